python -m pytest tests/test_gmm_gpu.py tests/test_plugin_gpu.py -x -q -m gpu 2>&1 | tail -15 > gpurun_out/s45_pytest.log
for V in 3 2; do
  (cd probability-of-collision-for-safe-planning_b200/csrc; touch mcsim_cuda.cu; make libmcsim_cuda.so NVFLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-Wall -I../../include -I. -DGMM_FUSED_CTAS=$V" > /dev/null 2>&1; grep -A3 "gmm_member_run" ptxas.log | grep -i "registers")
  echo "CTAS=$V"
  python tools/ncu_gmm_target.py 1332
  python tools/ncu_gmm_target.py 1024
  python tools/ncu_gmm_target.py 888
  python tools/ncu_gmm_target.py 256
  python tools/ncu_gmm_target.py 1
done > gpurun_out/s45_plain.log 2>&1

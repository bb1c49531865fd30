set -x
cd probability-of-collision-for-safe-planning_b200/csrc
for V in 5 4 6; do
  touch mcsim_cuda.cu
  make libmcsim_cuda.so NVFLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-Wall -I../../include -I. -DMC_P2_CTAS_PER_SM=$V" > /dev/null 2>&1
  grep -A3 "mc_philox2_kernelILb0ELi0" ptxas.log | grep -i "registers\|spill"
  (cd ../..; echo "CTAS=$V"; python tools/ncu_target.py 20000000; python tools/ncu_target.py 20000000)
done

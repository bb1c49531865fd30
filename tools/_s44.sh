python -m pytest tests/test_gmm_gpu.py tests/test_plugin_gpu.py -x -q -m gpu 2>&1 | tail -15 > gpurun_out/s44_pytest.log
MCSIM_TRACE=1 python tools/ncu_gmm_target.py 1024 > gpurun_out/s44_plain.log 2>&1
python tools/ncu_gmm_target.py 1332 >> gpurun_out/s44_plain.log 2>&1
python tools/ncu_gmm_target.py 256 >> gpurun_out/s44_plain.log 2>&1

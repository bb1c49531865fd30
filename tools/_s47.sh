python -m pytest tests/test_gmm_gpu.py tests/test_plugin_gpu.py -x -q -m gpu 2>&1 | tail -5 > gpurun_out/s47_pytest.log
python tools/ncu_gmm_target.py 1 > gpurun_out/s47_plain.log 2>&1
python tools/ncu_gmm_target.py 7 >> gpurun_out/s47_plain.log 2>&1
python tools/ncu_gmm_target.py 1332 >> gpurun_out/s47_plain.log 2>&1

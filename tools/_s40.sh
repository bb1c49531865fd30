python -m pytest tests/test_gmm_gpu.py tests/test_plugin_gpu.py -x -q -m gpu 2>&1 | tail -5 > gpurun_out/s40_pytest.log
python tools/ncu_gmm_target.py 1024 > gpurun_out/s40_plain.log 2>&1
python tools/ncu_gmm_target.py 256 >> gpurun_out/s40_plain.log 2>&1
python tools/ncu_gmm_target.py 1 >> gpurun_out/s40_plain.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/s40_launches.csv python tools/ncu_gmm_target.py 1024 > gpurun_out/s40_ncu.log 2>&1

python -m pytest tests/test_gmm_gpu.py tests/test_plugin_gpu.py tests/test_mc_gpu.py -x -q -s -m gpu 2>&1 | grep -v "^$" | tail -30 > gpurun_out/s38_pytest.log
python bench.py --steps 3 --warmup 3 > gpurun_out/s38_bench.json 2> gpurun_out/s38_bench.err

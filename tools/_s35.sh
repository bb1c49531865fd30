python -m pytest tests/test_mc_gpu.py tests/test_score_gpu.py tests/test_plugin_gpu.py -x -q -s -m gpu 2>&1 | grep -v "^$" | tail -40 > gpurun_out/s35_pytest.log
python bench.py --steps 3 --warmup 3 --no-extras > gpurun_out/s35_bench.json 2> gpurun_out/s35_bench.err

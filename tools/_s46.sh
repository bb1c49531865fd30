python -m pytest tests -x -q -m gpu 2>&1 | tail -5 > gpurun_out/s46_pytest.log
python bench.py > gpurun_out/s46_bench.json 2> gpurun_out/s46_bench.err
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s46_smoke.log 2>&1

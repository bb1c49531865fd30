python -m pytest tests -x -q -m gpu 2>&1 | tail -5 > gpurun_out/s48_pytest.log
python bench.py > gpurun_out/s48_bench.json 2> gpurun_out/s48_bench.err
python tools/ncu_gmm_target.py 1 > gpurun_out/s48_plain.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/s48_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/s48_ncu.log 2>&1

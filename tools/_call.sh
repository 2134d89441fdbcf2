mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x 2>&1 | tail -6 > gpurun_out/r2x_pytest.log; tail -4 gpurun_out/r2x_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
ORB_BENCH_CTA_CAP=0 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2x_launches.csv python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r2x_ncu.log 2>&1; wc -l gpurun_out/r2x_launches.csv

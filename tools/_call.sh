mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x 2>&1 | tail -6 > gpurun_out/r2r_pytest.log; tail -4 gpurun_out/r2r_pytest.log
bash tools/check_bounds_gpu.sh r2r

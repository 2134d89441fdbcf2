mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build_library()" 2>&1 | tail -2
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -15 > gpurun_out/r2m_noalloc_alone.log; tail -3 gpurun_out/r2m_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -q -x --durations=15 2>&1 | tail -40 > gpurun_out/r2m_pytest.log; tail -25 gpurun_out/r2m_pytest.log
timeout 600 python bench.py > gpurun_out/r2m_bench_ratio.json 2> gpurun_out/r2m_bench_ratio.err; tail -c 3000 gpurun_out/r2m_bench_ratio.json

# scratch driver for `gpurun -- 'bash tools/_call.sh'`: the end-of-round validation (GPU suite, smoke, one bench line)
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build_library()" 2>&1 | tail -2
timeout 1800 python -m pytest tests -m gpu -q -x 2>&1 | tail -6
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/bench_ratio.json 2> gpurun_out/bench_ratio.err; tail -c 400 gpurun_out/bench_ratio.json

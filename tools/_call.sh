mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build_library()" 2>&1 | tail -2
L=r2i_s2 bash tools/bench_multi_gpu.sh 8 ORB_BENCH_SLOTS=2
L=r2i_s1 bash tools/bench_multi_gpu.sh 8 ORB_BENCH_SLOTS=1

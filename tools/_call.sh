mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build_library()" 2>&1 | tail -2
timeout 600 python -m pytest tests/test_gpu_frames.py -m gpu -q -x -k "checked or handoff or slots" 2>&1 | tail -3
L=r2g bash tools/bench_multi_gpu.sh 2
L=r2g_nohalo bash tools/bench_multi_gpu.sh 2 ORB_BENCH_HALO=0
L=r2g_s1 bash tools/bench_multi_gpu.sh 2 ORB_BENCH_SLOTS=1

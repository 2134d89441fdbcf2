mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build_library()" 2>&1 | tail -2
timeout 600 python -m pytest tests/test_gpu_frames.py -m gpu -q -x -k "checked or handoff or slots" 2>&1 | tail -3
L=r2h_s2 bash tools/bench_multi_gpu.sh 4 ORB_BENCH_SLOTS=2
L=r2h_s3 bash tools/bench_multi_gpu.sh 4 ORB_BENCH_SLOTS=3
L=r2h_s4 bash tools/bench_multi_gpu.sh 4 ORB_BENCH_SLOTS=4

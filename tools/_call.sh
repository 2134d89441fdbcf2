mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build_library()" 2>&1 | tail -2   # only compiles when the sources are newer than the .so
timeout 1500 python -m pytest tests -m gpu -q -s 2>&1 | grep -v "^$" | grep -E "classified|passed|failed|FAILED|Error|error|assert|full-size" | cut -c1-1500 > gpurun_out/r2d_pytest.log; tail -3 gpurun_out/r2d_pytest.log
export ORB_TRACE_VARIANT=1
timeout 900 python tools/ab_trace2.py '' 'ORB_L2_PERSIST_MB=48' 'ORB_L2_PERSIST_MB=96' 'ORB_NODE_MIN=12' 'ORB_NODE_MIN=20' 'ORB_NODE_MIN=24' 'ORB_REFILL_MIN=8' 'ORB_REFILL_MIN=16' 'ORB_REFILL_MIN=20' 'ORB_PATH_NODE_MIN=12 ORB_PATH_SHADE_MIN=16' 'ORB_PATH_NODE_MIN=16 ORB_PATH_SHADE_MIN=24' 'ORB_TRACE_VARIANT=3 ORB_L2_PERSIST_MB=48' > gpurun_out/r2d_ab2.jsonl 2> gpurun_out/r2d_ab2.err; echo "ab2 rc=$?"
ORB200_LIB=$PWD/optix_renderer_b200/liborb200_tri1.so timeout 300 python tools/ab_trace2.py '' > gpurun_out/r2d_ab2_tri1.jsonl 2>> gpurun_out/r2d_ab2.err; echo "tri1 rc=$?"

mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s 2>&1 | grep -v "^$" | grep -E "classified|passed|failed|FAILED|Error|error|cfg2 frame|assert|full-size" | cut -c1-900 > gpurun_out/r2c_pytest.log; tail -3 gpurun_out/r2c_pytest.log
timeout 900 python tools/ab_trace.py > gpurun_out/r2c_ab_trace.jsonl 2> gpurun_out/r2c_ab_trace.err; echo "ab rc=$?"
ORB200_LIB=$PWD/optix_renderer_b200/liborb200_minb10.so timeout 600 python tools/ab_trace.py 3,12 3,0 > gpurun_out/r2c_ab_trace_minb10.jsonl 2>> gpurun_out/r2c_ab_trace.err; echo "ab10 rc=$?"

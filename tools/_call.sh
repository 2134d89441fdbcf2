mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build_library()" 2>&1 | tail -2
timeout 1500 python -m pytest tests -m gpu -q -s 2>&1 | grep -v "^$" | grep -E "classified|passed|failed|FAILED|Error|error|assert" | cut -c1-1200 > gpurun_out/r2f_pytest.log; tail -3 gpurun_out/r2f_pytest.log
timeout 900 python tools/ab_trace2.py 'ORB_SORT_RAYS=0' 'ORB_SORT_RAYS=1' 'ORB_SORT_CHUNK=256' 'ORB_SORT_CHUNK=512' 'ORB_SORT_CHUNK=2048' 'ORB_SORT_CHUNK=4096' 'ORB_SORT_CHUNK=16384' 'ORB_SORT_CHUNK=1024 ORB_NODE_MIN=20' 'ORB_SORT_CHUNK=1024 ORB_NODE_MIN=24 ORB_REFILL_MIN=16' 'ORB_SORT_CHUNK=1024 ORB_TRACE_VARIANT=3' > gpurun_out/r2f_ab_sort.jsonl 2> gpurun_out/r2f_ab_sort.err; echo "ab rc=$?"
ORB_BENCH_SLOTS=2 timeout 600 python bench.py --steps 60 --no-cpu-baseline > gpurun_out/r2f_ratio_s2.json 2> gpurun_out/r2f_ratio_s2.err; echo "ratio rc=$?"
ORB_BENCH_SLOTS=2 timeout 600 python bench.py --steps 60 --workload path --no-cpu-baseline > gpurun_out/r2f_path_s2.json 2> gpurun_out/r2f_path_s2.err; echo "path rc=$?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2f_*_s2.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["ms_per_step"], 3), "serial", round(d["e2e"].get("serial_ms_per_step", 0), 3), d["e2e"].get("readback_equals_device_frame"))
    except Exception as e:
        print(f, "ERR", e)
PY

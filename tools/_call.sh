mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build_library()" 2>&1 | tail -2
timeout 1500 python -m pytest tests -m gpu -q -s 2>&1 | grep -v "^$" | grep -E "classified|passed|failed|FAILED|Error|error|assert|full-size" | cut -c1-2500 > gpurun_out/r2e_pytest.log; tail -3 gpurun_out/r2e_pytest.log
export ORB_TRACE_VARIANT=1
for s in 1 2; do
ORB_BENCH_SLOTS=$s timeout 600 python bench.py --steps 60 --no-cpu-baseline > gpurun_out/r2e_ratio_s$s.json 2> gpurun_out/r2e_ratio_s$s.err; echo "ratio slots $s rc=$?"
ORB_BENCH_SLOTS=$s timeout 600 python bench.py --steps 60 --workload path --no-cpu-baseline > gpurun_out/r2e_path_s$s.json 2> gpurun_out/r2e_path_s$s.err; echo "path slots $s rc=$?"
done
for c in 4 6 7; do
ORB_TRACE_CTAS_PER_SM=$c timeout 600 python bench.py --steps 60 --no-cpu-baseline > gpurun_out/r2e_ratio_s2_c$c.json 2> gpurun_out/r2e_ratio_s2_c$c.err; echo "ratio ctas $c rc=$?"
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2e_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["ms_per_step"], 3), "serial", round(d["e2e"].get("serial_ms_per_step", 0), 3), d["e2e"].get("readback_equals_device_frame"))
    except Exception as e:
        print(f, "ERR", e)
PY
bash tools/ncu_full.sh r2e tools/ncu_target2.py 'k_trace_stream_spec|k_path_persistent' 6

mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build_library()" 2>&1 | tail -2
timeout 1500 python -m pytest tests -m gpu -q -s 2>&1 | grep -v "^$" | grep -E "classified|passed|failed|FAILED|Error|error|assert" | cut -c1-1500 > gpurun_out/r2l_pytest.log; tail -3 gpurun_out/r2l_pytest.log
timeout 600 python tools/ab_trace2.py '' 'ORB_LTC_PAIR_QUADS=0' > gpurun_out/r2l_ab_ltc.jsonl 2> gpurun_out/r2l_ab.err; echo "ab ltc rc=$?"
for m in 6 7 8; do ORB200_LIB=$PWD/optix_renderer_b200/liborb200_pmb$m.so timeout 300 python tools/ab_trace2.py '' >> gpurun_out/r2l_ab_pmb.jsonl 2>> gpurun_out/r2l_ab.err; done; echo "pmb rc=$?"
python - <<'PY'
import json
for f in ("gpurun_out/r2l_ab_ltc.jsonl", "gpurun_out/r2l_ab_pmb.jsonl"):
    for l in open(f):
        d = json.loads(l); print(d["knobs"], d["lib"], d["ratio_full"], "band", d["ratio_band135"].get("ltc"), "path", d["path_full"], d["path_band135"])
PY
bash tools/sanitize_gpu.sh r2l
timeout 900 python tools/ab_band.py 'SLOTS=1' 'SLOTS=2' 'SLOTS=2 ORB_STREAM_BLOCK=64' 'SLOTS=2 ORB_STREAM_BLOCK=32' 'SLOTS=2 ORB_TRACE_CTAS_PER_SM=6' 'SLOTS=2 ORB_STREAM_BLOCK=64 ORB_TRACE_CTAS_PER_SM=12' 'SLOTS=2 ORB_STREAM_BLOCK=32 ORB_TRACE_CTAS_PER_SM=24' 'SLOTS=3 ORB_STREAM_BLOCK=32' 'SLOTS=1 ORB_STREAM_BLOCK=32' > gpurun_out/r2l_ab_band.jsonl 2>> gpurun_out/r2l_ab.err; cat gpurun_out/r2l_ab_band.jsonl
for b in 128 64 32; do
ORB_STREAM_BLOCK=$b timeout 600 python bench.py --steps 60 --no-cpu-baseline > gpurun_out/r2l_ratio_b$b.json 2> gpurun_out/r2l_ratio_b$b.err
ORB_STREAM_BLOCK=$b timeout 600 python bench.py --steps 60 --workload path --no-cpu-baseline > gpurun_out/r2l_path_b$b.json 2> gpurun_out/r2l_path_b$b.err
done
ORB_DL_INPLACE=1 timeout 600 python bench.py --steps 60 --workload path --no-cpu-baseline > gpurun_out/r2l_path_inplace.json 2> gpurun_out/r2l_path_inplace.err
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2l_ratio_*.json") + glob.glob("gpurun_out/r2l_path_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["ms_per_step"], 3), "serial", round(d["e2e"].get("serial_ms_per_step", 0), 3), d["e2e"].get("readback_equals_device_frame"), {k: round(v["ms_per_frame"], 3) for k, v in d["passes"].items() if k in ("ltc", "trace", "path")})
    except Exception as e:
        print(f, "ERR", e)
PY

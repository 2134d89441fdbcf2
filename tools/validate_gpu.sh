# usage: tools/validate_gpu.sh [label] [ncu]   (label prefixes the files written under gpurun_out/)
L=${1:-r1q}
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/${L}_bench_ratio.json 2> gpurun_out/${L}_bench_ratio.err; echo "ratio rc=$?"
timeout 900 python bench.py --workload path > gpurun_out/${L}_bench_path.json 2> gpurun_out/${L}_bench_path.err; echo "path rc=$?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${L}_bench_ref.json 2> gpurun_out/${L}_bench_ref.err; echo "ref rc=$?"
L=$L python - <<'PY'
import json, os
L = os.environ['L']
for n in ("ratio","path","ref"):
    d=json.loads(open("gpurun_out/%s_bench_%s.json" % (L, n)).read().strip().splitlines()[-1])
    print(n, round(d["ms_per_step"],3), round(d["value"],2), "e2e", round(d["e2e"]["value"],2), d.get("gpu_launches"), (d.get("clocks") or {}).get("reasons"))
    if "passes" in d: print({k:(round(v["ms_per_frame"],3), round(v["frac_of_hbm_peak"] or 0,3)) for k,v in d["passes"].items()})
    if "roofline" in d: print({k:v for k,v in d["roofline"].items() if k not in ("note",)})
PY
if [ "$2" = "ncu" ]; then
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/${L}_short.json 2>/dev/null && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${L}_launches.csv \
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/${L}_ncu.log 2>&1; echo "ncu rc=$?"
wc -l gpurun_out/${L}_launches.csv
fi

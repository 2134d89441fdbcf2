timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
for wl in ratio; do
timeout 600 python bench.py --steps 60 --warmup 5 --no-cpu-baseline --workload $wl > gpurun_out/c_$wl.json 2> gpurun_out/c_$wl.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/c_$wl.json").read().strip().splitlines()[-1])
print(round(d["ms_per_step"],3), round(d["value"],1), "e2e", round(d["e2e"]["value"],1), d["roofline"]["frac"], d["roofline"]["traffic"])
print({k:(round(v["ms_per_frame"],3), round(v["frac_of_hbm_peak"] or 0,3)) for k,v in d["passes"].items()})
PY
done

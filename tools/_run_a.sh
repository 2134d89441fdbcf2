# GPU validation pass: tests, default bench, then the ncu launch list of a short bench run
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 600 python bench.py > gpurun_out/r1m_bench_ratio.json 2> gpurun_out/r1m_bench_ratio.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r1m_bench_ratio.json").read().strip().splitlines()[-1])
print(round(d["ms_per_step"],3), round(d["value"],1), "e2e", round(d["e2e"]["value"],1), d["roofline"], d.get("cpu_baseline"))
print({k:(round(v["ms_per_frame"],3), round(v["frac_of_hbm_peak"] or 0,3)) for k,v in d["passes"].items()})
PY
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r1m_short.json 2>/dev/null && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r1m_launches.csv \
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r1m_ncu.log 2>&1; echo "ncu rc=$?"
wc -l gpurun_out/r1m_launches.csv

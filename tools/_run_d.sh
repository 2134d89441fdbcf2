timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
run() { timeout 600 python bench.py --steps 40 --warmup 5 --no-cpu-baseline > gpurun_out/$1.json 2>gpurun_out/$1.err; python - <<PY
import json
d=json.loads(open("gpurun_out/$1.json").read().strip().splitlines()[-1])
print("$1", round(d["ms_per_step"],3), round(d["value"],1), d["roofline"].get("per_ray"), {k:round(v["ms_per_frame"],3) for k,v in d["passes"].items()})
PY
}
run d_wide
ORB_TRACE_VARIANT=1 run d_spec
ORB_WIDE_NODE_MIN=8 run d_wide_nm8
ORB_WIDE_NODE_MIN=24 run d_wide_nm24

run() { timeout 600 python bench.py --steps 40 --warmup 5 --no-cpu-baseline > gpurun_out/$1.json 2>gpurun_out/$1.err; python - <<PY
import json
d=json.loads(open("gpurun_out/$1.json").read().strip().splitlines()[-1])
print("$1", round(d["ms_per_step"],3), round(d["value"],1), {k:round(v["ms_per_frame"],3) for k,v in d["passes"].items()})
PY
}
ORB_POP_SKIP=0 run e_skip0
ORB_POP_SKIP=1 run e_skip1
ORB_POP_SKIP=2 run e_skip2
ORB_POP_SKIP=4 run e_skip4
ORB_POP_SKIP=2 ORB_NODE_MIN=12 run e_skip2_nm12
ORB_POP_SKIP=2 ORB_NODE_MIN=20 run e_skip2_nm20
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4

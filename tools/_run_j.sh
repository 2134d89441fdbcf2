run() { n=$1; shift; timeout 600 python bench.py --steps 40 --warmup 5 --no-cpu-baseline "$@" > gpurun_out/$n.json 2>gpurun_out/$n.err; python - <<PY
import json
d=json.loads(open("gpurun_out/$n.json").read().strip().splitlines()[-1])
print("$n", round(d["ms_per_step"],3), round(d["value"],1), {k:round(v["ms_per_frame"],3) for k,v in d["passes"].items()})
PY
}
ORB_Q_STACK=0 run j_q0
ORB_Q_STACK=1 run j_q1
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3

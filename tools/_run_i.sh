timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
run() { n=$1; shift; timeout 600 python bench.py --steps 40 --warmup 5 --no-cpu-baseline "$@" > gpurun_out/$n.json 2>gpurun_out/$n.err; python - <<PY
import json
d=json.loads(open("gpurun_out/$n.json").read().strip().splitlines()[-1])
print("$n", round(d["ms_per_step"],3), round(d["value"],1), {k:round(v["ms_per_frame"],3) for k,v in d["passes"].items()}, d.get("ltc_pass_one_quad",{}).get("ms"))
PY
}
run i_lut8
run i_lut64 --lut 64

mkdir -p gpurun_out
N=8
run() { L=$1; shift; for kv in "$@"; do export "$kv"; done
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus $N --steps 60 --warmup 5 --workload ratio > gpurun_out/r2q_n8_$L.json 2> gpurun_out/r2q_n8_$L.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r2q_n8_$L.json").read().strip().splitlines()[-1])
print("$L", round(d["ms_per_step"],3), round(d["e2e"]["ms_per_step"],3), d["host_enqueue_ms_per_step"], d["rank_kernel_ms_per_frame"][:3], {k:round(v["ms_per_frame"],3) for k,v in d["passes"].items()})
PY
for kv in "$@"; do unset "${kv%%=*}"; done; }
run tl ORB_BENCH_TIMELINE=gpurun_out/r2q_timeline_n8_s6.json
run cap3 ORB_BENCH_CTA_CAP=3
run s8cap2 ORB_BENCH_SLOTS=8

mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build_library()" 2>&1 | tail -2
run() { L=$1; shift; for kv in "$@"; do export "$kv"; done
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 40 --warmup 5 > gpurun_out/${L}.json 2> gpurun_out/${L}.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${L}.json").read().strip().splitlines()[-1])
    print("${L}", "dev", round(d["ms_per_step"],3), "host", round(d["host_enqueue_ms_per_step"],3), "e2e", round(d["e2e"]["ms_per_step"],3), "e2e host", d["e2e"].get("host_enqueue_ms_per_step"), d["e2e"].get("host_gather"))
except Exception as e:
    print("${L} ERR", e)
PY
  for kv in "$@"; do unset "${kv%%=*}"; done; }
nproc; free -g | head -2
run r2k_a ORB_BENCH_SLOTS=2 ORB_BENCH_TIMELINE=gpurun_out/r2k_timeline_n8_s2.json
run r2k_b ORB_BENCH_SLOTS=2 ORB_DL_INPLACE=0
run r2k_c ORB_BENCH_SLOTS=2 ORB_BENCH_HOST_GATHER=0
run r2k_d ORB_BENCH_SLOTS=1 ORB_BENCH_TIMELINE=gpurun_out/r2k_timeline_n8_s1.json

# usage: L=<label> tools/bench_multi_gpu.sh N [extra env assignments...]   -- bench.py on N GPUs of this box, both workloads
N=$1; shift
for kv in "$@"; do export "$kv"; done
mkdir -p gpurun_out
for wl in ratio path; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 60 --warmup 5 --workload $wl > gpurun_out/${L:-r2}_mg_${wl}_n$N.json 2> gpurun_out/${L:-r2}_mg_${wl}_n$N.err
  python - <<PY
import json
s=open("gpurun_out/${L:-r2}_mg_${wl}_n$N.json").read().strip().splitlines()
print("lines", len(s))
if s:
    d=json.loads(s[-1]); print(d.get("collectives_ms_per_frame_rank0"), d.get("assembled_frame_equals_single_gpu_frame"), d.get("rank_kernel_ms_per_frame"), d.get("halo")); print("$wl", d["n_gpus"], "ms/step", round(d["ms_per_step"],3), "Mrays/s", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), round(d["e2e"]["ms_per_step"],3), (d.get("bands") or d["config"].get("bands")), {k:round(v["ms_per_frame"],3) for k,v in d["passes"].items()})
PY
  grep -v "OMP_NUM_THREADS\|^\*\*\*\|^$" gpurun_out/${L:-r2}_mg_${wl}_n$N.err | tail -8
done

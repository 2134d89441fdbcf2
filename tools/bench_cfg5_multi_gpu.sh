N=$1
for wl in ratio path; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N bench.py --gpus $N --steps 64 --warmup 3 --workload $wl --scene cfg5 --no-cpu-baseline > gpurun_out/cfg5_${wl}_n$N.json 2> gpurun_out/cfg5_${wl}_n$N.err
  python - <<PY
import json
s=open("gpurun_out/cfg5_${wl}_n$N.json").read().strip().splitlines()
d=json.loads(s[-1]); print(d.get("collectives_ms_per_frame_rank0"), d.get("assembled_frame_equals_single_gpu_frame"), d.get("rank_kernel_ms_per_frame")); print("$wl", d["n_gpus"], round(d["ms_per_step"],3), round(d["value"],1), round(d["e2e"]["value"],1), (d.get("bands") or d["config"].get("bands")), {k:round(v["ms_per_frame"],3) for k,v in d["passes"].items()})
PY
  grep -v "OMP_NUM_THREADS\|^\*\*\*\|^$" gpurun_out/cfg5_${wl}_n$N.err | tail -4
done

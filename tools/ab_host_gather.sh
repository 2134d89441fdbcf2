# usage: tools/ab_host_gather.sh N [modes]  -- e2e of the N-GPU bench with the per-rank band read-back (1) and without (0)
N=$1
for hg in ${2:-1 0}; do for wl in ratio path; do
  ORB_BENCH_HOST_GATHER=$hg timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N bench.py --gpus $N --steps 30 --warmup 5 --workload $wl > gpurun_out/hg${hg}_${wl}_n$N.json 2> gpurun_out/hg${hg}_${wl}_n$N.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/hg${hg}_${wl}_n$N.json").read().strip().splitlines()[-1])
print("host_gather=$hg $wl N=$N", "device ms", round(d["ms_per_step"],3), "e2e ms", round(d["e2e"]["ms_per_step"],3), "same", d.get("assembled_frame_equals_single_gpu_frame"), d["e2e"].get("host_gather"))
PY
  grep -i "error\|Traceback\|differs" gpurun_out/hg${hg}_${wl}_n$N.err | head -5
done; done

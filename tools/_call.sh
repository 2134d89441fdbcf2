mkdir -p gpurun_out
N=4
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $N --steps 40 --warmup 5 --workload ratio > gpurun_out/r2y_n4_ratio.json 2> gpurun_out/r2y_n4_ratio.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2y_n4_ratio.json").read().strip().splitlines()[-1])
    print("ratio", round(d["ms_per_step"],3), round(d["e2e"]["ms_per_step"],3), d["host_enqueue_ms_per_step"], d["assembled_frame_equals_single_gpu_frame"], d.get("assembly","")[:30], d["rank_kernel_ms_per_frame"])
except Exception as e:
    print("ERR", e); print(open("gpurun_out/r2y_n4_ratio.err").read()[-1500:])
PY

mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_frames.py -m gpu -q -x -k "one_sided" 2>&1 | tail -5
N=2
for WL in ratio path; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 60 --warmup 5 --workload $WL > gpurun_out/r2y_n2_$WL.json 2> gpurun_out/r2y_n2_$WL.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2y_n2_$WL.json").read().strip().splitlines()[-1])
    print("$WL", round(d["ms_per_step"],3), round(d["e2e"]["ms_per_step"],3), d["host_enqueue_ms_per_step"], d["assembled_frame_equals_single_gpu_frame"], d.get("assembly","")[:30], d["rank_kernel_ms_per_frame"])
except Exception as e:
    print("$WL ERR", e); print(open("gpurun_out/r2y_n2_$WL.err").read()[-1500:])
PY
done

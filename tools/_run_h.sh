run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/dbg_$1.json 2> gpurun_out/dbg_$1.err; echo "== $1"; grep -A1 "assembled frame" gpurun_out/dbg_$1.err | cut -c1-200; python -c "
import json;d=json.loads(open('gpurun_out/dbg_$1.json').read().strip().splitlines()[-1]);print(d['ms_per_step'], d.get('assembled_frame_equals_single_gpu_frame'))"; }
ORB_BENCH_SYNC_ASSEMBLE=1 run sync
run async
ORB_RESOLVE_LANES=0 run async_nolanes

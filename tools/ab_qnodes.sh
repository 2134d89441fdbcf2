# usage: tools/ab_qnodes.sh   -- the 32-byte quantised nodes (ORB_TRACE_VARIANT=3, bvhq.cuh) in the ray-stream kernel:
# hit parity under the variant, then the bench's trace stage time with it (compare profiles/*_bench_ratio.json: 1.69 ms)
mkdir -p gpurun_out
ORB_TRACE_VARIANT=3 timeout 60 python -m pytest tests/test_gpu_frames.py -m gpu -q -x -k "brute_force_and_oracle or refit_after" 2>&1 | tail -2
ORB_TRACE_VARIANT=3 timeout 60 python bench.py --steps 30 --warmup 3 --no-cpu-baseline 2>/dev/null > gpurun_out/qnodes_bench.json
python -c "
import json
d=json.loads(open('gpurun_out/qnodes_bench.json').read().strip().splitlines()[-1])
print('variant 3: frame', round(d['ms_per_step'],4), 'trace', round(d['passes']['trace']['ms_per_frame'],4), 'nodes/ray', round(d['roofline']['per_ray']['nodes'],2))"

for q in 0 2 0 2; do
  ORB_Q_STACK=$q python bench.py --steps 60 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('q=$q', round(d['ms_per_step'],4), 'trace', round(d['passes']['trace']['ms_per_frame'],4))"
done
ORB_Q_STACK=2 python -m pytest tests/test_gpu_frames.py -m gpu -q -k "brute or refit or band or tile" 2>&1 | tail -2

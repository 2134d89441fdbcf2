run() { timeout 300 python bench.py --steps 40 --warmup 5 --no-cpu-baseline $2 > gpurun_out/$1.json 2>>gpurun_out/ab_err.log; python -c "
import json;d=json.loads(open('gpurun_out/$1.json').read().strip().splitlines()[-1]);print('$1', round(d['ms_per_step'],3),round(d['value'],1),{k:round(v['ms_per_frame'],3) for k,v in d['passes'].items()})"; }
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
run ab_ratio
ORB_LTC_OVERLAP=0 run ab_ratio_noov
tail -3 gpurun_out/ab_err.log

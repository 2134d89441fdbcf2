run() { timeout 300 python bench.py --steps 40 --warmup 5 --no-cpu-baseline $2 > gpurun_out/$1.json 2>>gpurun_out/ab_err.log; python -c "
import json;d=json.loads(open('gpurun_out/$1.json').read().strip().splitlines()[-1]);print('$1', round(d['ms_per_step'],3),round(d['value'],1),{k:round(v['ms_per_frame'],3) for k,v in d['passes'].items()})"; }
for sm in 20 24 28 32; do for nm in 8 12 16; do ORB_PATH_SHADE_MIN=$sm ORB_PATH_NODE_MIN=$nm run ab_path_s${sm}_n${nm} "--workload path"; done; done
tail -3 gpurun_out/ab_err.log

run() { python bench.py --steps 40 --warmup 5 --no-cpu-baseline $2 > gpurun_out/$1.json 2>>gpurun_out/r1h_err.log; python -c "
import json;d=json.load(open('gpurun_out/$1.json'));print('$1', round(d['ms_per_step'],3),round(d['value'],1),{k:round(v['ms_per_frame'],3) for k,v in d['passes'].items()})"; }
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
run r1h_ratio
ORB_MAX_LEAF=2 run r1h_ratio_leaf2
ORB_MAX_LEAF=1 run r1h_ratio_leaf1
ORB_MAX_LEAF=8 run r1h_ratio_leaf8
run r1h_path "--workload path"
ORB_MAX_LEAF=2 run r1h_path_leaf2 "--workload path"
tail -3 gpurun_out/r1h_err.log

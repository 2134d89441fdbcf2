mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_frames.py -m gpu -q -x -k "atrous or denoise or two_frame_slots or band_handoff" 2>&1 | tail -5
ORB200_LIB=$PWD/optix_renderer_b200/liborb200_mb3.so timeout 600 python bench.py --steps 60 --no-cpu-baseline > gpurun_out/r2n_ratio_mb3.json 2> gpurun_out/r2n_ratio_mb3.err
timeout 600 python bench.py --steps 60 --no-cpu-baseline > gpurun_out/r2n_ratio_mb4.json 2> gpurun_out/r2n_ratio_mb4.err
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2n_ratio_mb*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["ms_per_step"], 3), {k: round(v["ms_per_frame"], 4) for k, v in d["passes"].items()})
    except Exception as e:
        print(f, "ERR", e)
PY

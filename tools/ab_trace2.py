"""Second traversal A/B: environment knobs of the stream / path kernels on the bench workload, one JSON line each.
   python tools/ab_trace2.py 'ORB_NODE_MIN=12' 'ORB_NODE_MIN=20 ORB_REFILL_MIN=8' ...   ('' = defaults)"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optix_renderer_b200 import api, scenes

sc = scenes.scene_synthetic(1_000_000, 16, 1920, 1080)
base = dict(os.environ)
for spec in (sys.argv[1:] or [""]):
    os.environ.clear(); os.environ.update(base)
    for kv in spec.split():
        k, v = kv.split("=")
        os.environ[k] = v
    r = api.Renderer(sc, quantise_lut=1)
    out = {"knobs": spec, "lib": os.environ.get("ORB200_LIB", "default")[-14:]}
    for name, rows in (("full", None), ("band135", (472, 607))):
        for i in range(3):
            r.render(api.RATIO, accum_id=i, mode=1, shadow_spp=4, rows=rows)
        r.set_profiling(True); r.stage_times()
        for i in range(20):
            r.render(api.RATIO, accum_id=3 + i, mode=1, shadow_spp=4, rows=rows, sync=False)
        st = r.stage_times()
        out["ratio_" + name] = {k: round(v[0] / 20, 4) for k, v in st.items() if v[1] and k in ("trace", "primary", "ltc", "shadow")}
        for i in range(2):
            r.render(api.PATH, accum_id=i, mode=1, max_depth=8, rows=rows)
        r.stage_times()
        for i in range(10):
            r.render(api.PATH, accum_id=2 + i, mode=1, max_depth=8, rows=rows, sync=False)
        st = r.stage_times()
        out["path_" + name] = round(st["path"][0] / 10, 4)
        r.set_profiling(False)
    print(json.dumps(out), flush=True)
    r.close()

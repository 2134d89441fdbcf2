"""A/B of the traversal kernels' knobs on the bench workload (cfg3: 1080p, 1 M triangles, 32 TriLights):
   ORB_TRACE_VARIANT (1 = 64-byte fp32 nodes, 3 = 32-byte quantised nodes) x ORB_STACK_SMEM (stack entries per lane in
   shared memory).  Per configuration: stream-kernel hits == brute force on 20 k random rays, then CUDA-event stage
   times of 20 RATIO frames and 10 PATH frames (full frame and a 135-row band).  One JSON line per configuration.
       python tools/ab_trace.py [variant,smem ...]          (ORB200_LIB=<other build> for compile-time knobs)"""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optix_renderer_b200 import api, scenes

configs = [tuple(int(x) for x in a.split(",")) for a in sys.argv[1:]] or [(1, 0), (3, 0), (3, 8), (3, 12), (3, 16), (1, 12)]
sc = scenes.scene_synthetic(1_000_000, 16, 1920, 1080)
rng = np.random.default_rng(11)
n = 20000
org = (rng.uniform(-1, 1, (n, 3)) * 9).astype(np.float32)
org[:, 2] = np.abs(org[:, 2]) * 0.5 + 0.2
d = rng.normal(size=(n, 3)).astype(np.float32)
d /= np.linalg.norm(d, axis=1, keepdims=True)
d[:100, 1] = 0.0
ref_frame = None
for variant, sd in configs:
    os.environ["ORB_TRACE_VARIANT"] = str(variant)
    os.environ["ORB_STACK_SMEM"] = str(sd)
    r = api.Renderer(sc, quantise_lut=1)
    want = r.trace_rays(org, d, brute_force=True)
    got = r.trace_rays(org, d, brute_force=2)
    ok = all(np.array_equal(a, b) for a, b in zip(want, got))
    out = {"variant": variant, "stack_smem": sd, "lib": os.environ.get("ORB200_LIB", "default"), "hits_equal_brute_force": ok}
    for name, rows in (("full", None), ("band135", (472, 607))):
        for i in range(3):
            r.render(api.RATIO, accum_id=i, mode=1, shadow_spp=4, rows=rows)
        r.set_profiling(True); r.stage_times()
        for i in range(20):
            r.render(api.RATIO, accum_id=3 + i, mode=1, shadow_spp=4, rows=rows, sync=False)
        st = r.stage_times()
        out["ratio_" + name] = {k: round(v[0] / 20, 4) for k, v in st.items() if v[1]}
        for i in range(2):
            r.render(api.PATH, accum_id=i, mode=1, max_depth=8, rows=rows)
        r.stage_times()
        for i in range(10):
            r.render(api.PATH, accum_id=2 + i, mode=1, max_depth=8, rows=rows, sync=False)
        st = r.stage_times()
        out["path_" + name] = {k: round(v[0] / 10, 4) for k, v in st.items() if v[1]}
        r.set_profiling(False)
    r.render(api.RATIO, accum_id=0, mode=1, shadow_spp=4)
    fr = {k: r.download(k) for k in ("sto_direct", "sto_no_vis", "ltc")}
    if ref_frame is None:
        ref_frame = fr
    out["frame_equals_first_config"] = all(np.array_equal(fr[k], ref_frame[k], equal_nan=True) for k in fr)
    print(json.dumps(out), flush=True)
    r.close()

"""GPU check of BASELINE configs[4] on ONE GPU: 10M triangles, 64 quad lights (128 TriLights), 3840x2160,
RATIO 4 spp + denoise + combine and an 8-bounce PATH frame; prints build time, stage times, memory."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from optix_renderer_b200 import api, scenes

t0 = time.time()
sc = scenes.scene_synthetic(10_000_000, 64, 3840, 2160)
t1 = time.time()
r = api.Renderer(sc)
t2 = time.time()
c = r.counters()
out = {"scene_gen_s": round(t1 - t0, 1), "upload_commit_s": round(t2 - t1, 1), "triangles": c["bvh_triangles"],
       "bvh_nodes": c["bvh_nodes"], "bvh_height": c["bvh_height"], "bvh_build_ms": c["bvh_build_us"] / 1e3}
for rt, name in ((api.RATIO, "ratio"), (api.PATH, "path")):
    for i in range(2):
        r.render(rt, accum_id=i, mode=1, shadow_spp=4, max_depth=8)
    r.set_profiling(True)
    r.stage_times()
    n = 5
    for i in range(n):
        r.render(rt, accum_id=2 + i, mode=1, shadow_spp=4, max_depth=8, sync=False)
        if rt == api.RATIO:
            r.denoise()
            r.ratio_combine(use_denoised=True)
    st = r.stage_times()
    r.set_profiling(False)
    c = r.counters()
    out[name] = {"active": c["active_pixels"], "nan_pixels": c["nan_pixels"],
                 "ms": {k: round(v[0] / n, 3) for k, v in st.items() if v[1]},
                 "frame_ms": round(sum(v[0] for v in st.values()) / n, 3)}
    if rt == api.RATIO:
        fin = r.download("ratio_final", rows=(1000, 1010))
        out[name]["finite_frac_rows_1000_1010"] = float(np.isfinite(fin[..., :3]).mean())
import torch
free, total = torch.cuda.mem_get_info()
out["hbm_used_gb"] = round((total - free) / 2**30, 2)
print(json.dumps(out))

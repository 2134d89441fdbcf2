"""G0: the reference's device source as a literal one-thread-per-pixel megakernel on the
B200 (oracle/_ref/libref_gpu_intended.so; traceRay hook = this repository's BVH) on the bench scene, timed
beside the product's kernels for the same frames.  Kernel-only CUDA-event times, no copies.  TEST INFRASTRUCTURE
is what is timed on the G0 side; nothing here is a product path."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _oracle
from optix_renderer_b200 import api, scenes

W, H = 1920, 1080
sc = scenes.scene_synthetic(1_000_000, 16, W, H)
out = {"scene": "bench scene: 1920x1080, 1 M triangles, 32 TriLights; intended mode"}
g = _oracle.Oracle("reference-gpu-intended")
g.load_scene(sc)
r = api.Renderer(sc)
r.set_profiling(True)
for name, rt, spp, depth in (("ratio_4spp", _oracle.RATIO, 4, 1), ("path_8_bounces", _oracle.PATH, 1, 8),
                             ("ltc_baseline", _oracle.LTC_BASELINE, 4, 1)):
    g.set_options(1, spp, depth)
    ms, rays = [], 0
    for i in range(4):
        rays = g.render(W, H, rt, i)["rays"]
        ms.append(g.last_ms())
    r.stage_times()
    for i in range(8):
        r.render(rt, accum_id=i, mode=1, shadow_spp=spp, max_depth=depth, sync=False)
    r.sync()
    ours = sum(t for t, _ in r.stage_times().values()) / 8
    out[name] = {"megakernel_ms": float(np.median(ms[1:])), "rays_per_frame": int(rays), "ours_ms": ours,
                 "speedup": float(np.median(ms[1:])) / ours}
print(json.dumps(out))

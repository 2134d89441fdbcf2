"""A few frames of the bench workload (cfg3: 1080p, 1M triangles, 32 TriLights) for ncu captures:
   3 RATIO frames + denoise + combine, then 2 PATH frames (8 bounces).  Run it once without ncu first."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optix_renderer_b200 import api, scenes

sc = scenes.scene_synthetic(1_000_000, 16, 1920, 1080)
r = api.Renderer(sc)
for i in range(3):
    r.render(api.RATIO, accum_id=i, mode=api.INTENDED, shadow_spp=4)
    r.denoise(iterations=5)
    r.ratio_combine(use_denoised=True)
for i in range(2):
    r.render(api.PATH, accum_id=i, mode=api.INTENDED, max_depth=8)
print("ok", r.counters()["rays_shadow"], r.counters()["rays_bounce"])
r.close()

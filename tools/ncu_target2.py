"""Round-2 ncu target: the bench workload (cfg3: 1080p, 1 M triangles, 32 TriLights) in the regimes that matter:
   frame 0 (warm-up, BVH pages in), frame 1 = full RATIO frame + denoise + combine, a 135-row band of the same
   frame (what one of 8 ranks renders), one PATH frame of 8 bounces, the PATH band, and the stand-alone LTC pass
   with ONE quad light.  Run it once without ncu first."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optix_renderer_b200 import api, scenes

sc = scenes.scene_synthetic(1_000_000, 16, 1920, 1080)
r = api.Renderer(sc, quantise_lut=1)
for i in range(2):
    r.render(api.RATIO, accum_id=i, mode=api.INTENDED, shadow_spp=4)
    r.denoise(iterations=5)
    r.ratio_combine(use_denoised=True)
band = (472, 607)
r.render(api.RATIO, accum_id=2, mode=api.INTENDED, shadow_spp=4, rows=band)
r.render(api.PATH, accum_id=0, mode=api.INTENDED, max_depth=8)
r.render(api.PATH, accum_id=1, mode=api.INTENDED, max_depth=8, rows=band)
r.set_lights(sc["tri_lights"][:2])
r.ltc_gbuffer_pass()
print("ok", r.counters()["rays_shadow"], r.counters()["rays_bounce"])
r.close()

"""ncu target for the denoise stage: the bench frame (cfg3: 1080p, 1 M triangles, 32 TriLights, RATIO 4 spp), then two
rounds of pack + 5-pass a-trous (+ fused combine).  ORB_ATROUS4=0 gives the one-pixel-per-thread kernel."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optix_renderer_b200 import api, scenes

sc = scenes.scene_synthetic(1_000_000, 16, 1920, 1080)
r = api.Renderer(sc, quantise_lut=1)
r.set_fused_combine(True)
r.render(api.RATIO, accum_id=0, mode=api.INTENDED, shadow_spp=4)
for i in range(2):
    r.denoise(iterations=5)
r.sync()
print("ok")
r.close()

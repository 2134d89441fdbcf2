"""One GPU: a renderer that has only ever rendered BANDS of the frame then renders the whole frame; every buffer
must equal a fresh renderer's whole frame (nothing may survive from earlier frames)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optix_renderer_b200 import api, scenes

sc = scenes.scene_synthetic(1_000_000, 16, 1920, 1080)
a = api.Renderer(sc)
b = api.Renderer(sc)
fid = 38
def full(r):
    r.render(api.RATIO, accum_id=fid, mode=1, shadow_spp=4)
    r.denoise(iterations=5)
    r.ratio_combine(use_denoised=True)
def band(r, i, rows):
    r.render(api.RATIO, accum_id=i, mode=1, shadow_spp=4, rows=rows)
    r.denoise_pack(rows=rows)
    r.denoise_filter(iterations=5, rows=rows)
    r.ratio_combine(use_denoised=True, rows=rows)
full(a)
band(b, 0, (0, 540))
for i in range(1, 6):
    band(b, i, (0, 369))
full(b)
for n in ["position", "normal", "materialID", "ltc", "sto_direct", "sto_no_vis", "dn_guide", "dn_val", "sto_direct_dn", "sto_no_vis_dn", "ratio_final"]:
    x, y = a.download(n), b.download(n)
    bad = (x.view(np.int32) != y.view(np.int32)).any(axis=2)
    rows = np.nonzero(bad.any(axis=1))[0]
    print(n, "pixels differing:", int(bad.sum()), "rows", rows[:8].tolist(), "...", rows[-4:].tolist())

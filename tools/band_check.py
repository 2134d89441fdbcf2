"""One GPU, bench scene: a band rendered alone (rows y0..y1) against the same rows of the whole frame, buffer by
buffer, then the denoise / combine of the band with packed rows of the whole frame around it."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optix_renderer_b200 import api, scenes

y0, y1 = int(sys.argv[1]), int(sys.argv[2])
sc = scenes.scene_synthetic(1_000_000, 16, 1920, 1080)
dev_b = int(sys.argv[3]) if len(sys.argv) > 3 else 0   # second renderer on another GPU of the box
a = api.Renderer(sc)
b = api.Renderer(sc, device=dev_b)
fid = 38
a.render(api.RATIO, accum_id=fid, mode=1, shadow_spp=4)
a.denoise(iterations=5)
a.ratio_combine(use_denoised=True)
b.render(api.RATIO, accum_id=fid, mode=1, shadow_spp=4, rows=(y0, y1))
names = ["position", "normal", "albedo", "alpha", "materialID", "ltc", "sto_direct", "sto_no_vis"]
for n in names:
    x, y = a.download(n)[y0:y1], b.download(n)[y0:y1]
    bad = (x.view(np.int32) != y.view(np.int32)).any(axis=2)
    print(n, "rows differing:", int(bad.any(axis=1).sum()), "pixels:", int(bad.sum()))
# denoise the band with the whole frame's packed rows around it (as the halo exchange would deliver them)
a.denoise_pack()
g, v = a.download("dn_guide"), a.download("dn_val")
b.denoise_pack(rows=(y0, y1))
g2, v2 = b.download("dn_guide"), b.download("dn_val")
print("pack rows differing:", int((g.view(np.int32)[y0:y1] != g2.view(np.int32)[y0:y1]).any(axis=(1, 2)).sum()),
      int((v.view(np.int32)[y0:y1] != v2.view(np.int32)[y0:y1]).any(axis=(1, 2)).sum()))
b.upload("dn_guide", g)
b.upload("dn_val", v)
b.denoise_filter(iterations=5, rows=(y0, y1))
b.ratio_combine(use_denoised=True, rows=(y0, y1))
for n in ["sto_direct_dn", "sto_no_vis_dn", "ratio_final"]:
    x, y = a.download(n)[y0:y1], b.download(n)[y0:y1]
    bad = (x.view(np.int32) != y.view(np.int32)).any(axis=2)
    print(n, "rows differing:", int(bad.any(axis=1).sum()), "pixels:", int(bad.sum()))

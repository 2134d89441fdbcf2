"""Where the pipelined read-back's time goes: wall clock per frame of the bench scene for render only, and with the
asynchronous download of either or both result buffers.  Usage: python tools/e2e_probe.py [ratio|path]"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optix_renderer_b200 import api, scenes

wl = sys.argv[1] if len(sys.argv) > 1 else "path"
W, H = 1920, 1080
sc = scenes.scene_synthetic(n_tris=1_000_000, n_quad_lights=16, width=W, height=H)
r = api.Renderer(sc, quantise_lut=1)
f4 = torch.empty((2, H, W, 4), dtype=torch.float32, pin_memory=True).numpy()
u32 = torch.empty((2, H, W), dtype=torch.int32, pin_memory=True).numpy().view(np.uint32)
name = "ratio_final" if wl == "ratio" else "accum"


def frame(i):
    if wl == "ratio":
        r.render(api.RATIO, accum_id=i, mode=api.INTENDED, shadow_spp=4, sync=False)
        r.denoise(iterations=5)
        r.ratio_combine(use_denoised=True)
    else:
        r.render(api.PATH, accum_id=i, mode=api.INTENDED, max_depth=8, sync=False)


def run(what, n=40):
    for i in range(5):
        frame(i)
    r.sync()
    t = time.perf_counter()
    for i in range(n):
        frame(5 + i)
        if "f" in what:
            r.download_async(name, f4[i & 1])
        if "u" in what:
            r.download_async("rgba8", u32[i & 1])
    r.sync()
    r.download_wait()
    return (time.perf_counter() - t) / n * 1e3


out = {"workload": wl, "snapshot_kernel": os.environ.get("ORB_DL_SNAPSHOT_KERNEL", "default")}
for what in ("", "u", "f", "fu", ""):
    out["ms_per_frame[%s]" % (what or "render only")] = round(run(what), 4)
print(json.dumps(out))

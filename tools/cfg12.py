"""BASELINE configs[0] and configs[1], measured on the GPU box.

cfg1  ratio estimator on a saved 512x512 G-buffer, one quad light: the inputs (ltc, stoDirect, stoNoVis as
      float32[512,512,3], save.py's layout) come from a RATIO frame of the floor + blocker + one quad scene;
      timed: the reference's expression  ltc * (stoDirect / stoNoVis)  (ltc_ratio_estimator.py:8) in numpy on
      one host thread, with and without EXR file I/O (cv2; imageio is not installed), against the same arrays
      through orb_ratio_estimator_host (host buffers, copies included) and the device-resident kernel.
      Results must be bit-identical.
cfg2  Cornell box 1024x1024, one quad light (2 TriLights), RATIO with 1 spp, accumId 0..15: frame time, Mrays/s
      and the HBM fraction of the LTC pass.
Writes one JSON object to stdout."""
import json
import os
import sys
import tempfile
import time

os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")
import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optix_renderer_b200 import api, scenes

PEAK = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"] \
    if os.path.exists(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) else 6650.0
out = {"hbm_peak_gbs": PEAK}

# ---------------------------------------------------------------- cfg1
sc = scenes.scene_floor_blocker(512, 512)
r = api.Renderer(sc)
r.render(api.RATIO, accum_id=0, mode=api.INTENDED, shadow_spp=4)
flip = lambda a: np.ascontiguousarray(np.flipud(a)[..., :3])  # save.py: reshape, flipud, drop w
ltc, sd, sn = flip(r.download("ltc")), flip(r.download("sto_direct")), flip(r.download("sto_no_vis"))
n = 30
with np.errstate(all="ignore"):
    want = ltc * (sd / sn)
    t = time.perf_counter()
    for _ in range(n):
        want = ltc * (sd / sn)
    cpu_ms = (time.perf_counter() - t) / n * 1e3
io_ms = None
try:
    import cv2

    d = tempfile.mkdtemp()
    for name, a in (("ltc", ltc), ("stoDirect", sd), ("stoNoVis", sn)):
        cv2.imwrite(os.path.join(d, name + ".exr"), a[..., ::-1])
    t = time.perf_counter()
    for _ in range(5):
        a, b, c = (cv2.imread(os.path.join(d, k + ".exr"), cv2.IMREAD_UNCHANGED)[..., ::-1] for k in ("ltc", "stoDirect", "stoNoVis"))
        with np.errstate(all="ignore"):
            res = a * (b / c)
        cv2.imwrite(os.path.join(d, "final.exr"), res[..., ::-1])
    io_ms = (time.perf_counter() - t) / 5 * 1e3
except Exception as e:  # EXR codec missing
    io_ms = "unavailable: %s" % e
got = r.ratio_estimator_host(ltc, sd, sn)
t = time.perf_counter()
for _ in range(n):
    got = r.ratio_estimator_host(ltc, sd, sn)
host_ms = (time.perf_counter() - t) / n * 1e3
r.set_profiling(True)
r.stage_times()
for _ in range(50):
    r.ratio_combine(use_denoised=False)
r.sync()
dev_ms = r.stage_times()["combine"][0] / 50
r.set_profiling(False)
out["cfg1"] = {"what": "ltc * (stoDirect / stoNoVis) on float32[512,512,3] (ltc_ratio_estimator.py:8)",
               "cpu_numpy_ms": cpu_ms, "cpu_cores": 1, "cpu_numpy_with_exr_io_ms": io_ms,
               "gpu_host_buffers_ms": host_ms, "gpu_device_resident_ms": dev_ms,
               "gpu_device_gbs": 512 * 512 * 64 / (dev_ms * 1e-3) / 1e9,
               "note": "3 MB working set: L2-resident on the GPU, launch-bound (a 512x512 pass is far below the HBM roof)",
               "bit_identical_to_numpy": bool(np.array_equal(got, want, equal_nan=True))}
r.close()

# ---------------------------------------------------------------- cfg2
sc = scenes.scene_cornell(1024, 1024)
r = api.Renderer(sc)
for i in range(4):
    r.render(api.RATIO, accum_id=i, mode=api.INTENDED, shadow_spp=1)
r.set_profiling(True)
r.stage_times()
c0 = r.counters()
t = time.perf_counter()
for i in range(16):
    r.render(api.RATIO, accum_id=i, mode=api.INTENDED, shadow_spp=1, sync=False)
r.sync()
wall_ms = (time.perf_counter() - t) / 16 * 1e3
st = r.stage_times()
c1 = r.counters()
for _ in range(20):
    r.ltc_gbuffer_pass()
r.sync()
ltc_ms = r.stage_times()["ltc_gbuffer"][0] / 20
rays = (c1["rays_primary"] - c0["rays_primary"]) + (c1["rays_shadow_total"] - c0["rays_shadow_total"])
active = (c1["active_pixels_total"] - c0["active_pixels_total"]) / 16
px = 1024 * 1024
b = px * 32 + active * 48
out["cfg2"] = {"what": "Cornell box 1024x1024, one quad light, RATIO intended 1 spp, accumId 0..15",
               "ms_per_frame": wall_ms, "frames_per_s": 1e3 / wall_ms, "mrays_per_s": rays / 16 / (wall_ms * 1e-3) / 1e6,
               "stages_ms": {k: v[0] / 16 for k, v in st.items() if v[1]}, "active_pixels": active,
               "ltc_gbuffer_pass": {"ms": ltc_ms, "algorithmic_bytes": b, "gbs": b / (ltc_ms * 1e-3) / 1e9,
                                    "frac_of_hbm_peak": b / (ltc_ms * 1e-3) / 1e9 / PEAK},
               "nan_pixels": c1["nan_pixels"]}
r.close()
print(json.dumps(out))

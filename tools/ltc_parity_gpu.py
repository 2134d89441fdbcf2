"""GPU diagnostic: LTC radiance of the CUDA path against the oracle on three scenes -- fraction of
pixels within 1e-4 relative, p99.9 and max relative error, and the LTC stage time on the bench scene."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _oracle
import _parity as P
from optix_renderer_b200 import api, scenes

kind = "reference-intended" if _oracle.available("reference-intended") else "restatement"
o = _oracle.Oracle(kind)
out = {"lib": os.environ.get("ORB200_LIB", "liborb200.so"), "oracle": kind}
for name, sc, W, H in (("synthetic20k", scenes.scene_synthetic(20000, 16, 192, 128), 192, 128),
                       ("cornell", scenes.scene_cornell(128, 128), 128, 128),
                       ("floor", scenes.scene_floor_blocker(128, 128), 128, 128)):
    o.load_scene(sc)
    o.set_options(1, 4, 1)
    ref = o.render(W, H, _oracle.LTC_BASELINE, 0, nthreads=os.cpu_count())
    r = api.Renderer(sc)
    r.render(api.LTC_BASELINE, accum_id=0, mode=1)
    got = r.download("accum")
    prim, _, _ = r.hit_records()
    same = prim == ref["hit_prim"]
    e = P.rel_err(got, ref["accum"])[same]
    e = e[np.isfinite(e)]
    out[name] = {"pixels": int(same.sum()), "hit_mismatch": int((~same).sum()), "frac_le_1e-4": float((e <= 1e-4).mean()),
                 "frac_le_1e-5": float((e <= 1e-5).mean()), "p99.9": float(np.quantile(e, 0.999)), "max": float(e.max())}
    r.close()
sc = scenes.scene_synthetic(1_000_000, 16, 1920, 1080)
r = api.Renderer(sc)
for i in range(3):
    r.render(api.LTC_BASELINE, accum_id=i, mode=1)
r.set_profiling(True)
r.stage_times()
for i in range(10):
    r.render(api.LTC_BASELINE, accum_id=3 + i, mode=1, sync=False)
st = r.stage_times()
out["ltc_ms_1080p_32lights"] = st["ltc"][0] / 10
out["primary_ms"] = st["primary"][0] / 10
print(json.dumps(out))

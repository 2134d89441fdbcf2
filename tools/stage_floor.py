"""GPU diagnostic: stage times of a RATIO / PATH frame restricted to a few rows of the bench scene --
what a stage costs when it has almost no work (launch floors, tails)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optix_renderer_b200 import api, scenes

sc = scenes.scene_synthetic(1_000_000, 16, 1920, 1080)
r = api.Renderer(sc)
out = {}
for rows in ((135, 270) if os.environ.get("FLOOR_BANDS") else (1, 4, 16, 64, 256, 1080)):
    y0 = 300
    band = (y0, min(1080, y0 + rows)) if rows < 1080 else (0, 1080)
    for rt, name in ((api.RATIO, "ratio"), (api.PATH, "path")):
        for i in range(3):
            r.render(rt, accum_id=i, mode=1, shadow_spp=4, max_depth=8, rows=band)
        r.set_profiling(True)
        r.stage_times()
        for i in range(10):
            r.render(rt, accum_id=3 + i, mode=1, shadow_spp=4, max_depth=8, rows=band, sync=False)
        st = r.stage_times()
        r.set_profiling(False)
        c = r.counters()
        out["%s_rows%d" % (name, rows)] = {"active": c["active_pixels"], **{k: round(v[0] / 10, 4) for k, v in st.items() if v[1]}}
print(json.dumps(out, indent=1))

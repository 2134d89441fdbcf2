"""One GPU playing ONE RANK of an 8-way row-band partition: its 135-row band of the bench frame (RATIO 4 spp + denoise +
combine, or PATH 8 bounces) rendered back to back with n frame slots -- frames per second of the pipelined band, the
regime that limits strong scaling.  Knobs through the environment, one JSON line per setting:
    python tools/ab_band.py 'ORB_STREAM_BLOCK=64' 'ORB_STREAM_BLOCK=32 ORB_TRACE_CTAS_PER_SM=6' ...   (SLOTS=n in a spec)"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optix_renderer_b200 import api, scenes

sc = scenes.scene_synthetic(1_000_000, 16, 1920, 1080)
base = dict(os.environ)
band = (472, 607)
for spec in (sys.argv[1:] or [""]):
    os.environ.clear(); os.environ.update(base)
    slots = 2
    for kv in spec.split():
        k, v = kv.split("=")
        if k == "SLOTS":
            slots = int(v)
        elif k == "BAND":  # BAND=y0:y1 (default: the 135 rows an 8-way partition gives one rank)
            band = tuple(int(x) for x in v.split(":"))
        else:
            os.environ[k] = v
    r = api.Renderer(sc, quantise_lut=1)
    r.set_frame_slots(slots)
    out = {"knobs": spec, "slots": slots, "band": list(band), "lib": os.environ.get("ORB200_LIB", "default")[-14:]}
    for wl in ("ratio", "path"):
        def frame(i):
            if wl == "ratio":
                r.render(api.RATIO, accum_id=i, mode=1, shadow_spp=4, rows=band, sync=False)
                r.denoise(iterations=5, rows=band)
                r.ratio_combine(use_denoised=True, rows=band)
            else:
                r.render(api.PATH, accum_id=i, mode=1, max_depth=8, rows=band, sync=False)
        for i in range(8):
            frame(i)
        r.sync()
        n = 120
        t = time.perf_counter()
        for i in range(n):
            frame(8 + i)
        r.sync()
        out[wl + "_ms_per_frame"] = round((time.perf_counter() - t) * 1e3 / n, 4)
    print(json.dumps(out), flush=True)
    r.close()

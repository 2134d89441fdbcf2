"""Refit against rebuild on the bench scene (1 M triangles, 32 TriLights): device time of orb_scene_commit's
build, of orb_scene_refit after the lights and two surface patches moved, and the RATIO frame time over the
refitted tree against a tree built from scratch for the moved scene.  Writes gpurun_out/<label>_refit.json.
usage: python tools/refit_bench.py [label] [n_tris]"""
import copy
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optix_renderer_b200 import api, scenes  # noqa: E402

label = sys.argv[1] if len(sys.argv) > 1 else "r1q"
n_tris = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000


def moved(sc, amount):
    mv = copy.deepcopy(sc)
    c, s = np.float32(np.cos(0.35 * amount)), np.float32(np.sin(0.35 * amount))
    rot = np.array([[1, 0, 0], [0, c, -s], [0, s, c]], np.float32)
    changed = []
    for k, m in enumerate(mv["meshes"]):
        if m["is_light"]:
            ctr = m["vtx"].mean(0, dtype=np.float32)
            m["vtx"] = ((m["vtx"] - ctr) @ rot.T + ctr + amount * np.array([0.4, -0.3, -1.7], np.float32)).astype(np.float32)
            m["nrm"] = (m["nrm"] @ rot.T).astype(np.float32)
            changed.append(k)
        elif k in (0, 27):
            v = m["vtx"].copy()
            v[:, 2] += (amount * 0.6 * np.sin(1.9 * v[:, 0] + 0.7) * np.cos(1.1 * v[:, 1])).astype(np.float32)
            m["vtx"] = v
            changed.append(k)
    mv["tri_lights"] = scenes.make_tri_lights(mv["meshes"])
    return mv, changed


def frame_ms(r, n=30):
    for i in range(5):
        r.render(api.RATIO, accum_id=i, mode=1, shadow_spp=4)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(n):
        r.render(api.RATIO, accum_id=5 + i, mode=1, shadow_spp=4)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3


sc = scenes.scene_synthetic(n_tris=n_tris, n_quad_lights=16, width=1920, height=1080)
r = api.Renderer(sc)
out = {"triangles": scenes.n_triangles(sc), "build_ms": r.counters()["bvh_build_us"] / 1e3, "frame_ms_built": frame_ms(r)}
mv, changed = moved(sc, 1.0)
t0 = time.perf_counter()
for k in changed:
    m = mv["meshes"][k]
    r.update_mesh(k, m["vtx"], m["nrm"], m["idx"])
t1 = time.perf_counter()
r.refit()
t2 = time.perf_counter()
r.set_lights(mv["tri_lights"])
out["meshes_moved"] = len(changed)
out["triangles_moved"] = int(sum(mv["meshes"][k]["idx"].shape[0] for k in changed))
out["update_mesh_host_ms"] = (t1 - t0) * 1e3
out["refit_ms_device"] = r.counters()["bvh_refit_us"] / 1e3
out["refit_ms_wall"] = (t2 - t1) * 1e3
out["frame_ms_refitted"] = frame_ms(r)


def clean_frame(q):
    # rayGen leaves ltc / sto_* of miss pixels untouched (deviceCode.cu:91-95): start both contexts from zeros
    z = np.zeros((q.H, q.W, 4), np.float32)
    for k in ("ltc", "sto_direct", "sto_no_vis"):
        q.upload(k, z)
    q.render(api.RATIO, accum_id=99, mode=1, shadow_spp=4)
    return {k: q.download(k) for k in ("ltc", "sto_direct", "sto_no_vis", "position")}


a = clean_frame(r)
fresh = api.Renderer(mv)
out["rebuild_ms"] = fresh.counters()["bvh_build_us"] / 1e3
out["frame_ms_rebuilt"] = frame_ms(fresh)
b = clean_frame(fresh)
out["frames_bit_identical"] = bool(all(np.array_equal(a[k], b[k], equal_nan=True) for k in a))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/%s_refit.json" % label, "w"))
print(json.dumps(out))

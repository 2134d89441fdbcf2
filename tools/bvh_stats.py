"""GPU diagnostic: BVH traversal work per ray on the bench scene (instrumented trace through the
C-ABI, orb_trace_rays_host): nodes visited and triangles tested for camera rays and for
shadow-like rays (surface point -> random point on a random light)."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optix_renderer_b200 import api, scenes

leaf = int(sys.argv[1]) if len(sys.argv) > 1 else 0
sc = scenes.scene_synthetic(1_000_000, 16, 1920, 1080)
r = api.Renderer(sc, max_leaf_tris=leaf)
r.render(api.NORMALS)
pos, nrm = r.download("position")[..., :3], r.download("normal")[..., :3]
prim, _, _ = r.hit_records()
rng = np.random.default_rng(0)
ys, xs = rng.integers(0, 1080, 400000), rng.integers(0, 1920, 400000)
m = prim[ys, xs] >= 0
ys, xs = ys[m], xs[m]
cam = np.asarray(sc["camera"]["pos"], np.float32)
p = pos[ys, xs]
d = p - cam
d /= np.linalg.norm(d, axis=1, keepdims=True)
out = {"max_leaf_tris": leaf, "counters": {k: r.counters()[k] for k in ("bvh_nodes", "bvh_triangles", "bvh_height", "bvh_build_us")}}
r.trace_rays(np.repeat(cam[None], len(d), 0), d)
c = r.counters()
out["camera_rays"] = {"n": len(d), "nodes_per_ray": c["nodes_visited"] / len(d), "tris_per_ray": c["tris_tested"] / len(d)}
L = sc["tri_lights"].reshape(-1, 19)
li = rng.integers(0, len(L), len(p))
u = rng.uniform(0, 1, (len(p), 2)).astype(np.float32)
su = np.sqrt(u[:, :1])
q = (1 - su) * L[li, 0:3] + su * ((1 - u[:, 1:]) * L[li, 3:6] + u[:, 1:] * L[li, 6:9])
o = p + 1e-3 * nrm[ys, xs]
d2 = q - o
d2 /= np.linalg.norm(d2, axis=1, keepdims=True)
r.trace_rays(o, d2)
c = r.counters()
out["shadow_rays"] = {"n": len(p), "nodes_per_ray": c["nodes_visited"] / len(p), "tris_per_ray": c["tris_tested"] / len(p)}
print(json.dumps(out))

"""Scene construction on the host: the data contracts either side of the hot path.

Everything here mirrors what the reference's host code hands to OWL:
  * un-indexed meshes, 3 vertices per face, index = (3k, 3k+1, 3k+2)  -- src/Model.cpp:186-201
  * light meshes appended after the surface meshes with isLight = true, materialID = 0
                                                                        -- src/scene.cpp:5-13
  * the TriLight list                                                   -- include/viewer.hpp:236-265
  * the camera vectors                                                  -- include/viewer.hpp:635-641
  * the 8x8 LTC lookup table                                            -- include/ltc/ltc_isotropic.h
A scene is a plain dict: {"meshes": [...], "tri_lights": (n,19) f32, "lut": (m1,m2,m3),
"camera": {"pos","d00","du","dv"}, "width", "height"}.
"""
import os

import numpy as np

F = np.float32
_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")


def load_ltc_lut():
    """(m1, m2, m3), each (8, 8, 4) float32 -- rows = alpha, columns = theta (viewer.hpp:321-327)."""
    a = np.load(os.path.join(_DATA, "ltc_isotropic_8x8.npy"))
    return a[0].copy(), a[1].copy(), a[2].copy()


def upsample_lut(lut, n):
    """Bilinear (texture-unit style, texel centres at +0.5, clamp) resampling of the LUT to n x n.
    Only for the synthetic 64x64 perf variant named in BASELINE.json; parity runs use 8x8."""
    out = []
    for m in lut:
        h, w = m.shape[:2]
        xs = (np.arange(n, dtype=F) + F(0.5)) / F(n)
        xb = xs * F(w) - F(0.5)
        yb = xs * F(h) - F(0.5)
        i0 = np.floor(xb).astype(int)
        j0 = np.floor(yb).astype(int)
        a = (xb - i0).astype(F)[None, :, None]
        b = (yb - j0).astype(F)[:, None, None]
        i1 = np.clip(i0 + 1, 0, w - 1)
        j1 = np.clip(j0 + 1, 0, h - 1)
        i0 = np.clip(i0, 0, w - 1)
        j0 = np.clip(j0, 0, h - 1)
        r = ((1 - a) * (1 - b) * m[j0][:, i0] + a * (1 - b) * m[j0][:, i1] + (1 - a) * b * m[j1][:, i0]
             + a * b * m[j1][:, i1])
        out.append(np.ascontiguousarray(r, dtype=F))
    return tuple(out)


def _norm(v):
    """owl's normalize: v * (1 / sqrt(dot(v, v))), the dot product added left to right in fp32 (np.dot goes through BLAS,
    whose summation order / FMA use is not defined: one ulp off the C++ host's camera for some inputs)."""
    v = np.asarray(v, F)
    d = F(F(F(v[0] * v[0]) + F(v[1] * v[1])) + F(v[2] * v[2]))
    return (v * (F(1) / np.sqrt(d, dtype=F))).astype(F)


def _cross(a, b):
    """cross(a, b) in fp32, each component one multiply-subtract pair as the C++ host writes it."""
    a, b = np.asarray(a, F), np.asarray(b, F)
    return np.array([F(F(a[1] * b[2]) - F(a[2] * b[1])), F(F(a[2] * b[0]) - F(a[0] * b[2])), F(F(a[0] * b[1]) - F(a[1] * b[0]))], F)


def camera_from_lookat(frm, at, up, cos_fovy, width, height):
    """Viewer::cameraChanged, include/viewer.hpp:635-641, in fp32 and in the same order."""
    frm, at, up = np.asarray(frm, F), np.asarray(at, F), np.asarray(up, F)
    d00 = _norm(at - frm)
    aspect = F(width / float(height))
    ddu = (F(F(cos_fovy) * aspect) * _norm(_cross(d00, up))).astype(F)
    ddv = (F(cos_fovy) * _norm(_cross(ddu, d00))).astype(F)
    d00 = (d00 - F(0.5) * ddu).astype(F)
    d00 = (d00 - F(0.5) * ddv).astype(F)
    return {"pos": frm.copy(), "d00": d00, "du": ddu, "dv": ddv,
            "lookat": {"from": [float(x) for x in frm], "to": [float(x) for x in at], "up": [float(x) for x in up],
                       "cos_fovy": float(cos_fovy)}}


def make_mesh(tri_pos, tri_nrm, tri_uv, diffuse, alpha, emit=(0, 0, 0), is_light=False, material_id=1, tex=None):
    """tri_*: (nTri, 3, k) per-face corner attributes -> the reference's un-indexed mesh."""
    tri_pos = np.asarray(tri_pos, F)
    n = tri_pos.shape[0]
    return {
        "vtx": np.ascontiguousarray(tri_pos.reshape(-1, 3)),
        "nrm": np.ascontiguousarray(np.asarray(tri_nrm, F).reshape(-1, 3)),
        "uv": np.ascontiguousarray(np.asarray(tri_uv, F).reshape(-1, 2)),
        "idx": np.arange(3 * n, dtype=np.int32).reshape(n, 3),
        "diffuse": np.asarray(diffuse, F),
        "alpha": float(alpha),
        "emit": np.asarray(emit, F),
        "is_light": bool(is_light),
        "material_id": int(material_id),
        "tex": tex,
    }


def quad_tris(a, b, c, d, normal):
    """Two triangles (a,b,c), (a,c,d) with a constant vertex normal and unit-square uvs."""
    a, b, c, d = [np.asarray(x, F) for x in (a, b, c, d)]
    pos = np.array([[a, b, c], [a, c, d]], F)
    nrm = np.broadcast_to(np.asarray(normal, F), (2, 3, 3)).copy()
    uv = np.array([[[0, 0], [1, 0], [1, 1]], [[0, 0], [1, 1], [0, 1]]], F)
    return pos, nrm, uv


def quad_mesh(a, b, c, d, normal, **kw):
    return make_mesh(*quad_tris(a, b, c, d, normal), **kw)


def make_tri_lights(meshes):
    """include/viewer.hpp:236-265: one TriLight (19 floats) per triangle of every light mesh."""
    out = []
    for m in meshes:
        if not m["is_light"]:
            continue
        v, nr = m["vtx"], m["nrm"]
        for i0, i1, i2 in m["idx"]:
            v1, v2, v3 = v[i0], v[i1], v[i2]
            cg = ((v1 + v2 + v3) / F(3)).astype(F)
            nsum = (nr[i0] + nr[i1] + nr[i2]).astype(F)
            normal = _norm(nsum)
            cr = np.cross((v1 - v2).astype(F), (v3 - v2).astype(F)).astype(F)
            area = F(0.5) * np.sqrt(np.dot(cr, cr), dtype=F)
            out.append(np.concatenate([v1, v2, v3, cg, normal, m["emit"], [area]]).astype(F))
    return np.array(out, F).reshape(-1, 19)


def finish(meshes, camera, width, height, lut=None):
    meshes = [m for m in meshes if not m["is_light"]] + [m for m in meshes if m["is_light"]]
    for m in meshes:
        if m["is_light"]:
            m["material_id"] = 0  # src/scene.cpp:10
    return {"meshes": meshes, "tri_lights": make_tri_lights(meshes), "lut": lut or load_ltc_lut(),
            "camera": camera, "width": width, "height": height}


def n_triangles(scene):
    return int(sum(m["idx"].shape[0] for m in scene["meshes"]))


# ------------------------------------------------------------------------------------------------
def scene_floor_blocker(width=64, height=64):
    """The tiny z-up probe scene of SURVEY.md Appendix B: floor, blocker, one quad light."""
    up = (0, 0, 1)
    floor = quad_mesh((-2, -2, 0), (2, -2, 0), (2, 2, 0), (-2, 2, 0), up, diffuse=(.8, .6, .4), alpha=.3,
                      material_id=1)
    blocker = quad_mesh((-.3, -.3, .5), (.3, -.3, .5), (.3, .3, .5), (-.3, .3, .5), up, diffuse=(.2, .7, .2),
                        alpha=.5, material_id=2)
    light = quad_mesh((-.5, -.5, 1), (-.5, .5, 1), (.5, .5, 1), (.5, -.5, 1), (0, 0, -1), diffuse=(0, 0, 0),
                      alpha=1., emit=(10, 10, 10), is_light=True, material_id=0)
    cam = camera_from_lookat((3, -3, 2.5), (0, 0, .3), (0, 0, 1), .66, width, height)
    return finish([floor, blocker, light], cam, width, height)


def checker_texture(w, h, seed=3):
    """(h, w, 4) uint8 RGBA: every texel a different colour (a wrong texel pick changes the albedo), with texels at
    both ends of the unorm8 range so the /255 scale is visible, and an alpha channel that must be ignored."""
    rng = np.random.default_rng(seed)
    t = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    t[0, 0] = (0, 255, 1, 7)
    t[-1, -1] = (255, 0, 254, 200)
    return t


def scene_textured(width=48, height=40):
    """The floor / blocker / light scene with RGBA8 diffuse textures (hit_miss.cuh:39-43; nearest, clamp,
    include/viewer.hpp:436-447): the floor's texture coordinates run from -1.3 to 2.4, so abs(fmodf(uv, 1))
    wraps, mirrors the negative part and lands on texel borders; the blocker carries a 3x2 texture."""
    up = (0, 0, 1)
    fp, fn, fuv = quad_tris((-2, -2, 0), (2, -2, 0), (2, 2, 0), (-2, 2, 0), up)
    fuv = (fuv * F(3.7) - F(1.3)).astype(F)
    floor = make_mesh(fp, fn, fuv, diffuse=(.8, .6, .4), alpha=.3, material_id=1, tex=checker_texture(8, 4, 3))
    bp, bn, buv = quad_tris((-.3, -.3, .5), (.3, -.3, .5), (.3, .3, .5), (-.3, .3, .5), up)
    blocker = make_mesh(bp, bn, buv, diffuse=(.2, .7, .2), alpha=.5, material_id=2, tex=checker_texture(3, 2, 4))
    wall = quad_mesh((-2, 2, 0), (2, 2, 0), (2, 2, 1.5), (-2, 2, 1.5), (0, -1, 0), diffuse=(.5, .5, .7), alpha=.7,
                     material_id=3)  # untextured beside textured meshes
    light = quad_mesh((-.5, -.5, 1), (-.5, .5, 1), (.5, .5, 1), (.5, -.5, 1), (0, 0, -1), diffuse=(0, 0, 0),
                      alpha=1., emit=(10, 10, 10), is_light=True, material_id=0)
    cam = camera_from_lookat((3, -3, 2.5), (0, 0, .3), (0, 0, 1), .66, width, height)
    return finish([floor, blocker, wall, light], cam, width, height)


def _box_meshes(lo, hi, rot_deg, **kw):
    """Axis-aligned box [lo,hi] rotated about the y axis through its centre: 5 visible faces + bottom."""
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    c = (lo + hi) / 2
    t = np.deg2rad(rot_deg)
    R = np.array([[np.cos(t), 0, np.sin(t)], [0, 1, 0], [-np.sin(t), 0, np.cos(t)]])

    def P(x, y, z):
        return (R @ (np.array([x, y, z]) - c) + c).astype(F)

    x0, y0, z0 = lo
    x1, y1, z1 = hi
    faces = [
        ((x0, y1, z0), (x0, y1, z1), (x1, y1, z1), (x1, y1, z0), (0, 1, 0)),
        ((x0, y0, z0), (x1, y0, z0), (x1, y0, z1), (x0, y0, z1), (0, -1, 0)),
        ((x0, y0, z0), (x0, y0, z1), (x0, y1, z1), (x0, y1, z0), (-1, 0, 0)),
        ((x1, y0, z0), (x1, y1, z0), (x1, y1, z1), (x1, y0, z1), (1, 0, 0)),
        ((x0, y0, z0), (x0, y1, z0), (x1, y1, z0), (x1, y0, z0), (0, 0, -1)),
        ((x0, y0, z1), (x1, y0, z1), (x1, y1, z1), (x0, y1, z1), (0, 0, 1)),
    ]
    pos, nrm, uv = [], [], []
    for a, b, cc, d, n in faces:
        p, nn, u = quad_tris(P(*a), P(*b), P(*cc), P(*d), (R @ np.array(n, np.float64)).astype(F))
        pos.append(p)
        nrm.append(nn)
        uv.append(u)
    return make_mesh(np.concatenate(pos), np.concatenate(nrm), np.concatenate(uv), **kw)


def scene_cornell(width=1024, height=1024):
    """BASELINE.json config 2: Cornell box (classic 555-unit dimensions scaled to metres),
    one quad ceiling light (2 TriLights).  y-up, open towards -z, camera outside the opening."""
    s = 1.0 / 555.0
    W = 1.0
    white, red, green = (.73, .73, .73), (.65, .05, .05), (.12, .45, .15)
    meshes = [
        quad_mesh((0, 0, 0), (W, 0, 0), (W, 0, W), (0, 0, W), (0, 1, 0), diffuse=white, alpha=.6, material_id=1),
        quad_mesh((0, W, 0), (0, W, W), (W, W, W), (W, W, 0), (0, -1, 0), diffuse=white, alpha=.6, material_id=1),
        quad_mesh((0, 0, W), (W, 0, W), (W, W, W), (0, W, W), (0, 0, -1), diffuse=white, alpha=.6, material_id=1),
        quad_mesh((0, 0, 0), (0, 0, W), (0, W, W), (0, W, 0), (1, 0, 0), diffuse=red, alpha=.5, material_id=2),
        quad_mesh((W, 0, 0), (W, W, 0), (W, W, W), (W, 0, W), (-1, 0, 0), diffuse=green, alpha=.5, material_id=3),
        _box_meshes((130 * s, 0, 65 * s), (295 * s, 165 * s, 230 * s), -18, diffuse=white, alpha=.25,
                    material_id=4),
        _box_meshes((265 * s, 0, 295 * s), (430 * s, 330 * s, 460 * s), 15, diffuse=white, alpha=.1, material_id=5),
        quad_mesh((213 * s, W - 1e-3, 227 * s), (343 * s, W - 1e-3, 227 * s), (343 * s, W - 1e-3, 332 * s),
                  (213 * s, W - 1e-3, 332 * s), (0, -1, 0), diffuse=(0, 0, 0), alpha=1., emit=(17, 12, 4),
                  is_light=True, material_id=0),
    ]
    cam = camera_from_lookat((.5, .5, -1.44), (.5, .5, 0), (0, 1, 0), .66, width, height)
    return finish(meshes, cam, width, height)


def scene_synthetic(n_tris=1_000_000, n_quad_lights=16, width=1920, height=1080, seed=1234, patches=8):
    """BASELINE.json configs 3-5: deterministic height-field soup of ~n_tris triangles split into
    patches x patches meshes (alpha in U[0.05,1], diffuse in U[0.2,0.9]^3 per mesh), smooth vertex
    normals, with a sqrt(n) x sqrt(n) grid of constant-colour quad lights above (the reference has
    no textured lights: one emit per light mesh, cuda_include/common.cuh:64-72).  z-up."""
    rng = np.random.default_rng(seed)
    n = max(patches, int(round(np.sqrt(n_tris / 2.0))))
    n = (n // patches) * patches
    xs = np.linspace(-10, 10, n + 1)
    X, Y = np.meshgrid(xs, xs, indexing="xy")
    ph = rng.uniform(0, 2 * np.pi, 6)
    Z = (0.9 * np.sin(0.55 * X + ph[0]) * np.cos(0.45 * Y + ph[1]) + 0.35 * np.sin(1.7 * X + ph[2])
         * np.sin(1.3 * Y + ph[3]) + 0.08 * np.sin(6.1 * X + ph[4]) * np.cos(5.3 * Y + ph[5]))
    # box-like towers: raise a few random rectangular plateaus
    for _ in range(24):
        cx, cy = rng.uniform(-8, 8, 2)
        sx, sy = rng.uniform(0.3, 1.0, 2)
        hgt = rng.uniform(0.5, 2.5)
        Z = np.where((np.abs(X - cx) < sx) & (np.abs(Y - cy) < sy), Z + hgt, Z)
    P = np.stack([X, Y, Z], -1)
    # smooth normals from central differences of the height field
    dzdx = np.gradient(Z, xs, axis=1)
    dzdy = np.gradient(Z, xs, axis=0)
    N = np.stack([-dzdx, -dzdy, np.ones_like(Z)], -1)
    N /= np.linalg.norm(N, axis=-1, keepdims=True)
    UV = np.stack([(X + 10) / 20 * 8, (Y + 10) / 20 * 8], -1)
    meshes = []
    step = n // patches
    mid = 1
    for py in range(patches):
        for px in range(patches):
            j0, i0 = py * step, px * step
            sl = (slice(j0, j0 + step), slice(i0, i0 + step))
            sl_r = (slice(j0, j0 + step), slice(i0 + 1, i0 + step + 1))
            sl_u = (slice(j0 + 1, j0 + step + 1), slice(i0, i0 + step))
            sl_ur = (slice(j0 + 1, j0 + step + 1), slice(i0 + 1, i0 + step + 1))

            def corners(A):
                a, b, c, d = A[sl], A[sl_r], A[sl_ur], A[sl_u]
                t1 = np.stack([a, b, c], -2)
                t2 = np.stack([a, c, d], -2)
                return np.stack([t1, t2], 2).reshape(-1, 3, A.shape[-1])

            meshes.append(make_mesh(corners(P), corners(N), corners(UV), diffuse=rng.uniform(0.2, 0.9, 3),
                                    alpha=rng.uniform(0.05, 1.0), material_id=mid))
            mid += 1
    g = max(1, int(round(np.sqrt(n_quad_lights))))
    cnt = 0
    for ly in range(g):
        for lx in range(g):
            if cnt >= n_quad_lights:
                break
            cx = -7.5 + 15.0 * (lx + 0.5) / g
            cy = -7.5 + 15.0 * (ly + 0.5) / g
            h, z = 0.6, 6.0
            meshes.append(quad_mesh((cx - h, cy - h, z), (cx - h, cy + h, z), (cx + h, cy + h, z), (cx + h, cy - h, z),
                                    (0, 0, -1), diffuse=(0, 0, 0), alpha=1., emit=rng.uniform(1, 10, 3),
                                    is_light=True, material_id=0))
            cnt += 1
    cam = camera_from_lookat((13, -13, 9), (0, 0, 0.5), (0, 0, 1), .66, width, height)
    return finish(meshes, cam, width, height)


# ------------------------------------------------------------------------------------------------
# Reference file formats: scene JSON (src/scene.cpp:38-98) + OBJ/MTL as src/Model.cpp:130-252 reads
# them.  Used to feed the headless C++ host (host/orb_render) the same scenes the Python harness builds.
# ------------------------------------------------------------------------------------------------
def _g(x):
    return "%.9g" % float(x)


def _write_obj(path, meshes, light_file):
    """One `o` shape + one material per mesh; material k of the .mtl gets materialID k + 1 in the
    loader (Model.cpp:222), so surface materials are written at index material_id - 1."""
    mtl = os.path.splitext(path)[0] + ".mtl"
    if light_file:
        mats = list(range(len(meshes)))
        slot = {i: i for i in mats}
    else:
        n = max(m["material_id"] for m in meshes)
        mats = [None] * n
        slot = {}
        for i, m in enumerate(meshes):
            k = m["material_id"] - 1
            assert k >= 0, "surface meshes need material_id >= 1"
            if mats[k] is not None:
                o = meshes[mats[k]]
                assert np.array_equal(o["diffuse"], m["diffuse"]) and o["alpha"] == m["alpha"], "material id reused"
            else:
                mats[k] = i
            slot[i] = k
    with open(mtl, "w") as f:
        for k, i in enumerate(mats):
            f.write("newmtl m%d\n" % k)
            if i is None:
                f.write("Kd 0 0 0\nNs 1\nKe 0 0 0\n")
                continue
            m = meshes[i]
            f.write("Kd %s %s %s\nNs %s\nKe %s %s %s\n" % (*map(_g, m["diffuse"]), _g(m["alpha"]), *map(_g, m["emit"])))
            if m.get("tex") is not None:
                tex = np.asarray(m["tex"], np.uint8)
                name = os.path.basename(os.path.splitext(path)[0]) + "_tex%d.ppm" % k
                with open(os.path.join(os.path.dirname(path), name), "wb") as t:
                    t.write(b"P6\n%d %d\n255\n" % (tex.shape[1], tex.shape[0]))
                    t.write(np.ascontiguousarray(tex[::-1, :, :3]).tobytes())  # file rows top-down; loader mirrors y
                f.write("map_Kd %s\n" % name)
    with open(path, "w") as f:
        f.write("mtllib %s\n" % os.path.basename(mtl))
        base = 0
        for i, m in enumerate(meshes):
            f.write("o mesh%d\nusemtl m%d\n" % (i, slot[i]))
            for v in m["vtx"]:
                f.write("v %s %s %s\n" % tuple(map(_g, v)))
            for v in m["nrm"]:
                f.write("vn %s %s %s\n" % tuple(map(_g, v)))
            for v in m["uv"]:
                f.write("vt %s %s\n" % tuple(map(_g, v)))
            for a, b, c in m["idx"]:
                f.write("f %d/%d/%d %d/%d/%d %d/%d/%d\n" % ((base + a + 1,) * 3 + (base + b + 1,) * 3 + (base + c + 1,) * 3))
            base += len(m["vtx"])


def export_scene(scene, directory, name="scene", spp=1, renderers=(8,)):
    """Write `scene` as name.json + name_surface.obj/.mtl + name_lights.obj/.mtl; returns the JSON path."""
    import json

    os.makedirs(directory, exist_ok=True)
    surf = [m for m in scene["meshes"] if not m["is_light"]]
    lights = [m for m in scene["meshes"] if m["is_light"]]
    s_obj = os.path.join(directory, name + "_surface.obj")
    l_obj = os.path.join(directory, name + "_lights.obj")
    _write_obj(s_obj, surf, False)
    _write_obj(l_obj, lights, True)
    la = scene["camera"]["lookat"]
    doc = {"spp": spp, "width": scene["width"], "height": scene["height"], "renderers": list(renderers),
           "cameras": [{"from": la["from"], "to": la["to"], "up": la["up"], "cos_fovy": la["cos_fovy"]}],
           "surface_geometry": os.path.basename(s_obj), "area_lights": os.path.basename(l_obj)}
    path = os.path.join(directory, name + ".json")
    with open(path, "w") as f:
        json.dump(doc, f, indent=1)
    return path

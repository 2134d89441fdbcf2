"""Generate tests/golden/obj_ingest.json and scene_ingest.json: OBJ / MTL / texture files (and scene JSON files) and what the REFERENCE's own ingest makes of each --
src/Model.cpp loadOBJ / loadTexture with the tiny_obj_loader.h and stb_image.h it vendors, compiled where they lie as
oracle/_ref/ref_obj_dump, and src/scene.cpp parseScene / syncLights as oracle/_ref/ref_scene_dump (oracle/Makefile
target ref_host).  host/scene_io.cpp load_obj is pinned against these
(tests/test_host.py::test_host_obj_ingest_equals_the_reference_loader): mesh order, triangulation, un-indexed
vertex / normal / texcoord arrays bit for bit, material mapping, textures.

Run in the build container (needs /root/reference):
    make -C oracle ref_host && python tests/golden/make_golden_obj.py
"""
import base64
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.path.join(ROOT, "oracle", "_ref", "ref_obj_dump")


def cases():
    rng = np.random.default_rng(77)
    c = {}
    mtl = ("newmtl red\nKd 1 0 0\nNs 0.5\nKe 0 0 0\n"
           "newmtl green\nKd 0 1 0.25\nNs 0.25\nKe 2 3 4\nmap_Kd t.ppm\n"
           "newmtl blue\nKa 0.1 0.1 0.1\nKd 0.125 0.25 0.75\nKs 1 1 1\nNs 96.078431\nd 0.5\nillum 2\nmap_Kd sub\\\\t.ppm\nmap_Ks t.ppm\nmap_bump t.ppm\n")
    ppm = b"P6\n2 3\n255\n" + bytes(rng.integers(0, 256, 18, dtype=np.uint8))
    verts = "".join("v %.6f %.6f %.6f\n" % tuple(v) for v in rng.uniform(-2, 2, (12, 3)))
    norms = "".join("vn %.6f %.6f %.6f\n" % tuple(v) for v in rng.normal(size=(5, 3)))
    uvs = "".join("vt %.6f %.6f\n" % tuple(v) for v in rng.uniform(-1.5, 2.5, (7, 2)))
    head = "mtllib a.mtl\n" + verts + norms + uvs
    c["tri_quad_pentagon"] = {"a.obj": head + "usemtl red\nf 1/1/1 2/2/2 3/3/3\nf 4/4/4 5/5/5 6/6/1 7/7/2\nf 8/1/1 9/2/2 10/3/3 11/4/4 12/5/5\n"}
    c["negative_indices"] = {"a.obj": head + "usemtl green\nf -12/-7/-5 -11/-6/-4 -10/-5/-3\nf -1/-1/-1 -2/-2/-2 -3/-3/-3 -4/-4/-4\n"}
    c["materials_interleaved_in_one_shape"] = {"a.obj": head + "o s\nusemtl green\nf 1/1/1 2/2/1 3/3/1 4/4/1\nusemtl red\nf 5/5/2 6/6/2 7/7/2\n"
                                                               "usemtl green\nf 8/1/3 9/2/3 10/3/3\nusemtl blue\nf 10/1/4 11/2/4 12/3/4\n"}
    c["several_shapes_and_groups"] = {"a.obj": head + "o first\nusemtl blue\nf 1/1/1 2/2/2 3/3/3\ng grp\nusemtl red\nf 4/4/4 5/5/5 6/6/1\n"
                                                      "o second\nusemtl red\nf 7/7/2 8/1/3 9/2/4\nf 10/3/5 11/4/1 12/5/2\n"}
    c["shared_vertices_are_not_merged"] = {"a.obj": head + "usemtl red\nf 1/1/1 2/2/2 3/3/3\nf 1/1/1 3/3/3 4/4/4\nf 1/1/1 4/4/4 5/5/5\n"}
    c["concave_quad_and_hexagon"] = {"a.obj": "mtllib a.mtl\nv 0 0 0\nv 2 0 0\nv 0.5 0.5 0\nv 0 2 0\nv 3 0 0\nv 4 0 0\nv 5 1 0\nv 4 2 0\nv 3 2 0\nv 2.5 1 0\n"
                                              "vn 0 0 1\nvt 0 0\nvt 1 0\nvt 1 1\nvt 0 1\nusemtl red\nf 1/1/1 2/2/1 3/3/1 4/4/1\nf 5/1/1 6/2/1 7/3/1 8/4/1 9/1/1 10/2/1\n"}
    c["crlf_tabs_comments_scientific"] = {"a.obj": "# comment\r\nmtllib a.mtl\r\nv\t1e-3 -2.5E+1 .5\r\nv 1 0 0\r\nv 0 1 0 # trailing\r\nvn 0 0 1\r\nvt 0.25 0.75 0\r\n"
                                                   "usemtl red\r\nf 1/1/1  2/1/1\t3/1/1\r\n"}
    # ear clipping, stressed: many polygons of 4 .. 9 corners -- random star-shaped (concave) outlines in a random plane,
    # non-planar ones, ones with collinear and duplicated corners, a zero-area sliver
    lines, faces, nv = ["mtllib a.mtl", "vn 0 0 1", "vt 0.5 0.5"], [], 0
    for i in range(80):
        n = int(rng.integers(4, 10))
        ang = np.sort(rng.uniform(0, 2 * np.pi, n))
        rad = rng.uniform(0.2, 1.0, n) if i % 3 else np.ones(n)          # every third one convex
        pts = np.stack([rad * np.cos(ang), rad * np.sin(ang), np.zeros(n)], 1)
        if i % 4 == 1:
            pts[:, 2] = rng.normal(0, 0.3, n)                             # non-planar
        if i % 7 == 2:
            pts[1] = (pts[0] + pts[2]) / 2                                # a collinear corner
        if i % 11 == 3:
            pts[2] = pts[1]                                               # a duplicated corner
        if i % 13 == 4:
            pts[:, 1] = 0                                                 # zero area
        q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
        pts = pts @ q.T + rng.uniform(-3, 3, 3)
        if i % 2:
            pts = pts[::-1]                                               # both windings
        lines += ["v %.6f %.6f %.6f" % tuple(v) for v in pts]
        faces.append("f " + " ".join("%d/1/1" % (nv + k + 1) for k in range(n)))
        nv += n
    c["ear_clipping_80_polygons"] = {"a.obj": "\n".join(lines + ["usemtl red"] + faces) + "\n"}
    c["material_defaults_and_extras"] = {
        "a.obj": "mtllib b.mtl\nmtllib a.mtl\nv 0 0 0 1 0.5 0.25\nv 1 0 0 0 1 0\nv 0 1 0 0 0 1\nv 1 1 0\nvn 0 0 1\nvt 0.25 0.5 0.75\n"
                 "s 1\nusemtl bare\nf 1/1/1 2/1/1 3/1/1\ns off\nl 1 2\nf 1/1/1 2/1/1\nusemtl only_ns\nf 2/1/1 4/1/1 3/1/1\n"
                 "usemtl red\nf 1/1/1 4/1/1 3/1/1\n",
        "b.mtl": "# a material with nothing but a name, and one with a few fields\nnewmtl bare\n\nnewmtl only_ns\nNs 250\nKe 1 2 3\nTr 0.2\nNi 1.5\n"}
    for k in c:
        c[k]["a.mtl"] = mtl
        c[k]["t.ppm"] = ppm
        c[k]["sub/t.ppm"] = b"P6\n1 1\n255\n" + bytes([9, 8, 7])
    return c


def reference_dump(files):
    with tempfile.TemporaryDirectory() as t:
        for name, data in files.items():
            p = os.path.join(t, name)
            os.makedirs(os.path.dirname(p), exist_ok=True)
            open(p, "wb").write(data if isinstance(data, bytes) else data.encode())
        out = subprocess.run([REF, os.path.join(t, "a.obj")], capture_output=True, text=True)
        if out.returncode != 0:
            return {"error": out.stderr.strip().splitlines()[-1]}
        return json.loads(out.stdout.split("@@DUMP@@", 1)[1])


REF_SCENE = os.path.join(ROOT, "oracle", "_ref", "ref_scene_dump")


def scene_cases():
    obj = "mtllib a.mtl\nv 0 0 0\nv 1 0 0\nv 0 1 0\nv 1 1 0\nvn 0 0 1\nvt 0 0\nusemtl red\nf 1/1/1 2/1/1 3/1/1\nusemtl green\nf 2/1/1 4/1/1 3/1/1\n"
    rng = np.random.default_rng(5)   # two light meshes, scalene triangles, a different normal at every corner
    lobj = ("mtllib a.mtl\n" + "".join("v %.6f %.6f %.6f\n" % tuple(v) for v in rng.uniform(-2, 2, (7, 3)) + [0, 0, 4])
            + "".join("vn %.6f %.6f %.6f\n" % tuple(v) for v in rng.normal(size=(6, 3)) + [0, 0, -2]) + "vt 0 0\n"
            + "usemtl green\nf 1/1/1 2/1/2 3/1/3\nf 2/1/4 4/1/5 3/1/6\nusemtl red\nf 5/1/1 6/1/3 7/1/5\n")
    mtl = "newmtl red\nKd 1 0 0\nNs 0.5\nnewmtl green\nKd 0 1 0\nNs 0.25\nKe 2 3 4\n"
    base = {"spp": 3, "width": 48, "height": 32, "renderers": [8, 9, 7],
            "cameras": [{"from": [0, 0, 3], "to": [0, 0, 0], "up": [0, 1, 0], "cos_fovy": 0.66},
                        {"from": [1, 2, 3], "to": [0.5, 0, 0], "up": [0, 0, 1], "cos_fovy": 0.25}],
            "surface_geometry": "a.obj", "area_lights": "l.obj", "extra": {"k": [1, 2]}}

    def var(**kw):
        d = dict(base)
        for k, v in kw.items():
            if v is None:
                d.pop(k, None)
            else:
                d[k] = v
        return json.dumps(d)

    one_cam = [{"from": [0, 0, 3], "to": [0, 0, 0], "up": [0, 1, 0], "cos_fovy": 1}]
    texts = {"base": var(), "numbers_as_floats": var(spp=2.7, width=64.0, height=3.2e1, renderers=[8.0, 9.9]),
             "ints_for_floats": var(cameras=one_cam), "no_renderers_key": var(renderers=None), "empty_renderers": var(renderers=[]),
             "renderers_a_number": var(renderers=8), "empty_cameras": var(cameras=[]), "no_cameras_key": var(cameras=None),
             "camera_without_up": var(cameras=[{"from": [0, 0, 3], "to": [0, 0, 0], "cos_fovy": 0.5}]),
             "no_surface": var(surface_geometry=None), "no_lights": var(area_lights=None), "no_spp": var(spp=None),
             "malformed": '{"spp": 1,', "string_spp": var(spp="3")}
    return {k: {"a.obj": obj, "l.obj": lobj, "a.mtl": mtl, "s.json": t} for k, t in texts.items()}


def reference_scene(files):
    with tempfile.TemporaryDirectory() as t:
        for name, data in files.items():
            open(os.path.join(t, name), "w").write(data)
        out = subprocess.run([REF_SCENE, "s.json"], capture_output=True, text=True, cwd=t)   # paths as the reference opens them: cwd
        if "@@DUMP@@" not in out.stdout:
            return {"crashed": out.returncode}
        return json.loads(out.stdout.split("@@DUMP@@", 1)[1])


def main():
    if os.path.exists(REF_SCENE):
        sc = {name: {"files": files, "reference": reference_scene(files)} for name, files in scene_cases().items()}
        json.dump(sc, open(os.path.join(HERE, "scene_ingest.json"), "w"), indent=0, sort_keys=True)
        print("scene_ingest.json:", {k: ("ok" if v["reference"].get("ok") else "refused / threw") for k, v in sc.items()})
    if not os.path.exists(REF):
        sys.exit("oracle/_ref/ref_obj_dump is not built: make -C oracle ref_host (needs /root/reference)")
    out = {}
    for name, files in cases().items():
        enc = {k: base64.b64encode(v if isinstance(v, bytes) else v.encode()).decode() for k, v in files.items()}
        out[name] = {"files": enc, "reference": reference_dump(files)}
        print(name, "->", "error" if "error" in out[name]["reference"] else "%d meshes" % len(out[name]["reference"]["meshes"]))
    json.dump(out, open(os.path.join(HERE, "obj_ingest.json"), "w"), indent=0, sort_keys=True)


if __name__ == "__main__":
    main()

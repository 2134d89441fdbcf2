"""The headless C++ host (host/orb_render): reference file formats in, reference dump formats out.

CPU: the host parses scene JSON + OBJ/MTL written by scenes.export_scene() and must arrive at exactly the
arrays the Python harness hands to the C-ABI (mesh order, un-indexed corners, material mapping of
src/Model.cpp:204-223, light list of include/viewer.hpp:236-265, camera of :635-641), and must fail the
way parseScene does (src/scene.cpp:20-101).  GPU: `.btc` dumps of the host equal the frames rendered
through the ctypes binding, byte for byte."""
import json
import os
import subprocess

import numpy as np
import pytest

from optix_renderer_b200 import scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "host", "orb_render")


def fnv1a(a):
    h = 1469598103934665603
    for b in np.ascontiguousarray(a).view(np.uint8).ravel().tolist():
        h = ((h ^ b) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return "%016x" % h


@pytest.fixture(scope="module")
def host():
    if not os.path.exists(HOST):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "host")])
    return HOST


def dump(host, path, *extra):
    o = subprocess.run([host, path, "--dump-scene", *extra], capture_output=True, text=True)
    return o.returncode, (json.loads(o.stdout) if o.returncode == 0 else None), o.stderr


@pytest.mark.parametrize("make", [scenes.scene_floor_blocker, scenes.scene_cornell])
def test_host_parses_exported_scene(host, tmp_path, make):
    sc = make(48, 32)
    path = scenes.export_scene(sc, str(tmp_path), spp=3, renderers=(8, 9))
    rc, d, err = dump(host, path)
    assert rc == 0, err
    assert (d["spp"], d["width"], d["height"], d["renderers"]) == (3, 48, 32, [8, 9])
    assert len(d["meshes"]) == len(sc["meshes"])
    assert d["n_surface_meshes"] == sum(not m["is_light"] for m in sc["meshes"])
    for got, m in zip(d["meshes"], sc["meshes"]):
        assert got["triangles"] == len(m["idx"])
        assert got["material_id"] == m["material_id"] and got["is_light"] == m["is_light"]
        assert np.array_equal(np.float32(got["diffuse"]), m["diffuse"])
        assert np.float32(got["alpha"]) == np.float32(m["alpha"])
        assert np.array_equal(np.float32(got["emit"]), m["emit"])
        assert got["vertex_hash"] == fnv1a(m["vtx"]) and got["normal_hash"] == fnv1a(m["nrm"])
        assert got["texcoord_hash"] == fnv1a(m["uv"])
    assert d["n_tri_lights"] == len(sc["tri_lights"])
    assert d["tri_lights_hash"] == fnv1a(sc["tri_lights"])
    for k in ("pos", "d00", "du", "dv"):
        assert np.array_equal(np.float32(d["camera"][k]), sc["camera"][k]), k


def test_host_camera_follows_size(host, tmp_path):
    sc = scenes.scene_floor_blocker(64, 64)
    path = scenes.export_scene(sc, str(tmp_path))
    rc, d, err = dump(host, path, "--size", "128x32")
    assert rc == 0, err
    ref = scenes.camera_from_lookat(**{k: v for k, v in zip(("frm", "at", "up", "cos_fovy"),
                                                            [sc["camera"]["lookat"][x] for x in ("from", "to", "up", "cos_fovy")])},
                                    width=128, height=32)
    for k in ("pos", "d00", "du", "dv"):
        assert np.array_equal(np.float32(d["camera"][k]), ref[k]), k


def test_host_scene_errors(host, tmp_path):
    sc = scenes.scene_floor_blocker(16, 16)
    path = scenes.export_scene(sc, str(tmp_path))
    doc = json.load(open(path))
    # missing file
    rc, _, err = dump(host, str(tmp_path / "nope.json"))
    assert rc != 0 and "Could not load scene" in err
    # no cameras (src/scene.cpp:56-74)
    bad = dict(doc, cameras=[])
    p = tmp_path / "nocam.json"
    p.write_text(json.dumps(bad))
    rc, _, err = dump(host, str(p))
    assert rc != 0 and "No cameras defined." in err
    # renderers that are not numbers (:44-54); a MISSING list is an empty one, as in the reference (scene_ingest.json)
    bad = dict(doc, renderers=["ratio"])
    p = tmp_path / "norend.json"
    p.write_text(json.dumps(bad))
    rc, _, err = dump(host, str(p))
    assert rc != 0 and "No renderers defined." in err
    # OBJ without normals: the reference dereferences normal_index unconditionally (Model.cpp:193-201)
    (tmp_path / "flat.obj").write_text("mtllib scene_surface.mtl\nusemtl m0\nv 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n")
    bad = dict(doc, surface_geometry="flat.obj")
    p = tmp_path / "flat.json"
    p.write_text(json.dumps(bad))
    rc, _, err = dump(host, str(p))
    assert rc != 0 and "v/vt/vn" in err
    # malformed JSON
    p = tmp_path / "broken.json"
    p.write_text("{\"spp\": 1,")
    rc, _, err = dump(host, str(p))
    assert rc != 0 and "Could not load scene" in err


def test_host_obj_features(host, tmp_path):
    """Negative (relative) indices, polygon fan, several materials in one shape -> one mesh per
    (shape, material) in ascending material order, texture through map_Kd."""
    (tmp_path / "a.mtl").write_text("newmtl red\nKd 1 0 0\nNs 0.5\nnewmtl green\nKd 0 1 0\nNs 0.25\nKe 2 3 4\nmap_Kd t.ppm\n")
    (tmp_path / "t.ppm").write_bytes(b"P6\n2 1\n255\n" + bytes([255, 0, 0, 0, 255, 0]))
    (tmp_path / "a.obj").write_text(
        "mtllib a.mtl\nv 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nvn 0 0 1\nvt 0 0\nvt 1 0\nvt 1 1\nvt 0 1\n"
        "o s\nusemtl green\nf 1/1/1 2/2/1 3/3/1 4/4/1\nusemtl red\nf -4/-4/-1 -3/-3/-1 -2/-2/-1\n")
    (tmp_path / "l.obj").write_text("mtllib a.mtl\nv 0 0 1\nv 1 0 1\nv 0 1 1\nvn 0 0 -1\nvt 0 0\nusemtl green\nf 1/1/1 2/1/1 3/1/1\n")
    (tmp_path / "s.json").write_text(json.dumps({
        "spp": 1, "width": 8, "height": 8, "renderers": [9],
        "cameras": [{"from": [0, 0, 3], "to": [0, 0, 0], "up": [0, 1, 0], "cos_fovy": 0.66}],
        "surface_geometry": "a.obj", "area_lights": "l.obj"}))
    rc, d, err = dump(host, str(tmp_path / "s.json"))
    assert rc == 0, err
    red, green, light = d["meshes"]
    assert (red["material_id"], red["triangles"], red["diffuse"]) == (1, 1, [1, 0, 0])
    assert (green["material_id"], green["triangles"], green["alpha"], green["tex"]) == (2, 2, 0.25, [2, 1])
    assert light["is_light"] and light["material_id"] == 0 and light["emit"] == [2, 3, 4]
    assert d["n_tri_lights"] == 1
    fan = np.float32([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 0, 0], [1, 1, 0], [0, 1, 0]])
    assert green["vertex_hash"] == fnv1a(fan)


def test_host_obj_ingest_equals_the_reference_loader(host, tmp_path):
    """host/scene_io.cpp load_obj against the reference's OWN ingest -- src/Model.cpp loadOBJ with the tiny_obj_loader.h
    it vendors, compiled where it lies as oracle/_ref/ref_obj_dump and recorded in tests/golden/obj_ingest.json by
    tests/golden/make_golden_obj.py: mesh order (shape, then ascending material), polygon triangulation (tinyobj's ear
    clipping: 80 concave / non-planar / degenerate polygons), the un-indexed vertex / normal / texcoord arrays BIT FOR
    BIT, Kd / Ns / Ke mapping, materialID = mtl index + 1, map_Kd texels (backslash paths too)."""
    import base64
    G = json.load(open(os.path.join(ROOT, "tests", "golden", "obj_ingest.json")))
    assert len(G) >= 8
    for name, case in sorted(G.items()):
        d0 = tmp_path / name
        for k, v in case["files"].items():
            (d0 / k).parent.mkdir(parents=True, exist_ok=True)
            (d0 / k).write_bytes(base64.b64decode(v))
        (d0 / "l.obj").write_text("mtllib a.mtl\nv 0 0 1\nv 1 0 1\nv 0 1 1\nvn 0 0 -1\nvt 0 0\nusemtl green\nf 1/1/1 2/1/1 3/1/1\n")
        (d0 / "s.json").write_text(json.dumps({
            "spp": 1, "width": 8, "height": 8, "renderers": [9],
            "cameras": [{"from": [0, 0, 3], "to": [0, 0, 0], "up": [0, 1, 0], "cos_fovy": 0.66}],
            "surface_geometry": "a.obj", "area_lights": "l.obj"}))
        rc, d, err = dump(host, str(d0 / "s.json"))
        assert rc == 0, (name, err)
        got, want = d["meshes"][:d["n_surface_meshes"]], case["reference"]["meshes"]
        assert len(got) == len(want), name
        for i, (g, w) in enumerate(zip(got, want)):
            for k in ("triangles", "material_id", "tex", "tex_hash", "vertex_hash", "normal_hash", "texcoord_hash"):
                assert g[k] == w[k], (name, i, k)
            for k in ("diffuse", "alpha", "emit"):
                assert np.array_equal(np.float32(g[k]), np.float32(w[k])), (name, i, k)


def test_host_scene_ingest_equals_the_reference_parser(host, tmp_path):
    """host/scene_io.cpp parse_scene against the reference's OWN parseScene + Scene::syncLights (src/scene.cpp with the
    json.h it vendors, compiled where it lies as oracle/_ref/ref_scene_dump; tests/golden/scene_ingest.json): spp / size /
    renderers (floats truncate, a missing list is empty, a lone number is a list of one), every camera, the light
    meshes appended with is_light and materialID 0 -- and against the two pieces of host arithmetic of include/viewer.hpp
    (which cannot be compiled whole: OpenGL / OWL), cut out of the file where it lies: the TriLight list and the camera
    vectors, bit for bit.  Where the reference refuses or throws, the host refuses; the one
    deliberate difference: a scene WITHOUT cameras, which the reference accepts and then dereferences cameras[0]."""
    G = json.load(open(os.path.join(ROOT, "tests", "golden", "scene_ingest.json")))
    assert len(G) >= 14
    seen_ok = 0
    for name, case in sorted(G.items()):
        d0 = tmp_path / name
        d0.mkdir()
        for k, v in case["files"].items():
            (d0 / k).write_text(v)
        o = subprocess.run([host, "s.json", "--dump-scene"], capture_output=True, text=True, cwd=str(d0))
        ref = case["reference"]
        if not ref.get("ok"):
            assert o.returncode != 0, name
            continue
        if not ref["cameras"]:
            assert o.returncode != 0 and "No cameras defined." in o.stderr, name
            continue
        assert o.returncode == 0, (name, o.stderr)
        d = json.loads(o.stdout)
        seen_ok += 1
        assert (d["spp"], d["width"], d["height"], d["renderers"]) == (ref["spp"], ref["width"], ref["height"], ref["renderers"]), name
        assert d["n_cameras"] == len(ref["cameras"]) and len(d["meshes"]) == len(ref["meshes"]), name
        for g, w in zip(d["meshes"], ref["meshes"]):
            for k in ("triangles", "material_id", "is_light", "vertex_hash"):
                assert g[k] == w[k], (name, k)
            assert np.array_equal(np.float32(g["emit"]), np.float32(w["emit"])), name
        # the TriLight list (include/viewer.hpp:236-265, cut out of the file where it lies and compiled in the harness): 76-byte
        # records v1 v2 v3 cg normal emit area, bit for bit
        assert (d["n_tri_lights"], d["tri_lights_hash"]) == (ref["n_tri_lights"], ref["tri_lights_hash"]), name
        for i, cam in enumerate(ref["cameras"]):      # the camera vectors (viewer.hpp:635-641, compiled the same way), bit for bit
            o = subprocess.run([host, "s.json", "--dump-scene", "--camera", str(i)], capture_output=True, text=True, cwd=str(d0))
            assert o.returncode == 0, (name, o.stderr)
            want = scenes.camera_from_lookat(np.float32(cam["from"]), np.float32(cam["to"]), np.float32(cam["up"]),
                                             np.float32(cam["cos_fovy"]), width=ref["width"], height=ref["height"])
            got = json.loads(o.stdout)["camera"]
            flat = np.float32(got["pos"] + got["d00"] + got["du"] + got["dv"])
            assert np.array_equal(flat, np.float32(ref["camera_vectors"][i])), (name, i)
            for k in ("pos", "d00", "du", "dv"):   # ... and so does the Python restatement the GPU tests and bench.py use
                assert np.array_equal(np.float32(got[k]), want[k]), (name, i, k)
    assert seen_ok >= 6


def _png(path, arr, ctype, depth=8, palette=None, trns=None):
    """Minimal PNG writer (zlib + CRC), filter type chosen per row to exercise all five."""
    import struct, zlib
    h, w = arr.shape[:2]
    raw = bytearray()
    a = np.ascontiguousarray(arr).reshape(h, -1).astype(np.uint8)
    bpp = max(1, a.shape[1] // w)
    prev = np.zeros(a.shape[1], np.int32)
    for y in range(h):
        cur = a[y].astype(np.int32)
        ft = y % 5
        left = np.concatenate([np.zeros(bpp, np.int32), cur[:-bpp]])
        ul = np.concatenate([np.zeros(bpp, np.int32), prev[:-bpp]])
        if ft == 0:
            enc = cur
        elif ft == 1:
            enc = cur - left
        elif ft == 2:
            enc = cur - prev
        elif ft == 3:
            enc = cur - ((left + prev) >> 1)
        else:
            p = left + prev - ul
            pa, pb, pc = np.abs(p - left), np.abs(p - prev), np.abs(p - ul)
            pred = np.where((pa <= pb) & (pa <= pc), left, np.where(pb <= pc, prev, ul))
            enc = cur - pred
        raw += bytes([ft]) + bytes((enc & 255).astype(np.uint8).tolist())
        prev = cur
    def chunk(t, d):
        return struct.pack(">I", len(d)) + t + d + struct.pack(">I", zlib.crc32(t + d) & 0xffffffff)
    out = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, ctype, 0, 0, 0))
    if palette is not None:
        out += chunk(b"PLTE", bytes(np.asarray(palette, np.uint8).ravel().tolist()))
    if trns is not None:
        out += chunk(b"tRNS", bytes(trns))
    comp = zlib.compress(bytes(raw))
    out += chunk(b"IDAT", comp[:len(comp) // 2]) + chunk(b"IDAT", comp[len(comp) // 2:]) + chunk(b"IEND", b"")
    open(path, "wb").write(out)


def test_host_texture_formats(host, tmp_path):
    """map_Kd through PNG (RGB, RGBA, grey, palette + tRNS; all five row filters, split IDAT) and TGA (raw and RLE):
    the texels the host hands to orb_scene_add_mesh are the file's, as RGBA8 with the BOTTOM row first -- what the
    reference gets from stbi_load(..., STBI_rgb_alpha) plus its y mirror (src/Model.cpp:81-128)."""
    rng = np.random.default_rng(9)
    w, h = 7, 6
    rgb = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    rgba = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    grey = rng.integers(0, 256, (h, w, 1), dtype=np.uint8)
    pal = rng.integers(0, 256, (5, 3), dtype=np.uint8)
    idx = rng.integers(0, 5, (h, w, 1), dtype=np.uint8)
    cases = {}
    _png(tmp_path / "rgb.png", rgb, 2)
    cases["rgb.png"] = np.concatenate([rgb, np.full((h, w, 1), 255, np.uint8)], -1)
    _png(tmp_path / "rgba.png", rgba, 6)
    cases["rgba.png"] = rgba
    _png(tmp_path / "grey.png", grey, 0)
    cases["grey.png"] = np.concatenate([grey, grey, grey, np.full((h, w, 1), 255, np.uint8)], -1)
    _png(tmp_path / "pal.png", idx, 3, palette=pal, trns=[10, 20, 30])
    a = np.array([10, 20, 30, 255, 255], np.uint8)[idx[..., 0]][..., None]
    cases["pal.png"] = np.concatenate([pal[idx[..., 0]], a], -1)
    bgr = rgb[::-1, :, ::-1]  # TGA: bottom-up rows, BGR
    (tmp_path / "raw.tga").write_bytes(bytes([0, 0, 2, 0, 0, 0, 0, 0, 0, 0, 0, 0, w, 0, h, 0, 24, 0]) + bgr.tobytes())
    cases["raw.tga"] = cases["rgb.png"]
    rle = bytearray()
    for row in bgr.reshape(h, w, 3):
        rle += bytes([128 | 1]) + row[0].tobytes()          # a run of 2 x the first pixel ...
        rle += bytes([w - 3]) + row[2:].tobytes()           # ... then w - 2 literal pixels
    (tmp_path / "rle.tga").write_bytes(bytes([0, 0, 10, 0, 0, 0, 0, 0, 0, 0, 0, 0, w, 0, h, 0, 24, 0]) + bytes(rle))
    t = rgb.copy()
    t[:, 1] = t[:, 0]
    cases["rle.tga"] = np.concatenate([t, np.full((h, w, 1), 255, np.uint8)], -1)
    for name, want in cases.items():
        (tmp_path / "a.mtl").write_text("newmtl m\nKd 1 0 0\nNs 0.5\nmap_Kd %s\n" % name)
        (tmp_path / "a.obj").write_text("mtllib a.mtl\nv 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nvt 0 0\nusemtl m\nf 1/1/1 2/1/1 3/1/1\n")
        (tmp_path / "s.json").write_text(json.dumps({
            "spp": 1, "width": 8, "height": 8, "renderers": [9],
            "cameras": [{"from": [0, 0, 3], "to": [0, 0, 0], "up": [0, 1, 0], "cos_fovy": 0.66}],
            "surface_geometry": "a.obj", "area_lights": "a.obj"}))
        rc, d, err = dump(host, str(tmp_path / "s.json"))
        assert rc == 0, err
        m = d["meshes"][0]
        assert m["tex"] == [w, h], (name, err)
        assert m["tex_hash"] == fnv1a(want[::-1]), name      # bottom row first


def test_host_exr_writer(host, tmp_path):
    """--exr writes what save.py makes of a .btc dump (flipud, RGB, float32): read back with OpenCV's EXR codec
    (imageio, which the reference uses, is not installed)."""
    os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")
    cv2 = pytest.importorskip("cv2")
    p = tmp_path / "t.exr"
    assert subprocess.run([host, "--selftest-exr", str(p)]).returncode == 0
    a = cv2.imread(str(p), cv2.IMREAD_UNCHANGED)
    if a is None:
        pytest.skip("this OpenCV build has no EXR codec")
    h, w = 5, 7
    y, x, c = np.meshgrid(np.arange(h), np.arange(w), np.arange(3), indexing="ij")
    want = (1000 * (h - 1 - y) + 10 * x + c + 0.25).astype(np.float32)   # row 0 of the buffer is the bottom image row
    assert a.dtype == np.float32 and np.array_equal(a[..., ::-1], want)


def test_host_textures_equal_the_reference_decoder(host, tmp_path):
    """map_Kd through JPEG (baseline AND progressive; 4:4:4 / 4:2:2 / 4:2:0 / 4:4:0 / 4:1:1, grey, optimised Huffman tables,
    restart intervals, 1x1 and odd sizes), PNG (every colour type x bit depth, plain and Adam7, tRNS on every type) and
    TGA (15 / 16 / 24 / 32-bit colour, grey, grey + alpha, colour-mapped with 15 ... 32-bit palettes, RLE, top-down) and
    BMP (OS/2 and Windows headers, palettes, 16 / 32-bit default and explicit masks, all-zero alpha, top-down): the texels the host hands to orb_scene_add_mesh equal, BIT FOR BIT, what the reference gets
    from the same file -- its vendored stb_image.h behind loadTexture's stbi_load + y mirror (src/Model.cpp:99-118),
    recorded in tests/golden/textures.npz by tests/golden/make_golden_textures.py.  A JPEG's texels are decoder
    defined (IDCT rounding, chroma upsampling filter, fixed-point colour conversion): host/image_io.cpp follows the
    reference's decoder in all three."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "textures.npz"))
    names = [str(n) for n in g["names"]]
    assert len(names) >= 275 and sum(n.endswith(".jpg") for n in names) >= 85 and sum(n.endswith(".tga") for n in names) >= 15
    assert sum(n.endswith(".bmp") for n in names) >= 45 and sum(n.endswith(".png") for n in names) >= 120
    (tmp_path / "a.obj").write_text("mtllib a.mtl\nv 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nvt 0 0\nusemtl m\nf 1/1/1 2/1/1 3/1/1\n")
    (tmp_path / "s.json").write_text(json.dumps({
        "spp": 1, "width": 8, "height": 8, "renderers": [9],
        "cameras": [{"from": [0, 0, 3], "to": [0, 0, 0], "up": [0, 1, 0], "cos_fovy": 0.66}],
        "surface_geometry": "a.obj", "area_lights": "a.obj"}))
    for i, name in enumerate(names):
        (tmp_path / name).write_bytes(g["file_%d" % i].tobytes())
        (tmp_path / "a.mtl").write_text("newmtl m\nKd 1 0 0\nNs 0.5\nmap_Kd %s\n" % name)
        rc, d, err = dump(host, str(tmp_path / "s.json"), "--dump-texels")
        assert rc == 0, err
        m = d["meshes"][0]
        want = g["rgba_%d" % i]
        assert m["tex"] == [want.shape[1], want.shape[0]], (name, err)
        got = np.frombuffer(bytes.fromhex(m["texels"]), np.uint8).reshape(want.shape)
        assert np.array_equal(got, want), (name, int(np.abs(got.astype(int) - want.astype(int)).max()))
    # arithmetic coding (SOF9): refused -- the scene renders with the material's Kd and a message -- never mis-decoded
    bad = bytearray(g["file_%d" % names.index(next(n for n in names if n.endswith("444_std.jpg") and n.startswith("c")))].tobytes())
    k = bad.find(b"\xff\xc0")
    assert k > 0
    bad[k + 1] = 0xC9
    (tmp_path / "ari.jpg").write_bytes(bytes(bad))
    (tmp_path / "a.mtl").write_text("newmtl m\nKd 1 0 0\nNs 0.5\nmap_Kd ari.jpg\n")
    rc, d, err = dump(host, str(tmp_path / "s.json"), "--dump-texels")
    assert rc == 0 and d["meshes"][0]["tex"] == [0, 0] and "Could not load texture" in err
    # a truncated file: refused or decoded to SOMETHING of the right size, never a crash
    for cut in (200, len(bad) // 2):
        (tmp_path / "cut.jpg").write_bytes(g["file_%d" % names.index("noise_420.jpg")].tobytes()[:cut])
        (tmp_path / "a.mtl").write_text("newmtl m\nKd 1 0 0\nNs 0.5\nmap_Kd cut.jpg\n")
        rc, d, err = dump(host, str(tmp_path / "s.json"), "--dump-texels")
        assert rc == 0 and d["meshes"][0]["tex"] in ([0, 0], [40, 24])


@pytest.mark.gpu
def test_host_checkpoint_resume_and_exr(host, tmp_path):
    """2 frames + checkpoint + 3 more frames == 5 frames in one go, bit for bit (accum and accumId travel in the
    checkpoint); the EXR dumps run through the reference's offline expression (ltc_ratio_estimator.py:8)."""
    os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")
    sc = scenes.scene_cornell(96, 64)
    path = scenes.export_scene(sc, str(tmp_path))
    a, b, ck = tmp_path / "a", tmp_path / "b", tmp_path / "state.bin"
    run = lambda *x: subprocess.run([host, path, "--renderer", "path", "--depth", "3", *x], capture_output=True, text=True)
    o = run("--out", str(a), "--frames", "5")
    assert o.returncode == 0, o.stderr
    o = run("--out", str(tmp_path / "b0"), "--frames", "2", "--save-state", str(ck))
    assert o.returncode == 0, o.stderr
    o = run("--out", str(b), "--frames", "3", "--resume", str(ck))
    assert o.returncode == 0 and "resumed at accumId 2" in o.stderr, o.stderr
    assert (a / "path.btc").read_bytes() == (b / "path.btc").read_bytes()
    assert (a / "rendered_output.ppm").read_bytes() == (b / "rendered_output.ppm").read_bytes()
    o = run("--out", str(b), "--frames", "1", "--resume", str(ck), "--size", "32x32")
    assert o.returncode != 0 and "cannot resume" in o.stderr
    # EXR dumps of a denoised RATIO run -> the reference's offline combine
    cv2 = pytest.importorskip("cv2")
    r = tmp_path / "r"
    o = subprocess.run([host, path, "--renderer", "ratio", "--frames", "1", "--denoise", "--exr", "--out", str(r)],
                       capture_output=True, text=True)
    assert o.returncode == 0, o.stderr
    rd = lambda n: cv2.imread(str(r / (n + ".exr")), cv2.IMREAD_UNCHANGED)
    ltc, sd, sn, fin = rd("ltc"), rd("stodirectDenoise"), rd("stoNoVisDenoise"), rd("final_ratio_estimator")
    if ltc is None:
        pytest.skip("this OpenCV build has no EXR codec")
    with np.errstate(all="ignore"):
        assert np.array_equal(ltc * (sd / sn), fin, equal_nan=True)
    btc = np.fromfile(str(r / "ltc.btc"), np.float32).reshape(64, 96, 4)
    assert np.array_equal(np.flipud(btc)[..., :3], ltc[..., ::-1])          # save.py:9-11


@pytest.mark.gpu
@pytest.mark.parametrize("renderer,frames", [("ratio", 2), ("path", 3), ("ltc_baseline", 1)])
def test_host_dumps_equal_binding(host, tmp_path, renderer, frames):
    from optix_renderer_b200 import api

    sc = scenes.scene_cornell(96, 64)
    path = scenes.export_scene(sc, str(tmp_path))
    out = tmp_path / "saves"
    args = [host, path, "--out", str(out), "--renderer", renderer, "--frames", str(frames), "--depth", "3"]
    if renderer == "ratio":
        args.append("--denoise")
    o = subprocess.run(args, capture_output=True, text=True)
    assert o.returncode == 0, o.stderr
    r = api.Renderer(sc, quantise_lut=1)  # the host fetches the LTC table like the texture unit by default
    rt = {"ratio": api.RATIO, "path": api.PATH, "ltc_baseline": api.LTC_BASELINE}[renderer]
    for f in range(frames):
        r.render(rt, accum_id=f, mode=api.INTENDED, shadow_spp=4, max_depth=3)

    def btc(name):
        return np.fromfile(str(out / (name + ".btc")), np.float32).reshape(64, 96, 4)

    if renderer == "ratio":
        r.denoise()
        r.ratio_combine(use_denoised=True)
        pairs = [("stoDirect", "sto_direct", 1), ("stoNoVis", "sto_no_vis", 1), ("ltc", "ltc", 1),
                 ("normal_screen_buffer", "normal", 1), ("materialID_screen_buffer", "materialID", 1),
                 ("stodirectDenoise", "sto_direct_dn", 1), ("stoNoVisDenoise", "sto_no_vis_dn", 1),
                 ("final_ratio_estimator", "ratio_final", 1)]
    elif renderer == "path":
        pairs = [("path", "accum", frames), ("position_screen_buffer", "position", 1), ("albedo_screen_buffer", "albedo", 1)]
    else:
        pairs = [("ltc_baseline", "accum", 1)]
    for fname, buf, avg in pairs:
        want = (r.download(buf) / np.float32(avg)).astype(np.float32)
        assert np.array_equal(btc(fname), want, equal_nan=True), fname
    # PPM = the RGBA8 frame, top row first
    raw = (out / "rendered_output.ppm").read_bytes()
    assert raw.startswith(b"P6\n96 64\n255\n")
    rgb = np.frombuffer(raw[len(b"P6\n96 64\n255\n"):], np.uint8).reshape(64, 96, 3)
    px = r.download("rgba8")
    want = np.stack([px & 255, (px >> 8) & 255, (px >> 16) & 255], -1).astype(np.uint8)[::-1]
    assert np.array_equal(rgb, want)
    # the reference's offline pipeline on the dumps: save.py's reshape/flip, ltc_ratio_estimator.py:8
    if renderer == "ratio":
        ltc, sd, sn = (np.flipud(btc(n))[..., :3] for n in ("ltc", "stodirectDenoise", "stoNoVisDenoise"))
        with np.errstate(all="ignore"):
            final = ltc * (sd / sn)
        assert np.array_equal(final, np.flipud(btc("final_ratio_estimator"))[..., :3], equal_nan=True)
    r.close()


@pytest.mark.gpu
def test_host_move_lights_goes_through_refit(host, tmp_path):
    """--move-lights: the host moves the light meshes of the COMMITTED scene (orb_scene_update_mesh +
    orb_scene_refit + orb_scene_set_lights, no rebuild).  Its dumps equal, bit for bit, those of a scene that was
    built with the lights already there."""
    import copy

    from optix_renderer_b200 import api

    sc = scenes.scene_cornell(96, 64)
    path = scenes.export_scene(sc, str(tmp_path))
    out = tmp_path / "saves"
    o = subprocess.run([host, path, "--out", str(out), "--renderer", "ratio", "--frames", "2", "--move-lights",
                        "0.125,-0.25,0.0625"], capture_output=True, text=True)
    assert o.returncode == 0, o.stderr
    assert "update + refit" in o.stderr
    mv = copy.deepcopy(sc)
    for m in mv["meshes"]:
        if m["is_light"]:
            m["vtx"] = (m["vtx"] + np.array([0.125, -0.25, 0.0625], np.float32)).astype(np.float32)
    mv["tri_lights"] = scenes.make_tri_lights(mv["meshes"])
    r = api.Renderer(mv, quantise_lut=1)
    for f in range(2):
        r.render(api.RATIO, accum_id=f, mode=api.INTENDED, shadow_spp=4)
    for fname, buf in (("stoDirect", "sto_direct"), ("stoNoVis", "sto_no_vis"), ("ltc", "ltc"),
                       ("normal_screen_buffer", "normal"), ("materialID_screen_buffer", "materialID")):
        got = np.fromfile(str(out / (fname + ".btc")), np.float32).reshape(64, 96, 4)
        assert np.array_equal(got, r.download(buf), equal_nan=True), fname
    still = api.Renderer(sc, quantise_lut=1)
    still.render(api.RATIO, accum_id=1, mode=api.INTENDED, shadow_spp=4)
    assert not np.array_equal(still.download("ltc"), r.download("ltc"), equal_nan=True)  # the lights did move
    still.close()
    r.close()

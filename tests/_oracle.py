"""ctypes binding for the CPU checkers declared in oracle/oracle_api.h.

TEST INFRASTRUCTURE: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs load this.  Three interchangeable libraries:
  kind="restatement"        oracle/liboracle.so            (this repo's CPU restatement)
  kind="reference-literal"  oracle/_ref/libref_literal.so  (reference TU compiled unmodified)
  kind="reference-intended" oracle/_ref/libref_intended.so (reference TU + patches P1-P3)
"""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_PATHS = {
    "restatement": os.path.join(ROOT, "oracle", "liboracle.so"),
    "reference-literal": os.path.join(ROOT, "oracle", "_ref", "libref_literal.so"),
    "reference-intended": os.path.join(ROOT, "oracle", "_ref", "libref_intended.so"),
    # the same reference TU compiled by nvcc for sm_100a: the literal megakernel ON THE GPU (oracle/ref_shim_gpu);
    # scene / frame entry points only, whole frames only
    "reference-gpu-literal": os.path.join(ROOT, "oracle", "_ref", "libref_gpu_literal.so"),
    "reference-gpu-intended": os.path.join(ROOT, "oracle", "_ref", "libref_gpu_intended.so"),
}

BUFFER_NAMES = ["position", "normal", "uv", "albedo", "alpha", "materialID", "bounce0", "bounce1", "bounce2",
                "ltc", "sto_direct", "sto_no_vis", "accum"]

# RendererType, cuda_include/common.cuh:17-29
DIFFUSE, ALPHA, NORMALS, SHADE_NORMALS, POSITION, MASK, MATERIAL_ID, LTC_BASELINE, RATIO, PATH = range(10)

_f = C.POINTER(C.c_float)
_i = C.POINTER(C.c_int32)
_u = C.POINTER(C.c_uint32)
_b = C.POINTER(C.c_uint8)


class OraBuffers(C.Structure):
    _fields_ = [(n, _f) for n in BUFFER_NAMES] + [("rgba8", _u), ("hit_prim", _i), ("hit_uv", _f)]


def available(kind):
    return os.path.exists(_PATHS[kind])


def _fp(a):
    return a.ctypes.data_as(_f)


def f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


class Oracle:
    """One loaded checker library (global state inside the library: one scene at a time)."""

    def __init__(self, kind="restatement"):
        self.lib = C.CDLL(_PATHS[kind])
        L = self.lib
        L.ora_kind.restype = C.c_char_p
        L.ora_add_mesh.argtypes = [_f, _f, _f, C.c_int, _i, C.c_int, _f, C.c_float, _f, C.c_int, C.c_uint32, _b,
                                   C.c_int, C.c_int]
        L.ora_set_lights.argtypes = [_f, C.c_int]
        L.ora_set_ltc_lut.argtypes = [_f, _f, _f, C.c_int, C.c_int, C.c_int]
        L.ora_commit.argtypes = [C.c_int]
        L.ora_set_camera.argtypes = [_f, _f, _f, _f]
        L.ora_set_options.argtypes = [C.c_int, C.c_int, C.c_int]
        L.ora_render.argtypes = [C.c_int] * 6 + [C.POINTER(OraBuffers), C.c_int]
        L.ora_render.restype = C.c_longlong
        self.kind = L.ora_kind().decode()
        self._keep = []
        if self.kind.startswith("reference-gpu"):
            L.ora_last_ms.restype = C.c_float
            return
        L.ora_ltc_gbuffer.argtypes = [C.c_int, _f, _f, _f, _f, _f, _f, C.c_int]
        L.ora_u_rng.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, _u, _u, _f]
        L.ora_u_clip.argtypes = [C.c_int, _f]
        L.ora_u_integrate_edge.argtypes = [_f, _f]
        L.ora_u_integrate_edge.restype = C.c_float
        L.ora_u_fetch_ltc.argtypes = [C.c_float, C.c_float, _f]
        L.ora_u_mat_inverse.argtypes = [_f, _f]
        L.ora_u_onb.argtypes = [_f, _f]
        L.ora_u_sample_direction.argtypes = [_f, C.c_float, C.c_float, _f, C.c_float, C.c_float, _f]
        L.ora_u_evaluate.argtypes = [_f, _f, _f, C.c_float, _f]
        L.ora_u_pdf.argtypes = [_f, _f, _f, C.c_float, C.c_float]
        L.ora_u_pdf.restype = C.c_float
        L.ora_u_sample_tri.argtypes = [_f, _f, _f, C.c_float, C.c_float, _f]
        L.ora_u_pdf_a2w.argtypes = [C.c_float] * 3
        L.ora_u_pdf_a2w.restype = C.c_float
        L.ora_u_make_rgba.argtypes = [_f]
        L.ora_u_make_rgba.restype = C.c_uint32
        L.ora_ratio_combine.argtypes = [C.c_longlong, _f, _f, _f, _f]
        self.kind = L.ora_kind().decode()
        if self.kind == "restatement":
            L.ora_ltc_margin.argtypes = [C.c_int, _f, _f, _f, _f, _f, C.c_int]
            L.ora_denoise.argtypes = [C.c_int, C.c_int, _f, _f, _f, _f, _f, _f, C.c_int, C.c_float, C.c_float,
                                      C.c_float, _f, _f, C.c_int]
        self._keep = []

    def last_ms(self):
        """reference-gpu kinds: CUDA-event time of the last render()'s megakernel."""
        return float(self.lib.ora_last_ms())

    # ---- scene ----
    def reset(self):
        self.lib.ora_reset()
        self._keep = []

    def add_mesh(self, vtx, nrm, uv, idx, diffuse, alpha, emit, is_light, material_id, tex=None):
        vtx, nrm, uv = f32(vtx), f32(nrm), f32(uv)
        idx = np.ascontiguousarray(idx, dtype=np.int32)
        d, e = f32(diffuse), f32(emit)
        tp, tw, th = None, 0, 0
        if tex is not None:
            tex = np.ascontiguousarray(tex, dtype=np.uint8)
            th, tw = tex.shape[:2]
            tp = tex.ctypes.data_as(_b)
        return self.lib.ora_add_mesh(_fp(vtx), _fp(nrm), _fp(uv), vtx.shape[0], idx.ctypes.data_as(_i), idx.shape[0],
                                     _fp(d), float(alpha), _fp(e), int(is_light), int(material_id), tp, tw, th)

    def set_lights(self, tri_lights):
        t = f32(tri_lights).reshape(-1, 19)
        self.lib.ora_set_lights(_fp(t), t.shape[0])

    def set_ltc_lut(self, m1, m2, m3, quantise=0):
        m1, m2, m3 = f32(m1), f32(m2), f32(m3)
        h, w = m1.shape[:2]
        self.lib.ora_set_ltc_lut(_fp(m1), _fp(m2), _fp(m3), w, h, int(quantise))

    def commit(self, use_bvh=True):
        self.lib.ora_commit(int(use_bvh))

    def set_camera(self, pos, d00, du, dv):
        a = [f32(x) for x in (pos, d00, du, dv)]
        self.lib.ora_set_camera(*[_fp(x) for x in a])

    def set_options(self, mode, shadow_spp=4, max_depth=1):
        return self.lib.ora_set_options(int(mode), int(shadow_spp), int(max_depth))

    def load_scene(self, scene, quantise=0, use_bvh=True):
        """scene: the dict made by optix_renderer_b200.scenes (meshes, lights, lut, camera)."""
        self.reset()
        for m in scene["meshes"]:
            self.add_mesh(m["vtx"], m["nrm"], m["uv"], m["idx"], m["diffuse"], m["alpha"], m["emit"], m["is_light"],
                          m["material_id"], m.get("tex"))
        self.set_lights(scene["tri_lights"])
        self.set_ltc_lut(*scene["lut"], quantise=quantise)
        self.commit(use_bvh)
        cam = scene["camera"]
        self.set_camera(cam["pos"], cam["d00"], cam["du"], cam["dv"])

    # ---- frames ----
    def render(self, W, H, renderer, accum_id=0, accum=None, rows=None, nthreads=1):
        out = {n: np.zeros((H, W, 4), np.float32) for n in BUFFER_NAMES}
        if accum is not None:
            out["accum"][...] = accum
        out["rgba8"] = np.zeros((H, W), np.uint32)
        out["hit_prim"] = np.full((H, W), -1, np.int32)
        out["hit_uv"] = np.zeros((H, W, 2), np.float32)
        b = OraBuffers()
        for n in BUFFER_NAMES:
            setattr(b, n, _fp(out[n]))
        b.rgba8 = out["rgba8"].ctypes.data_as(_u)
        b.hit_prim = out["hit_prim"].ctypes.data_as(_i)
        b.hit_uv = _fp(out["hit_uv"])
        y0, y1 = rows if rows else (0, H)
        out["rays"] = self.lib.ora_render(W, H, int(renderer), int(accum_id), y0, y1, C.byref(b), int(nthreads))
        return out

    def ltc_gbuffer(self, position, normal, albedo, alpha, cam_pos, nthreads=1):
        position, normal, albedo, alpha = f32(position), f32(normal), f32(albedo), f32(alpha)
        n = position.size // 4
        out = np.zeros_like(position)
        c = f32(cam_pos)
        self.lib.ora_ltc_gbuffer(n, _fp(position), _fp(normal), _fp(albedo), _fp(alpha), _fp(c), _fp(out), nthreads)
        return out

    def ltc_margin(self, position, normal, alpha, cam_pos, nthreads=1):
        """liboracle.so only: distance of every pixel's LTC evaluation to its nearest sign decision (see
        oracle_api.h); the scene's lights and LUT must be loaded."""
        position, normal, alpha = f32(position), f32(normal), f32(alpha)
        n = position.size // 4
        out = np.zeros(position.shape[:-1], np.float32)
        c = f32(cam_pos)
        self.lib.ora_ltc_margin(n, _fp(position), _fp(normal), _fp(alpha), _fp(c), _fp(out), nthreads)
        return out

    def ratio_combine(self, ltc, sd, snv):
        ltc, sd, snv = f32(ltc), f32(sd), f32(snv)
        out = np.empty_like(ltc)
        self.lib.ora_ratio_combine(ltc.size, _fp(ltc), _fp(sd), _fp(snv), _fp(out))
        return out

    def denoise(self, normal, position, material_id, sto_direct, sto_no_vis, cam_pos, iterations=5,
                sigma_normal=32.0, sigma_position=0.25, sigma_colour=0.0, nthreads=1):
        """Own definition (no reference denoiser exists); liboracle.so only."""
        a = [f32(x) for x in (normal, position, material_id, sto_direct, sto_no_vis)]
        H, W = a[0].shape[:2]
        c = f32(cam_pos)
        od, on = np.empty_like(a[3]), np.empty_like(a[4])
        self.lib.ora_denoise(W, H, *[_fp(x) for x in a], _fp(c), int(iterations), sigma_normal, sigma_position,
                             sigma_colour, _fp(od), _fp(on), int(nthreads))
        return od, on

    # ---- units ----
    def u_rng(self, frame, px, py, W, H, n):
        s = np.zeros(1, np.uint32)
        raw = np.zeros(n, np.uint32)
        flt = np.zeros(n, np.float32)
        self.lib.ora_u_rng(frame, px, py, W, H, n, s.ctypes.data_as(_u), raw.ctypes.data_as(_u), _fp(flt))
        return int(s[0]), raw, flt

    def u_clip(self, vc, v):
        v = f32(v).reshape(5, 3).copy()
        n = self.lib.ora_u_clip(vc, _fp(v))
        return n, v

    def u_integrate_edge(self, a, b):
        a, b = f32(a), f32(b)
        return self.lib.ora_u_integrate_edge(_fp(a), _fp(b))

    def u_fetch_ltc(self, alpha, theta):
        o = np.zeros(10, np.float32)
        self.lib.ora_u_fetch_ltc(alpha, theta, _fp(o))
        return o[:9].reshape(3, 3).copy(), float(o[9])

    def u_mat_inverse(self, m):
        m = f32(m)
        o = np.zeros(9, np.float32)
        self.lib.ora_u_mat_inverse(_fp(m), _fp(o))
        return o.reshape(3, 3)

    def u_onb(self, n):
        n = f32(n)
        o = np.zeros(18, np.float32)
        self.lib.ora_u_onb(_fp(n), _fp(o))
        return o[:9].reshape(3, 3).copy(), o[9:].reshape(3, 3).copy()

    def u_sample_direction(self, wo, u1, u2, base, metal, alpha):
        wo, base = f32(wo), f32(base)
        o = np.zeros(4, np.float32)
        self.lib.ora_u_sample_direction(_fp(wo), u1, u2, _fp(base), metal, alpha, _fp(o))
        return o[:3].copy(), float(o[3])

    def u_evaluate(self, wi, wo, base, alpha):
        wi, wo, base = f32(wi), f32(wo), f32(base)
        o = np.zeros(3, np.float32)
        self.lib.ora_u_evaluate(_fp(wi), _fp(wo), _fp(base), alpha, _fp(o))
        return o

    def u_pdf(self, wi, wo, base, metal, alpha):
        wi, wo, base = f32(wi), f32(wo), f32(base)
        return self.lib.ora_u_pdf(_fp(wi), _fp(wo), _fp(base), metal, alpha)

    def u_sample_tri(self, v1, v2, v3, u1, u2):
        v1, v2, v3 = f32(v1), f32(v2), f32(v3)
        o = np.zeros(3, np.float32)
        self.lib.ora_u_sample_tri(_fp(v1), _fp(v2), _fp(v3), u1, u2, _fp(o))
        return o

    def u_pdf_a2w(self, pdf, d2, c):
        return self.lib.ora_u_pdf_a2w(pdf, d2, c)

    def u_make_rgba(self, c):
        c = f32(c)
        return self.lib.ora_u_make_rgba(_fp(c))

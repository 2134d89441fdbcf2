"""ctypes binding for tests/hostcheck/libhostcheck.so -- the product's __host__ __device__ headers
compiled with g++ (TEST INFRASTRUCTURE; see tests/hostcheck/hostcheck.cpp)."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "hostcheck", "hostcheck.cpp")
LIB = os.path.join(HERE, "hostcheck", "libhostcheck.so")
CSRC = os.path.join(os.path.dirname(HERE), "optix_renderer_b200", "csrc")

BUFFER_NAMES = ["position", "normal", "uv", "albedo", "alpha", "materialID", "bounce0", "bounce1", "bounce2",
                "ltc", "sto_direct", "sto_no_vis", "accum"]
_f = C.POINTER(C.c_float)
_i = C.POINTER(C.c_int32)
_u = C.POINTER(C.c_uint32)
_b = C.POINTER(C.c_uint8)


class HcBuffers(C.Structure):
    _fields_ = [(n, _f) for n in BUFFER_NAMES] + [("rgba8", _u), ("hit_prim", _i), ("hit_uv", _f)]


def build(force=False):
    deps = [SRC] + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cuh")]
    if not force and os.path.exists(LIB) and all(os.path.getmtime(LIB) >= os.path.getmtime(d) for d in deps):
        return LIB
    cuda_inc = os.environ.get("CUDA_INC", "/usr/local/cuda/include")
    subprocess.check_call(["g++", "-O2", "-mfma", "-ffp-contract=off", "-std=c++17", "-shared", "-fPIC",
                           "-I" + cuda_inc, "-o", LIB, SRC])
    return LIB


def f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _fp(a):
    return a.ctypes.data_as(_f)


class HostCheck:
    def __init__(self):
        self.lib = L = C.CDLL(build())
        L.hc_add_mesh.argtypes = [_f, _f, _f, C.c_int, _i, C.c_int, _f, C.c_float, _f, C.c_int, C.c_uint32, _b, C.c_int,
                                  C.c_int]
        L.hc_set_lights.argtypes = [_f, C.c_int]
        L.hc_set_ltc_lut.argtypes = [_f, _f, _f, C.c_int, C.c_int, C.c_int]
        L.hc_commit.argtypes = [C.c_int]
        L.hc_set_camera.argtypes = [_f, _f, _f, _f]
        L.hc_trace.argtypes = [_f, _f, C.c_int, C.c_int, _i, _f, _f, _f]
        L.hc_trace_counted.argtypes = [_f, _f, C.c_int, C.c_int, _i, _f, _f, _f, C.POINTER(C.c_ulonglong),
                                       C.POINTER(C.c_ulonglong)]
        L.hc_render.argtypes = [C.c_int] * 7 + [C.POINTER(HcBuffers)]
        L.hc_render.restype = C.c_longlong
        L.hc_ltc_gbuffer.argtypes = [C.c_int, _f, _f, _f, _f, _f, _f]
        L.hc_u_rng.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, _u, _u, _f]
        L.hc_u_rng_skip.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int]
        L.hc_u_rng_skip.restype = C.c_uint32
        L.hc_u_clip.argtypes = [_f, C.c_int, _i]
        L.hc_u_clip.restype = C.c_float
        L.hc_u_integrate_edge.argtypes = [_f, _f]
        L.hc_u_integrate_edge.restype = C.c_float
        L.hc_u_fetch_ltc.argtypes = [C.c_float, C.c_float, _f]
        L.hc_u_mat_inverse.argtypes = [_f, _f]
        L.hc_u_onb.argtypes = [_f, _f]
        L.hc_u_sample_direction.argtypes = [_f, C.c_float, C.c_float, _f, C.c_float, _f]
        L.hc_u_evaluate.argtypes = [_f, _f, _f, C.c_float, _f]
        L.hc_u_pdf.argtypes = [_f, _f, _f, C.c_float]
        L.hc_u_pdf.restype = C.c_float
        L.hc_u_make_rgba.argtypes = [_f]
        L.hc_u_make_rgba.restype = C.c_uint32
        L.hc_u_sample_tri.argtypes = [_f, C.c_float, C.c_float, _f]
        L.hc_u_pdf_a2w.argtypes = [C.c_float] * 3
        L.hc_u_pdf_a2w.restype = C.c_float

    def load_scene(self, scene, quantise=0, max_leaf=4):
        L = self.lib
        L.hc_reset()
        for m in scene["meshes"]:
            vtx, nrm, uv = f32(m["vtx"]), f32(m["nrm"]), f32(m["uv"])
            idx = np.ascontiguousarray(m["idx"], np.int32)
            d, e = f32(m["diffuse"]), f32(m["emit"])
            tex = m.get("tex")
            tp, tw, th = None, 0, 0
            if tex is not None:
                tex = np.ascontiguousarray(tex, np.uint8)
                th, tw = tex.shape[:2]
                tp = tex.ctypes.data_as(_b)
            L.hc_add_mesh(_fp(vtx), _fp(nrm), _fp(uv), vtx.shape[0], idx.ctypes.data_as(_i), idx.shape[0], _fp(d),
                          float(m["alpha"]), _fp(e), int(m["is_light"]), int(m["material_id"]), tp, tw, th)
        t = f32(scene["tri_lights"]).reshape(-1, 19)
        L.hc_set_lights(_fp(t), t.shape[0])
        m1, m2, m3 = [f32(x) for x in scene["lut"]]
        L.hc_set_ltc_lut(_fp(m1), _fp(m2), _fp(m3), m1.shape[1], m1.shape[0], int(quantise))
        L.hc_commit(max_leaf)
        cam = scene["camera"]
        a = [f32(cam[k]) for k in ("pos", "d00", "du", "dv")]
        L.hc_set_camera(*[_fp(x) for x in a])

    def render(self, W, H, renderer, accum_id=0, accum=None, mode=0, spp=4, depth=1):
        out = {n: np.zeros((H, W, 4), np.float32) for n in BUFFER_NAMES}
        if accum is not None:
            out["accum"][...] = accum
        out["rgba8"] = np.zeros((H, W), np.uint32)
        out["hit_prim"] = np.full((H, W), -1, np.int32)
        out["hit_uv"] = np.zeros((H, W, 2), np.float32)
        b = HcBuffers()
        for n in BUFFER_NAMES:
            setattr(b, n, _fp(out[n]))
        b.rgba8 = out["rgba8"].ctypes.data_as(_u)
        b.hit_prim = out["hit_prim"].ctypes.data_as(_i)
        b.hit_uv = _fp(out["hit_uv"])
        out["rays"] = self.lib.hc_render(W, H, int(renderer), int(accum_id), int(mode), int(spp), int(depth), C.byref(b))
        return out

    def trace(self, org, dirs, brute=False):
        org, dirs = f32(org).reshape(-1, 3), f32(dirs).reshape(-1, 3)
        n = org.shape[0]
        prim = np.zeros(n, np.int32)
        t, u, v = (np.zeros(n, np.float32) for _ in range(3))
        self.lib.hc_trace(_fp(org), _fp(dirs), n, int(brute), prim.ctypes.data_as(_i), _fp(t), _fp(u), _fp(v))
        return prim, t, u, v

    def trace_counted(self, org, dirs, mode=0):
        """(prim, t, u, v), node records fetched, triangles tested; mode 0 = fp32 nodes, 5 = 32-byte quantised nodes."""
        org, dirs = f32(org).reshape(-1, 3), f32(dirs).reshape(-1, 3)
        n = org.shape[0]
        prim = np.zeros(n, np.int32)
        t, u, v = (np.zeros(n, np.float32) for _ in range(3))
        nn, nt = C.c_ulonglong(), C.c_ulonglong()
        self.lib.hc_trace_counted(_fp(org), _fp(dirs), n, int(mode), prim.ctypes.data_as(_i), _fp(t), _fp(u), _fp(v),
                                  C.byref(nn), C.byref(nt))
        return (prim, t, u, v), nn.value, nt.value

    def ltc_gbuffer(self, position, normal, albedo, alpha, cam_pos):
        position, normal, albedo, alpha = f32(position), f32(normal), f32(albedo), f32(alpha)
        out = np.zeros_like(position)
        c = f32(cam_pos)
        self.lib.hc_ltc_gbuffer(position.size // 4, _fp(position), _fp(normal), _fp(albedo), _fp(alpha), _fp(c), _fp(out))
        return out

    # ---- unit wrappers: same names / meaning as optix_renderer_b200.api.Renderer.u_* ----
    def u_rng(self, frame, px, py, W, n):
        s = np.zeros(1, np.uint32)
        raw, flt = np.zeros(n, np.uint32), np.zeros(n, np.float32)
        self.lib.hc_u_rng(frame, px, py, W, n, s.ctypes.data_as(_u), raw.ctypes.data_as(_u), _fp(flt))
        return int(s[0]), raw, flt

    def u_rng_skip(self, frame, px, py, W, n):
        return int(self.lib.hc_u_rng_skip(frame, px, py, W, n))

    def u_clip(self, abc, quad=False):
        a = f32(abc).reshape(9)
        n = np.zeros(1, np.int32)
        ff = self.lib.hc_u_clip(_fp(a), int(quad), n.ctypes.data_as(_i))
        return int(n[0]), float(ff)

    def u_integrate_edge(self, a, b):
        a, b = f32(a), f32(b)
        return float(self.lib.hc_u_integrate_edge(_fp(a), _fp(b)))

    def u_fetch_ltc(self, alpha, theta):
        o = np.zeros(10, np.float32)
        self.lib.hc_u_fetch_ltc(alpha, theta, _fp(o))
        return o[:9].reshape(3, 3).copy(), float(o[9])

    def u_mat_inverse(self, m):
        m, o = f32(m).reshape(9), np.zeros(9, np.float32)
        self.lib.hc_u_mat_inverse(_fp(m), _fp(o))
        return o.reshape(3, 3)

    def u_onb(self, n):
        n, o = f32(n), np.zeros(9, np.float32)
        self.lib.hc_u_onb(_fp(n), _fp(o))
        return o.reshape(3, 3)

    def u_sample_direction(self, wo, u1, u2, base, alpha):
        wo, base, o = f32(wo), f32(base), np.zeros(4, np.float32)
        self.lib.hc_u_sample_direction(_fp(wo), u1, u2, _fp(base), alpha, _fp(o))
        return o[:3].copy(), float(o[3])

    def u_evaluate(self, wi, wo, base, alpha):
        wi, wo, base, o = f32(wi), f32(wo), f32(base), np.zeros(3, np.float32)
        self.lib.hc_u_evaluate(_fp(wi), _fp(wo), _fp(base), alpha, _fp(o))
        return o

    def u_pdf(self, wi, wo, base, alpha):
        wi, wo, base = f32(wi), f32(wo), f32(base)
        return float(self.lib.hc_u_pdf(_fp(wi), _fp(wo), _fp(base), alpha))

    def u_make_rgba(self, c):
        c = f32(c)
        return int(self.lib.hc_u_make_rgba(_fp(c)))

    def u_sample_tri(self, v9, u1, u2):
        v, o = f32(v9).reshape(9), np.zeros(3, np.float32)
        self.lib.hc_u_sample_tri(_fp(v), u1, u2, _fp(o))
        return o

    def u_pdf_a2w(self, pdf, d2, c):
        return float(self.lib.hc_u_pdf_a2w(pdf, d2, c))

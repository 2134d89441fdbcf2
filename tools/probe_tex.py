"""How does the B200 texture unit quantise bilinear weights?  (oracle/_ref/libref_gpu_intended.so, ora_probe_tex)"""
import ctypes as C, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
L = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_gpu_intended.so"))
fp = C.POINTER(C.c_float)
L.ora_probe_tex.argtypes = [fp, C.c_int, C.c_int, fp]
w = 8
rng = np.random.default_rng(1)
xs = np.concatenate([np.linspace(0, 1, 200001), rng.uniform(0, 1, 200000)]).astype(np.float32)
out = np.zeros_like(xs)
L.ora_probe_tex(xs.ctypes.data_as(fp), len(xs), w, out.ctypes.data_as(fp))
xb = xs.astype(np.float32) * np.float32(w) - np.float32(0.5)
cands = {}
def model(q):  # q = fixed-point coordinate in 1/256 units
    i = np.floor(q / 256.0)
    a = (q - i * 256) / 256.0
    i0 = np.clip(i, 0, w - 1); i1 = np.clip(i + 1, 0, w - 1)
    return ((1 - a) * i0 + a * i1).astype(np.float32)
cands["exact"] = model(xb.astype(np.float64) * 256.0)
cands["round_nearest_1/256"] = model(np.floor(xb.astype(np.float64) * 256.0 + 0.5))
cands["truncate_1/256"] = model(np.floor(xb.astype(np.float64) * 256.0))
xw = xs.astype(np.float32) * np.float32(w)
cands["round(x*w*256) then -128"] = model(np.floor(xw.astype(np.float64) * 256.0 + 0.5) - 128)
cands["trunc(x*w*256) then -128"] = model(np.floor(xw.astype(np.float64) * 256.0) - 128)
res = {k: {"max_abs_err": float(np.abs(v - out).max()), "frac_exact": float((v == out).mean())} for k, v in cands.items()}

# ---- 2-D, the real LTC tables -------------------------------------------------------------------
sys.path.insert(0, ROOT)
from optix_renderer_b200 import scenes
lut = scenes.load_ltc_lut()
L.ora_probe_tex2d.argtypes = [fp, C.c_int, C.c_int, fp, C.c_int, fp]
n = 300000
xy = rng.uniform(0, 1, (n, 2)).astype(np.float32)
m = np.ascontiguousarray(lut[0], np.float32)  # (8, 8, 4): rows = y
h, w = m.shape[:2]
got = np.zeros((n, 4), np.float32)
L.ora_probe_tex2d(m.ctypes.data_as(fp), w, h, xy.ctypes.data_as(fp), n, got.ctypes.data_as(fp))
def wq(c, size):
    b = c.astype(np.float32) * np.float32(size) - np.float32(0.5)
    f = np.floor(b)
    a = np.floor((b - f).astype(np.float32) * np.float32(256) + np.float32(0.5)) / np.float32(256)
    i0 = np.clip(f.astype(int), 0, size - 1); i1 = np.clip(f.astype(int) + 1, 0, size - 1)
    return a.astype(np.float32), i0, i1
a, i0, i1 = wq(xy[:, 0], w)
b, j0, j1 = wq(xy[:, 1], h)
t00, t10, t01, t11 = m[j0, i0], m[j0, i1], m[j1, i0], m[j1, i1]
A, B = a[:, None], b[:, None]
one = np.float32(1)
forms = {
    "fp32 weights w00*t00+w10*t10+w01*t01+w11*t11": ((one - A) * (one - B)) * t00 + (A * (one - B)) * t10 + ((one - A) * B) * t01 + (A * B) * t11,
    "fp32 nested lerp in x then y": (one - B) * ((one - A) * t00 + A * t10) + B * ((one - A) * t01 + A * t11),
    "fp64 exact with the quantised weights": ((1 - A.astype(np.float64)) * (1 - B) * t00 + A.astype(np.float64) * (1 - B) * t10
                                               + (1 - A.astype(np.float64)) * B * t01 + A.astype(np.float64) * B * t11),
}
# the rule that fits: fractions rounded to 1/256, w11 = round(a b) in 1/256 units, the rest by subtraction
qa, qb = np.floor(a * np.float32(256) + np.float32(0.5)).astype(np.float64), np.floor(b * np.float32(256) + np.float32(0.5)).astype(np.float64)
q11 = np.floor(qa * qb / 256.0 + 0.5)
W4 = [(256 - qa - qb + q11) / 256.0, (qa - q11) / 256.0, (qb - q11) / 256.0, q11 / 256.0]
forms["MEASURED RULE: w11 = round(a b), others by subtraction (1/256 units)"] = sum(wk[:, None] * tk for wk, tk in zip(W4, (t00, t10, t01, t11)))
res2 = {}
for k, v in forms.items():
    v32 = v.astype(np.float32)
    dd = np.abs(v32.astype(np.float64) - got)
    nz = np.abs(got) > 1e-6
    res2[k] = {"frac_bit_exact_nonzero_channels": float((v32 == got)[:, [0, 2]].mean()), "max_abs_err": float(dd.max()),
               "p99_abs_err": float(np.quantile(dd[nz], 0.99))}
print(json.dumps({"gpu": "B200 texture unit, float4 texture, cudaFilterModeLinear, normalised coordinates, clamp",
                  "one_dimensional_weight": res, "two_dimensional_on_ltc_table": res2, "samples": [len(xs), n]}, indent=1))

"""The reference's own device source running ON THE GPU (oracle/_ref/libref_gpu_*.so: cuda_src/deviceCode.cu
compiled by nvcc for sm_100a behind oracle/ref_shim_gpu, traceRay hook = this repository's BVH, tex2D = the
texture unit) against the product's kernels -- at sizes the CPU oracle cannot reach in test time, up to
BASELINE's full 1080p / 1 M-triangle frame.  TEST INFRASTRUCTURE on one side, product on the other."""
import numpy as np
import pytest

import _oracle
import _parity as P
from optix_renderer_b200 import api, scenes
from test_gpu_frames import BUFS, check_frame, gpu_frame

pytestmark = pytest.mark.gpu


def check_vs_megakernel(got, ref, rt, tag=""):
    """Bars for GPU-against-GPU frames.  nvcc contracts the reference's barycentric sums and dot products into
    fmas, the product rounds every step of the geometry path explicitly (to match the CPU checker bit for
    bit), so G-buffer values differ in the last bits: absolute RMSE <= 1e-5 (scene units are metres), and
    <= 1e-4 relative on >= 98 % of the values (the rest are components next to zero, e.g. of smooth normals).
    LTC / ratio terms: <= 1e-4 relative on >= 99.5 % of pixels (the product fetches the table with the texture
    unit's measured weight rule, tools/probe_tex.py); stochastic images RMSE <= 1e-3 (north_star's tolerances)."""
    rep = P.frame_report({k: got[k] for k in BUFS}, ref, got["hit_prim"], ref["hit_prim"])
    n = rep["pixels"]
    keys = ("hit_mismatch", "position", "normal") + (("accum",) if rt != _oracle.RATIO else ("ltc", "sto_direct", "sto_no_vis"))
    print(tag, {k: (rep[k] if not isinstance(rep[k], dict) else {m: rep[k][m] for m in ("frac_le_1e-4", "rmse", "nan_mismatch")})
                for k in keys})
    assert rep["hit_mismatch"] <= max(2, n // 2000), rep["hit_mismatch"]
    for k in ("position", "normal", "uv", "albedo", "alpha", "materialID"):
        assert rep[k]["rmse"] <= 1e-5 and rep[k]["frac_le_1e-4"] >= 0.98, (k, rep[k])
    if rt in (_oracle.LTC_BASELINE, _oracle.RATIO):
        k = "accum" if rt == _oracle.LTC_BASELINE else "ltc"
        assert rep[k]["frac_le_1e-4"] >= 0.995, (k, rep[k])
        assert rep[k]["nan_mismatch"] <= max(1, n // 5000)
    if rt == _oracle.RATIO:
        for k in ("sto_direct", "sto_no_vis"):
            assert rep[k]["rmse"] <= 1e-3, (k, rep[k])
    if rt == _oracle.PATH:
        assert rep["accum"]["rmse"] <= 1e-3, rep["accum"]
    return rep


def ref_gpu(mode):
    kind = "reference-gpu-intended" if mode else "reference-gpu-literal"
    if not _oracle.available(kind):
        pytest.skip("oracle/_ref/%s not built (make -C oracle ref_gpu needs the reference tree)" % kind)
    return _oracle.Oracle(kind)


@pytest.mark.parametrize("mode", [0, 1])
def test_megakernel_vs_cpu_reference(mode):
    """The shim itself: the reference source on the GPU against the same source on the CPU (texture unit vs
    emulated bilinear fetch with 1/256 weights, libdevice vs glibc, fma contraction)."""
    sc = scenes.scene_cornell(96, 80)
    g = ref_gpu(mode)
    c = _oracle.Oracle("reference-intended" if mode else "reference-literal")
    for o in (g, c):
        o.load_scene(sc, quantise=1)
        assert o.set_options(mode, 4, 4 if mode else 1) == 0
    for rt in (_oracle.LTC_BASELINE, _oracle.RATIO, _oracle.PATH):
        a, b = g.render(96, 80, rt, 0), c.render(96, 80, rt, 0, nthreads=8)
        rep = P.frame_report({k: a[k] for k in BUFS}, b, a["hit_prim"], b["hit_prim"])
        assert rep["hit_mismatch"] == 0
        for k in ("position", "normal", "albedo"):  # nvcc contracts the barycentric sums into fmas, g++ was told not to
            assert rep[k]["frac_le_1e-4"] >= 0.99 and rep[k]["rmse"] <= 1e-5, (k, rep[k])
        k = "accum" if rt == _oracle.LTC_BASELINE else ("ltc" if rt == _oracle.RATIO else None)
        if k:
            assert rep[k]["frac_le_1e-4"] >= 0.995, (k, rep[k])
        if rt == _oracle.RATIO:
            assert rep["sto_direct"]["rmse"] <= 1e-3 and rep["sto_no_vis"]["rmse"] <= 1e-3


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("scene_fn,W,H", [(lambda w, h: scenes.scene_cornell(w, h), 512, 384),
                                          (lambda w, h: scenes.scene_synthetic(200_000, 8, w, h), 640, 360)])
def test_product_vs_megakernel(scene_fn, W, H, mode):
    sc = scene_fn(W, H)
    g = ref_gpu(mode)
    g.load_scene(sc)
    depth = 8 if mode else 1
    assert g.set_options(mode, 4, depth) == 0
    r = api.Renderer(sc, quantise_lut=1)  # the texture unit filters with 1/256 weights
    for rt in (_oracle.LTC_BASELINE, _oracle.RATIO, _oracle.PATH):
        ref = g.render(W, H, rt, 0)
        got = gpu_frame(r, rt, 0, mode, 4, depth)
        check_vs_megakernel(got, ref, rt, "%dx%d mode %d rt %d" % (W, H, mode, rt))
        rays = W * H + got["counters"]["rays_shadow"] + got["counters"]["rays_bounce"]
        assert abs(rays - ref["rays"]) <= max(4, ref["rays"] // 500), (rays, ref["rays"])
    r.close()


def test_product_vs_megakernel_full_size():
    """BASELINE configs[2] at full size: 1920x1080, 1 M triangles, 32 TriLights, RATIO intended 4 spp."""
    W, H = 1920, 1080
    sc = scenes.scene_synthetic(1_000_000, 16, W, H)
    g = ref_gpu(1)
    g.load_scene(sc)
    assert g.set_options(1, 4, 1) == 0
    r = api.Renderer(sc, quantise_lut=1)
    ref = g.render(W, H, _oracle.RATIO, 0)
    got = gpu_frame(r, _oracle.RATIO, 0, 1, 4, 1)
    rep = check_vs_megakernel(got, ref, _oracle.RATIO, "full size")
    assert abs(got["counters"]["rays_shadow"] + W * H - ref["rays"]) <= ref["rays"] // 2000
    print("full-size parity:", {k: rep[k] for k in ("hit_mismatch", "ltc", "sto_direct", "sto_no_vis")}, "megakernel ms", g.last_ms())
    r.close()

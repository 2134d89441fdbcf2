"""GPU: whole frames of the CUDA wavefront path through the C-ABI against the oracle (the
reference's own device source compiled for the CPU when oracle/_ref is present, else the
restatement pinned to it), plus size-independent properties at full BASELINE sizes.

Tolerances: hit primitive ids exact except ties (counted); G-buffers 1e-5 relative; LTC radiance
<= 1e-4 relative on >= 99.9 % of pixels (case-flip pixels counted); stochastic
images RMSE <= 1e-3; integer outputs (ray counts, queue lengths) exact."""
import os

import numpy as np
import pytest

import _oracle
import _parity as P
from conftest import best_oracle
from optix_renderer_b200 import api, scenes

pytestmark = pytest.mark.gpu

BUFS = ("position", "normal", "uv", "albedo", "alpha", "materialID", "ltc", "sto_direct", "sto_no_vis", "accum")


def gpu_frame(r, rt, accum_id=0, mode=0, spp=4, depth=1, rows=None):
    # rayGen leaves ltc / sto_* of MISS pixels untouched (deviceCode.cu:91-95,117-145), so they keep
    # whatever an earlier frame with another jitter wrote; the oracle starts from zeroed buffers
    z = np.zeros((r.H, r.W, 4), np.float32)
    for k in ("ltc", "sto_direct", "sto_no_vis"):
        r.upload(k, z)
    r.render(rt, accum_id=accum_id, mode=mode, shadow_spp=spp, max_depth=depth, rows=rows)
    out = {k: r.download(k) for k in BUFS}
    out["rgba8"] = r.download("rgba8")
    out["hit_prim"], out["hit_uv"], _ = r.hit_records()
    out["counters"] = r.counters()
    return out


def check_frame(got, ref, rt, strict_hits=True):
    rep = P.frame_report({k: got[k] for k in BUFS}, ref, got["hit_prim"], ref["hit_prim"])
    n = rep["pixels"]
    assert rep["hit_mismatch"] <= (0 if strict_hits else max(2, n // 2000)), rep["hit_mismatch"]
    for k in ("position", "normal", "uv", "albedo", "alpha", "materialID"):
        assert rep[k]["max_rel"] <= 1e-5, (k, rep[k])
    if rt in (_oracle.LTC_BASELINE, _oracle.RATIO):
        k = "accum" if rt == _oracle.LTC_BASELINE else "ltc"
        assert rep[k]["frac_le_1e-4"] >= 0.999, (k, rep[k])
        assert rep[k]["nan_mismatch"] <= max(1, n // 5000)
    if rt == _oracle.RATIO:
        for k in ("sto_direct", "sto_no_vis"):
            assert rep[k]["rmse"] <= 1e-3, (k, rep[k])
            assert rep[k]["nan_mismatch"] <= max(1, n // 5000)
    if rt == _oracle.PATH:
        assert rep["accum"]["rmse"] <= 1e-3, rep["accum"]
        assert rep["accum"]["nan_mismatch"] <= max(2, n // 1000)
    return rep


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("scene_fn,W,H", [(scenes.scene_floor_blocker, 64, 64), (scenes.scene_cornell, 96, 80),
                                          (scenes.scene_textured, 96, 80)])
def test_frames_vs_oracle(renderer_factory, scene_fn, W, H, mode):
    sc = scene_fn(W, H)
    r = renderer_factory(sc)
    o = best_oracle(mode)
    o.load_scene(sc)
    depth = 8 if mode else 1
    assert o.set_options(mode, 4, depth) == 0
    for rt in (_oracle.LTC_BASELINE, _oracle.RATIO, _oracle.PATH):
        ref = o.render(W, H, rt, 0, nthreads=8)
        got = gpu_frame(r, rt, 0, mode, 4, depth)
        check_frame(got, ref, rt)
        c = got["counters"]
        assert c["rays_primary"] > 0
        rays = W * H + c["rays_shadow"] + c["rays_bounce"]                  # device-counted
        if rt == _oracle.PATH and mode == 1:
            # a bounce ray that grazes an edge can end a path one bounce earlier or later when the
            # direction differs in the last ulp (fma contraction, libdevice vs glibc sin/cos)
            assert abs(rays - ref["rays"]) <= max(4, ref["rays"] // 500), (rays, ref["rays"])
        else:
            assert rays == ref["rays"]                                      # exact


def test_cfg2_cornell_1024_one_shadow_sample(renderer_factory):
    """BASELINE configs[1] as stated: the Cornell box at 1024x1024, one quad light, LTC direct + ONE stochastic
    shadow ray per pixel (RATIO, shadow_spp = 1), accumId 0..3, against the reference's own source on the host
    cores; mismatching pixels classified."""
    W = H = 1024
    sc = scenes.scene_cornell(W, H)
    r = renderer_factory(sc)
    o = best_oracle(1)
    o.load_scene(sc)
    assert o.set_options(1, 1, 1) == 0
    acc = None
    for i in range(4):
        ref = o.render(W, H, _oracle.RATIO, i, accum=acc, nthreads=os.cpu_count() or 8)
        acc = ref["accum"]
        z = np.zeros((H, W, 4), np.float32)
        for k in ("ltc", "sto_direct", "sto_no_vis"):
            r.upload(k, z)
        r.render(api.RATIO, accum_id=i, mode=1, shadow_spp=1)
        got = {k: r.download(k) for k in BUFS}
        got["hit_prim"], got["hit_uv"], _ = r.hit_records()
        c = r.counters()
        rep = check_frame(got, ref, _oracle.RATIO, strict_hits=False)
        assert W * H + c["rays_shadow"] == ref["rays"] and c["rays_shadow"] == c["active_pixels"]   # 1 spp, exact
        hits = P.classify_hit_mismatches(got, ref, sc["camera"]["pos"], scene=sc)
        assert hits["other"] == 0, hits
        print("cfg2 frame %d:" % i, hits, {k: rep[k]["frac_le_1e-4"] for k in ("ltc", "accum")},
              {k: rep[k]["rmse"] for k in ("sto_direct", "sto_no_vis")})


def test_lut_64x64_vs_oracle(renderer_factory):
    """A 64x64 LTC table (north_star; too large for the kernel's shared-memory copy: the global-memory path)
    against the reference source with the same table."""
    sc = scenes.scene_cornell(96, 80)
    sc = dict(sc, lut=scenes.upsample_lut(sc["lut"], 64))
    r = renderer_factory(sc)
    o = best_oracle(1)
    o.load_scene(sc)
    assert o.set_options(1, 4, 1) == 0
    for rt in (_oracle.LTC_BASELINE, _oracle.RATIO):
        ref = o.render(96, 80, rt, 0, nthreads=8)
        check_frame(gpu_frame(r, rt, 0, 1, 4, 1), ref, rt)


def test_debug_aov_modes(renderer_factory):
    sc = scenes.scene_floor_blocker(48, 48)
    r = renderer_factory(sc)
    o = best_oracle(0)
    o.load_scene(sc)
    for rt in range(7):
        ref = o.render(48, 48, rt, 0)
        got = gpu_frame(r, rt)
        assert P.rel_err(got["accum"], ref["accum"]).max() <= 1e-5
        assert (got["rgba8"] != ref["rgba8"]).mean() < 0.01


def test_golden_frames(renderer_factory):
    """The committed fixtures (reference source output), no oracle library involved."""
    g = np.load(os.path.join(P.G, "frames_floor32.npz"))
    r = renderer_factory(scenes.scene_floor_blocker(32, 32))
    for mode in (0, 1):
        for rt, name in ((_oracle.LTC_BASELINE, "ltc"), (_oracle.RATIO, "ratio"), (_oracle.PATH, "path")):
            pre = "%s_m%d_a0_" % (name, mode)
            ref = {k: g[pre + k] for k in BUFS + ("hit_prim",)}
            got = gpu_frame(r, rt, 0, mode, 4, 4 if mode else 1)
            check_frame(got, ref, rt)
            r.render(rt, accum_id=1, mode=mode, shadow_spp=4, max_depth=4 if mode else 1)
            acc = r.download("accum")
            ref1 = g["%s_m%d_a1_accum" % (name, mode)]
            m = np.isfinite(acc) & np.isfinite(ref1)
            assert np.sqrt(np.mean((acc[m] - ref1[m]) ** 2)) <= 1e-3
            assert P.rel_err(acc, ref1)[np.isfinite(ref1).all(-1) & np.isfinite(acc).all(-1)].mean() < 1e-4


def test_progressive_accumulation_and_rgba8(renderer_factory):
    sc = scenes.scene_cornell(64, 64)
    r = renderer_factory(sc)
    o = best_oracle(1)
    o.load_scene(sc)
    o.set_options(1, 4, 5)
    acc = None
    for i in range(4):
        ref = o.render(64, 64, _oracle.PATH, i, accum=acc, nthreads=8)
        acc = ref["accum"]
        r.render(api.PATH, accum_id=i, mode=1, max_depth=5)
    got = r.download("accum")
    m = np.isfinite(got).all(-1) & np.isfinite(acc).all(-1)
    assert (np.isnan(got).any(-1) != np.isnan(acc).any(-1)).sum() <= 4
    assert np.sqrt(np.mean((got[m] - acc[m]) ** 2)) <= 1e-3
    assert (r.download("rgba8")[m] != ref["rgba8"][m]).mean() < 0.02


def test_row_tiles_are_bit_identical_to_full_frame(renderer_factory):
    """8(e): pixels do not depend on the partition (RNG keyed by global pixel and dims)."""
    sc = scenes.scene_cornell(64, 48)
    r = renderer_factory(sc)
    full = gpu_frame(r, api.RATIO, 2, 1, 4, 1)
    r2 = renderer_factory(sc)
    for rows in ((0, 13), (13, 32), (32, 48)):
        r2.render(api.RATIO, accum_id=2, mode=1, shadow_spp=4, rows=rows)
    for k in ("ltc", "sto_direct", "sto_no_vis", "position", "accum"):
        a, b = full[k], r2.download(k)
        assert np.array_equal(a, b, equal_nan=True), k
    r3 = renderer_factory(sc)
    r3.render(api.PATH, accum_id=0, mode=1, max_depth=6)
    a = r3.download("accum")
    for rows in ((0, 20), (20, 48)):
        r2.render(api.PATH, accum_id=0, mode=1, max_depth=6, rows=rows)
    assert np.array_equal(a, r2.download("accum"), equal_nan=True)


def test_trace_hit_ids_bvh_vs_brute_force_and_oracle(renderer_factory):
    sc = scenes.scene_synthetic(n_tris=20000, n_quad_lights=4, width=32, height=32)
    r = renderer_factory(sc)
    rng = np.random.default_rng(11)
    n = 20000
    org = (rng.uniform(-1, 1, (n, 3)) * 3).astype(np.float32)
    org[:, 2] = np.abs(org[:, 2]) + 0.3
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    p1, t1, u1, v1 = r.trace_rays(org, d, brute_force=False)
    c = r.counters()
    p2, t2, u2, v2 = r.trace_rays(org, d, brute_force=True)
    assert np.array_equal(p1, p2) and np.array_equal(t1, t2) and np.array_equal(u1, u2) and np.array_equal(v1, v2)
    assert 0 < c["tris_tested"] < 0.05 * n * c["bvh_triangles"]      # the BVH actually prunes
    # against the oracle's stand-in closest hit through a rendered frame
    o = best_oracle(0)
    o.load_scene(sc)
    ref = o.render(32, 32, _oracle.NORMALS, 0, nthreads=8)
    got = gpu_frame(r, api.NORMALS)
    ties = int((got["hit_prim"] != ref["hit_prim"]).sum())
    assert ties <= 2, ties


@pytest.mark.parametrize("n_tris,leaf", [(1, 0), (3, 4), (5, 4), (2000, 1), (20000, 2), (200000, 8)])
def test_sah_and_lbvh_builders_agree_with_brute_force(renderer_factory, n_tris, leaf):
    """Both GPU builders (binned SAH, the default; Morton LBVH) against the brute-force closest hit, from
    one triangle (root is a leaf) to 200 k, at several leaf sizes; frames rendered over the two trees
    are bit-identical (the result does not depend on the hierarchy)."""
    sc = scenes.scene_synthetic(n_tris=n_tris, n_quad_lights=2, width=48, height=32, patches=1 if n_tris < 200 else 8)
    rng = np.random.default_rng(5)
    n = 6000
    org = (rng.uniform(-1, 1, (n, 3)) * 9).astype(np.float32)
    org[:, 2] = np.abs(org[:, 2]) * 0.5 + 0.2
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[:50, 0] = 0.0                      # axis-parallel components (infinite slab distances)
    d[50:100, 2] = 0.0
    frames = []
    for builder in (api.BVH_SAH, api.BVH_LBVH):
        r = renderer_factory(sc, max_leaf_tris=leaf, builder=builder)
        p1, t1, u1, v1 = r.trace_rays(org, d, brute_force=False)
        c = r.counters()
        p2, t2, u2, v2 = r.trace_rays(org, d, brute_force=True)
        assert np.array_equal(p1, p2) and np.array_equal(t1, t2) and np.array_equal(u1, u2) and np.array_equal(v1, v2)
        assert c["bvh_triangles"] == scenes.n_triangles(sc)
        if n_tris >= 20000:
            assert c["tris_tested"] < 0.05 * n * c["bvh_triangles"]
        # the ray-stream kernel (persistent warps with refill) finds the same hits, bit for bit
        p3, t3, u3, v3 = r.trace_rays(org, d, brute_force=2)
        assert np.array_equal(p3, p2) and np.array_equal(t3, t2) and np.array_equal(u3, u2) and np.array_equal(v3, v2)
        c3 = r.counters()
        assert c3["tris_tested"] >= (p3 >= 0).sum()
        r.render(api.RATIO, accum_id=1, mode=1, shadow_spp=2)
        frames.append({k: r.download(k) for k in ("ltc", "sto_direct", "sto_no_vis", "position")})
        frames[-1]["hit"] = r.hit_records()[0]
    for k in frames[0]:
        assert np.array_equal(frames[0][k], frames[1][k], equal_nan=True), k


def test_ltc_gbuffer_pass_and_ratio_combine(renderer_factory):
    """BASELINE config 1 shape: saved G-buffer in, LTC pass + ratio arithmetic out."""
    g = np.load(os.path.join(P.G, "gbuffer48.npz"))
    sc = scenes.scene_floor_blocker(48, 48)
    r = renderer_factory(sc)
    for k in ("position", "normal", "albedo", "alpha"):
        r.upload(k, g[k])
    r.ltc_gbuffer_pass()
    out = r.download("ltc")
    e = P.rel_err(out, g["ltc_gbuffer"])
    assert (e <= 1e-4).mean() >= 0.999 and np.array_equal(out[..., 3], np.ones((48, 48), np.float32))
    r.upload("ltc", g["ltc_frame"])
    r.upload("sto_direct", g["sto_direct"])
    r.upload("sto_no_vis", g["sto_no_vis"])
    r.ratio_combine(use_denoised=False)
    with np.errstate(all="ignore"):
        want = g["ltc_frame"][..., :3] * (g["sto_direct"][..., :3] / g["sto_no_vis"][..., :3])
    got = r.download("ratio_final")[..., :3]
    assert np.array_equal(got, want, equal_nan=True)        # IEEE mul/div: bit-exact
    # host-buffer entry point, ragged length (not a multiple of 4), zeros -> inf/nan like numpy
    rng = np.random.default_rng(5)
    for n in (0, 1, 3, 4, 1001, 48 * 48 * 3):
        a, b, c = (rng.uniform(0, 2, n).astype(np.float32) for _ in range(3))
        if n > 10:
            c[::7] = 0
            b[::14] = 0
        with np.errstate(all="ignore"):
            want = a * (b / c)
        assert np.array_equal(r.ratio_estimator_host(a, b, c), want, equal_nan=True)


def test_denoise_vs_definition(renderer_factory, restatement):
    sc = scenes.scene_cornell(96, 64)
    r = renderer_factory(sc)
    r.render(api.RATIO, accum_id=0, mode=1, shadow_spp=4)
    f = {k: r.download(k) for k in ("normal", "position", "materialID", "sto_direct", "sto_no_vis")}
    for iters, sc_ in ((1, 0.0), (5, 0.0), (3, 0.5)):
        r.denoise(iterations=iters, sigma_colour=sc_)
        gd, gn = r.download("sto_direct_dn"), r.download("sto_no_vis_dn")
        od, on = restatement.denoise(f["normal"], f["position"], f["materialID"], f["sto_direct"], f["sto_no_vis"],
                                     sc["camera"]["pos"], iterations=iters, sigma_colour=sc_, nthreads=8)
        for a, b in ((gd, od), (gn, on)):
            m = np.isfinite(a) & np.isfinite(b)
            assert (np.isnan(a) != np.isnan(b)).sum() == 0
            np.testing.assert_allclose(a[m], b[m], rtol=2e-4, atol=1e-6)
    r.ratio_combine(use_denoised=True)
    with np.errstate(all="ignore"):
        want = r.download("ltc")[..., :3] * (gd[..., :3] / gn[..., :3])
    assert np.array_equal(r.download("ratio_final")[..., :3], want, equal_nan=True)
    # NaN values are skipped as taps (and replaced where neighbours exist); a pixel with a NaN guide is passed
    # through and never used as a tap -- same as the definition, where every weight it enters is NaN
    rng = np.random.default_rng(5)
    ok = np.argwhere(f["materialID"][..., 0] != 0)
    pick = ok[rng.choice(len(ok), 12, replace=False)]
    for k, (y, x) in enumerate(pick):
        if k < 5:
            f["sto_direct"][y, x, :3] = np.nan
        elif k < 9:
            f["sto_no_vis"][y, x, :3] = np.nan
        else:
            f["normal"][y, x, k % 3] = np.nan
    for k in ("normal", "sto_direct", "sto_no_vis"):
        r.upload(k, f[k])
    for iters, sc_ in ((1, 0.0), (4, 0.0), (2, 0.5)):
        r.denoise(iterations=iters, sigma_colour=sc_)
        gd, gn = r.download("sto_direct_dn"), r.download("sto_no_vis_dn")
        od, on = restatement.denoise(f["normal"], f["position"], f["materialID"], f["sto_direct"], f["sto_no_vis"],
                                     sc["camera"]["pos"], iterations=iters, sigma_colour=sc_, nthreads=8)
        for a, b in ((gd, od), (gn, on)):
            assert np.array_equal(np.isnan(a[..., 0]), np.isnan(b[..., 0])), (iters, sc_)
            m = np.isfinite(a) & np.isfinite(b)
            np.testing.assert_allclose(a[m], b[m], rtol=2e-4, atol=1e-6)


def test_edge_cases(renderer_factory):
    # a scene with no surface hit at all: every pixel is a miss
    sc = scenes.scene_floor_blocker(16, 16)
    cam = scenes.camera_from_lookat((0, 0, 50), (0, 0, 100), (0, 1, 0), 0.66, 16, 16)
    sc2 = dict(sc, camera=cam)
    r = renderer_factory(sc2)
    for rt in (api.LTC_BASELINE, api.RATIO, api.PATH):
        r.render(rt, mode=1, max_depth=4)
        assert not r.download("accum")[..., :3].any()
        assert np.array_equal(r.download("position")[..., 3], np.ones((16, 16), np.float32))
        c = r.counters()
        assert c["active_pixels"] == 0 and c["rays_shadow"] == 0 and c["rays_bounce"] == 0
    # 1x1 frame and a non-multiple-of-tile frame
    for W, H in ((1, 1), (17, 9)):
        scx = scenes.scene_floor_blocker(W, H)
        rr = renderer_factory(scx)
        o = best_oracle(1)
        o.load_scene(scx)
        o.set_options(1, 4, 3)
        ref = o.render(W, H, _oracle.RATIO, 0)
        got = gpu_frame(rr, api.RATIO, 0, 1, 4, 3)
        check_frame(got, ref, _oracle.RATIO)
    # error behaviour: bad arguments return status codes with a message, never crash
    with pytest.raises(api.OrbError):
        r.render(api.RATIO, rows=(5, 3))
    with pytest.raises(api.OrbError):
        r.render(api.RATIO, mode=7)
    with pytest.raises(api.OrbError):
        api.Renderer().render(api.RATIO)


def test_full_size_properties(renderer_factory):
    """BASELINE config 3 size (1920x1080, ~1M triangles, 16 quad lights): properties that need
    no oracle -- exact ray accounting, idempotence, tile independence, finite LTC."""
    sc = scenes.scene_synthetic(n_tris=1_000_000, n_quad_lights=16, width=1920, height=1080)
    r = renderer_factory(sc)
    r.render(api.RATIO, accum_id=0, mode=1, shadow_spp=4)
    c = r.counters()
    assert c["rays_primary"] == 1920 * 1080
    assert c["rays_shadow"] == 4 * c["active_pixels"]
    assert c["bvh_triangles"] == scenes.n_triangles(sc)
    ltc = r.download("ltc")
    band = r.download("ltc", rows=(500, 540))
    sd = r.download("sto_direct", rows=(500, 540))
    sn = r.download("sto_no_vis", rows=(500, 540))
    fin = np.isfinite(sd[..., 0]) & np.isfinite(sn[..., 0])
    hit = r.download("materialID", rows=(500, 540))[..., 0] != 0
    assert (sd[..., 0][fin & hit] <= sn[..., 0][fin & hit] * (1 + 1e-5) + 1e-7).all()    # shadowed <= unshadowed
    assert np.isfinite(ltc[..., :3]).mean() > 0.999
    r.render(api.RATIO, accum_id=0, mode=1, shadow_spp=4)                      # idempotent
    assert np.array_equal(ltc, r.download("ltc"), equal_nan=True)
    r.render(api.RATIO, accum_id=0, mode=1, shadow_spp=4, rows=(500, 540))     # tile == full
    assert np.array_equal(band, r.download("ltc", rows=(500, 540)), equal_nan=True)
    assert np.array_equal(sd, r.download("sto_direct", rows=(500, 540)), equal_nan=True)
    # hit ids of the BVH against brute force on a sample of camera rays
    prim, _, _ = r.hit_records()
    ys, xs = np.random.default_rng(1).integers(0, 1080, 300), np.random.default_rng(2).integers(0, 1920, 300)
    pos = r.download("position")
    cam = np.asarray(sc["camera"]["pos"], np.float32)
    m = prim[ys, xs] >= 0
    d = pos[ys, xs, :3][m] - cam
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    o = np.repeat(cam[None], d.shape[0], 0)
    p1, t1, _, _ = r.trace_rays(o, d, brute_force=False)
    p2, t2, _, _ = r.trace_rays(o, d, brute_force=True)
    assert np.array_equal(p1, p2) and np.array_equal(t1, t2)


def test_resolve_per_sample_lanes_equal_per_pixel_kernel(renderer_factory, monkeypatch):
    """k_shadow_resolve_lanes (one thread per sample, in-order sum over the pixel's lanes) is bit-identical to
    the one-thread-per-pixel resolve kernel, for every lane-parallel spp; odd spp fall back to the latter."""
    sc = scenes.scene_synthetic(n_tris=50000, n_quad_lights=5, width=161, height=95)
    a = renderer_factory(sc)
    monkeypatch.setenv("ORB_RESOLVE_LANES", "0")
    b = renderer_factory(sc)
    monkeypatch.delenv("ORB_RESOLVE_LANES")
    for spp in (2, 4, 8, 5):
        for r in (a, b):
            r.render(api.RATIO, accum_id=3, mode=1, shadow_spp=spp)
        for k in ("sto_direct", "sto_no_vis"):
            assert np.array_equal(a.download(k), b.download(k), equal_nan=True), (spp, k)
        assert a.counters()["rays_shadow"] == b.counters()["rays_shadow"] == spp * a.counters()["active_pixels"]


def test_quantised_node_stream_kernel(renderer_factory, monkeypatch):
    """ORB_TRACE_VARIANT=3: the ray-stream kernel over the 32-byte nodes of bvhq.cuh (child boxes on a 16-bit grid,
    conservative) returns the brute-force hits bit for bit, renders the same RATIO / PATH-wavefront frames as the
    fp32-node kernel, and fetches about as many nodes."""
    sc = scenes.scene_synthetic(n_tris=60000, n_quad_lights=4, width=128, height=72)
    a = renderer_factory(sc)
    monkeypatch.setenv("ORB_TRACE_VARIANT", "3")
    b = renderer_factory(sc)
    monkeypatch.delenv("ORB_TRACE_VARIANT")
    rng = np.random.default_rng(12)
    n = 20000
    org = (rng.uniform(-1, 1, (n, 3)) * 9).astype(np.float32)
    org[:, 2] = np.abs(org[:, 2]) * 0.5 + 0.2
    org[:300] *= 6                      # outside the grid
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[:100, 1] = 0.0
    d[100:200, 2] = 0.0
    want = b.trace_rays(org, d, brute_force=True)
    got = b.trace_rays(org, d, brute_force=2)
    q_nodes = b.counters()["nodes_visited"]
    for x, y in zip(got, want):
        assert np.array_equal(x, y)
    a.trace_rays(org, d, brute_force=2)
    f_nodes = a.counters()["nodes_visited"]
    assert 0.95 * f_nodes < q_nodes < 1.05 * f_nodes, (q_nodes, f_nodes)
    for r in (a, b):
        r.render(api.RATIO, accum_id=2, mode=1, shadow_spp=4)
    for k in ("sto_direct", "sto_no_vis", "ltc"):
        assert np.array_equal(a.download(k), b.download(k), equal_nan=True), k
    for r in (a, b):
        r.render(api.PATH, accum_id=0, mode=1, max_depth=4, path_wavefront=True)
    assert np.array_equal(a.download("accum"), b.download("accum"), equal_nan=True)


def test_pipelined_download_equals_blocking_download(renderer_factory):
    """orb_buffer_download_rows_async + orb_download_wait: frame i is read back while frames i+1.. render; each
    host frame equals a blocking download taken right after the same frame, bit for bit (full frame and a band)."""
    import torch
    sc = scenes.scene_synthetic(n_tris=50000, n_quad_lights=5, width=160, height=96)
    a, b = renderer_factory(sc), renderer_factory(sc)
    n = 5
    pin_f = torch.empty((n, 96, 160, 4), dtype=torch.float32, pin_memory=True).numpy()
    pin_u = torch.empty((n, 96, 160), dtype=torch.int32, pin_memory=True).numpy().view(np.uint32)
    pin_band = torch.empty((n, 30, 160, 4), dtype=torch.float32, pin_memory=True).numpy()
    want = []
    for i in range(n):
        a.render(api.RATIO, accum_id=i, mode=1, shadow_spp=4)
        want.append((a.download("accum"), a.download("rgba8"), a.download("sto_direct", rows=(11, 41))))
        b.render(api.RATIO, accum_id=i, mode=1, shadow_spp=4, sync=False)
        b.download_async("accum", pin_f[i])
        b.download_async("rgba8", pin_u[i])
        b.download_async("sto_direct", pin_band[i], rows=(11, 41))
    b.download_wait()
    for i in range(n):
        assert np.array_equal(pin_f[i], want[i][0], equal_nan=True), i
        assert np.array_equal(pin_u[i], want[i][1]), i
        assert np.array_equal(pin_band[i], want[i][2], equal_nan=True), i
    with pytest.raises(api.OrbError):
        b.download_async("accum", pin_f[0], rows=(5, 500))


def test_instrumented_traversal_counts(renderer_factory):
    """ORB_RENDER_COUNT_TRAVERSAL: the instrumented twin of the shadow-ray kernel writes the same frame and
    reports node / triangle fetch counts (bench.py's algorithmic bytes); the TRACE stage timer covers it."""
    sc = scenes.scene_synthetic(n_tris=50000, n_quad_lights=5, width=160, height=96)
    r = renderer_factory(sc)
    r.render(api.RATIO, accum_id=0, mode=1, shadow_spp=4)
    a = {k: r.download(k) for k in ("sto_direct", "sto_no_vis", "ltc", "accum")}
    c0 = r.counters()
    assert c0["nodes_visited"] == 0 and c0["tris_tested"] == 0
    r.set_profiling(True)
    r.stage_times()
    r.render(api.RATIO, accum_id=0, mode=1, shadow_spp=4, count_traversal=True)
    st = r.stage_times()
    r.set_profiling(False)
    c1 = r.counters()
    for k in a:
        assert np.array_equal(a[k], r.download(k), equal_nan=True), k
    rays = c1["rays_shadow"]
    assert rays == 4 * c1["active_pixels"] > 0
    assert c1["nodes_visited"] >= rays            # every ray fetches at least the root
    assert 0 < c1["tris_tested"] < 64 * rays
    assert st["trace"][1] == 1 and st["trace"][0] > 0 and st["shadow"][1] == 2
    # the persistent path kernel has the same twin
    r.render(api.PATH, accum_id=0, mode=1, max_depth=4)
    want = r.download("accum")
    assert r.counters()["nodes_visited"] == 0
    r.render(api.PATH, accum_id=0, mode=1, max_depth=4, count_traversal=True)
    c2 = r.counters()
    assert np.array_equal(want, r.download("accum"), equal_nan=True)
    assert c2["nodes_visited"] >= c2["rays_bounce"] > 0 and c2["tris_tested"] > 0


def test_path_persistent_equals_wavefront(renderer_factory):
    """PATH: the persistent all-bounces kernel (default) against the per-bounce ray-stream wavefront: the
    same device functions in another schedule.  fma contraction may differ between the two kernels, so
    compared like the fused form below."""
    sc = scenes.scene_synthetic(n_tris=50000, n_quad_lights=5, width=160, height=96)
    r = renderer_factory(sc)
    for depth in (1, 3, 8):
        r.render(api.PATH, accum_id=0, mode=1, max_depth=depth, path_wavefront=True)
        a, ca = r.download("accum"), r.counters()
        r.render(api.PATH, accum_id=0, mode=1, max_depth=depth)
        b, cb = r.download("accum"), r.counters()
        assert (np.isnan(a).any(-1) != np.isnan(b).any(-1)).sum() <= 4
        m = np.isfinite(a).all(-1) & np.isfinite(b).all(-1)
        d = (a[m] - b[m])[:, :3] / np.maximum(1.0, np.abs(a[m][:, :3]))
        assert np.sqrt((d * d).mean()) <= 1e-3
        assert (np.abs(d).max(-1) > 1e-4).mean() < 2e-3
        assert abs(ca["rays_bounce"] - cb["rays_bounce"]) <= max(4, ca["rays_bounce"] // 500)
        assert cb["rays_bounce"] >= cb["active_pixels"]
        assert cb["kernel_launches"] - ca["kernel_launches"] == 3        # primary, all bounces, accumulate


def test_path_stream_equals_fused_bounce_kernel(renderer_factory):
    """PATH: per-bounce ray streams vs the fused bounce kernel.  Same arithmetic, different fma
    contraction, so a handful of grazing bounce rays may end a path differently: compared by
    normalised RMSE and ray count, not bit for bit."""
    sc = scenes.scene_synthetic(n_tris=50000, n_quad_lights=5, width=160, height=96)
    r = renderer_factory(sc)
    r.render(api.PATH, accum_id=0, mode=1, max_depth=8, fused_shadow=True)
    a, ca = r.download("accum"), r.counters()
    r.render(api.PATH, accum_id=0, mode=1, max_depth=8, fused_shadow=False, path_wavefront=True)
    b, cb = r.download("accum"), r.counters()
    assert (np.isnan(a).any(-1) != np.isnan(b).any(-1)).sum() <= 4
    m = np.isfinite(a).all(-1) & np.isfinite(b).all(-1)
    d = (a[m] - b[m])[:, :3] / np.maximum(1.0, np.abs(a[m][:, :3]))
    assert np.sqrt((d * d).mean()) <= 1e-3
    assert (np.abs(d).max(-1) > 1e-4).mean() < 2e-3
    assert abs(ca["rays_bounce"] - cb["rays_bounce"]) <= max(4, ca["rays_bounce"] // 500)
    assert cb["rays_bounce"] > cb["active_pixels"]


@pytest.mark.parametrize("W,H", [(96, 64), (203, 157)])
def test_atrous_four_pixels_per_thread_equals_one_pixel_kernel(renderer_factory, monkeypatch, W, H):
    """k_denoise_atrous4 (four pixels per thread sharing tap rows, shared weight sum, NaN fall-back) against the
    one-pixel kernel (ORB_ATROUS4=0): bit for bit -- every pass count up to 8 (run-time steps 32 .. 128), frames that are
    not a multiple of the CTA tile, bands, NaN values and guides; and the combine fused into the final pass."""
    sc = scenes.scene_cornell(W, H)
    a = renderer_factory(sc)
    monkeypatch.setenv("ORB_ATROUS4", "0")
    b = renderer_factory(sc)
    monkeypatch.delenv("ORB_ATROUS4")
    a.set_fused_combine(True)
    names = ("sto_direct_dn", "sto_no_vis_dn", "ratio_final")
    for r in (a, b):
        r.render(api.RATIO, accum_id=0, mode=1, shadow_spp=4)
    f = {k: a.download(k) for k in ("normal", "sto_direct", "sto_no_vis", "materialID")}
    rng = np.random.default_rng(9)
    for with_nan in (False, True):
        if with_nan:
            ok = np.argwhere(f["materialID"][..., 0] != 0)
            for k, (y, x) in enumerate(ok[rng.choice(len(ok), 40, replace=False)]):
                f[("sto_direct", "sto_no_vis", "normal")[k % 3]][y, x, :3] = np.nan if k % 5 else np.inf
            for r in (a, b):
                for k in ("normal", "sto_direct", "sto_no_vis"):
                    r.upload(k, f[k])
        for iters in (1, 2, 3, 5, 8):
            for rows in (None, (0, 37), (37, 38), (38, H)):
                for r in (a, b):
                    z = np.zeros((H, W, 4), np.float32)
                    for k in names:
                        r.upload(k, z)
                    r.denoise_pack()
                    r.denoise_filter(iterations=iters, rows=rows)
                b.ratio_combine(use_denoised=True, rows=rows)
                for k in names:
                    assert np.array_equal(a.download(k), b.download(k), equal_nan=True), (k, iters, rows, with_nan)
    assert a.counters()["kernel_launches"] < b.counters()["kernel_launches"]


def test_denoise_halves_and_row_counts(renderer_factory):
    """orb_denoise == orb_denoise_pack + orb_denoise_filter; a band filtered from packed rows that were
    written band by band (as ranks exchange them) equals the same rows of the full-frame filter; the
    per-row active counts equal what the G-buffer says."""
    sc = scenes.scene_cornell(96, 160)
    r = renderer_factory(sc)
    r.render(api.RATIO, accum_id=0, mode=1, shadow_spp=4)
    r.denoise(iterations=4)
    want = {k: r.download(k) for k in ("sto_direct_dn", "sto_no_vis_dn")}
    z = np.zeros((160, 96, 4), np.float32)
    for k in want:
        r.upload(k, z)
    for rows in ((0, 50), (50, 51), (51, 160)):     # every "rank" packs its own rows
        r.denoise_pack(rows=rows)
    g, v = r.download("dn_guide"), r.download("dn_val")
    assert g.shape == (160, 96, 4) and v.shape == (160, 96, 2)
    use = g[..., 3] >= 0                               # lights / misses carry the skip sentinel and a zero value
    assert np.array_equal(v[..., 0][use], r.download("sto_direct")[..., 0][use], equal_nan=True)
    assert use.any() and (~use).any() and np.all(g[..., 3][~use] == np.float32(-1e30)) and not v[~use].any()
    for rows in ((51, 160), (0, 50), (50, 51)):
        for name, a in (("dn_guide", g), ("dn_val", v)):  # the filter's ping-pong overwrote dn_val: restore the gathered rows
            r.upload(name, a)
        r.denoise_filter(iterations=4, rows=rows)
        for k in want:
            assert np.array_equal(r.download(k, rows=rows), want[k][rows[0]:rows[1]], equal_nan=True), (k, rows)
    mid = r.download("materialID")[..., 0]
    prim, _, _ = r.hit_records()
    assert np.array_equal(r.row_active_counts(), ((prim >= 0) & (mid != 0)).sum(1).astype(np.int32))


def _moved_scene(sc):
    """The same meshes after an animation step: every light quad swings down and sideways (rotation about
    x plus a translation), two surface patches get a travelling bump.  Triangle counts are unchanged."""
    import copy
    mv = copy.deepcopy(sc)
    c, s = np.float32(np.cos(0.35)), np.float32(np.sin(0.35))
    rot = np.array([[1, 0, 0], [0, c, -s], [0, s, c]], np.float32)
    changed = []
    for k, m in enumerate(mv["meshes"]):
        if m["is_light"]:
            ctr = m["vtx"].mean(0, dtype=np.float32)
            m["vtx"] = ((m["vtx"] - ctr) @ rot.T + ctr + np.array([0.4, -0.3, -1.7], np.float32)).astype(np.float32)
            m["nrm"] = (m["nrm"] @ rot.T).astype(np.float32)
            changed.append(k)
        elif k in (0, 5):
            v = m["vtx"].copy()
            v[:, 2] += (0.6 * np.sin(1.9 * v[:, 0] + 0.7) * np.cos(1.1 * v[:, 1])).astype(np.float32)
            m["vtx"] = v
            changed.append(k)
    mv["tri_lights"] = scenes.make_tri_lights(mv["meshes"])
    return mv, changed


@pytest.mark.parametrize("builder", ["sah", "lbvh"])
@pytest.mark.parametrize("n_tris", [6, 20000])
def test_refit_after_moving_meshes(renderer_factory, builder, n_tris):
    """orb_scene_update_mesh + orb_scene_refit (animated lights without a rebuild): the
    refitted hierarchy finds the brute-force closest hits bit for bit, and hits and RATIO frames equal those of
    a context that built the moved scene from scratch (closest t, ties to the lowest primitive id: the result
    does not depend on the hierarchy).  A stale BVH refuses to render."""
    bl = api.BVH_SAH if builder == "sah" else api.BVH_LBVH
    sc = scenes.scene_synthetic(n_tris=n_tris, n_quad_lights=4, width=48, height=32, patches=1 if n_tris < 200 else 8)
    mv, changed = _moved_scene(sc)
    r = renderer_factory(sc, builder=bl, max_leaf_tris=8 if n_tris < 200 else 0)
    r.render(api.RATIO, accum_id=0, mode=1, shadow_spp=2)
    before = r.download("position")
    for k in changed:
        m = mv["meshes"][k]
        r.update_mesh(k, m["vtx"], m["nrm"], m["idx"])
    with pytest.raises(api.OrbError, match="orb_scene_refit"):
        r.render(api.RATIO, accum_id=0, mode=1, shadow_spp=2)
    with pytest.raises(api.OrbError, match="triangle count"):
        m = mv["meshes"][changed[0]]
        r.update_mesh(changed[0], m["vtx"], m["nrm"], m["idx"][:-1])
    with pytest.raises(api.OrbError, match="no such mesh"):
        r.update_mesh(len(mv["meshes"]), m["vtx"], m["nrm"], m["idx"])
    r.refit()
    r.set_lights(mv["tri_lights"])
    c = r.counters()
    assert c["bvh_triangles"] == scenes.n_triangles(sc)
    fresh = renderer_factory(mv, builder=bl, max_leaf_tris=8 if n_tris < 200 else 0)
    rng = np.random.default_rng(23)
    n = 8000
    org = (rng.uniform(-1, 1, (n, 3)) * 9).astype(np.float32)
    org[:, 2] = np.abs(org[:, 2]) * 0.5 + 0.2
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[:40, 1] = 0.0
    want = r.trace_rays(org, d, brute_force=True)
    assert (want[0] >= 0).sum() > n // 4
    for got in (r.trace_rays(org, d), r.trace_rays(org, d, brute_force=2), fresh.trace_rays(org, d)):
        for a, b in zip(got, want):
            assert np.array_equal(a, b)
    frames = []
    for q in (r, fresh):
        z = np.zeros((q.H, q.W, 4), np.float32)
        for k in ("ltc", "sto_direct", "sto_no_vis"):
            q.upload(k, z)
        q.render(api.RATIO, accum_id=3, mode=1, shadow_spp=2)
        frames.append({k: q.download(k) for k in ("position", "normal", "ltc", "sto_direct", "sto_no_vis")})
        frames[-1]["hit"] = q.hit_records()[0]
    for k in frames[0]:
        assert np.array_equal(frames[0][k], frames[1][k], equal_nan=True), k
    assert not np.array_equal(before, frames[0]["position"])     # the geometry did move
    # a second refit without changes is idempotent
    r.refit()
    got = r.trace_rays(org, d)
    for a, b in zip(got, want):
        assert np.array_equal(a, b)


@pytest.mark.parametrize("slots,cap", [(2, 0), (6, 3)])
def test_two_frame_slots_equal_one(renderer_factory, slots, cap):
    """orb_set_frame_slots(n): consecutive frames cycle through n buffer sets and n streams (with an optional cap on the
    persistent kernels' CTAs per SM, orb_set_persistent_ctas_per_sm); accum / RGBA8 are
    shared and ordered.  Frames, the running sum and the ray totals equal those of a one-slot context bit for bit,
    for RATIO (+ denoise + combine on the frame's slot) and PATH, with pipelined read-backs of every frame."""
    import torch
    sc = scenes.scene_synthetic(n_tris=50000, n_quad_lights=5, width=160, height=96)
    a, b = renderer_factory(sc), renderer_factory(sc)
    b.set_frame_slots(slots)
    b.set_persistent_ctas_per_sm(cap)
    assert b.frame_slots == slots and a.frame_slots == 1
    n = 6 if slots == 2 else 14
    pin = torch.empty((n, 96, 160, 4), dtype=torch.float32, pin_memory=True).numpy()
    pin_u = torch.empty((n, 96, 160), dtype=torch.int32, pin_memory=True).numpy().view(np.uint32)
    want = []
    for i in range(n):
        a.render(api.RATIO, accum_id=i, mode=1, shadow_spp=4)
        a.denoise(iterations=3)
        a.ratio_combine(use_denoised=True)
        want.append((a.download("ratio_final"), a.download("rgba8"), a.download("accum"), (a.download("normal")[..., :3] != 0).any(-1)))
        b.render(api.RATIO, accum_id=i, mode=1, shadow_spp=4, sync=False)
        assert b.frame_slot == (i + 1) % slots
        b.denoise(iterations=3)
        b.ratio_combine(use_denoised=True)
        b.download_async("ratio_final", pin[i])
        b.download_async("rgba8", pin_u[i])
    b.sync()
    b.download_wait()
    for i in range(n):
        # compared on the pixels frame i WRITES: a pixel whose primary ray misses keeps what an earlier frame of the
        # same buffer set left in ltc / sto_* (as in the reference, deviceCode.cu:90-94) -- frame i-1 with one slot,
        # frame i-2 with two
        w = want[i][3]
        assert np.array_equal(pin[i][w], want[i][0][w], equal_nan=True), i
        assert np.array_equal(pin_u[i], want[i][1]), i
    assert np.array_equal(b.download("accum"), want[-1][2], equal_nan=True)
    ca, cb = a.counters(), b.counters()
    for k in ("rays_primary", "rays_shadow_total", "active_pixels_total"):
        assert ca[k] == cb[k], k
    for i in range(4):
        a.render(api.PATH, accum_id=i, mode=1, max_depth=5)
        b.render(api.PATH, accum_id=i, mode=1, max_depth=5, sync=False)
    b.sync()
    assert np.array_equal(a.download("accum"), b.download("accum"), equal_nan=True)
    assert a.counters()["rays_bounce_total"] == b.counters()["rays_bounce_total"]
    b.set_frame_slots(1)      # back to one slot: buffers re-allocated, frames still right
    b.render(api.RATIO, accum_id=0, mode=1, shadow_spp=4)
    a.render(api.RATIO, accum_id=0, mode=1, shadow_spp=4)
    w = (a.download("normal")[..., :3] != 0).any(-1)
    assert np.array_equal(a.download("sto_direct")[w], b.download("sto_direct")[w], equal_nan=True)


@pytest.mark.parametrize("slots", [1, 2, 6])
def test_band_handoff_through_peer_stores(renderer_factory, slots):
    """The fused tile hand-off (csrc/halo.cuh, orb_denoise_band): three contexts of ONE process play the ranks of a
    row-band partition -- bands of 50 / 8 / 102 rows, the middle one narrower than the filter's reach so its
    neighbours' halos span two peers -- each packing its band and storing the rows the others' filters read straight
    into their buffers, ordered by epoch flags only.  Denoised bands equal the same rows of a whole-frame denoise bit
    for bit over several frames (both frame slots in use), and no flag wait times out."""
    from optix_renderer_b200 import tiles
    W, H, iters = 96, 160, 4
    sc = scenes.scene_cornell(W, H)
    bands = [(0, 50), (50, 58), (58, 160)]
    ranks = [renderer_factory(sc) for _ in bands]
    for r in ranks:
        r.set_frame_slots(slots)
    ptrs = [[r.halo_local(k) for k in range(slots)] for r in ranks]
    for i, ri in enumerate(ranks):
        for j in range(len(ranks)):
            if i != j:
                for k in range(slots):
                    ri.halo_connect(i, j, k, *ptrs[j][k])
    whole = renderer_factory(sc)
    for frame in range(5):
        whole.render(api.RATIO, accum_id=frame, mode=1, shadow_spp=4)
        whole.denoise(iterations=iters)
        want = {k: whole.download(k) for k in ("sto_direct_dn", "sto_no_vis_dn")}
        wrote = (whole.download("normal")[..., :3] != 0).any(-1)   # miss pixels keep an earlier frame's values (per buffer set)
        for i, r in enumerate(ranks):
            r.render(api.RATIO, accum_id=frame, mode=1, shadow_spp=4, rows=bands[i], sync=False)
            plan = tiles.halo_plan(bands, i, H, iters)
            sends = [(q, a, b) for q, (a, b), kind in plan if kind == "send"]
            recv = sorted({q for q, _, kind in plan if kind == "recv"})
            assert sends and recv
            r.denoise_band(bands[i], sends, recv, iterations=iters)
        for i, r in enumerate(ranks):
            for k in want:
                m = wrote[bands[i][0]:bands[i][1]]
                assert np.array_equal(r.download(k, rows=bands[i])[m], want[k][bands[i][0]:bands[i][1]][m], equal_nan=True), (frame, i, k)
    for r in ranks:
        assert r.halo_status() == 0
    with pytest.raises(api.OrbError, match="not connected"):
        ranks[0].denoise_band(bands[0], [(5, 10, 20)], [])
    for r in ranks:
        r.halo_disconnect()
    with pytest.raises(api.OrbError, match="no peers"):
        ranks[0].denoise_band(bands[0], [], [])


def test_one_sided_frame_assembly(renderer_factory):
    """orb_copy_rows_to_peer: three contexts of ONE process play the ranks of a row-band partition with two frame slots;
    every rank copies its finished band of ratio_final (per slot) and rgba8 (shared by the slots) into the other two
    ranks' buffers.  After a sync every rank holds the whole frame: equal to the frame one context renders whole,
    bit for bit, for several frames in a row."""
    W, H, slots = 96, 120, 2
    sc = scenes.scene_cornell(W, H)
    bands = [(0, 50), (50, 58), (58, 120)]
    ranks = [renderer_factory(sc) for _ in bands]
    for r in ranks:
        r.set_frame_slots(slots)
        r.set_fused_combine(True)
    whole = renderer_factory(sc)
    for frame in range(5):
        whole.render(api.RATIO, accum_id=frame, mode=1, shadow_spp=4)
        whole.denoise(iterations=2)
        whole.ratio_combine(use_denoised=True)
        want = whole.download("ratio_final")
        want_u = whole.download("rgba8")
        wrote = (whole.download("normal")[..., :3] != 0).any(-1)
        for i, r in enumerate(ranks):
            r.render(api.RATIO, accum_id=frame, mode=1, shadow_spp=4, rows=bands[i], sync=False)
        for i, r in enumerate(ranks):   # the packed rows of every band first: a band's filter reads its neighbours' rows
            r.denoise_pack(rows=bands[i])
        packed = [(r.download("dn_guide"), r.download("dn_val")) for r in ranks]
        for i, r in enumerate(ranks):
            g, v = packed[i][0].copy(), packed[i][1].copy()
            for j, (a, b) in enumerate(bands):
                g[a:b], v[a:b] = packed[j][0][a:b], packed[j][1][a:b]
            r.upload("dn_guide", g)
            r.upload("dn_val", v)
            r.denoise_filter(iterations=2, rows=bands[i])
            k = r.frame_slot
            ops = [(n, ranks[j].slot_buffer_ptr(k, n)) for j in range(len(ranks)) if j != i for n in ("ratio_final", "rgba8")]
            for q, (n, ptr) in enumerate(ops):
                r.copy_rows_to_peer(n, ptr, bands[i], last=(q == len(ops) - 1))
        for r in ranks:
            r.sync()
        for i, r in enumerate(ranks):
            assert r.frame_slot == (frame + 1) % slots
            got = r.download("ratio_final")
            assert np.array_equal(got[wrote], want[wrote], equal_nan=True), (frame, i)
            assert np.array_equal(r.download("rgba8"), want_u), (frame, i)
    with pytest.raises(api.OrbError, match="bad row range"):
        ranks[0].copy_rows_to_peer("ratio_final", ranks[1].slot_buffer_ptr(0, "ratio_final"), (10, H + 1))


def test_checked_render_counts_ties_and_near_decisions(renderer_factory, restatement):
    """ORB_RENDER_CHECKED: OrbCounters.hit_ties = primary hits within 1e-5 (barycentric) of a triangle edge -- the
    pixels where another ray/triangle test may report the neighbouring triangle -- and OrbCounters.ltc_near_decision
    = active pixels within 1e-4 of an LTC sign decision: the counts the hit records and the oracle's ltc_margin give
    on the frame's own buffers.  The frame is the unchecked one."""
    W, H = 512, 384
    sc = scenes.scene_synthetic(n_tris=200000, n_quad_lights=6, width=W, height=H)
    r = renderer_factory(sc)
    r.render(api.RATIO, accum_id=0, mode=1, shadow_spp=2)
    want = {k: r.download(k) for k in ("ltc", "sto_direct", "accum")}
    c0 = r.counters()
    assert c0["hit_ties"] == 0 and c0["ltc_near_decision"] == 0
    r.render(api.RATIO, accum_id=0, mode=1, shadow_spp=2, checked=True)
    c = r.counters()
    for k in want:
        assert np.array_equal(want[k], r.download(k), equal_nan=True), k
    prim, uv, t = r.hit_records()
    u, v = uv[..., 0], uv[..., 1]
    edge = (prim >= 0) & (np.minimum(np.minimum(u, v), np.float32(1) - u - v) <= np.float32(1e-5))
    assert c["hit_ties"] == int(edge.sum())
    pos = r.download("position")
    cam = np.asarray(sc["camera"]["pos"], np.float32)
    restatement.reset()
    restatement.set_lights(sc["tri_lights"])
    restatement.set_ltc_lut(*sc["lut"], quantise=0)
    m = restatement.ltc_margin(pos, r.download("normal"), r.download("alpha"), cam, nthreads=8)
    active = (prim >= 0) & (r.download("materialID")[..., 0] != 0)
    lo, hi = int((active & (m <= 0.8e-4)).sum()), int((active & (m <= 1.25e-4)).sum())
    assert hi > 0 and lo <= c["ltc_near_decision"] <= hi, (lo, c["ltc_near_decision"], hi)
    r.render(api.PATH, accum_id=0, mode=1, max_depth=2, checked=True)
    assert r.counters()["hit_ties"] == c["hit_ties"] and r.counters()["ltc_near_decision"] == 0

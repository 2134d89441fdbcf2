"""CPU: the product's own __host__ __device__ headers (optix_renderer_b200/csrc/*.cuh), compiled
with g++ by tests/hostcheck, against the oracle.  This is NOT the product path (that is the
CUDA library, tested under -m gpu); it lets the arithmetic be checked where there is no GPU."""
import numpy as np
import pytest

import _oracle
import _parity as P
from conftest import best_oracle
from optix_renderer_b200 import scenes


@pytest.fixture(scope="module")
def hc(hostcheck):
    hostcheck.load_scene(scenes.scene_floor_blocker(8, 8))
    return hostcheck


@pytest.fixture(scope="module")
def ora(restatement):
    restatement.load_scene(scenes.scene_floor_blocker(8, 8))
    return restatement


def test_units(hc, ora):
    u = P.units()
    P.check_rng(hc, u)
    P.check_clip(hc, ora, u)
    P.check_scalar_units(hc, u)
    P.check_bsdf(hc, u)


def test_rng_skip_equals_sequential_draws(hc):
    for n in (0, 1, 2, 7, 42, 1000):
        s, raw, _ = hc.u_rng(10011, 17, 3, 640, n)
        want = s if n == 0 else int(raw[-1])
        assert hc.u_rng_skip(10011, 17, 3, 640, n) == want


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("scene_fn,W,H", [(scenes.scene_floor_blocker, 48, 40), (scenes.scene_cornell, 40, 40),
                                          (scenes.scene_textured, 48, 40)])
def test_frames(hostcheck, scene_fn, W, H, mode):
    sc = scene_fn(W, H)
    hostcheck.load_scene(sc)
    o = best_oracle(mode)
    o.load_scene(sc)
    depth = 6 if mode else 1
    assert o.set_options(mode, 4, depth) == 0
    for rt in (_oracle.LTC_BASELINE, _oracle.RATIO, _oracle.PATH, _oracle.NORMALS, _oracle.MATERIAL_ID):
        a = o.render(W, H, rt, 0)
        b = hostcheck.render(W, H, rt, 0, mode=mode, spp=4, depth=depth)
        rep = P.frame_report({k: b[k] for k in ("accum", "ltc", "sto_direct", "sto_no_vis", "position", "normal", "albedo", "uv")},
                             a, b["hit_prim"], a["hit_prim"])
        assert rep["hit_mismatch"] == 0
        assert a["rays"] == b["rays"]
        for k in ("albedo", "uv"):  # the texel pick (hit_miss.cuh:39-43) and uv.z = diffuse.z (deviceCode.cu:167): exact
            assert np.array_equal(a[k], b[k]), (rt, k)
        for k in ("accum", "ltc", "position", "normal"):
            assert rep[k]["max_rel"] <= 1e-4, (rt, k, rep[k])
            assert rep[k]["nan_mismatch"] == 0
        for k in ("sto_direct", "sto_no_vis"):
            assert rep[k]["rmse"] <= 1e-3 and rep[k]["nan_mismatch"] == 0, (rt, k, rep[k])
        assert np.array_equal(a["rgba8"] == b["rgba8"], np.ones_like(a["rgba8"], bool)) or \
            (a["rgba8"] != b["rgba8"]).mean() < 0.01  # 8-bit flips from last-ulp differences


def test_lut_64x64(hostcheck):
    """north_star names a 64x64 LTC table; the reference ships 8x8.  Any size goes through the same fetch
    (fetchLtcMat, ltc_utils.cuh:10-23): the product against the reference source with a 64x64 table."""
    sc = scenes.scene_cornell(40, 40)
    sc = dict(sc, lut=scenes.upsample_lut(sc["lut"], 64))
    hostcheck.load_scene(sc)
    o = best_oracle(1)
    o.load_scene(sc)
    assert o.set_options(1, 4, 1) == 0
    a = o.render(40, 40, _oracle.RATIO, 0)
    b = hostcheck.render(40, 40, _oracle.RATIO, 0, mode=1, spp=4, depth=1)
    rep = P.frame_report({k: b[k] for k in ("accum", "ltc", "sto_direct", "sto_no_vis", "position", "normal")},
                         a, b["hit_prim"], a["hit_prim"])
    assert rep["ltc"]["max_rel"] <= 1e-4 and rep["ltc"]["nan_mismatch"] == 0, rep["ltc"]
    assert np.abs(a["ltc"][..., :3]).max() > 0


def test_accumulation(hostcheck):
    sc = scenes.scene_floor_blocker(32, 32)
    hostcheck.load_scene(sc)
    o = best_oracle(1)
    o.load_scene(sc)
    o.set_options(1, 4, 4)
    acc_a = acc_b = None
    for i in range(3):
        a = o.render(32, 32, _oracle.PATH, i, accum=acc_a)
        b = hostcheck.render(32, 32, _oracle.PATH, i, accum=acc_b, mode=1, spp=4, depth=4)
        acc_a, acc_b = a["accum"], b["accum"]
    m = np.isfinite(acc_a) & np.isfinite(acc_b)
    assert np.abs(acc_a[m] - acc_b[m]).max() <= 1e-4 * 3
    assert (np.isnan(acc_a) != np.isnan(acc_b)).sum() == 0


def test_bvh_traversal_equals_brute_force(hostcheck):
    sc = scenes.scene_synthetic(n_tris=3000, n_quad_lights=2, width=16, height=16)
    hostcheck.load_scene(sc)
    rng = np.random.default_rng(3)
    org = rng.uniform(-1, 1, (4000, 3)).astype(np.float32) * np.float32(3)
    org[:, 2] = np.abs(org[:, 2]) + 0.5
    d = rng.normal(size=(4000, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[:40, 0] = 0.0   # axis-parallel components: slab distances must stay finite
    d[40:80, 2] = 0.0
    d[80:100, 1] = -0.0
    p1, t1, u1, v1 = hostcheck.trace(org, d, brute=False)
    p2, t2, u2, v2 = hostcheck.trace(org, d, brute=True)
    assert np.array_equal(p1, p2) and np.array_equal(t1, t2) and np.array_equal(u1, u2) and np.array_equal(v1, v2)
    assert (p1 >= 0).mean() > 0.2


def test_ltc_gbuffer_pass(hostcheck, restatement):
    import os
    g = np.load(os.path.join(P.G, "gbuffer48.npz"))
    hostcheck.load_scene(scenes.scene_floor_blocker(48, 48))
    out = hostcheck.ltc_gbuffer(g["position"], g["normal"], g["albedo"], g["alpha"], g["cam"])
    e = P.rel_err(out, g["ltc_gbuffer"])
    assert e.max() <= 1e-4


@pytest.mark.parametrize("n_tris,leaf", [(5, 4), (60, 1), (3000, 4), (40000, 4)])
def test_quantised_nodes_find_the_same_hits(hostcheck, n_tris, leaf):
    """bvhq.cuh: 32-byte nodes (child boxes on a 16-bit grid, decoded inside the slab FMA) are conservative -- the
    closest hits equal the brute-force ones bit for bit, also for axis-parallel rays and rays that start outside
    the scene box -- and cost only a few per cent more node visits than the fp32 nodes they were encoded from."""
    sc = scenes.scene_synthetic(n_tris=n_tris, n_quad_lights=2, width=16, height=16, patches=1 if n_tris < 200 else 4)
    hostcheck.load_scene(sc, max_leaf=leaf)
    rng = np.random.default_rng(100 + n_tris)
    n = 6000
    org = rng.uniform(-1, 1, (n, 3)).astype(np.float32) * np.float32(9)
    org[:, 2] = np.abs(org[:, 2]) * np.float32(0.5) + np.float32(0.2)
    org[:500] *= np.float32(8)                      # far outside the grid
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[:60, 0] = 0.0
    d[60:120, 2] = 0.0
    d[120:160, 1] = -0.0
    tgt = rng.uniform(-9, 9, (1500, 3)).astype(np.float32)   # rays aimed into the terrain from above and from the side
    tgt[:, 2] = 0.0
    d[1000:2500] = tgt - org[1000:2500]
    d[1000:2500] /= np.linalg.norm(d[1000:2500], axis=1, keepdims=True)
    want = hostcheck.trace(org, d, brute=True)
    got32, nodes32, tris32 = hostcheck.trace_counted(org, d, mode=0)
    gotq, nodesq, trisq = hostcheck.trace_counted(org, d, mode=5)
    for a, b, c in zip(want, got32, gotq):
        assert np.array_equal(a, b) and np.array_equal(a, c)
    # the hybrid stack of the GPU traversal kernels (first entries in a shared-memory column, deeper ones in local
    # memory; here 3 + local) walks both hierarchies to the same hits with the same number of node visits
    goth, nodesh, _ = hostcheck.trace_counted(org, d, mode=6)
    gothq, nodeshq, _ = hostcheck.trace_counted(org, d, mode=7)
    for a, b, c in zip(want, goth, gothq):
        assert np.array_equal(a, b) and np.array_equal(a, c)
    assert nodesh == nodes32 and nodeshq == nodesq
    assert (want[0] >= 0).mean() > 0.15
    if n_tris >= 3000:
        assert nodes32 > 0 and nodesq >= nodes32
        assert nodesq <= 1.06 * nodes32, (nodesq, nodes32)
        assert trisq <= 1.06 * tris32, (trisq, tris32)
        print("quantised nodes: %d tris: node visits x%.4f, triangle tests x%.4f" % (n_tris, nodesq / nodes32, trisq / tris32))


def test_quantised_nodes_large_extent_axis_parallel_rays(hostcheck):
    """bvhq.cuh with a grid step far above 40 (scene extent 1e7): rays with exactly-zero direction components
    must not lose every node to an inf - inf in the grid-space slab constants (trav_begin_q clamps the
    substituted reciprocal).  Hits equal brute force and the fp32-node traversal."""
    import copy
    sc = copy.deepcopy(scenes.scene_synthetic(n_tris=3000, n_quad_lights=2, width=16, height=16, patches=4))
    k = np.float32(5e5)
    for m in sc["meshes"]:
        m["vtx"] = (m["vtx"] * k).astype(np.float32)
    hostcheck.load_scene(sc)
    rng = np.random.default_rng(77)
    n = 3000
    org = (rng.uniform(-1, 1, (n, 3)) * 9).astype(np.float32)
    org[:, 2] = np.abs(org[:, 2]) * np.float32(0.5) + np.float32(1.5)
    org *= k
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d[:, 2] = -np.abs(d[:, 2])
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[:800] = (0, 0, -1)              # straight down: two exactly-zero components
    d[800:1200, 0] = 0.0
    d[1200:1600, 1] = -0.0
    d[800:1600] /= np.linalg.norm(d[800:1600], axis=1, keepdims=True)
    want = hostcheck.trace(org, d, brute=True)
    got32, _, _ = hostcheck.trace_counted(org, d, mode=0)
    gotq, _, _ = hostcheck.trace_counted(org, d, mode=5)
    assert (want[0][:800] >= 0).mean() > 0.9          # the vertical rays do hit the terrain
    for a, b, c in zip(want, got32, gotq):
        assert np.array_equal(a, b) and np.array_equal(a, c)

"""orb_trace_rays_no_alloc (include/orb200.h): the ray-stream kernel with its scratch in the library's static
device array, through the C-ABI, against the existing entry point orb_trace_rays_host(brute_force=2) -- bit for
bit (integer ids and the same fp32 arithmetic), on two calls in a row that share the scratch array.

Collected last on purpose; independent of the other tests: own Renderer, own
scene, device buffers from torch (plumbing: the caller owns inputs and outputs)."""
import numpy as np
import pytest

from optix_renderer_b200 import api, scenes

pytestmark = pytest.mark.gpu


def _rays(seed, n):
    rng = np.random.default_rng(seed)
    org = (rng.uniform(-1, 1, (n, 3)) * 3).astype(np.float32)
    org[:, 2] = np.abs(org[:, 2]) + 0.3
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[:16, 0] = 0.0  # axis-parallel components
    return org, d


def _no_alloc(r, torch, org, d):
    n = org.shape[0]
    dev = torch.device("cuda:0")
    d_o = torch.from_numpy(org).to(dev).contiguous()
    d_d = torch.from_numpy(d).to(dev).contiguous()
    d_p = torch.full((n,), -77, dtype=torch.int32, device=dev)
    d_t, d_u, d_v = (torch.full((n,), -5.0, dtype=torch.float32, device=dev) for _ in range(3))
    torch.cuda.synchronize()
    r.trace_rays_no_alloc(d_o.data_ptr(), d_d.data_ptr(), n, d_p.data_ptr(), d_t.data_ptr(), d_u.data_ptr(), d_v.data_ptr())
    r.sync()
    return d_p.cpu().numpy(), d_t.cpu().numpy(), d_u.cpu().numpy(), d_v.cpu().numpy()


def test_no_alloc_trace_matches_the_host_entry_point_twice_in_a_row():
    torch = pytest.importorskip("torch")
    assert torch.cuda.is_available()
    sc = scenes.scene_synthetic(n_tris=20000, n_quad_lights=2, width=32, height=32)
    r = api.Renderer(sc)
    try:
        # second call: fewer rays than the first, so stale rays / hits / cursor of call 1 would show
        for seed, n in ((3, 50000), (4, 12345), (5, int(api.NO_ALLOC_MAX_RAYS))):
            org, d = _rays(seed, n)
            want = r.trace_rays(org, d, brute_force=2)
            got = _no_alloc(r, torch, org, d)
            assert (want[0] >= 0).any() and (want[0] < 0).any()
            for w, g, name in zip(want, got, ("prim", "t", "u", "v")):
                assert np.array_equal(w, g), name
        # beyond the fixed scratch: refused with a clear error, nothing launched
        org, d = _rays(6, 8)
        with pytest.raises(api.OrbError, match="ORB_NO_ALLOC_MAX_RAYS"):
            r.trace_rays_no_alloc(1, 1, int(api.NO_ALLOC_MAX_RAYS) + 1, 1, 1, 1, 1)
    finally:
        r.close()

"""CPU: the multi-GPU host logic (row bands, denoise halo, frame assembly) with world_size 2 and
3 over gloo.  Bands are produced by the oracle so the check is end to end: tile-partitioned
frame == full frame, bit for bit."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import _oracle
from optix_renderer_b200 import scenes, tiles


def test_row_bands_cover_the_frame():
    for H in (1, 7, 48, 1080, 2160):
        for n in (1, 2, 3, 4, 8):
            b = tiles.row_bands(H, n)
            assert len(b) == n and b[0][0] == 0 and max(y1 for _, y1 in b) == H
            assert all(b[i][1] == b[i + 1][0] or b[i + 1][0] == H for i in range(n - 1))
            assert sum(y1 - y0 for y0, y1 in b) == H
    assert tiles.row_bands(1080, 8)[3] == (405, 540)
    assert tiles.denoise_reach(5) == 62 and tiles.denoise_reach(0) == 0
    assert tiles.halo_bands((405, 540), 1080, 5) == [(343, 405), (540, 602)]
    assert tiles.halo_bands((0, 135), 1080, 5) == [(135, 197)]


def test_denoise_halo_is_sufficient(restatement):
    """Filtering a band from (band + halo) rows only gives the same pixels as filtering the frame."""
    W, H, iters = 40, 96, 3
    sc = scenes.scene_floor_blocker(W, H)
    restatement.load_scene(sc)
    restatement.set_options(1, 4, 1)
    f = restatement.render(W, H, _oracle.RATIO, 0, nthreads=4)
    args = lambda s: [f[k][s] for k in ("normal", "position", "materialID", "sto_direct", "sto_no_vis")]
    full_d, full_n = restatement.denoise(*args(slice(None)), sc["camera"]["pos"], iterations=iters)
    r = tiles.denoise_reach(iters)
    for band in tiles.row_bands(H, 3):
        lo, hi = max(0, band[0] - r), min(H, band[1] + r)
        d, n = restatement.denoise(*args(slice(lo, hi)), sc["camera"]["pos"], iterations=iters)
        assert np.array_equal(d[band[0] - lo:band[1] - lo], full_d[band[0]:band[1]], equal_nan=True)
        assert np.array_equal(n[band[0] - lo:band[1] - lo], full_n[band[0]:band[1]], equal_nan=True)
    # one row less is NOT enough (the reach is tight)
    band = tiles.row_bands(H, 3)[1]
    lo, hi = band[0] - r + 1, band[1] + r - 1
    d, _ = restatement.denoise(*args(slice(lo, hi)), sc["camera"]["pos"], iterations=iters)
    assert not np.array_equal(d[band[0] - lo:band[1] - lo], full_d[band[0]:band[1]], equal_nan=True)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, W, H, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sc = scenes.scene_floor_blocker(W, H)
        o = _oracle.Oracle("restatement")
        o.load_scene(sc)
        o.set_options(1, 4, 1)
        band = tiles.row_bands(H, world)[rank]
        part = o.render(W, H, _oracle.RATIO, 3, rows=band) if band[1] > band[0] else None
        res = {}
        for name in ("ltc", "sto_direct", "accum"):
            full = torch.zeros((H, W, 4), dtype=torch.float32)
            if part is not None:
                full[band[0]:band[1]] = torch.from_numpy(part[name][band[0]:band[1]])
            tiles.assemble(full, band, world)
            res[name] = full.numpy()
        rg = torch.zeros((H, W), dtype=torch.int32)
        if part is not None:
            rg[band[0]:band[1]] = torch.from_numpy(part["rgba8"][band[0]:band[1]].view(np.int32))
        tiles.assemble(rg, band, world)
        res["rgba8"] = rg.numpy()
        np.savez(os.path.join(out_dir, "rank%d.npz" % rank), **res)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,W,H", [(2, 24, 32), (3, 16, 20), (2, 8, 1)])
def test_assembly_over_gloo_equals_full_frame(tmp_path, world, W, H):
    mp.spawn(_worker, args=(world, _free_port(), W, H, str(tmp_path)), nprocs=world, join=True)
    sc = scenes.scene_floor_blocker(W, H)
    o = _oracle.Oracle("restatement")
    o.load_scene(sc)
    o.set_options(1, 4, 1)
    full = o.render(W, H, _oracle.RATIO, 3)
    for rank in range(world):
        got = np.load(os.path.join(str(tmp_path), "rank%d.npz" % rank))
        for name in ("ltc", "sto_direct", "accum"):
            assert np.array_equal(got[name], full[name], equal_nan=True), (rank, name)
        assert np.array_equal(got["rgba8"].view(np.uint32), full["rgba8"])

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


def test_balanced_bands():
    for H, n in ((1080, 8), (1080, 2), (37, 4), (8, 8), (5, 3)):
        rng = np.random.default_rng(H + n)
        cost = rng.uniform(0, 1, H) * (np.arange(H) < H // 2)      # all the work in the lower half
        b = tiles.balanced_bands(cost, n)
        assert len(b) == n and b[0][0] == 0 and b[-1][1] == H
        assert all(b[i][1] == b[i + 1][0] for i in range(n - 1))
        assert all(y1 > y0 for y0, y1 in b)                          # nobody is left without rows
        if H >= 100:
            sums = [cost[y0:y1].sum() for y0, y1 in b]
            assert max(sums) <= cost.sum() / n + cost.max() * 2       # equal to within a row or two
    # uniform cost -> equal bands
    assert tiles.balanced_bands(np.ones(1080), 8) == tiles.row_bands(1080, 8)
    # the cost model: a row of sky costs W, a row full of geometry (1 + ACTIVE_PIXEL_COST) W
    c = tiles.row_costs([0, 1920], 1920)
    assert c[0] == 1920 and c[1] == 1920 * (1 + tiles.ACTIVE_PIXEL_COST)
    assert tiles.balanced_bands([1.0], 1) == [(0, 1)]


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
        bands = tiles.row_bands(H, world) if H < 8 else tiles.balanced_bands(np.arange(H) % 5 + 1.0, world)
        band = bands[rank]
        part = o.render(W, H, _oracle.RATIO, 3, rows=band) if band[1] > band[0] else None
        res = {}
        for name in ("ltc", "sto_direct", "accum"):
            full = torch.zeros((H, W, 4), dtype=torch.float32)
            if part is not None:
                full[band[0]:band[1]] = torch.from_numpy(part[name][band[0]:band[1]])
            tiles.assemble(full, bands, rank)
            res[name] = full.numpy()
        rg = torch.zeros((H, W), dtype=torch.int32)
        if part is not None:
            rg[band[0]:band[1]] = torch.from_numpy(part["rgba8"][band[0]:band[1]].view(np.int32))
        tiles.assemble(rg, bands, rank)
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


def test_halo_plan_is_symmetric():
    for H, n, iters in ((1080, 8, 5), (96, 3, 3), (40, 4, 2), (200, 8, 5)):
        bands = tiles.balanced_bands(np.arange(H) % 7 + 1.0, n)
        plans = [tiles.halo_plan(bands, r, H, iters) for r in range(n)]
        for r in range(n):
            need = tiles.halo_bands(bands[r], H, iters)
            got = sorted(o for _, o, k in plans[r] if k == "recv")
            # the received pieces tile the halo exactly
            assert sum(b - a for a, b in got) == sum(b - a for a, b in need)
            for q, o, k in plans[r]:
                other = "send" if k == "recv" else "recv"
                assert (r, o, other) in plans[q]


def _halo_worker(rank, world, port, H, iters, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        bands = tiles.balanced_bands(np.arange(H) % 5 + 1.0, world)
        y0, y1 = bands[rank]
        a = torch.full((H, 3, 4), -1.0)
        b = torch.full((H, 3, 2), -1.0)
        rows = torch.arange(H, dtype=torch.float32)
        a[y0:y1] = rows[y0:y1, None, None]
        b[y0:y1] = rows[y0:y1, None, None] * 2
        tiles.exchange_halos([a, b], bands, rank, iters)
        np.savez(os.path.join(out_dir, "halo%d.npz" % rank), a=a.numpy(), b=b.numpy(), bands=np.array(bands))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,H,iters", [(2, 40, 2), (3, 64, 3)])
def test_halo_exchange_over_gloo(tmp_path, world, H, iters):
    mp.spawn(_halo_worker, args=(world, _free_port(), H, iters, str(tmp_path)), nprocs=world, join=True)
    reach = tiles.denoise_reach(iters)
    for rank in range(world):
        g = np.load(os.path.join(str(tmp_path), "halo%d.npz" % rank))
        y0, y1 = g["bands"][rank]
        lo, hi = max(0, y0 - reach), min(H, y1 + reach)
        rows = np.arange(H, dtype=np.float32)
        assert np.array_equal(g["a"][lo:hi, 0, 0], rows[lo:hi])       # band + halo present
        assert np.array_equal(g["b"][lo:hi, 2, 1], rows[lo:hi] * 2)
        assert (g["a"][:lo] == -1).all() and (g["a"][hi:] == -1).all()  # nothing else was sent


def _shared_worker(rank, world, port, W, H, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sh = tiles.SharedHostFrames(rank, world, {"ratio_final": ((H, W, 4), np.float32), "rgba8": ((H, W), np.uint32)},
                                    pin=False)
        bands = tiles.balanced_bands(np.arange(H) % 3 + 1.0, world)
        y0, y1 = bands[rank]
        for slot in (0, 1):
            # what TiledRenderer.read_back_bands_async does through the C-ABI: this rank's rows, nobody else's
            sh["ratio_final"][slot][y0:y1] = _pattern(H, W, slot)[y0:y1]
            sh["rgba8"][slot][y0:y1] = (_pattern(H, W, slot)[..., 0] * 7).astype(np.uint32)[y0:y1]
        dist.barrier()
        if rank == 0:
            assert not os.path.exists(sh.path)  # unlinked once every rank had mapped it
            np.savez(os.path.join(out_dir, "shared.npz"), f=np.array(sh["ratio_final"]), g=np.array(sh["rgba8"]))
        dist.barrier()
        sh.close()
    finally:
        dist.destroy_process_group()


def _pattern(H, W, slot):
    return (np.arange(H * W * 4, dtype=np.float32).reshape(H, W, 4) + 1000.0 * slot)


@pytest.mark.parametrize("world,W,H", [(2, 16, 24), (3, 8, 10)])
def test_shared_host_frames_over_gloo(tmp_path, world, W, H):
    """N > 1 end-to-end read-back: every rank writes ITS band into one host frame mapped by all ranks of the box;
    the frame rank 0 reads afterwards is whole, in both slots."""
    mp.spawn(_shared_worker, args=(world, _free_port(), W, H, str(tmp_path)), nprocs=world, join=True)
    got = np.load(os.path.join(str(tmp_path), "shared.npz"))
    for slot in (0, 1):
        assert np.array_equal(got["f"][slot], _pattern(H, W, slot))
        assert np.array_equal(got["g"][slot], (_pattern(H, W, slot)[..., 0] * 7).astype(np.uint32))


class _FakeRenderer:
    """The part of api.Renderer that TiledRenderer's one-sided frame assembly uses, with file-backed arrays as the
    "device" buffers: ipc_export hands out the file's path, ipc_open maps it, copy_rows_to_peer copies rows."""

    def __init__(self, out_dir, rank, W, H, slots):
        self.W, self.H, self.frame_slots, self.frame_slot = W, H, slots, 0
        self.rank, self.dir = rank, out_dir
        self._reg = {}          # "pointer" -> (path, array)
        self._next = 0x1000
        self._bufs = {}
        self.copies = []        # (name, rows, last) in call order
        for k in range(slots):
            self._bufs[(k, "ratio_final")] = self._alloc("r%d_ratio_%d" % (rank, k), (H, W, 4), np.float32)
        shared = self._alloc("r%d_rgba8" % rank, (H, W), np.uint32)   # one buffer for every slot
        for k in range(slots):
            self._bufs[(k, "rgba8")] = shared

    def _alloc(self, tag, shape, dt):
        path = os.path.join(self.dir, tag + ".bin")
        a = np.lib.format.open_memmap(path, mode="w+", dtype=dt, shape=shape)
        ptr = self._next
        self._next += 0x1000
        self._reg[ptr] = (path, a)
        return ptr

    def set_fused_combine(self, on):
        pass

    def slot_buffer_ptr(self, k, name):
        return self._bufs[(k, name)]

    def array(self, k, name):
        return self._reg[self._bufs[(k, name)]][1]

    def ipc_export(self, ptr):
        return self._reg[ptr][0].encode().ljust(64, b"\0")[:64] if len(self._reg[ptr][0]) <= 64 else self._reg[ptr][0].encode()

    def ipc_open(self, handle):
        path = bytes(handle).rstrip(b"\0").decode()
        assert all(path != p for p, _ in self._reg.values()), "a handle was opened twice (or is this rank's own)"
        ptr = self._next
        self._next += 0x1000
        self._reg[ptr] = (path, np.load(path, mmap_mode="r+"))
        return ptr

    def copy_rows_to_peer(self, name, ptr, rows, last=False):
        y0, y1 = rows
        dst = self._reg[ptr][1]
        dst[y0:y1] = self.array(self.frame_slot, name)[y0:y1]
        dst.flush()
        self.copies.append((name, (y0, y1), bool(last)))


def _pat2(H, W, frame):
    return np.arange(H * W, dtype=np.float32).reshape(H, W) + 1000.0 * frame


def _peer_asm_worker(rank, world, port, W, H, slots, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        r = _FakeRenderer(out_dir, rank, W, H, slots)
        T = tiles.TiledRenderer(r, rank, world)
        T.bands = tiles.balanced_bands(np.arange(H) % 7 + 1.0, world)
        names = ["ratio_final", "rgba8"]
        assert T.connect_assembly(names)
        res = {}
        for frame in range(slots + 1):                       # every slot, and the first one a second time
            r.frame_slot = (frame + 1) % slots
            y0, y1 = T.band
            r.array(r.frame_slot, "ratio_final")[y0:y1] = _pat2(H, W, frame)[y0:y1, :, None] + np.arange(4, dtype=np.float32)
            r.array(r.frame_slot, "rgba8")[y0:y1] = _pat2(H, W, frame)[y0:y1].astype(np.uint32) + 7
            n0 = len(r.copies)
            T.assemble_async(names)
            mine = r.copies[n0:]
            assert len(mine) == 2 * (world - 1) and [c[2] for c in mine] == [False] * (len(mine) - 1) + [True]
            assert all(c[1] == (y0, y1) for c in mine)
            dist.barrier()                                   # every rank's copies have landed
            res["ratio_%d" % frame] = np.array(r.array(r.frame_slot, "ratio_final"))
            res["rgba8_%d" % frame] = np.array(r.array(r.frame_slot, "rgba8"))
            dist.barrier()                                   # nobody writes the next frame before everyone has read
        np.savez(os.path.join(out_dir, "rank%d.npz" % rank), **res)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,W,H,slots", [(2, 12, 40, 2), (3, 8, 33, 3)])
def test_one_sided_assembly_protocol_over_gloo(tmp_path, world, W, H, slots):
    """TiledRenderer.connect_assembly / assemble_async (the orb_copy_rows_to_peer path) with a file-backed stand-in for
    the C-ABI: every rank maps every peer's buffers once (the buffer shared by the slots through ONE handle), sends its
    band to every peer with `last` on the final copy only, and after a barrier every rank holds the whole frame in
    the buffers of the frame's slot -- for every slot."""
    mp.spawn(_peer_asm_worker, args=(world, _free_port(), W, H, slots, str(tmp_path)), nprocs=world, join=True)
    for rank in range(world):
        got = np.load(os.path.join(str(tmp_path), "rank%d.npz" % rank))
        for frame in range(slots + 1):
            want = _pat2(H, W, frame)
            assert np.array_equal(got["ratio_%d" % frame], want[:, :, None] + np.arange(4, dtype=np.float32)), (rank, frame)
            assert np.array_equal(got["rgba8_%d" % frame], want.astype(np.uint32) + 7), (rank, frame)

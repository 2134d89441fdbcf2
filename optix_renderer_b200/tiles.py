"""Image-tile partition of a frame across the GPUs of one box.

The reference is single-GPU (owlContextCreate(nullptr, 1), include/viewer.hpp:177).  Here a
frame is split into contiguous row bands, one per rank; scene, BVH, lights and LUT are
replicated; every pixel's RNG is keyed by its GLOBAL index and the GLOBAL frame size
(cuda_include/lcg_random.cuh:54-62), so a band renders bit-identically to the same rows of a
full frame.  Rendering needs no exchange.  The only collective is frame assembly: one
all-gather of each rank's finished band (NCCL over NVLink on the GPU box; gloo in CPU tests).
The denoise filter reaches 2*(2^iters - 1) rows beyond a band; the owner renders those halo
rows redundantly (their rays are not counted) instead of exchanging them.
"""
import numpy as np


def row_bands(height, world_size):
    """Contiguous bands of ceil(H / N) rows; trailing ranks may get fewer (or zero) rows."""
    rows = -(-height // world_size)
    return [(min(r * rows, height), min((r + 1) * rows, height)) for r in range(world_size)]


def denoise_reach(iterations):
    """Rows an a-trous filter of `iterations` passes (5 taps, step 2^i) reaches beyond a band."""
    return 2 * ((1 << iterations) - 1) if iterations > 0 else 0


def halo_bands(band, height, iterations):
    """The (up to two) row ranges outside `band` that its denoise pass reads."""
    y0, y1 = band
    if y1 <= y0:
        return []
    r = denoise_reach(iterations)
    out = []
    if y0 > 0 and r:
        out.append((max(0, y0 - r), y0))
    if y1 < height and r:
        out.append((y1, min(height, y1 + r)))
    return out


class _CudaView:
    """Zero-copy torch view of a library-owned device buffer (__cuda_array_interface__)."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2, "strides": None}


def device_tensor(renderer, name):
    """torch tensor aliasing one of the renderer's frame buffers ((H, W, 4) f32 or (H, W) i32)."""
    import torch

    ptr = renderer.device_ptr(name)
    if name == "rgba8":
        v = _CudaView(ptr, (renderer.H, renderer.W), "<i4")
    else:
        v = _CudaView(ptr, (renderer.H, renderer.W, 4), "<f4")
    return torch.as_tensor(v, device="cuda")


def assemble(full, band, world_size, group=None):
    """All-gather every rank's band of `full` (a (H, ...) tensor, rows [y0, y1) valid on the rank
    that owns them) so that all ranks end with the whole frame.  Bands come from row_bands()."""
    import torch.distributed as dist

    H = full.shape[0]
    bands = row_bands(H, world_size)
    rows = bands[0][1] - bands[0][0]
    y0, y1 = band
    if H == rows * world_size:  # the common case (1080 / 8, 2160 / 8 ...): gather in place
        dist.all_gather_into_tensor(full, full[y0:y1], group=group)
        return full
    import torch

    mine = torch.zeros((rows,) + tuple(full.shape[1:]), dtype=full.dtype, device=full.device)
    mine[: y1 - y0] = full[y0:y1]
    parts = [torch.empty_like(mine) for _ in range(world_size)]
    dist.all_gather(parts, mine, group=group)
    for (a, b), p in zip(bands, parts):
        if b > a:
            full[a:b] = p[: b - a]
    return full


class TiledRenderer:
    """One rank's share of a tile-partitioned frame: renders its band (+ denoise halo) with the
    rank's Renderer and assembles finished buffers across ranks."""

    def __init__(self, renderer, rank, world_size, denoise_iterations=5):
        self.r = renderer
        self.rank, self.world = rank, world_size
        self.iters = denoise_iterations
        self.band = row_bands(renderer.H, world_size)[rank]
        self.halos = halo_bands(self.band, renderer.H, denoise_iterations) if world_size > 1 else []

    def render_ratio_frame(self, accum_id, mode=1, shadow_spp=4, denoise=True):
        """RATIO frame + denoise + combine for this rank's band; asynchronous on the renderer's stream."""
        from . import api

        y0, y1 = self.band
        if y1 <= y0:
            return
        rows = None if self.world == 1 else (y0, y1)
        self.r.render(api.RATIO, accum_id=accum_id, mode=mode, shadow_spp=shadow_spp, rows=rows, sync=False)
        if denoise:
            for h in self.halos:
                self.r.render(api.RATIO, accum_id=accum_id, mode=mode, shadow_spp=shadow_spp, rows=h, sync=False,
                              count_rays=False)
            self.r.denoise(iterations=self.iters, rows=rows)
        self.r.ratio_combine(use_denoised=denoise, rows=rows)

    def render_path_frame(self, accum_id, mode=1, max_depth=8):
        from . import api

        y0, y1 = self.band
        if y1 <= y0:
            return
        rows = None if self.world == 1 else (y0, y1)
        self.r.render(api.PATH, accum_id=accum_id, mode=mode, max_depth=max_depth, rows=rows, sync=False)

    def assemble(self, names, group=None):
        if self.world == 1:
            return
        for n in names:
            assemble(device_tensor(self.r, n), self.band, self.world, group)

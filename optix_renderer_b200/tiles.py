"""Image-tile partition of a frame across the GPUs of one box.

The reference is single-GPU (owlContextCreate(nullptr, 1), include/viewer.hpp:177).  Here a
frame is split into contiguous row bands, one per rank; scene, BVH, lights and LUT are
replicated; every pixel's RNG is keyed by its GLOBAL index and the GLOBAL frame size
(cuda_include/lcg_random.cuh:54-62), so a band renders bit-identically to the same rows of a
full frame.  Rendering needs no exchange.

Bands are BALANCED BY COST, not by row count: sky rows cost one cheap primary ray per pixel, rows
full of geometry cost the LTC sum and four shadow rays per pixel on top (about 10x).  After a first
frame every rank counts its active pixels per row, the counts are all-gathered, and all ranks cut
the frame where the cumulative cost crosses k/N (same input, same cuts everywhere).

Exchanges (NVLink on the GPU box; gloo in the CPU tests):
  * denoise input: each rank packs ITS rows (guide float4 + value float2, 24 B/px) and the
    2*(2^iters - 1) rows either side of a cut reach the neighbours that filter them, so nobody renders a
    halo twice.  After connect_halo() the pack kernel itself stores those rows into the neighbours'
    buffers through peer-mapped pointers (csrc/halo.cuh: cudaIpc handles exchanged once, epoch flags, no
    NCCL call on the frame's critical path); without it they go point to point through torch.distributed;
  * frame assembly: one all-gather of each rank's finished band per output buffer.
Bands are uneven, so the gathers are lists of per-rank row views of the destination buffer.
"""
import os

import numpy as np

ACTIVE_PIXEL_COST = 10.0  # cost of an active pixel relative to a primary ray (bench scene: 2.0 ns vs 0.19 ns)


def row_bands(height, world_size):
    """Contiguous bands of ceil(H / N) rows; trailing ranks may get fewer (or zero) rows."""
    rows = -(-height // world_size)
    return [(min(r * rows, height), min((r + 1) * rows, height)) for r in range(world_size)]


def balanced_bands(row_cost, world_size):
    """Contiguous bands of (nearly) equal summed cost.  row_cost: per-row non-negative costs.
    Deterministic in its input; every band of a frame with H >= N rows gets at least one row."""
    c = np.asarray(row_cost, np.float64)
    H = len(c)
    if world_size <= 1:
        return [(0, H)]
    c = np.maximum(c, 0.0) + 1e-9
    cum = np.concatenate([[0.0], np.cumsum(c)])
    cuts = [0]
    for k in range(1, world_size):
        y = int(np.searchsorted(cum, cum[-1] * k / world_size, side="left"))
        lo = min(cuts[-1] + 1, H)                # at least one row for the previous band ...
        hi = max(H - (world_size - k), lo)      # ... and one for each band still to come
        cuts.append(min(max(y, lo), hi))
    cuts.append(H)
    return [(cuts[i], cuts[i + 1]) for i in range(world_size)]


def row_costs(active_per_row, width):
    """Cost model of a row: one primary ray per pixel + ACTIVE_PIXEL_COST per active pixel."""
    return float(width) + ACTIVE_PIXEL_COST * np.asarray(active_per_row, np.float64)


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
    """torch tensor aliasing one of the renderer's frame buffers: (H, W, 4) f32, (H, W, 2) f32 for
    the packed denoise values, (H, W) i32 for RGBA8."""
    import torch

    ptr = renderer.device_ptr(name)
    if name == "rgba8":
        v = _CudaView(ptr, (renderer.H, renderer.W), "<i4")
    else:
        v = _CudaView(ptr, (renderer.H, renderer.W, 2 if name == "dn_val" else 4), "<f4")
    return torch.as_tensor(v, device="cuda")


def _contiguous_equal(bands, height):
    return (len({b[1] - b[0] for b in bands}) == 1 and bands[0][0] == 0 and bands[-1][1] == height
            and all(bands[i][1] == bands[i + 1][0] for i in range(len(bands) - 1)))


def assemble(full, bands, rank, group=None):
    """All-gather every rank's band of `full` (a (H, ...) tensor whose rows bands[rank] are valid on
    this rank) so that all ranks end with the whole frame.  `bands`: the same list on every rank."""
    import torch
    import torch.distributed as dist

    world = len(bands)
    if world == 1:
        return full
    y0, y1 = bands[rank]
    if _contiguous_equal(bands, full.shape[0]):
        dist.all_gather_into_tensor(full, full[y0:y1], group=group)  # equal bands: gather in place
        return full
    if dist.get_backend(group) == "nccl" and all(b[1] > b[0] for b in bands):
        # uneven bands: the output list are row views of the destination itself (NCCL runs this as one
        # group of broadcasts, no staging copy)
        dist.all_gather([full[a:b] for a, b in bands], full[y0:y1], group=group)
        return full
    rows = max(max(b[1] - b[0] for b in bands), 1)  # gloo, or empty bands: pad to a common size
    mine = torch.zeros((rows,) + tuple(full.shape[1:]), dtype=full.dtype, device=full.device)
    mine[: y1 - y0] = full[y0:y1]
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine, group=group)
    for (a, b), p in zip(bands, parts):
        if b > a:
            full[a:b] = p[: b - a]
    return full


def _overlap(a, b):
    lo, hi = max(a[0], b[0]), min(a[1], b[1])
    return (lo, hi) if hi > lo else None


def halo_plan(bands, rank, height, iterations):
    """Row ranges to exchange so that `rank` holds the rows its denoise pass reads: a list of
    (peer, (y0, y1), 'send' | 'recv'), in an order both sides agree on."""
    plan = []
    mine = bands[rank]
    my_halos = halo_bands(mine, height, iterations)
    for q, bq in enumerate(bands):
        if q == rank or bq[1] <= bq[0]:
            continue
        for h in my_halos:                       # rows of q's band that I read
            o = _overlap(bq, h)
            if o:
                plan.append((q, o, "recv"))
        if mine[1] > mine[0]:
            for h in halo_bands(bq, height, iterations):  # rows of my band that q reads
                o = _overlap(mine, h)
                if o:
                    plan.append((q, o, "send"))
    return plan


def exchange_halos(tensors, bands, rank, iterations, group=None):
    """Point-to-point exchange of the denoise halo rows of every tensor in `tensors` ((H, ...) each):
    at most 2*(2^iters - 1) rows per side instead of the whole frame."""
    import torch.distributed as dist

    H = tensors[0].shape[0]
    plan = halo_plan(bands, rank, H, iterations)
    ops = []
    for peer, (a, b), kind in plan:
        for t in tensors:
            ops.append(dist.P2POp(dist.isend if kind == "send" else dist.irecv, t[a:b], peer, group))
    if not ops:
        return
    for w in dist.batch_isend_irecv(ops):
        w.wait()


class SharedHostFrames:
    """Whole-frame host buffers shared by the ranks of one box: a /dev/shm file that every rank maps and
    page-locks (cudaHostRegister).  Every rank then copies ITS band of a finished frame straight into the
    frame over its own PCIe link (TiledRenderer.read_back_bands_async), instead of rank 0 reading the whole
    assembled frame back alone: the host frame is complete after wait_read_back_bands().  The device-side
    all-gather stays what puts the whole frame on every GPU; this is the host's copy of it.

    spec: name -> (per-frame shape, numpy dtype); `slots` frames per name (a consumer reads one slot while
    the next frame lands in the other).  pin=False maps without CUDA (CPU tests)."""

    def __init__(self, rank, world_size, spec, slots=2, group=None, pin=True):
        import torch.distributed as dist

        self.rank, self.world, self.group = rank, world_size, group
        self._pinned = []
        sizes = {n: int(slots * np.prod(shape) * np.dtype(dt).itemsize) for n, (shape, dt) in spec.items()}
        total = sum((b + 4095) // 4096 * 4096 for b in sizes.values())
        box = [None]
        if rank == 0:
            vfs = os.statvfs("/dev/shm")
            if vfs.f_bavail * vfs.f_frsize >= total + (64 << 20):
                # private to this user, never an existing file, and the pages RESERVED now: a sparse file on a short
                # tmpfs would end in SIGBUS on first touch (or a partial cudaHostRegister) long after this check
                name = "/dev/shm/orb200_frames_%d_%s" % (os.getpid(), os.urandom(8).hex())
                try:
                    fd = os.open(name, os.O_RDWR | os.O_CREAT | os.O_EXCL, 0o600)
                    try:
                        os.posix_fallocate(fd, 0, max(total, 4096))
                        box[0] = name
                    except OSError:  # ENOSPC: the same None answer on every rank below
                        os.unlink(name)
                    finally:
                        os.close(fd)
                except OSError:
                    pass
        if world_size > 1:
            dist.broadcast_object_list(box, src=0, group=group)
        self.path = box[0]
        if self.path is None:  # the same answer on every rank
            raise RuntimeError("SharedHostFrames: /dev/shm has less than %d MB free" % ((total >> 20) + 64))
        self._map = np.memmap(self.path, dtype=np.uint8, mode="r+", shape=(max(total, 4096),))
        if world_size > 1:
            dist.barrier(group=group)  # every rank holds its mapping: the name can go
        if rank == 0:
            os.unlink(self.path)
        self.frames = {}
        ofs = 0
        for n, (shape, dt) in spec.items():
            b = sizes[n]
            self.frames[n] = self._map[ofs:ofs + b].view(dt).reshape((slots,) + tuple(shape))
            ofs += (b + 4095) // 4096 * 4096
        if pin:
            import torch

            rt = torch.cuda.cudart()
            err = int(rt.cudaHostRegister(self._map.ctypes.data, self._map.nbytes, 0))
            if err == 0:
                self._pinned.append(self._map.ctypes.data)
            oks = [err == 0]
            if world_size > 1:  # every rank pins or none does: the callers' fallback must be the same everywhere
                oks = [None] * world_size
                dist.all_gather_object(oks, err == 0, group=group)
            if not all(oks):
                self.close()
                raise RuntimeError("SharedHostFrames: cudaHostRegister failed on ranks %s" %
                                   [i for i, ok in enumerate(oks) if not ok])

    def __getitem__(self, name):
        return self.frames[name]

    def close(self):
        """Unpins and unmaps.  Views handed out by [] must not be used afterwards."""
        if self._pinned:
            import torch

            for p in self._pinned:
                torch.cuda.cudart().cudaHostUnregister(p)
            self._pinned = []
        self.frames = {}
        m, self._map = getattr(self, "_map", None), None
        if m is not None and getattr(m, "_mmap", None) is not None:
            try:
                m._mmap.close()
            except (BufferError, ValueError):  # a caller still holds a view: the mapping goes with its last reference
                pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class TiledRenderer:
    """One rank's share of a tile-partitioned frame: renders its band with the rank's Renderer,
    exchanges packed denoise rows, and assembles finished buffers across ranks."""

    def __init__(self, renderer, rank, world_size, denoise_iterations=5, group=None):
        self.r = renderer
        self.rank, self.world = rank, world_size
        self.iters = denoise_iterations
        self.group = group
        self.bands = row_bands(renderer.H, world_size)
        self.row_cost = None
        self.floor_ms = 0.0  # fixed per-frame kernel time of a band of any height (see rebalance)
        self._views = {}
        self._streams = {}
        self._ipc = []        # peer mappings opened by connect_halo()
        self.halo = False     # True: the denoise halo is handed over by peer stores inside the pack kernel
        # the filter's final pass writes ratio_final too (orb_set_fused_combine); ORB_FUSED_COMBINE=0: separate kernel
        self.fused_combine = os.environ.get("ORB_FUSED_COMBINE", "1") != "0"
        renderer.set_fused_combine(self.fused_combine)
        self._plan_key, self._plan = None, None
        self.spans = None  # set to {} to collect CUDA-event spans of the collectives: name -> [(e0, e1), ...]
        self._stage_key, self._stage_in, self._stage_all, self._stage_meta, self._done_bands = None, None, None, {}, None
        self._comm = None  # side stream the frame assembly runs on, beside the next frame's kernels
        self._ring, self._ring_i, self._cur_ent = [], -1, None
        self._peer_asm, self._peer_names = None, ()  # connect_assembly(): (peer, slot, name) -> mapped pointer
        self._done = None
        self._copy = None  # stream the assembled frame is read back to the host on (read_back_async)
        self._read = None

    @property
    def band(self):
        return self.bands[self.rank]

    def _t(self, name):
        """torch view of the ACTIVE frame slot's buffer (the slot of the last render())."""
        key = (name, self.r.frame_slot)
        if key not in self._views:
            self._views[key] = device_tensor(self.r, name)
        return self._views[key]

    def slot_stream(self):
        """torch handle of the stream the active frame slot's kernels run on (the library's own stream)."""
        import torch

        return self.stream_of(self.r.frame_slot)

    def stream_of(self, k):
        import torch

        if k not in self._streams:
            self._streams[k] = torch.cuda.ExternalStream(self.r.slot_stream(k))
        return self._streams[k]

    def join(self, stream):
        """Orders `stream` behind everything enqueued so far on every frame slot and behind the last assembly."""
        import torch

        for k in range(self.r.frame_slots):
            e = torch.cuda.Event()
            e.record(self.stream_of(k))
            stream.wait_event(e)
        if self._done is not None:
            stream.wait_event(self._done)

    def connect_halo(self):
        """Map every other rank's packed denoise buffers and control block into this process (cudaIpc handles
        through one all_gather_object) so that render_ratio_frame() hands its halo rows over with peer stores
        from inside the pack kernel.  All ranks must call it, after set_frame_slots / resize."""
        if self.world == 1:
            return False
        import torch.distributed as dist

        mine = []
        for k in range(self.r.frame_slots):
            g, v, f = self.r.halo_local(k)
            mine.append((self.r.ipc_export(g), self.r.ipc_export(v)))
        mine.append(self.r.ipc_export(self.r.halo_local(0)[2]))
        everyone = [None] * self.world
        dist.all_gather_object(everyone, mine, group=self.group)
        for q, theirs in enumerate(everyone):
            if q == self.rank:
                continue
            flags = self.r.ipc_open(theirs[-1])
            self._ipc.append(flags)
            for k, (hg, hv) in enumerate(theirs[:-1]):
                g, v = self.r.ipc_open(hg), self.r.ipc_open(hv)
                self._ipc += [g, v]
                self.r.halo_connect(self.rank, q, k, g, v, flags)
        dist.barrier(group=self.group)
        self.halo = True
        return True

    def connect_assembly(self, names):
        """One-sided frame assembly (orb_copy_rows_to_peer): map every other rank's buffers `names` of every frame slot
        into this process, so that assemble_async() copies this rank's finished band straight into the peers' frames
        with the copy engines -- no NCCL kernel, no rendezvous.  All ranks must call it, after set_frame_slots /
        resize.  Returns False (and assemble_async keeps the NCCL all-gather) if any rank failed to map a peer."""
        if self.world == 1:
            return False
        import torch.distributed as dist

        slots = self.r.frame_slots
        mine, seen = {}, {}
        for k in range(slots):
            for n in names:
                ptr = self.r.slot_buffer_ptr(k, n)
                if ptr not in seen:  # accum / rgba8 are one buffer for every slot: one handle, opened once
                    seen[ptr] = self.r.ipc_export(ptr)
                mine[(k, n)] = seen[ptr]
        everyone = [None] * self.world
        dist.all_gather_object(everyone, mine, group=self.group)
        table, ok = {}, True
        try:
            for q, theirs in enumerate(everyone):
                if q == self.rank:
                    continue
                opened = {}
                for (k, n), h in sorted(theirs.items()):
                    if h not in opened:
                        opened[h] = self.r.ipc_open(h)
                        self._ipc.append(opened[h])
                    table[(q, k, n)] = opened[h]
        except Exception:  # noqa: BLE001 -- any failure: every rank falls back together
            ok = False
        oks = [None] * self.world
        dist.all_gather_object(oks, ok, group=self.group)
        if not all(oks):
            return False
        self._peer_asm, self._peer_names = table, tuple(names)
        return True

    def disconnect_halo(self):
        if not self.halo:
            return
        import torch.distributed as dist

        self.r.sync()
        dist.barrier(group=self.group)  # nobody may still be pushing into a mapping that is about to go
        self.r.halo_disconnect()
        for p in self._ipc:
            self.r.ipc_close(p)
        self._ipc = []
        self._peer_asm, self._peer_names = None, ()  # its mappings were in the same list
        self.halo = False

    def _band_plan(self):
        """(sends, recv_ranks) of this rank's band for orb_denoise_band, from the plan both sides agree on."""
        key = tuple(self.bands)
        if self._plan_key != key:
            plan = halo_plan(self.bands, self.rank, self.r.H, self.iters)
            self._plan = ([(q, a, b) for q, (a, b), kind in plan if kind == "send"],
                          sorted({q for q, _, kind in plan if kind == "recv"}))
            self._plan_key = key
        return self._plan

    def _span(self, name):
        """Context manager: CUDA-event pair around a collective when self.spans is a dict."""
        import contextlib

        import torch

        if self.spans is None:
            return contextlib.nullcontext()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.spans.setdefault(name, []).append((e0, e1))

        @contextlib.contextmanager
        def cm():
            e0.record()
            yield
            e1.record()
        return cm()

    def span_ms(self):
        """Summed milliseconds per collective since spans were switched on (synchronises)."""
        import torch

        torch.cuda.synchronize()
        return {k: sum(a.elapsed_time(b) for a, b in v) for k, v in (self.spans or {}).items()}

    def rebalance(self, measured_ms=None):
        """Cut the bands by cost (all ranks must call).  First call: the model of row_costs() from the
        active-pixel counts of the last frame.  With measured_ms (this rank's device time for its band
        in the last frames) the model is corrected by feedback: every row of a band is scaled by
        measured time / modelled cost of that band, then the frame is cut again."""
        if self.world == 1:
            return self.bands
        import torch
        import torch.distributed as dist

        nccl = dist.get_backend(self.group) == "nccl"
        self.r.sync()  # every frame in flight has left the shared accum and the peers' halo rows
        if nccl:
            with torch.cuda.stream(self.slot_stream()):
                assemble(self._t("accum"), self.bands, self.rank, self.group)  # progressive sums follow their rows
            self.slot_stream().synchronize()
        else:
            assemble(self._t("accum"), self.bands, self.rank, self.group)
        if self.row_cost is None or measured_ms is None:
            counts = self.r.row_active_counts()
            y0, y1 = self.band
            mine = np.zeros(self.r.H, np.int32)
            mine[y0:y1] = counts[y0:y1]  # only the rows this rank rendered are current
            t = torch.from_numpy(mine)
            t = t.cuda() if nccl else t
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
            self.row_cost = row_costs(t.cpu().numpy(), self.r.W)
        else:
            t = torch.zeros(self.world, dtype=torch.float64)
            t[self.rank] = float(measured_ms)
            t = t.cuda() if nccl else t
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
            ms = t.cpu().numpy()
            cost = self.row_cost.copy()
            # a band costs floor + sum of its rows: the kernels' fixed launch tails (FLOOR_MS, profiles/r1k_stage_floors)
            # do not shrink with the band, so only the part above the floor is spread over the rows
            for (y0, y1), m in zip(self.bands, ms):
                if y1 > y0 and m > 0:
                    cost[y0:y1] *= max(m - self.floor_ms, 0.1 * m) / max(cost[y0:y1].sum(), 1e-30)
            self.row_cost = cost
        self.bands = balanced_bands(self.row_cost, self.world)
        return self.bands

    def render_ratio_frame(self, accum_id, mode=1, shadow_spp=4, denoise=True):
        """RATIO frame + denoise + combine for this rank's band; asynchronous on the stream of the frame slot the
        render takes (torch-side exchanges are issued on that same stream, slot_stream())."""
        from . import api

        y0, y1 = self.band
        rows = None if self.world == 1 else (y0, y1)
        if y1 > y0:
            self.r.render(api.RATIO, accum_id=accum_id, mode=mode, shadow_spp=shadow_spp, rows=rows, sync=False)
        if denoise:
            if self.world == 1:
                self.r.denoise(iterations=self.iters)
            elif self.halo:
                sends, recv = self._band_plan()
                self.r.denoise_band((y0, max(y0, y1)), sends, recv, iterations=self.iters)  # every rank, every frame
            else:
                import torch

                with torch.cuda.stream(self.slot_stream()):
                    if y1 > y0:
                        self.r.denoise_pack(rows=rows)
                    with self._span("halo_exchange"):
                        exchange_halos([self._t("dn_guide"), self._t("dn_val")], self.bands, self.rank, self.iters, self.group)
                    if y1 > y0:
                        self.r.denoise_filter(iterations=self.iters, rows=rows)
        if y1 > y0 and not (denoise and self.fused_combine):
            self.r.ratio_combine(use_denoised=denoise, rows=rows)

    def render_path_frame(self, accum_id, mode=1, max_depth=8):
        from . import api

        y0, y1 = self.band
        if y1 <= y0:
            return
        rows = None if self.world == 1 else (y0, y1)
        self.r.render(api.PATH, accum_id=accum_id, mode=mode, max_depth=max_depth, rows=rows, sync=False)

    def assemble_async(self, names):
        """Frame assembly off the critical path.  This rank's finished band of every buffer in `names` is copied into
        ONE staging block (rows of [name0 row | name1 row | ...] bytes, padded to the tallest band) on the frame
        slot's stream, and a single all_gather_into_tensor of that block runs on a side stream while the slot goes on
        with its next frame: one NCCL call per frame whatever the number of ranks and buffers (the list-of-views
        all_gather of uneven bands it replaces issued N broadcasts per buffer, ~0.3 ms of host time per frame at
        N = 8).  Every rank then holds every band; assembled(name) lays them out as the whole frame.
        wait_assembled() / join() order a stream behind the last gather."""
        if self.world == 1:
            return
        import torch
        import torch.distributed as dist

        if self._peer_asm is not None and tuple(names) == self._peer_names:
            # one-sided: this rank's band into every peer's frame of the same slot, by the copy engines, behind the
            # slot's kernels.  join() / orb_sync cover the copies (the slot's stream waits for them); the frame is
            # whole on every rank once all ranks have synchronised (a barrier).
            y0, y1 = self.band
            k = self.r.frame_slot
            if y1 > y0:
                ops = [(n, self._peer_asm[(q, k, n)]) for q in range(self.world) if q != self.rank for n in names]
                for i, (n, ptr) in enumerate(ops):
                    self.r.copy_rows_to_peer(n, ptr, (y0, y1), last=(i == len(ops) - 1))
            self._done = None
            self._done_bands = list(self.bands)
            return
        cur = self.slot_stream()
        if self._comm is None:
            # high priority: the gather's CTAs go ahead of the queued CTAs of the frames in flight, whose persistent
            # kernels otherwise keep it waiting for SMs (measured at N = 2: 1.78 ms per frame with the gather starved
            # on ONE staging block, 1.36 ms without any assembly)
            self._comm = torch.cuda.Stream(priority=-1)
        y0, y1 = self.band
        key = (tuple(names), tuple(self.bands), self.r.frame_slots)
        if self._stage_key != key:
            torch.cuda.synchronize()
            views = [self._t(n) for n in names]
            row_bytes = [v[0].numel() * v.element_size() for v in views]
            offs = [sum(row_bytes[:i]) for i in range(len(names))]
            rows = max(max(b - a for a, b in self.bands), 1)
            # a ring of staging blocks, one per frame in flight: frame i+1 copies its band in while frame i's gather
            # still reads ITS block (one block made every frame wait for the previous frame's gather)
            self._ring = []
            for _ in range(max(2, self.r.frame_slots)):
                blk_in = torch.empty((rows, sum(row_bytes)), dtype=torch.uint8, device=views[0].device)
                blk_all = torch.empty((self.world,) + tuple(blk_in.shape), dtype=torch.uint8, device=views[0].device)
                self._ring.append({"in": blk_in, "all": blk_all, "done": None, "read": None})
            self._ring_i = -1
            self._stage_meta = {n: (o, rb, v.dtype, tuple(v.shape[1:])) for n, o, rb, v in zip(names, offs, row_bytes, views)}
            self._stage_key = key
        self._ring_i = (self._ring_i + 1) % len(self._ring)
        ent = self._ring[self._ring_i]
        if ent["done"] is not None:
            cur.wait_event(ent["done"])  # the gather that last used this block (a ring's length of frames ago)
        if ent["read"] is not None:
            cur.wait_event(ent["read"])  # ... and a read-back of that gathered block
            ent["read"] = None
        self._stage_in, self._stage_all, self._cur_ent = ent["in"], ent["all"], ent
        with torch.cuda.stream(cur):
            if y1 > y0:
                for n in names:
                    o, rb, dt, shp = self._stage_meta[n]
                    dst = self._stage_in[: y1 - y0, o:o + rb].view(dt).view((y1 - y0,) + shp)
                    dst.copy_(self._t(n)[y0:y1])
        ready = torch.cuda.Event()
        ready.record(cur)
        with torch.cuda.stream(self._comm):
            self._comm.wait_event(ready)
            dist.all_gather_into_tensor(self._stage_all.view(-1), self._stage_in.view(-1), group=self.group)
            self._done = torch.cuda.Event()
            self._done.record(self._comm)
            ent["done"] = self._done
            self._done_bands = list(self.bands)

    def assembled(self, name, stream=None):
        """The whole frame of buffer `name` as gathered by the last assemble_async: a new tensor, laid out from the
        N bands of the staging block on `stream` (default: the active frame slot's), behind the gather."""
        import torch

        st = stream or self.slot_stream()
        if self._peer_asm is not None and name in self._peer_names:
            # the peers stored their bands into this rank's own buffer of the frame's slot (valid once every rank has
            # synchronised behind its copies)
            with torch.cuda.stream(st):
                return self._t(name).clone()
        o, rb, dt, shp = self._stage_meta[name]
        with torch.cuda.stream(st):
            if self._done is not None:
                st.wait_event(self._done)
            out = torch.empty((self.r.H,) + shp, dtype=dt, device=self._stage_all.device)
            for q, (a, b) in enumerate(self._done_bands):
                if b > a:
                    out[a:b].copy_(self._stage_all[q, : b - a, o:o + rb].view(dt).view((b - a,) + shp))
        return out

    def read_back_async(self, host):
        """After assemble_async: copy the assembled frame to page-locked host tensors `host[name]` on a copy stream,
        behind the gather and beside the next frame's kernels.  The next assemble_async waits for it on the device
        before it touches the staging block; wait_read_back() blocks the host until the frame has landed."""
        import torch

        if self._copy is None:
            self._copy = torch.cuda.Stream()
        with torch.cuda.stream(self._copy):
            for n, t in host.items():
                t.copy_(self.assembled(n, self._copy), non_blocking=True)
            self._read = torch.cuda.Event()
            self._read.record(self._copy)
            if getattr(self, "_cur_ent", None) is not None:
                self._cur_ent["read"] = self._read

    def read_back_bands_async(self, shared, slot, names):
        """This rank's rows of the finished buffers `names` into the box-wide host frames `shared`
        (SharedHostFrames), through the C-ABI's pipelined read-back (orb_buffer_download_rows_async: a device
        snapshot on the render stream, the PCIe copy on the context's copy stream beside the next frame)."""
        y0, y1 = self.band
        if y1 <= y0:
            return
        for n in names:
            self.r.download_async(n, shared[n][slot][y0:y1], rows=(y0, y1))

    def wait_read_back_bands(self):
        """Blocks until every rank's band has landed: the shared host frames are whole on return."""
        self.r.download_wait()
        if self.world > 1:
            import torch.distributed as dist

            dist.barrier(group=self.group)

    def wait_read_back(self):
        if self._read is not None:
            self._read.synchronize()

    def wait_assembled(self, stream=None):
        """Orders `stream` (default: the active frame slot's) behind the last gather."""
        if self._done is not None:
            (stream or self.slot_stream()).wait_event(self._done)

    def assemble(self, names):
        if self.world == 1:
            return
        import torch

        with torch.cuda.stream(self.slot_stream()):
            with self._span("assemble"):
                for n in names:
                    assemble(self._t(n), self.bands, self.rank, self.group)

"""bench.py -- BASELINE.json's metric on the hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload ratio|path]

A step = one pass of the hot path over one frame of the synthetic scene:
  ratio (default; BASELINE configs[2], the config the "1080p frames/s" metric is quoted on):
      1920x1080, ~1M triangles, 16 quad lights (32 TriLights), RATIO renderer in intended mode
      -- primary rays, LTC sum over all lights, 4 shadow rays per pixel -- + 5-pass a-trous
      denoise of the two stochastic images + ratio combine.  accumId advances every step.
  path  (configs[3]): 8-bounce PATH frame of the same scene.
Metric: Mrays/s = device-counted rays (primary + shadow + bounce) / second, whole job.
N > 1 (torchrun, one rank per GPU): the frame is split into row bands (strong scaling), the
BVH is replicated, and the finished band of the output (float4 ratio image + RGBA8) is
assembled with one NCCL all-gather per buffer inside the timed region.

--impl reference times the reference's own CPU code for the same frame (oracle/_ref = the
reference's device source compiled as host C++; falls back to the oracle restatement) on all host
cores, each step a bounded sample of rows of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H = 1920, 1080
N_TRIS, N_QUADS = 1_000_000, 16
SPP, DEPTH, DN_ITERS = 4, 8, 5


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks and throttle reasons DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                o = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                    "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                if o.returncode == 0:
                    self.rows.append([x.strip() for x in o.stdout.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        self.stop_flag = True
        self.join(timeout=6)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "samples": len(sm),
                "power_w_max": max(float(r[2]) for r in self.rows), "reasons": reasons}


# ALGORITHMIC bytes per stage, float4 AoS as at the boundary.
def stage_bytes(stage, px, active, accum_id_gt0=True, trav=None):
    if stage == "path":           # all bounces in one kernel: the traversal fetches (counted) + per path 64 B of G-buffer
        if not trav:              # read and 16 B of colour written; no ray stream, no path state in HBM
            return 0
        return trav["nodes"] * 64 + trav["tris"] * 48 + active * 80
    if stage == "trace":          # nodes visited x 64 B + triangles tested x 48 B (counted by the
        if not trav:              # instrumented twin of the kernel) + the ray stream in (32 B) and hit records out (16 B)
            return 0
        return trav["nodes"] * 64 + trav["tris"] * 48 + trav["rays"] * 48
    if stage == "primary":        # 6 G-buffers + hit record written, flags
        return px * (6 * 16 + 16 + 4) + active * 4
    if stage in ("ltc", "ltc_gbuffer"):  # position, normal, albedo, alpha read; ltc written (+ queue index)
        return active * (80 + (4 if stage == "ltc" else 0))
    if stage == "shadow":         # gen: queue, position, normal read, SPP rays (32 B) written; resolve: SPP hit records
        return active * (4 + 32 + 32 * SPP) + active * (4 + 16 * SPP + 64 + 32)  # + 4 G-buffers read, 2 sto written
    if stage == "accumulate":     # colour + flags read, accum read/write, RGBA8 written
        return px * (16 + 4 + 16 + (16 if accum_id_gt0 else 0) + 4)
    if stage == "denoise_pack":   # normal, position, materialID, 2 sto read; guide + packed value written
        return px * (5 * 16 + 16 + 8)
    if stage == "denoise":        # per pass: guide + value read, value written; last pass writes 2 float4
        return px * ((DN_ITERS - 1) * (16 + 8 + 8) + (16 + 8 + 32))
    if stage == "combine":        # 3 float4 read, 1 written
        return px * 64
    return 0


def run_reference(args):
    """The reference's CPU implementation of the frame on the host cores (bounded sample)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle
    from optix_renderer_b200 import scenes

    kind = "reference-intended" if _oracle.available("reference-intended") else "restatement"
    o = _oracle.Oracle(kind)
    sc = scenes.scene_synthetic(N_TRIS, N_QUADS, W, H)
    o.load_scene(sc)
    cores = os.cpu_count() or 1
    rt = _oracle.RATIO if args.workload == "ratio" else _oracle.PATH
    o.set_options(1, SPP, DEPTH if args.workload == "path" else 1)
    # calibrate the sample: rows of the middle of the frame worth ~1.5 s per step
    t = time.perf_counter()
    o.render(W, H, rt, 0, rows=(536, 544), nthreads=cores)
    per_row = (time.perf_counter() - t) / 8
    budget = min(150.0, max(10.0, 2.0 * (args.steps + args.warmup))) / (args.steps + args.warmup)
    rows = int(max(8, min(H, budget / per_row)))
    y0 = (H - rows) // 2
    for i in range(args.warmup):
        o.render(W, H, rt, i, rows=(y0, y0 + rows), nthreads=cores)
    rays, t0 = 0, time.perf_counter()
    for i in range(args.steps):
        rays += o.render(W, H, rt, args.warmup + i, rows=(y0, y0 + rows), nthreads=cores)["rays"]
    dt = time.perf_counter() - t0
    v = rays / dt / 1e6
    sample = "%d of %d rows (y %d..%d) of the %s frame per step, %d threads" % (rows, H, y0, y0 + rows, args.workload, cores)
    line = {"impl": "reference", "metric": "Mrays/s", "value": v, "unit": "Mrays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "ms_per_full_frame_extrapolated": dt / args.steps * 1e3 * H / rows,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, args.gpus),  # the job's config as the GPU arm states it for this N
            "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": cores,
                             "kind": "reference" if kind.startswith("reference") else "port", "sample": sample},
            "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "CPU: reference device source (rayGen) compiled as host C++ behind the OWL shim; denoise not "
                    "included (the reference has none)"}
    print(json.dumps(line))


def workload_config(args, n):
    if args.scene == "cfg5":
        tag = "BASELINE configs[4] scene (4K, 10 M triangles, 64 quad lights), "
    else:
        tag = ""
    if args.workload == "ratio":
        desc = ("BASELINE configs[2]: synthetic %d-triangle scene %dx%d, %d quad lights (%d TriLights), RATIO intended mode "
                "%d spp shadow rays + %d-pass a-trous denoise + ratio combine" % (N_TRIS, W, H, N_QUADS, 2 * N_QUADS, SPP, DN_ITERS))
    else:
        desc = "BASELINE configs[3]: same scene, PATH intended mode, %d bounces, progressive" % DEPTH
    return {"workload": tag + desc,
            "width": W, "height": H, "triangles": N_TRIS, "tri_lights": 2 * N_QUADS, "shadow_spp": SPP,
            "max_depth": DEPTH if args.workload == "path" else 1,
            "lut": "8x8 (reference ltc_isotropic.h)" if args.lut == 8 else "%dx%d (bilinear resampling of the reference table)" % (args.lut, args.lut),
            "partition": ("%d cost-balanced row bands, BVH replicated, packed denoise halo rows stored into the neighbours' "
                          "buffers by the pack kernel (peer-mapped memory over NVLink, epoch flags), finished bands copied into every "
                          "GPU's frame beside the next frames (one-sided copy-engine transfers, or an NCCL all-gather: this "
                          "line's assembly key)" % n) if n > 1 else "single GPU",
            "frames_in_flight": "several per GPU (orb_set_frame_slots; the count is this line's frames_in_flight key): the "
                                "next frames' kernels run beside a frame's traversal tail; accum / RGBA8 shared and "
                                "accumulated in order",
            "l2": "working set per frame (13+ float4 frame buffers of 33 MB per frame in flight, 40 MB triangles, 37 MB BVH) exceeds the "
                  "126 MB L2; no explicit flush"}


def cpu_baseline_sample(workload, seconds=12.0):
    """Rank 0, N = 1: the reference's CPU code on a bounded sample of the same frame."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle
    from optix_renderer_b200 import scenes

    kind = "reference-intended" if _oracle.available("reference-intended") else "restatement"
    o = _oracle.Oracle(kind)
    o.load_scene(scenes.scene_synthetic(N_TRIS, N_QUADS, W, H))
    cores = os.cpu_count() or 1
    rt = _oracle.RATIO if workload == "ratio" else _oracle.PATH
    o.set_options(1, SPP, DEPTH if workload == "path" else 1)
    o.render(W, H, rt, 0, rows=(536, 544), nthreads=cores)  # warm-up (page-in, thread start)
    t = time.perf_counter()
    o.render(W, H, rt, 0, rows=(500, 580), nthreads=cores)
    per_row = (time.perf_counter() - t) / 80
    rows = int(max(8, min(H, 0.25 * seconds / per_row)))
    y0 = (H - rows) // 2
    rays, reps, t = 0, 0, time.perf_counter()
    while reps < 4 or (time.perf_counter() - t < seconds and reps < 64):
        rays += o.render(W, H, rt, 1 + reps, rows=(y0, y0 + rows), nthreads=cores)["rays"]
        reps += 1
    dt = time.perf_counter() - t
    return {"value": rays / dt / 1e6, "unit": "Mrays/s", "cores": cores,
            "kind": "reference" if kind.startswith("reference") else "port",
            "sample": "%d x (%d of %d rows, y %d..%d) of the same %s frame, %d threads, %.1f s; denoise "
                      "excluded (the reference has none)" % (reps, rows, H, y0, y0 + rows, workload, cores, dt),
            "frames_per_s_extrapolated": 1.0 / (dt / reps * H / rows)}


def run_ours(args):
    import torch
    import torch.distributed as dist

    from optix_renderer_b200 import api, scenes, tiles

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries exactly ONE JSON line: library chatter (NCCL prints its version there) goes to stderr
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # The only collective of the job is the per-frame all-gather of the finished bands (20 - 40 MB), beside the next
        # frames' kernels.  An NCCL kernel is a rendezvous: its CTAs spin until the peers' kernels arrive, and the
        # ranks of a 6-deep frame pipeline drift apart -- with NCCL's default channel count the spinning CTAs took
        # 0.2 - 0.4 ms per frame from the render kernels at N = 2 (1.36 ms without the gather, 1.72 - 1.78 with).
        # Two channels move 40 MB per frame many times over.
        # (one channel is too few: 1.87 ms, the ring of staging blocks backs up).  N = 8 gathers more bytes per frame
        # (bands padded to the tallest) and keeps NCCL's own choice: not measured otherwise.
        # N = 4: 0.927 ms per RATIO frame with 4 channels against 0.986 with NCCL's default (profiles/r2w_n4_*.json).
        nch = os.environ.get("ORB_BENCH_NCCL_CHANNELS", "2" if world == 2 else ("4" if world <= 4 else ""))
        if nch:
            os.environ.setdefault("NCCL_MAX_NCHANNELS", nch)
            os.environ.setdefault("NCCL_MIN_NCHANNELS", "1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ratio = args.workload == "ratio"

    sc = scenes.scene_synthetic(N_TRIS, N_QUADS, W, H)
    if args.lut != 8:
        sc["lut"] = scenes.upsample_lut(sc["lut"], args.lut)
    r = api.Renderer(sc, device=local, quantise_lut=1)  # table fetched with the texture unit's weights, as the reference's tex2D
    # frames in flight per GPU (orb_set_frame_slots) and the persistent kernels' share of an SM.  Measured on one GPU
    # rendering the band an N-way partition gives a rank (tools/ab_band.py, profiles/r2o_ab_band*.jsonl), ms per frame
    # RATIO / PATH:            2 slots, no cap    4 slots, 4 CTAs    6 slots
    #   whole frame            2.717 / 1.880      2.611 / 1.948*     2.600 / 1.802 (cap 4)        * cap 6
    #   270 rows (N = 4)       1.103 / 1.056      0.993 / 0.843      0.979 / 0.827 (cap 3)
    #   135 rows (N = 8)       0.669 / 0.737      0.538 / 0.527      0.521 / 0.466 (cap 2)
    # A persistent kernel on a full grid keeps the other frames' CTAs off the SMs until its last ray ends, and a band's
    # kernel is mostly tail: more frames in flight, each with a smaller share of every SM.
    SLOTS = max(1, min(8, int(os.environ.get("ORB_BENCH_SLOTS", "6"))))
    # 8 GPUs (profiles/r2q_n8_*.json): 6 slots cap 2 / cap 3 / 8 slots cap 2 = 0.672 / 0.638 / 0.694 ms per RATIO frame
    CTA_CAP = int(os.environ.get("ORB_BENCH_CTA_CAP", "4" if world == 1 else "3"))
    r.set_frame_slots(SLOTS)
    if "ORB_TRACE_CTAS_PER_SM" not in os.environ:
        r.set_persistent_ctas_per_sm(CTA_CAP)
    stream = torch.cuda.Stream()  # carries the timing events; ordered behind the frame slots' own streams by T.join()
    T = tiles.TiledRenderer(r, rank, world, DN_ITERS)
    HALO = world > 1 and ratio and os.environ.get("ORB_BENCH_HALO", "1") == "1" and T.connect_halo()
    out_names = ["ratio_final", "rgba8"] if ratio else ["accum", "rgba8"]
    # frame assembly: one-sided copies of the finished band into every peer's frame (orb_copy_rows_to_peer: copy engines,
    # no rendezvous) instead of an NCCL all-gather; ORB_BENCH_PEER_ASSEMBLE=0 or a failed mapping keeps the all-gather
    # (only with the e2e loop's per-rank host gather: a read-back of the ASSEMBLED frame right behind a frame would need
    # a cross-rank wait for the peers' copies, which the all-gather gives and one-sided copies do not)
    PEER_ASM = (world > 1 and os.environ.get("ORB_BENCH_PEER_ASSEMBLE", "1") == "1"
                and not os.environ.get("ORB_BENCH_SYNC_ASSEMBLE") and os.environ.get("ORB_BENCH_SYNC_DOWNLOAD", "0") != "1"
                and os.environ.get("ORB_BENCH_HOST_GATHER", "1") == "1" and T.connect_assembly(out_names))

    NO_ASSEMBLE = os.environ.get("ORB_BENCH_NO_ASSEMBLE") == "1"
    SYNC_ASSEMBLE = bool(os.environ.get("ORB_BENCH_SYNC_ASSEMBLE"))  # A/B: gather in place on the render stream
    SYNC_DOWNLOAD = os.environ.get("ORB_BENCH_SYNC_DOWNLOAD", "0") == "1"  # A/B: blocking read-back every frame

    def render(i):
        if ratio:
            T.render_ratio_frame(i, mode=api.INTENDED, shadow_spp=SPP)
        else:
            T.render_path_frame(i, mode=api.INTENDED, max_depth=DEPTH)

    def step(i):
        # the finished band is handed to a side stream that all-gathers it while this stream renders the next
        # frame; T.join(stream) (before the closing event / before a read) orders `stream` behind the last gather
        render(i)
        if NO_ASSEMBLE:  # diagnostic only (ORB_BENCH_NO_ASSEMBLE=1): what the frame assembly costs the pipeline
            return
        if SYNC_ASSEMBLE:
            T.assemble(out_names)
        else:
            T.assemble_async(out_names)

    def barrier():
        r.sync()
        torch.cuda.synchronize()  # this rank's copies into its peers' frames have landed before anyone passes the barrier
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(x, op):
        if world == 1:
            return x
        t = torch.tensor([float(x)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=op)
        return float(t.item())

    with torch.cuda.stream(stream):
        sampler = ClockSampler(local)
        sampler.start()
        T.floor_ms = float(os.environ.get("ORB_BAND_FLOOR_MS", "0.25" if ratio else "0.45"))
        n_calib = int(os.environ.get("ORB_BAND_CALIB", "6")) if world > 1 else 0
        # N > 1: after the last re-cut every frame slot renders one more untimed frame, so that whatever is sized by the
        # band (ray streams, staging) is re-allocated before the clock starts (a cudaFree / cudaMalloc costs ~10 ms
        # beside 7 peers' mapped memory)
        n_settle = SLOTS if world > 1 else 0
        base_id = args.warmup + n_calib + n_settle  # accumId of the first timed frame
        for i in range(args.warmup + n_calib + n_settle):  # N > 1: n_calib extra untimed frames first, to settle the band cuts
            if i >= args.warmup + n_calib:
                step(i)
                continue
            if world > 1:
                # cut the row bands by cost: a model from the first frame's active pixels, then corrected by
                # this rank's measured kernel time for its band (stage timers: the collectives and the
                # waiting for slower neighbours are not in it) over the following warm-up frames
                r.set_profiling(True)
                r.stage_times()
                step(i)
                mine_ms = sum(t for k, (t, _) in r.stage_times().items() if k != "halo")
                r.set_profiling(False)
                T.rebalance(None if i == 0 else mine_ms)
            else:
                step(i)
        barrier()
        c0 = r.counters()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        ev0.record(stream)          # every stream is idle: the clock starts here
        t_host = time.perf_counter()
        for i in range(args.steps):
            step(base_id + i)
        host_ms = (time.perf_counter() - t_host) * 1e3 / args.steps  # host time to ENQUEUE a frame (no device wait in it)
        T.join(stream)              # behind both frame slots and the last assembly
        ev1.record(stream)
        barrier()
        ms = ev0.elapsed_time(ev1)
        c1 = r.counters()
        # per-stage CUDA-event times, OUTSIDE the timed region: the same frames one at a time (a sync after each), so
        # the stages of two frames in flight do not overlap inside each other's event pairs
        P_FRAMES = max(3, min(10, args.steps))
        # ... and with the persistent kernels' CTA cap lifted: the cap leaves room for the OTHER frames in flight; a
        # kernel timed alone (the roofline's duration) runs on the whole SM
        if "ORB_TRACE_CTAS_PER_SM" not in os.environ:
            r.set_persistent_ctas_per_sm(0)
        r.set_profiling(True)
        r.stage_times()
        if world > 1:
            T.spans = {}
        for i in range(P_FRAMES):
            step(base_id + args.steps + i)
            r.sync()
        st = r.stage_times()
        coll_ms = T.span_ms() if world > 1 else {}
        T.spans = None
        r.set_profiling(False)
        if "ORB_TRACE_CTAS_PER_SM" not in os.environ:
            r.set_persistent_ctas_per_sm(CTA_CAP)
        st = {k: (t * args.steps / P_FRAMES, n * args.steps / P_FRAMES) for k, (t, n) in st.items()}  # per `steps` frames, as below
        coll_ms = {k: v * args.steps / P_FRAMES for k, v in coll_ms.items()}
        base_id += P_FRAMES
        if os.environ.get("ORB_BENCH_TIMELINE"):
            # a timeline of a few frames AS THEY RUN in the timed loop (frames in flight overlap): every stage's
            # CUDA-event start / end on rank 0, relative to the first -- where the overlap happens and where it does not
            r.set_profiling(True)
            r.stage_times()
            for i in range(6):
                step(base_id + i)
            spans = r.stage_spans()
            r.stage_times()
            r.set_profiling(False)
            base_id += 6
            if rank == 0:
                with open(os.environ["ORB_BENCH_TIMELINE"], "w") as fh:
                    json.dump({"n_gpus": world, "slots": SLOTS, "workload": args.workload, "bands": [list(b) for b in T.bands],
                               "spans": [[n, k, round(a, 4), round(b, 4)] for n, k, a, b in spans]}, fh)
        ms = reduce(ms, dist.ReduceOp.MAX) if world > 1 else ms
        rank_ms = None
        if world > 1:  # every rank's own kernel time per frame (stage timers): how even the bands are
            mine = torch.zeros(world, dtype=torch.float64, device="cuda")
            mine[rank] = sum(t for k, (t, _) in st.items() if k != "halo") / args.steps  # halo = waiting for the peers' flags
            dist.all_reduce(mine)
            rank_ms = [round(x, 4) for x in mine.tolist()]
        d = {k: c1[k] - c0[k] for k in ("rays_primary", "rays_shadow_total", "rays_bounce_total", "kernel_launches",
                                         "active_pixels_total")}
        rays = reduce(d["rays_primary"] + d["rays_shadow_total"] + d["rays_bounce_total"], dist.ReduceOp.SUM if world > 1 else None)
        launches = reduce(d["kernel_launches"], dist.ReduceOp.SUM if world > 1 else None)
        value = rays / (ms * 1e-3) / 1e6

        # ---- end to end through the C-ABI with host buffers: camera in (host), frame out (pinned host) ----
        cam = sc["camera"]
        y0, y1 = (0, H)
        t_f4 = torch.empty((2, H, W, 4), dtype=torch.float32, pin_memory=True)   # two host frames: a consumer reads one
        t_u32 = torch.empty((2, H, W), dtype=torch.int32, pin_memory=True)       # while the next one lands
        host_f4, host_u32 = t_f4.numpy(), t_u32.numpy().view(np.uint32)
        e_steps = max(3, min(args.steps, 30))
        # N > 1: every rank copies ITS band into one page-locked host frame shared by the ranks of the box (each
        # over its own PCIe link) instead of rank 0 reading the whole assembled frame back alone
        HOST_GATHER = world > 1 and not SYNC_ASSEMBLE and not SYNC_DOWNLOAD and os.environ.get("ORB_BENCH_HOST_GATHER", "1") == "1"
        shared = None
        if HOST_GATHER:
            try:
                shared = tiles.SharedHostFrames(rank, world, {out_names[0]: ((H, W, 4), np.float32), "rgba8": ((H, W), np.uint32)})
            except RuntimeError as e:  # raised on every rank or on none: rank 0 reads the assembled frame back instead
                if rank == 0:
                    print("bench.py: %s; e2e falls back to rank 0's read-back of the assembled frame" % e, file=sys.stderr)
                HOST_GATHER = False
                PEER_ASM = False
                T._peer_asm, T._peer_names = None, ()  # rank 0 reads the gathered frame back: the all-gather again

        e2e_host = [None]
        e2e_host_parts = {"set_camera": 0.0, "render": 0.0, "assemble": 0.0, "read_back": 0.0, "final_wait": 0.0, "frames": 0}

        def e2e_loop(pipelined, id0, n):
            barrier()
            t0 = time.perf_counter()
            hb = e2e_host_parts
            for k in hb:
                hb[k] = 0  # the last (timed) loop's numbers are the ones reported
            for i in range(n):
                if i == n - 1:
                    e2e_host[0] = (time.perf_counter() - t0) * 1e3 / max(n - 1, 1)  # host time per frame of the enqueue loop
                ta = time.perf_counter()
                r.set_camera(cam["pos"], cam["d00"], cam["du"], cam["dv"])  # host -> library (launch parameters)
                tb = time.perf_counter()
                render(id0 + i)
                tc = time.perf_counter()
                if SYNC_ASSEMBLE:
                    T.assemble(out_names)
                else:
                    T.assemble_async(out_names)
                td = time.perf_counter()
                hb["set_camera"] += tb - ta; hb["render"] += tc - tb; hb["assemble"] += td - tc; hb["frames"] += 1
                if rank == 0 and world == 1 and pipelined:
                    # D2H through the C-ABI, every frame, overlapped with the next frame's kernels
                    r.download_async(out_names[0], host_f4[i & 1], rows=(y0, y1))
                    r.download_async("rgba8", host_u32[i & 1], rows=(y0, y1))
                elif rank == 0 and world == 1:
                    r.download(out_names[0], rows=(y0, y1), out=host_f4[i & 1])  # synchronises every frame
                    r.download("rgba8", rows=(y0, y1), out=host_u32[i & 1])
                elif HOST_GATHER:
                    T.read_back_bands_async(shared, i & 1, out_names)
                elif rank == 0 and not SYNC_ASSEMBLE and not SYNC_DOWNLOAD:
                    # the gathered frame: read back behind the gather, beside the next frame
                    T.read_back_async({out_names[0]: t_f4[i & 1], "rgba8": t_u32[i & 1]})
                elif rank == 0:
                    T.join(stream)
                    stream.synchronize()
                    t_f4[i & 1].copy_(tiles.device_tensor(r, out_names[0]) if SYNC_ASSEMBLE else T.assembled(out_names[0]))
                    t_u32[i & 1].copy_(tiles.device_tensor(r, "rgba8") if SYNC_ASSEMBLE else T.assembled("rgba8"))
                    torch.cuda.current_stream().synchronize()
                hb["read_back"] += time.perf_counter() - td
            if HOST_GATHER:
                T.join(stream)
                stream.synchronize()
                tw = time.perf_counter()
                T.wait_read_back_bands()  # every band of the last frame is on the host when the clock stops
                hb["final_wait"] = (time.perf_counter() - tw) * hb["frames"]  # reported per frame like the others: x frames here
            elif world > 1 and rank == 0:
                T.join(stream)
                stream.synchronize()
                T.wait_read_back()  # the last frame is on the host when the clock stops
            if world == 1 and pipelined:
                r.sync()
                r.download_wait()  # the last frame is on the host when the clock stops
            barrier()
            t = (time.perf_counter() - t0) * 1e3
            return reduce(t, dist.ReduceOp.MAX) if world > 1 else t

        PIPE = world == 1 and not SYNC_DOWNLOAD
        serial_ms = e2e_loop(False, base_id + args.steps, e_steps) if PIPE else None
        if PIPE or HOST_GATHER:
            # untimed: the read-back staging is allocated on first use -- two staging buffers per output and frame slot,
            # so 2 x SLOTS frames touch them all (a cudaMalloc / cudaFree inside the timed loop costs milliseconds at N = 8)
            e2e_loop(PIPE, base_id + args.steps, max(3, 2 * SLOTS))
        e_ms = e2e_loop(PIPE, base_id + args.steps + (e_steps if PIPE else 0), e_steps)
        # the frame that arrived through the pipelined read-back is the frame the device holds
        readback_ok = bool(np.array_equal(host_f4[(e_steps - 1) & 1], r.download(out_names[0]), equal_nan=True)) if PIPE else None
        e_value = (rays / args.steps) * e_steps / (e_ms * 1e-3) / 1e6
        clocks = sampler.summary()
        clocks["window"] = "warm-up + timed region + e2e loop (one continuous load)"
        same = None
        if ratio and world > 1 and rank == 0:
            # parity of the partitioned job: the frame assembled from N bands (halo rows exchanged, bands gathered)
            # against the same frame rendered whole by this rank alone -- bit for bit, outside every timed region
            T.join(stream)
            stream.synchronize()
            got = (tiles.device_tensor(r, "ratio_final") if SYNC_ASSEMBLE else T.assembled("ratio_final")).clone()
            if HOST_GATHER:  # the frame as the host received it: N bands, each copied by its own rank
                got = torch.from_numpy(np.array(shared[out_names[0]][(e_steps - 1) & 1])).to("cuda")
            elif not SYNC_ASSEMBLE and not SYNC_DOWNLOAD:  # the frame as the host received it through the pipelined read-back
                got = t_f4[(e_steps - 1) & 1].to("cuda")
            last = base_id + args.steps + e_steps - 1
            r.render(api.RATIO, accum_id=last, mode=api.INTENDED, shadow_spp=SPP, count_rays=False)
            r.denoise(iterations=DN_ITERS)
            r.ratio_combine(use_denoised=True)
            torch.cuda.synchronize()
            want = tiles.device_tensor(r, "ratio_final")
            # compared on the pixels this frame WRITES: a pixel whose primary ray missed keeps whatever an earlier
            # frame left in ltc / sto_* (the reference's rayGen writes nothing there either, deviceCode.cu:90-94,
            # 118-143), and after the bands moved the ranks hold different histories for such silhouette pixels
            nrm = tiles.device_tensor(r, "normal")
            written = (nrm[..., :3] != 0).any(dim=2)
            px_bad = (got.view(torch.int32) != want.view(torch.int32)).any(dim=2) & written
            same = not bool(px_bad.any())
            if not same:
                rows_bad = torch.nonzero(px_bad.any(dim=1)).flatten().tolist()
                print("bench.py: assembled frame differs from the single-GPU frame in %d pixels, rows %s ... %s; bands %s" %
                      (int(px_bad.sum()), rows_bad[:12], rows_bad[-12:], T.bands), file=sys.stderr)
        if (not ratio) and world > 1:
            # parity of the partitioned PATH job: frame accumId = 0 (accum = that frame's colour, no history, so the
            # band cuts that moved during warm-up do not matter) rendered in N bands and gathered in place, against
            # the same frame rendered whole by rank 0 -- bit for bit, outside every timed region
            T.join(stream)
            stream.synchronize()
            render(0)
            T.assemble(out_names)
            torch.cuda.synchronize()
            if rank == 0:
                got = tiles.device_tensor(r, "accum").clone()
                r.render(api.PATH, accum_id=0, mode=api.INTENDED, max_depth=DEPTH, count_rays=False)
                torch.cuda.synchronize()
                want = tiles.device_tensor(r, "accum")
                px_bad = (got.view(torch.int32) != want.view(torch.int32)).any(dim=2)
                same = not bool(px_bad.any())
                if not same:
                    print("bench.py: assembled PATH frame differs from the single-GPU frame in %d pixels; bands %s" %
                          (int(px_bad.sum()), T.bands), file=sys.stderr)
            barrier()
        trav, one_quad = None, None
        if ratio and world == 1:
            # BASELINE's "HBM GB/s fraction on the LTC pass": the stand-alone LTC pass over the resident G-buffer
            # with ONE quad light (2 TriLights, the only light count near the HBM roof); outside
            # the timed region, 20 launches, working set 166 MB > L2
            r.set_lights(sc["tri_lights"][:2])
            r.stage_times()
            for _ in range(3):
                r.ltc_gbuffer_pass()
            torch.cuda.synchronize()
            r.set_profiling(True)
            r.stage_times()
            for _ in range(20):
                r.ltc_gbuffer_pass()
            torch.cuda.synchronize()
            one_quad = r.stage_times()["ltc_gbuffer"][0] / 20
            r.set_profiling(False)
            r.set_lights(sc["tri_lights"])
        if ratio and world == 1:
            # one extra frame, outside every timed region, through the instrumented twin of the traversal kernel:
            # node / triangle records per shadow ray of THIS workload (the algorithmic bytes of the roofline)
            r.render(api.RATIO, accum_id=base_id + args.steps, mode=api.INTENDED, shadow_spp=SPP, count_rays=False,
                     count_traversal=True)
            cc = r.counters()
            trav = {"nodes": cc["nodes_visited"], "tris": cc["tris_tested"], "rays": cc["rays_shadow"]}
        if not ratio and world == 1:
            r.render(api.PATH, accum_id=base_id + args.steps, mode=api.INTENDED, max_depth=DEPTH, count_rays=False,
                     count_traversal=True)
            cc = r.counters()
            trav = {"nodes": cc["nodes_visited"], "tris": cc["tris_tested"], "rays": cc["rays_bounce"]}

    if rank == 0:
        px = W * H // world  # per rank (bands); stage times are rank 0's
        active = d["active_pixels_total"] / args.steps
        peak, peak_src = peaks()
        passes = {}
        for name, (t_ms, n_launch) in st.items():
            if n_launch == 0:
                continue
            per = t_ms / args.steps
            b = stage_bytes(name, px, active, trav=trav)
            passes[name] = {"ms_per_frame": per, "launches_per_frame": n_launch / args.steps, "algorithmic_bytes": b,
                            "gbs": b / (per * 1e-3) / 1e9 if per > 0 else None,
                            "frac_of_hbm_peak": b / (per * 1e-3) / 1e9 / peak if per > 0 else None}
        top = max(passes, key=lambda k: passes[k]["ms_per_frame"])
        tp = passes[top]
        n_top = tp["launches_per_frame"]
        kernel_of = {"trace": "k_trace_stream_spec", "path": "k_path_persistent", "ltc": "k_ltc", "primary": "k_primary",
                     "denoise": "k_denoise_atrous"}
        # ---- roofline of the dominant kernel --------------------------------------------------------------------
        # The traversal kernels are not HBM-bound (their fetches are served by L1 / L2): the roofs they are under are
        # the SM's issue slots and the L1 data pipe (one wavefront per cycle and SM).  Numerators come from the
        # committed `ncu --set full` capture of this same workload (profiles/roofline_traffic.json, written by
        # tools/ncu_summary.py --json: warp instructions, L1 wavefronts and DRAM bytes per launch -- the workload is
        # deterministic), denominators from THIS run: the kernel's CUDA-event time and the SM clock sampled under load.
        kname = kernel_of.get(top, top)
        n_sm = torch.cuda.get_device_properties(local).multi_processor_count
        clk_hz = (clocks.get("sm_mhz") or 1965.0) * 1e6
        dur_s = tp["ms_per_frame"] / n_top * 1e-3
        cap = None
        tj = os.path.join(ROOT, "profiles", "roofline_traffic.json")
        if os.path.exists(tj) and world == 1 and args.scene == "cfg3":
            cap = json.load(open(tj)).get(kname)
        roofline = {"kernel": kname, "stage": top, "avg_launch_ms": dur_s * 1e3,
                    "share_of_step_serial": tp["ms_per_frame"] / sum(v["ms_per_frame"] for v in passes.values()),
                    "sm_clock_mhz": clk_hz / 1e6, "sms": n_sm,
                    "fetched_bytes_per_launch": tp["algorithmic_bytes"] / n_top if tp["algorithmic_bytes"] else None,
                    "fetched_gbs": tp["gbs"], "hbm_peak_gbs": peak, "peak_source": peak_src}
        if cap and cap.get("warp_inst") and top in ("trace", "path", "primary"):
            issue_rate, issue_peak = cap["warp_inst"] / dur_s, n_sm * 4 * clk_hz
            l1_rate, l1_peak = (cap.get("l1_wavefronts") or 0) / dur_s, n_sm * clk_hz
            l1_bound = l1_rate / l1_peak > issue_rate / issue_peak
            roofline.update({
                "bound": "l1-wavefronts" if l1_bound else "issue",
                "achieved": (l1_rate if l1_bound else issue_rate) / 1e9, "peak": (l1_peak if l1_bound else issue_peak) / 1e9,
                "unit": "G wavefronts/s" if l1_bound else "G warp-instructions/s",
                "frac": (l1_rate / l1_peak) if l1_bound else (issue_rate / issue_peak),
                "issue_frac": issue_rate / issue_peak, "l1_wavefront_frac": l1_rate / l1_peak if l1_rate else None,
                "lanes_per_inst": cap.get("threads_per_inst"), "traffic": cap.get("dram_bytes_per_launch"),
                "hbm_frac": cap["dram_bytes_per_launch"] / dur_s / 1e9 / peak if cap.get("dram_bytes_per_launch") else None,
                "counts_source": cap.get("source"),
                "note": "issue roof = SMs x 4 schedulers x SM clock; L1 roof = SMs x 1 wavefront per cycle x SM clock; "
                        "instruction / wavefront / DRAM-byte counts per launch from the committed ncu capture of this "
                        "workload, duration and clock measured in this run.  fetched_* = node and triangle records "
                        "fetched (counted by the instrumented twin) -- served by L1 / L2, not an HBM figure"})
        elif top in ("trace", "path", "primary"):
            roofline.update({"bound": "issue", "achieved": None, "peak": n_sm * 4 * clk_hz / 1e9, "unit": "G warp-instructions/s",
                             "frac": None, "traffic": None,
                             "note": "no ncu capture of this kernel on this partition / scene under profiles/roofline_traffic.json: "
                                     "counts unknown (the capture is the single-GPU full frame of configs[2] / configs[3])"})
        else:  # a streaming pass on top: HBM roof, algorithmic bytes over its own time
            roofline.update({"bound": "hbm", "achieved": tp["gbs"], "peak": peak, "unit": "GB/s", "frac": tp["frac_of_hbm_peak"],
                             "traffic": cap.get("dram_bytes_per_launch") if cap else None})
        if one_quad:
            b = px * 32 + active * 48  # normal read + ltc written everywhere; position, albedo, alpha where a surface was hit
            line_ltc = {"tri_lights": 2, "ms": one_quad, "algorithmic_bytes": b, "gbs": b / (one_quad * 1e-3) / 1e9,
                        "frac_of_hbm_peak": b / (one_quad * 1e-3) / 1e9 / peak,
                        "what": "orb_ltc_gbuffer_pass over the frame's resident G-buffer, one quad light: 32 B per pixel + 48 B per "
                                "surface pixel"}
        if trav:
            roofline["per_ray"] = {"nodes": trav["nodes"] / max(trav["rays"], 1), "tris": trav["tris"] / max(trav["rays"], 1),
                                   "rays_per_launch": trav["rays"]}
        line = {"metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "frames_per_s": args.steps / (ms * 1e-3),
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "config": workload_config(args, world), "rays_per_frame": rays / args.steps,
                "e2e": {"value": e_value, "unit": "Mrays/s", "ms_per_step": e_ms / e_steps,
                        "frames_per_s": e_steps / (e_ms * 1e-3), "h2d_bytes_per_step": 48 + 32,
                        "d2h_bytes_per_step": W * H * 20, "steps": e_steps, "host_enqueue_ms_per_step": e2e_host[0],
                        "host_ms_per_step_by_call": {k: round(v * 1e3 / max(1, e2e_host_parts["frames"]), 4)
                                                     for k, v in e2e_host_parts.items() if k != "frames"},
                        "what": "orb_set_camera + orb_render_async(+denoise+combine) + %s of the "
                                "float4 result and the RGBA8 frame into pinned host memory, every frame"
                                % ("orb_buffer_download_rows_async (read-back of frame i overlaps the kernels of frame i+1; "
                                   "orb_download_wait inside the timed region)" if PIPE else
                                   "orb_buffer_download_rows" if world == 1 else
                                   "every rank's orb_buffer_download_rows_async of ITS band into one page-locked host "
                                   "frame shared by the ranks of the box (tiles.SharedHostFrames; the device all-gather of "
                                   "the frame runs beside it as in the device-timed loop; orb_download_wait on every rank "
                                   "and a barrier inside the timed region)" if HOST_GATHER else
                                   "rank 0's copy of the assembled frame" + ("" if SYNC_ASSEMBLE or SYNC_DOWNLOAD else
                                   " on a copy stream, behind the all-gather and beside the next frame's kernels; the "
                                   "last copy is waited for inside the timed region"))},
                "gpu_launches": int(launches), "frames_in_flight": SLOTS, "host_enqueue_ms_per_step": host_ms, "roofline": roofline, "passes": passes,
                "persistent_ctas_per_sm": CTA_CAP, "passes_note": "per-stage CUDA-event times of %d extra frames run ONE AT A TIME (CTA cap lifted) after the timed region "
                               "(with %d frames in flight the stages of consecutive frames overlap, so their sum "
                               "exceeds ms_per_step)" % (P_FRAMES, SLOTS),
                "clocks": clocks}
        if PIPE:
            line["e2e"]["serial_ms_per_step"] = serial_ms / e_steps  # the same loop with a blocking download per frame
            line["e2e"]["readback_equals_device_frame"] = readback_ok
        if one_quad:
            line["ltc_pass_one_quad"] = line_ltc
        if world > 1:
            line["e2e"]["host_gather"] = bool(HOST_GATHER)
            line["bands"] = [list(b) for b in T.bands]  # beside config, so both arms print the same config at every N
            line["assembled_frame_equals_single_gpu_frame"] = same
            line["assembly"] = ("orb_copy_rows_to_peer: every rank's band into every peer's frame by the copy engines over "
                                "NVLink, no rendezvous") if PEER_ASM else "NCCL all_gather_into_tensor on a side stream"
            line["rank_kernel_ms_per_frame"] = rank_ms
            line["collectives_ms_per_frame_rank0"] = {k: v / args.steps for k, v in coll_ms.items()}
            line["halo"] = ("peer stores inside k_denoise_pack + epoch flags (orb_denoise_band); status %d" % r.halo_status()) if HALO \
                else ("torch.distributed point-to-point" if ratio else None)
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_sample(args.workload)
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        print(json.dumps(line), flush=True)
    r.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="ratio", choices=["ratio", "path"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--lut", type=int, default=8, help="LTC table size (8 = the reference's ltc_isotropic.h; 64 = north_star's)")
    ap.add_argument("--scene", default="cfg3", choices=["cfg3", "cfg5"],
                    help="cfg3: 1080p, 1 M triangles, 16 quad lights (default); cfg5: 4K, 10 M triangles, 64 quad lights")
    args = ap.parse_args()
    global W, H, N_TRIS, N_QUADS
    if args.scene == "cfg5":
        W, H, N_TRIS, N_QUADS = 3840, 2160, 10_000_000, 64
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

"""ctypes binding of liborb200.so (include/orb200.h) and a host-side mirror of the reference's
`Viewer` (include/viewer.hpp:163-845) for the per-pixel hot path.

The library is the product; this file is the thin Python view of its C-ABI that the tests and
bench.py use.  There is no CPU fallback: if liborb200.so is missing or no CUDA device is
visible, construction raises.

    Viewer ctor        include/viewer.hpp:163-551   -> Renderer(scene)
    Viewer::resize     :571-612                      -> Renderer.resize(W, H)
    cameraChanged      :621-657                      -> Renderer.set_camera(...)
    Viewer::render     :553-568  (owlLaunch2D)       -> Renderer.render(renderer_type, accum_id, ...)
    savebuffer         :695-713                      -> Renderer.download(buffer) / save_btc(...)
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ORB200_LIB") or os.path.join(HERE, "liborb200.so")  # env override: A/B builds in tools/

# RendererType, cuda_include/common.cuh:17-29
DIFFUSE, ALPHA, NORMALS, SHADE_NORMALS, POSITION, MASK, MATERIAL_ID, LTC_BASELINE, RATIO, PATH = range(10)
LITERAL, INTENDED = 0, 1

BUFFERS = {
    "position": 0, "normal": 1, "uv": 2, "albedo": 3, "alpha": 4, "materialID": 5, "bounce0": 6, "bounce1": 7,
    "bounce2": 8, "ltc": 9, "sto_direct": 10, "sto_no_vis": 11, "accum": 12, "rgba8": 13, "hit": 14,
    "sto_direct_dn": 15, "sto_no_vis_dn": 16, "ratio_final": 17, "dn_guide": 18, "dn_val": 19,
}

_f = C.POINTER(C.c_float)
_i = C.POINTER(C.c_int32)
_u = C.POINTER(C.c_uint32)
_b = C.POINTER(C.c_uint8)


class OrbError(RuntimeError):
    pass


class OrbMeshDesc(C.Structure):
    _fields_ = [("vertex", _f), ("normal", _f), ("texcoord", _f), ("n_vertices", C.c_int32), ("index", _i),
                ("n_triangles", C.c_int32), ("diffuse", C.c_float * 3), ("alpha", C.c_float),
                ("emit", C.c_float * 3), ("is_light", C.c_int32), ("material_id", C.c_uint32),
                ("diffuse_tex_rgba8", _b), ("tex_w", C.c_int32), ("tex_h", C.c_int32)]


BVH_SAH, BVH_LBVH = 0, 1
NO_ALLOC_MAX_RAYS = 65536  # ORB_NO_ALLOC_MAX_RAYS (include/orb200.h)


class OrbBvhOptions(C.Structure):
    _fields_ = [("max_leaf_tris", C.c_int32), ("builder", C.c_int32), ("reserved", C.c_int32 * 2)]


class OrbRenderConfig(C.Structure):
    _fields_ = [("renderer_type", C.c_int32), ("accum_id", C.c_int32), ("mode", C.c_int32),
                ("shadow_spp", C.c_int32), ("max_depth", C.c_int32), ("y0", C.c_int32), ("y1", C.c_int32),
                ("flags", C.c_int32)]


class OrbDenoiseConfig(C.Structure):
    _fields_ = [("iterations", C.c_int32), ("sigma_normal", C.c_float), ("sigma_position", C.c_float),
                ("sigma_colour", C.c_float)]


class OrbHaloSend(C.Structure):
    _fields_ = [("peer_rank", C.c_int32), ("y0", C.c_int32), ("y1", C.c_int32)]


class OrbCounters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("rays_primary", "rays_shadow", "rays_bounce", "kernel_launches",
                                          "nan_pixels", "bvh_nodes", "bvh_triangles", "bvh_height",
                                          "active_pixels", "nodes_visited", "tris_tested",
                                          "rays_shadow_total", "rays_bounce_total", "active_pixels_total", "bvh_build_us", "bvh_refit_us",
                                          "hit_ties", "ltc_near_decision")]


STAGES = ["primary", "ltc", "shadow", "path", "accumulate", "denoise_pack", "denoise", "combine", "ltc_gbuffer",
          "trace", "halo"]


class OrbStageSpan(C.Structure):
    _fields_ = [("stage", C.c_int32), ("slot", C.c_int32), ("t0_ms", C.c_float), ("t1_ms", C.c_float)]


class OrbStageTimes(C.Structure):
    _fields_ = [("ms", C.c_double * 11), ("launches", C.c_uint64 * 11)]


_lib = None


def load_library():
    """Load liborb200.so (built in-tree by __graft_entry__.build()).  Raises if it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise OrbError("liborb200.so not built (run __graft_entry__.build()); there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    L.orb_last_error.restype = C.c_char_p
    L.orb_create.argtypes = [C.c_int, C.POINTER(vp)]
    L.orb_destroy.argtypes = [vp]
    L.orb_destroy.restype = None
    L.orb_stream.argtypes = [vp]
    L.orb_stream.restype = vp
    L.orb_set_stream.argtypes = [vp, vp]
    L.orb_sync.argtypes = [vp]
    L.orb_set_profiling.argtypes = [vp, C.c_int]
    L.orb_get_stage_times.argtypes = [vp, C.POINTER(OrbStageTimes)]
    L.orb_get_stage_spans.argtypes = [vp, C.POINTER(OrbStageSpan), C.c_int32, C.POINTER(C.c_int32)]
    L.orb_scene_clear.argtypes = [vp]
    L.orb_scene_add_mesh.argtypes = [vp, C.POINTER(OrbMeshDesc)]
    L.orb_scene_set_lights.argtypes = [vp, _f, C.c_int32]
    L.orb_scene_set_ltc_lut.argtypes = [vp, _f, _f, _f, C.c_int32, C.c_int32, C.c_int32]
    L.orb_scene_commit.argtypes = [vp, C.POINTER(OrbBvhOptions)]
    L.orb_scene_update_mesh.argtypes = [vp, C.c_int32, _f, _f, C.c_int32, _i, C.c_int32]
    L.orb_scene_refit.argtypes = [vp]
    L.orb_set_camera.argtypes = [vp, _f, _f, _f, _f]
    L.orb_resize.argtypes = [vp, C.c_int32, C.c_int32]
    L.orb_render.argtypes = [vp, C.POINTER(OrbRenderConfig)]
    L.orb_render_async.argtypes = [vp, C.POINTER(OrbRenderConfig)]
    L.orb_set_frame_slots.argtypes = [vp, C.c_int32]
    L.orb_frame_slot.argtypes = [vp]
    L.orb_frame_slots.argtypes = [vp]
    L.orb_slot_stream.argtypes = [vp, C.c_int32]
    L.orb_slot_stream.restype = vp
    L.orb_ltc_gbuffer_pass.argtypes = [vp, C.c_int32, C.c_int32]
    L.orb_denoise.argtypes = [vp, C.POINTER(OrbDenoiseConfig), C.c_int32, C.c_int32]
    L.orb_denoise_pack.argtypes = [vp, C.c_int32, C.c_int32]
    L.orb_denoise_filter.argtypes = [vp, C.POINTER(OrbDenoiseConfig), C.c_int32, C.c_int32]
    L.orb_row_active_counts.argtypes = [vp, vp]
    L.orb_halo_local.argtypes = [vp, C.c_int32, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]
    L.orb_ipc_export.argtypes = [vp, vp, C.c_char_p]
    L.orb_ipc_open.argtypes = [vp, C.c_char_p, C.POINTER(vp)]
    L.orb_ipc_close.argtypes = [vp, vp]
    L.orb_halo_connect.argtypes = [vp, C.c_int32, C.c_int32, C.c_int32, vp, vp, vp]
    L.orb_halo_disconnect.argtypes = [vp]
    L.orb_halo_status.argtypes = [vp, C.POINTER(C.c_int32)]
    L.orb_denoise_band.argtypes = [vp, C.POINTER(OrbDenoiseConfig), C.c_int32, C.c_int32, C.POINTER(OrbHaloSend), C.c_int32,
                                   _i, C.c_int32]
    L.orb_ratio_combine.argtypes = [vp, C.c_int32, C.c_int32, C.c_int32]
    L.orb_set_fused_combine.argtypes = [vp, C.c_int32]
    L.orb_set_persistent_ctas_per_sm.argtypes = [vp, C.c_int32]
    L.orb_buffer_bytes.argtypes = [vp, C.c_int32]
    L.orb_buffer_bytes.restype = C.c_size_t
    L.orb_buffer_device_ptr.argtypes = [vp, C.c_int32]
    L.orb_buffer_device_ptr.restype = vp
    L.orb_slot_buffer_device_ptr.argtypes = [vp, C.c_int32, C.c_int32]
    L.orb_slot_buffer_device_ptr.restype = vp
    L.orb_copy_rows_to_peer.argtypes = [vp, C.c_int32, vp, C.c_int32, C.c_int32, C.c_int32]
    L.orb_buffer_download.argtypes = [vp, C.c_int32, vp, C.c_size_t]
    L.orb_buffer_download_rows.argtypes = [vp, C.c_int32, vp, C.c_int32, C.c_int32]
    L.orb_buffer_download_rows_async.argtypes = [vp, C.c_int32, vp, C.c_int32, C.c_int32]
    L.orb_download_wait.argtypes = [vp]
    L.orb_buffer_upload.argtypes = [vp, C.c_int32, vp, C.c_size_t]
    L.orb_get_counters.argtypes = [vp, C.POINTER(OrbCounters)]
    L.orb_ratio_estimator_host.argtypes = [vp, _f, _f, _f, _f, C.c_size_t]
    L.orb_trace_rays_host.argtypes = [vp, _f, _f, C.c_size_t, C.c_int32, _i, _f, _f, _f]
    L.orb_trace_rays_no_alloc.argtypes = [vp, vp, vp, C.c_uint32, vp, vp, vp, vp]
    L.orb_unit_rng.argtypes = [vp, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, _u, _u, _f]
    L.orb_unit_clip.argtypes = [vp, _f, C.c_int, _i, _f]
    L.orb_unit_integrate_edge.argtypes = [vp, _f, _f, _f]
    L.orb_unit_fetch_ltc.argtypes = [vp, C.c_float, C.c_float, _f]
    L.orb_unit_mat_inverse.argtypes = [vp, _f, _f]
    L.orb_unit_onb.argtypes = [vp, _f, _f]
    L.orb_unit_sample_direction.argtypes = [vp, _f, C.c_float, C.c_float, _f, C.c_float, _f]
    L.orb_unit_evaluate.argtypes = [vp, _f, _f, _f, C.c_float, _f]
    L.orb_unit_pdf.argtypes = [vp, _f, _f, _f, C.c_float, _f]
    L.orb_unit_make_rgba.argtypes = [vp, _f, _u]
    L.orb_unit_sample_tri.argtypes = [vp, _f, C.c_float, C.c_float, _f]
    L.orb_unit_pdf_a2w.argtypes = [vp, C.c_float, C.c_float, C.c_float, _f]
    L.orb_unit_ltc_pixel.argtypes = [vp, _f, _f, _f, _f, C.c_float, _f]
    _lib = L
    return L


def f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _fp(a):
    return a.ctypes.data_as(_f)


class Renderer:
    """One context on one GPU.  `scene` is the dict made by optix_renderer_b200.scenes."""

    def __init__(self, scene=None, device=0, quantise_lut=0, max_leaf_tris=0, builder=BVH_SAH):
        self.lib = load_library()
        h = C.c_void_p()
        self._h = None
        self._ck(self.lib.orb_create(int(device), C.byref(h)))
        self._h = h
        self.W = self.H = 0
        self._dl_keep = []
        if scene is not None:
            self.load_scene(scene, quantise_lut, max_leaf_tris, builder)

    def _ck(self, rc):
        if rc != 0:
            raise OrbError("orb status %d: %s" % (rc, self.lib.orb_last_error().decode()))

    def close(self):
        if self._h is not None:
            self.lib.orb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- scene (Viewer ctor, viewer.hpp:234-517) ----
    def add_mesh(self, m):
        vtx, nrm, uv = f32(m["vtx"]), f32(m["nrm"]), f32(m["uv"])
        idx = np.ascontiguousarray(m["idx"], np.int32)
        d = OrbMeshDesc()
        d.vertex, d.normal, d.texcoord, d.n_vertices = _fp(vtx), _fp(nrm), _fp(uv), vtx.shape[0]
        d.index, d.n_triangles = idx.ctypes.data_as(_i), idx.shape[0]
        d.diffuse = (C.c_float * 3)(*[float(x) for x in m["diffuse"]])
        d.alpha = float(m["alpha"])
        d.emit = (C.c_float * 3)(*[float(x) for x in m["emit"]])
        d.is_light = int(m["is_light"])
        d.material_id = int(m["material_id"])
        tex = m.get("tex")
        if tex is not None:
            tex = np.ascontiguousarray(tex, np.uint8)
            d.diffuse_tex_rgba8 = tex.ctypes.data_as(_b)
            d.tex_h, d.tex_w = tex.shape[:2]
        self._ck(self.lib.orb_scene_add_mesh(self._h, C.byref(d)))

    def set_lights(self, tri_lights):
        t = f32(tri_lights).reshape(-1, 19)
        self._ck(self.lib.orb_scene_set_lights(self._h, _fp(t), t.shape[0]))

    def set_ltc_lut(self, lut, quantise=0):
        m1, m2, m3 = [f32(x) for x in lut]
        self._ck(self.lib.orb_scene_set_ltc_lut(self._h, _fp(m1), _fp(m2), _fp(m3), m1.shape[1], m1.shape[0],
                                                int(quantise)))

    def commit(self, max_leaf_tris=0, builder=BVH_SAH):
        o = OrbBvhOptions()
        o.max_leaf_tris = int(max_leaf_tris)
        o.builder = int(builder)
        self._ck(self.lib.orb_scene_commit(self._h, C.byref(o)))

    def update_mesh(self, mesh_index, vtx, nrm, idx):
        """New corner positions / normals for a mesh of the committed scene (same triangle count); call refit()."""
        vtx, nrm = f32(vtx), f32(nrm)
        idx = np.ascontiguousarray(idx, np.int32)
        self._ck(self.lib.orb_scene_update_mesh(self._h, int(mesh_index), _fp(vtx), _fp(nrm), vtx.shape[0],
                                                idx.ctypes.data_as(_i), idx.shape[0]))

    def refit(self):
        """Recompute triangle records and BVH boxes on the GPU, topology unchanged (viewer.hpp:506-517 rebuilds)."""
        self._ck(self.lib.orb_scene_refit(self._h))

    def set_camera(self, pos, d00, du, dv):
        a = [f32(x) for x in (pos, d00, du, dv)]
        self._ck(self.lib.orb_set_camera(self._h, *[_fp(x) for x in a]))

    def load_scene(self, scene, quantise_lut=0, max_leaf_tris=0, builder=BVH_SAH):
        self._ck(self.lib.orb_scene_clear(self._h))
        for m in scene["meshes"]:
            self.add_mesh(m)
        self.set_lights(scene["tri_lights"])
        self.set_ltc_lut(scene["lut"], quantise_lut)
        self.commit(max_leaf_tris, builder)
        cam = scene["camera"]
        self.set_camera(cam["pos"], cam["d00"], cam["du"], cam["dv"])
        if "width" in scene:
            self.resize(scene["width"], scene["height"])

    def resize(self, W, H):
        self._ck(self.lib.orb_resize(self._h, int(W), int(H)))
        self.W, self.H = int(W), int(H)

    # ---- launch (Viewer::render, viewer.hpp:553-568) ----
    def _cfg(self, renderer_type, accum_id, mode, shadow_spp, max_depth, rows, count_rays=True, fused_shadow=False,
             path_wavefront=False, count_traversal=False, checked=False):
        c = OrbRenderConfig()
        # ORB_RENDER_NO_COUNT, _FUSED_SHADOW, _PATH_WAVEFRONT, _COUNT_TRAVERSAL
        c.flags = ((0 if count_rays else 1) | (2 if fused_shadow else 0) | (4 if path_wavefront else 0)
                   | (8 if count_traversal else 0) | (32 if checked else 0))
        c.renderer_type, c.accum_id, c.mode = int(renderer_type), int(accum_id), int(mode)
        c.shadow_spp, c.max_depth = int(shadow_spp), int(max_depth)
        c.y0, c.y1 = rows if rows else (0, 0)
        return c

    def render(self, renderer_type, accum_id=0, mode=LITERAL, shadow_spp=4, max_depth=1, rows=None, sync=True,
               count_rays=True, fused_shadow=False, path_wavefront=False, count_traversal=False, checked=False):
        c = self._cfg(renderer_type, accum_id, mode, shadow_spp, max_depth, rows, count_rays, fused_shadow, path_wavefront,
                      count_traversal, checked)
        self._ck((self.lib.orb_render if sync else self.lib.orb_render_async)(self._h, C.byref(c)))

    def set_frame_slots(self, n):
        """1 (default) or 2 frames in flight: with 2, consecutive render() calls alternate between two sets of frame
        buffers and two streams (accum / RGBA8 shared, accumulated in call order); everything called after a render()
        -- denoise, combine, download, device_ptr, stream -- refers to THAT frame until the next render()."""
        self._ck(self.lib.orb_set_frame_slots(self._h, int(n)))

    @property
    def frame_slot(self):
        return int(self.lib.orb_frame_slot(self._h))

    @property
    def frame_slots(self):
        return int(self.lib.orb_frame_slots(self._h))

    def slot_stream(self, k):
        """cudaStream_t (as int) of frame slot k."""
        return self.lib.orb_slot_stream(self._h, int(k))

    def ltc_gbuffer_pass(self, rows=None):
        y0, y1 = rows if rows else (0, 0)
        self._ck(self.lib.orb_ltc_gbuffer_pass(self._h, y0, y1))

    def denoise(self, iterations=5, sigma_normal=32.0, sigma_position=0.25, sigma_colour=0.0, rows=None):
        d = OrbDenoiseConfig(int(iterations), float(sigma_normal), float(sigma_position), float(sigma_colour))
        y0, y1 = rows if rows else (0, 0)
        self._ck(self.lib.orb_denoise(self._h, C.byref(d), y0, y1))

    def denoise_pack(self, rows=None):
        """First half of denoise(): this band's rows -> the packed guide / value buffers."""
        y0, y1 = rows if rows else (0, 0)
        self._ck(self.lib.orb_denoise_pack(self._h, y0, y1))

    def denoise_filter(self, iterations=5, sigma_normal=32.0, sigma_position=0.25, sigma_colour=0.0, rows=None):
        """Second half: the a-trous passes over rows; reads packed rows up to denoise_reach beyond them."""
        d = OrbDenoiseConfig(int(iterations), float(sigma_normal), float(sigma_position), float(sigma_colour))
        y0, y1 = rows if rows else (0, 0)
        self._ck(self.lib.orb_denoise_filter(self._h, C.byref(d), y0, y1))

    # ---- tile hand-off between the ranks of a row-band partition (include/orb200.h, csrc/halo.cuh) ----
    def halo_local(self, slot=0):
        """Device pointers (guide, val, control block) a peer needs to push its rows into this context's `slot`."""
        g, v, f = C.c_void_p(), C.c_void_p(), C.c_void_p()
        self._ck(self.lib.orb_halo_local(self._h, int(slot), C.byref(g), C.byref(v), C.byref(f)))
        return g.value, v.value, f.value

    def ipc_export(self, ptr):
        h = C.create_string_buffer(64)
        self._ck(self.lib.orb_ipc_export(self._h, C.c_void_p(ptr), h))
        return h.raw

    def ipc_open(self, handle):
        out = C.c_void_p()
        self._ck(self.lib.orb_ipc_open(self._h, C.create_string_buffer(bytes(handle), 64), C.byref(out)))
        return out.value

    def ipc_close(self, ptr):
        self._ck(self.lib.orb_ipc_close(self._h, C.c_void_p(ptr)))

    def halo_connect(self, my_rank, peer_rank, slot, guide, val, flags):
        self._ck(self.lib.orb_halo_connect(self._h, int(my_rank), int(peer_rank), int(slot), C.c_void_p(guide), C.c_void_p(val),
                                           C.c_void_p(flags)))

    def halo_disconnect(self):
        self._ck(self.lib.orb_halo_disconnect(self._h))

    def halo_status(self):
        s = C.c_int32()
        self._ck(self.lib.orb_halo_status(self._h, C.byref(s)))
        return int(s.value)

    def denoise_band(self, rows, sends=(), recv_ranks=(), iterations=5, sigma_normal=32.0, sigma_position=0.25, sigma_colour=0.0):
        """Denoise of this rank's band with the hand-off fused in: pack (+ peer stores of the rows in `sends` =
        [(peer_rank, y0, y1), ...]), flags, filter once the rows of `recv_ranks` have arrived.  Asynchronous."""
        d = OrbDenoiseConfig(int(iterations), float(sigma_normal), float(sigma_position), float(sigma_colour))
        sa = (OrbHaloSend * max(len(sends), 1))(*[OrbHaloSend(int(p), int(a), int(b)) for p, a, b in sends])
        ra = (C.c_int32 * max(len(recv_ranks), 1))(*[int(x) for x in recv_ranks])
        self._ck(self.lib.orb_denoise_band(self._h, C.byref(d), int(rows[0]), int(rows[1]), sa, len(sends), ra, len(recv_ranks)))

    def row_active_counts(self):
        """Active (non-light hit) pixels per row of the last frame, int32[H]."""
        out = np.zeros(self.H, np.int32)
        self._ck(self.lib.orb_row_active_counts(self._h, out.ctypes.data_as(C.c_void_p)))
        return out

    def set_persistent_ctas_per_sm(self, n):
        """Frame slots > 1: CTAs per SM the persistent traversal kernels may take (0 = all that fit)."""
        self._ck(self.lib.orb_set_persistent_ctas_per_sm(self._h, int(n)))

    def set_fused_combine(self, enable=True):
        """The filter's final pass also writes ratio_final (= ratio_combine(use_denoised=True), which is then left out)."""
        self._ck(self.lib.orb_set_fused_combine(self._h, int(bool(enable))))

    def ratio_combine(self, use_denoised=False, rows=None):
        y0, y1 = rows if rows else (0, 0)
        self._ck(self.lib.orb_ratio_combine(self._h, int(use_denoised), y0, y1))

    def sync(self):
        self._ck(self.lib.orb_sync(self._h))

    def stream(self):
        return self.lib.orb_stream(self._h)

    def set_stream(self, cuda_stream):
        self._ck(self.lib.orb_set_stream(self._h, C.c_void_p(cuda_stream)))

    def set_profiling(self, enable=True):
        self._ck(self.lib.orb_set_profiling(self._h, int(enable)))

    def stage_spans(self, max_spans=4096):
        """[(stage name, slot, t0_ms, t1_ms), ...] recorded since the last stage_times(): a timeline (synchronises)."""
        a = (OrbStageSpan * max_spans)()
        n = C.c_int32()
        self._ck(self.lib.orb_get_stage_spans(self._h, a, max_spans, C.byref(n)))
        return [(STAGES[a[i].stage], a[i].slot, a[i].t0_ms, a[i].t1_ms) for i in range(min(n.value, max_spans))]

    def stage_times(self):
        """{stage: (summed ms, kernel launches)} since the previous call (synchronises)."""
        t = OrbStageTimes()
        self._ck(self.lib.orb_get_stage_times(self._h, C.byref(t)))
        return {n: (t.ms[i], int(t.launches[i])) for i, n in enumerate(STAGES)}

    # ---- buffers (savebuffer, viewer.hpp:695-713) ----
    def device_ptr(self, name):
        return self.lib.orb_buffer_device_ptr(self._h, BUFFERS[name])

    def buffer_bytes(self, name):
        return self.lib.orb_buffer_bytes(self._h, BUFFERS[name])

    def slot_buffer_ptr(self, slot, name):
        """Device pointer of buffer `name` of frame slot `slot` (accum / rgba8: one buffer for every slot)."""
        return self.lib.orb_slot_buffer_device_ptr(self._h, int(slot), BUFFERS[name])

    def copy_rows_to_peer(self, name, peer_ptr, rows, last=False):
        """orb_copy_rows_to_peer: rows of the active slot's buffer into the same rows of a peer's buffer (copy engines,
        behind the slot's kernels); last=True closes the frame's batch."""
        self._ck(self.lib.orb_copy_rows_to_peer(self._h, BUFFERS[name], peer_ptr, int(rows[0]), int(rows[1]), int(bool(last))))

    def download(self, name, rows=None, out=None):
        bid = BUFFERS[name]
        y0, y1 = rows if rows else (0, self.H)
        if out is None:
            out = (np.empty((y1 - y0, self.W), np.uint32) if name == "rgba8"
                   else np.empty((y1 - y0, self.W, 2 if name == "dn_val" else 4), np.float32))
        self._ck(self.lib.orb_buffer_download_rows(self._h, bid, out.ctypes.data_as(C.c_void_p), y0, y1))
        return out

    def download_async(self, name, out, rows=None):
        """Pipelined read-back (orb_buffer_download_rows_async): `out` is a page-locked array that must stay
        alive and untouched until download_wait()."""
        y0, y1 = rows if rows else (0, self.H)
        elem = 4 if name == "rgba8" else (8 if name == "dn_val" else 16)
        need = (y1 - y0) * self.W * elem
        if not (isinstance(out, np.ndarray) and out.flags.c_contiguous and out.flags.writeable and out.nbytes == need):
            # the copy stream writes `need` bytes at out's address asynchronously: a strided view or a short array
            # would be overrun silently
            raise OrbError("download_async: `out` must be a writable C-contiguous array of exactly %d bytes "
                           "(rows %d..%d of %s)" % (need, y0, y1, name))
        self._ck(self.lib.orb_buffer_download_rows_async(self._h, BUFFERS[name], out.ctypes.data_as(C.c_void_p), y0, y1))
        self._dl_keep.append(out)  # alive until download_wait()
        return out

    def download_wait(self):
        self._ck(self.lib.orb_download_wait(self._h))
        self._dl_keep = []

    def upload(self, name, arr):
        a = np.ascontiguousarray(arr)
        self._ck(self.lib.orb_buffer_upload(self._h, BUFFERS[name], a.ctypes.data_as(C.c_void_p), a.nbytes))

    def hit_records(self):
        h = self.download("hit")
        return h[..., 0].copy().view(np.int32), h[..., 1:3].copy(), h[..., 3].copy()

    def counters(self):
        o = OrbCounters()
        self._ck(self.lib.orb_get_counters(self._h, C.byref(o)))
        return {n: int(getattr(o, n)) for n, _ in OrbCounters._fields_}

    def save_btc(self, name, path, avg_factor=1.0):
        """savebuffer(): raw float4 rows, bottom row first, every float divided by avg_factor."""
        a = self.download(name)
        (a / np.float32(avg_factor)).astype(np.float32).tofile(path)

    # ---- host-buffer entry points ----
    def ratio_estimator_host(self, ltc, sto_direct, sto_no_vis, out=None):
        """ltc_ratio_estimator.py:8 on host arrays (H2D, kernel, D2H inside the call)."""
        a, b, c = f32(ltc), f32(sto_direct), f32(sto_no_vis)
        if out is None:
            out = np.empty_like(a)
        self._ck(self.lib.orb_ratio_estimator_host(self._h, _fp(a), _fp(b), _fp(c), _fp(out), a.size))
        return out

    def trace_rays(self, origin, direction, brute_force=False):
        o, d = f32(origin).reshape(-1, 3), f32(direction).reshape(-1, 3)
        n = o.shape[0]
        prim = np.empty(n, np.int32)
        t, u, v = (np.empty(n, np.float32) for _ in range(3))
        self._ck(self.lib.orb_trace_rays_host(self._h, _fp(o), _fp(d), n, int(brute_force), prim.ctypes.data_as(_i),
                                              _fp(t), _fp(u), _fp(v)))
        return prim, t, u, v

    def trace_rays_no_alloc(self, d_origin, d_direction, n, d_prim, d_t, d_u, d_v):
        """orb_trace_rays_no_alloc: DEVICE pointers (ints) in and out, scratch in the library's static device array,
        no allocation; enqueued on stream() -- sync() before reading the outputs.  n <= NO_ALLOC_MAX_RAYS."""
        self._ck(self.lib.orb_trace_rays_no_alloc(self._h, d_origin, d_direction, int(n), d_prim, d_t, d_u, d_v))

    # ---- known-answer entry points ----
    def u_rng(self, frame, px, py, W, n):
        s = np.zeros(1, np.uint32)
        raw, flt = np.zeros(n, np.uint32), np.zeros(n, np.float32)
        self._ck(self.lib.orb_unit_rng(self._h, frame, px, py, W, n, s.ctypes.data_as(_u), raw.ctypes.data_as(_u),
                                       _fp(flt)))
        return int(s[0]), raw, flt

    def u_clip(self, abc, quad=False):
        a = f32(abc).reshape(9)
        n = np.zeros(1, np.int32)
        ff = np.zeros(1, np.float32)
        self._ck(self.lib.orb_unit_clip(self._h, _fp(a), int(quad), n.ctypes.data_as(_i), _fp(ff)))
        return int(n[0]), float(ff[0])

    def u_integrate_edge(self, a, b):
        a, b, o = f32(a), f32(b), np.zeros(1, np.float32)
        self._ck(self.lib.orb_unit_integrate_edge(self._h, _fp(a), _fp(b), _fp(o)))
        return float(o[0])

    def u_fetch_ltc(self, alpha, theta):
        o = np.zeros(10, np.float32)
        self._ck(self.lib.orb_unit_fetch_ltc(self._h, alpha, theta, _fp(o)))
        return o[:9].reshape(3, 3).copy(), float(o[9])

    def u_mat_inverse(self, m):
        m, o = f32(m).reshape(9), np.zeros(9, np.float32)
        self._ck(self.lib.orb_unit_mat_inverse(self._h, _fp(m), _fp(o)))
        return o.reshape(3, 3)

    def u_onb(self, n):
        n, o = f32(n), np.zeros(9, np.float32)
        self._ck(self.lib.orb_unit_onb(self._h, _fp(n), _fp(o)))
        return o.reshape(3, 3)

    def u_sample_direction(self, wo, u1, u2, base, alpha):
        wo, base, o = f32(wo), f32(base), np.zeros(4, np.float32)
        self._ck(self.lib.orb_unit_sample_direction(self._h, _fp(wo), u1, u2, _fp(base), alpha, _fp(o)))
        return o[:3].copy(), float(o[3])

    def u_evaluate(self, wi, wo, base, alpha):
        wi, wo, base, o = f32(wi), f32(wo), f32(base), np.zeros(3, np.float32)
        self._ck(self.lib.orb_unit_evaluate(self._h, _fp(wi), _fp(wo), _fp(base), alpha, _fp(o)))
        return o

    def u_pdf(self, wi, wo, base, alpha):
        wi, wo, base, o = f32(wi), f32(wo), f32(base), np.zeros(1, np.float32)
        self._ck(self.lib.orb_unit_pdf(self._h, _fp(wi), _fp(wo), _fp(base), alpha, _fp(o)))
        return float(o[0])

    def u_make_rgba(self, c):
        c, o = f32(c), np.zeros(1, np.uint32)
        self._ck(self.lib.orb_unit_make_rgba(self._h, _fp(c), o.ctypes.data_as(_u)))
        return int(o[0])

    def u_sample_tri(self, v9, u1, u2):
        v, o = f32(v9).reshape(9), np.zeros(3, np.float32)
        self._ck(self.lib.orb_unit_sample_tri(self._h, _fp(v), u1, u2, _fp(o)))
        return o

    def u_pdf_a2w(self, pdf, d2, c):
        o = np.zeros(1, np.float32)
        self._ck(self.lib.orb_unit_pdf_a2w(self._h, pdf, d2, c, _fp(o)))
        return float(o[0])

    def u_ltc_pixel(self, cam, p, n, diffuse, alpha):
        cam, p, n, diffuse, o = f32(cam), f32(p), f32(n), f32(diffuse), np.zeros(3, np.float32)
        self._ck(self.lib.orb_unit_ltc_pixel(self._h, _fp(cam), _fp(p), _fp(n), _fp(diffuse), alpha, _fp(o)))
        return o

/* orb200.h -- C-ABI of liborb200.so: the B200 (sm_100a) per-pixel hot path of the renderer.
 *
 * This library replaces what the reference reaches through the OWL host API in
 * include/viewer.hpp (context/module :177-178, buffers :181-225, launch params :271-364,
 * geometry + GAS/IAS build :375-517, miss/raygen programs :522-543, owlLaunch2D :565,
 * owlBufferGetPointer :701) together with the three OptiX device programs it launches
 * (cuda_src/deviceCode.cu:59-175 rayGen; cuda_include/hit_miss.cuh:14-63 triangleMeshCH, miss).
 * Plain pointers and sizes only; no C++ or torch types.  All `const T*` inputs are HOST
 * pointers borrowed for the duration of the call unless a name says `_dev`.  Every function
 * returns 0 on success or an OrbStatus code; orb_last_error() gives the text for the calling
 * thread.  A context is single-threaded (externally synchronised) and owns one CUDA stream.
 * There is no CPU fallback: without a CUDA device orb_create() fails.
 */
#ifndef ORB200_H
#define ORB200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORB_API_VERSION 2

typedef enum OrbStatus {
    ORB_OK = 0,
    ORB_ERR_INVALID = 1,   /* bad argument / call order */
    ORB_ERR_CUDA = 2,      /* a CUDA call failed; text in orb_last_error() */
    ORB_ERR_NO_DEVICE = 3, /* no usable CUDA device */
    ORB_ERR_NOMEM = 4
} OrbStatus;

/* RendererType, cuda_include/common.cuh:17-29 (same values). */
typedef enum OrbRendererType {
    ORB_DIFFUSE = 0,
    ORB_ALPHA = 1,
    ORB_NORMALS = 2,
    ORB_SHADE_NORMALS = 3,
    ORB_POSITION = 4,
    ORB_MASK = 5,
    ORB_MATERIAL_ID = 6,
    ORB_LTC_BASELINE = 7,
    ORB_RATIO = 8,
    ORB_PATH = 9
} OrbRendererType;

/* Frame buffers.  0..12 are the 13 float4 screen buffers of LaunchParams
 * (cuda_include/common.cuh:84-100) in declaration order, row 0 = bottom image row, w = 1 as
 * rayGen writes them (cuda_src/deviceCode.cu:158-168).  The rest are this library's own. */
typedef enum OrbBufferId {
    ORB_BUF_POSITION = 0,
    ORB_BUF_NORMAL = 1,
    ORB_BUF_UV = 2,
    ORB_BUF_ALBEDO = 3,
    ORB_BUF_ALPHA = 4,
    ORB_BUF_MATERIAL_ID = 5,
    ORB_BUF_BOUNCE0 = 6,
    ORB_BUF_BOUNCE1 = 7,
    ORB_BUF_BOUNCE2 = 8,
    ORB_BUF_LTC = 9,
    ORB_BUF_STO_DIRECT = 10,
    ORB_BUF_STO_NO_VIS = 11,
    ORB_BUF_ACCUM = 12,
    ORB_BUF_RGBA8 = 13,        /* uint32 per pixel: RayGenData.frameBuffer, owl::make_rgba        */
    ORB_BUF_HIT = 14,          /* int32 prim, float u, float v, float t per pixel (primary hit)    */
    ORB_BUF_STO_DIRECT_DN = 15, /* float4: denoised sto_direct  (stodirectDenoise.exr)             */
    ORB_BUF_STO_NO_VIS_DN = 16, /* float4: denoised sto_no_vis  (stoNoVisDenoise.exr)              */
    ORB_BUF_RATIO_FINAL = 17,  /* float4: ltc * (stoDirect / stoNoVis) (final_ratio_estimator.exr) */
    ORB_BUF_DN_GUIDE = 18,     /* float4: denoise guide (n.xyz, camera distance; -1e30 on lights / misses)       */
    ORB_BUF_DN_VAL = 19,       /* float2 (8 B/px): packed (sto_direct.x, sto_no_vis.x), the filter's input      */
    ORB_BUF_COUNT = 20
} OrbBufferId;

/* TriLight, cuda_include/common.cuh:64-72 -- 76 bytes, same field order. */
typedef struct OrbTriLight {
    float v1[3], v2[3], v3[3];
    float cg[3];
    float normal[3];
    float emit[3];
    float area;
} OrbTriLight;

/* One triangle mesh = one OWL geometry + GAS of the reference (TriangleMeshData,
 * cuda_include/common.cuh:130-151; upload include/viewer.hpp:421-501). */
typedef struct OrbMeshDesc {
    const float* vertex;   /* n_vertices * 3 */
    const float* normal;   /* n_vertices * 3 */
    const float* texcoord; /* n_vertices * 2 */
    int32_t n_vertices;
    const int32_t* index; /* n_triangles * 3, indices into this mesh's vertices */
    int32_t n_triangles;
    float diffuse[3];
    float alpha; /* clamped to [0.01, 1] at hit time, hit_miss.cuh:45-46 */
    float emit[3];
    int32_t is_light;
    uint32_t material_id;
    const uint8_t* diffuse_tex_rgba8; /* NULL = no texture; nearest / clamp, viewer.hpp:441-447 */
    int32_t tex_w, tex_h;
} OrbMeshDesc;

typedef struct OrbBvhOptions {
    int32_t max_leaf_tris; /* 0 = default (4) */
    int32_t builder;       /* ORB_BVH_SAH (default) or ORB_BVH_LBVH; both run on the GPU */
    int32_t reserved[2];
} OrbBvhOptions;
#define ORB_BVH_SAH 0  /* breadth-first binned surface-area heuristic, 16 bins x 3 axes per node */
#define ORB_BVH_LBVH 1 /* Morton codes + Karras hierarchy: ~10x faster to build, ~25 % more node visits per ray */

typedef struct OrbRenderConfig {
    int32_t renderer_type; /* OrbRendererType */
    int32_t accum_id;      /* LaunchParams.accumId */
    int32_t mode;          /* 0 literal (committed source), 1 intended (patches P1-P3) */
    int32_t shadow_spp;    /* RATIO: shadow samples per pixel; literal fixes 4 (deviceCode.cu:128) */
    int32_t max_depth;     /* PATH: bounces; literal fixes 1 (deviceCode.cu:151) */
    int32_t y0, y1;        /* rows [y0, y1) rendered by this context (tile partition); 0,0 = all */
    int32_t flags;         /* ORB_RENDER_* */
} OrbRenderConfig;
#define ORB_RENDER_FUSED_SHADOW 2 /* PATH: every bounce as one per-path kernel that traces its ray inline instead of
                                     through the ray stream (the form literal mode uses; a cross-check in the tests) */
#define ORB_RENDER_PATH_WAVEFRONT 4 /* PATH: one generate / trace / shade wavefront per bounce instead of the
                                       persistent all-bounces kernel (kept for A/B measurement) */
#define ORB_RENDER_COUNT_TRAVERSAL 8 /* RATIO / PATH: run the instrumented twin of the shadow-ray / path kernel; node
                                        records fetched / triangle records tested land in OrbCounters.nodes_visited
                                        / tris_tested (same results, slower; bench.py's roofline uses it once) */
#define ORB_RENDER_CHECKED 32 /* count what north_star wants reported with a frame: OrbCounters.hit_ties (primary hits within
                                 1e-5, barycentric, of an edge of their triangle: another test may report the neighbour)
                                 and OrbCounters.ltc_near_decision (LTC / RATIO pixels within 1e-4 of a clip / back-face /
                                 horizon decision, the "case flip" candidates).  One extra kernel; same frame. */
#define ORB_RENDER_NO_COUNT 1 /* rays of this launch are not added to the *_total counters (halo rows
                                 a tile owner renders redundantly for the denoise filter) */

typedef struct OrbDenoiseConfig {
    int32_t iterations; /* a-trous passes, step 1,2,4,...; default 5 */
    float sigma_normal; /* edge stop on normal (power of the dot product), default 32 */
    float sigma_position; /* edge stop on camera-distance difference, per unit of tap spacing; default 0.25 */
    float sigma_colour; /* edge stop on colour difference; <= 0 disables (default) */
} OrbDenoiseConfig;

typedef struct OrbCounters {
    uint64_t rays_primary;
    uint64_t rays_shadow;
    uint64_t rays_bounce;
    uint64_t kernel_launches; /* kernels launched by this context since creation */
    uint64_t nan_pixels;      /* pixels of the last frame whose colour had a NaN channel */
    uint64_t bvh_nodes;
    uint64_t bvh_triangles;
    uint64_t bvh_height;     /* inner-node levels of the hierarchy (sizes the traversal stack) */
    uint64_t active_pixels;  /* last frame: non-light hit pixels (length of the active-pixel queue) */
    uint64_t nodes_visited;  /* last orb_trace_rays_host call (instrumented traversal) */
    uint64_t tris_tested;
    uint64_t rays_shadow_total; /* since orb_create, summed on the device at the end of every frame */
    uint64_t rays_bounce_total;
    uint64_t active_pixels_total;
    uint64_t bvh_build_us; /* device time of the last orb_scene_commit's hierarchy build (CUDA events) */
    uint64_t bvh_refit_us; /* device time of the last orb_scene_refit (CUDA events) */
    uint64_t hit_ties;          /* last frame rendered with ORB_RENDER_CHECKED, else 0 */
    uint64_t ltc_near_decision; /* last frame rendered with ORB_RENDER_CHECKED, else 0 */
} OrbCounters;

/* Stages of a frame, for orb_get_stage_times(): one entry per kernel family. */
typedef enum OrbStage {
    ORB_STAGE_PRIMARY = 0,      /* camera rays + closest hit + G-buffer write + queue compaction */
    ORB_STAGE_LTC = 1,          /* LTC area-light sum over the active-pixel queue                */
    ORB_STAGE_SHADOW = 2,       /* ratio estimator: shadow-ray generation + resolve (traversal: TRACE) */
    ORB_STAGE_PATH = 3,         /* all bounce kernels of a PATH frame                            */
    ORB_STAGE_ACCUMULATE = 4,   /* accum += colour, RGBA8                                        */
    ORB_STAGE_DENOISE_PACK = 5, /* guide / value packing                                         */
    ORB_STAGE_DENOISE = 6,      /* a-trous iterations                                            */
    ORB_STAGE_COMBINE = 7,      /* ratio combine                                                 */
    ORB_STAGE_LTC_GBUFFER = 8,  /* stand-alone LTC pass over a saved G-buffer                    */
    ORB_STAGE_TRACE = 9,        /* RATIO: the shadow-ray stream traversal kernel alone           */
    ORB_STAGE_HALO = 10,        /* tile partition: flag kernels of the hand-off (time = waiting for the peers) */
    ORB_STAGE_COUNT = 11
} OrbStage;

typedef struct OrbStageTimes {
    double ms[ORB_STAGE_COUNT];        /* summed CUDA-event time of the stage's launches */
    uint64_t launches[ORB_STAGE_COUNT]; /* kernels launched in the stage                  */
} OrbStageTimes;

typedef struct OrbStageSpan {
    int32_t stage, slot; /* OrbStage, frame slot */
    float t0_ms, t1_ms;  /* CUDA-event times relative to the first recorded span */
} OrbStageSpan;

typedef struct OrbContext OrbContext;

/* ---- context ---- replaces owlContextCreate/ModuleCreate, viewer.hpp:177-178 */
int orb_api_version(void);
const char* orb_last_error(void);
int orb_create(int device, OrbContext** out);
void orb_destroy(OrbContext* ctx);
void* orb_stream(OrbContext* ctx); /* cudaStream_t the context launches on */
/* Launch on a caller-owned cudaStream_t instead (e.g. the stream a host framework times with). */
int orb_set_stream(OrbContext* ctx, void* cuda_stream);
int orb_sync(OrbContext* ctx);
/* Per-stage timing with cudaEvents recorded on the context's stream around each stage (off by
 * default).  orb_get_stage_times() synchronises, returns the sums since the previous call and
 * resets them.  New in this library (the reference has no timing code at all). */
int orb_set_profiling(OrbContext* ctx, int enable);
int orb_get_stage_times(OrbContext* ctx, OrbStageTimes* out);
/* the spans recorded since the last orb_get_stage_times, as a timeline (several frame slots overlap); synchronises */
int orb_get_stage_spans(OrbContext* ctx, OrbStageSpan* out, int32_t max, int32_t* n);

/* ---- scene ---- replaces owlGeom*, owlTriangles*, owlGroupBuildAccel, viewer.hpp:375-517 */
int orb_scene_clear(OrbContext* ctx);
int orb_scene_add_mesh(OrbContext* ctx, const OrbMeshDesc* mesh); /* returns mesh index via *mesh order */
int orb_scene_set_lights(OrbContext* ctx, const OrbTriLight* lights, int32_t n); /* viewer.hpp:336-338 */
/* Three RGBA32F tables of w*h texels, rows = alpha, columns = theta (viewer.hpp:321-331).
 * quantise_weights = 1 filters with 8-bit fractional weights like the CUDA texture unit. */
int orb_scene_set_ltc_lut(OrbContext* ctx, const float* m1, const float* m2, const float* m3, int32_t w,
                          int32_t h, int32_t quantise_weights);
int orb_scene_commit(OrbContext* ctx, const OrbBvhOptions* opt); /* builds the BVH on the GPU */
/* Animated geometry (moving lights or meshes) on a COMMITTED scene.  The reference can only rebuild:
 * owlTrianglesSetVertices + owlGroupBuildAccel of the GAS and the IAS again, viewer.hpp:494-517.
 * orb_scene_update_mesh replaces the corner positions and normals of mesh `mesh_index` (the order
 * of the orb_scene_add_mesh calls); arrays as in OrbMeshDesc, the triangle count must not change.
 * orb_scene_refit then recomputes the triangle records and every BVH box on the GPU, topology
 * unchanged; until it ran, orb_render* and orb_trace_rays_host fail with ORB_ERR_INVALID.
 * TriLights that moved are passed again through orb_scene_set_lights. */
int orb_scene_update_mesh(OrbContext* ctx, int32_t mesh_index, const float* vertex, const float* normal,
                          int32_t n_vertices, const int32_t* index, int32_t n_triangles);
int orb_scene_refit(OrbContext* ctx);

/* ---- camera / frame ---- replaces owlParamsSet3f camera.*, viewer.hpp:647-650, and resize :571-612 */
int orb_set_camera(OrbContext* ctx, const float pos[3], const float dir_00[3], const float dir_du[3],
                   const float dir_dv[3]);
int orb_resize(OrbContext* ctx, int32_t width, int32_t height);

/* ---- launch ---- replaces owlLaunch2D(rayGen, W, H, launchParams), viewer.hpp:553-568.
 * Blocking, like owlLaunch2D.  orb_render_async() enqueues on orb_stream() and returns. */
int orb_render(OrbContext* ctx, const OrbRenderConfig* cfg);
int orb_render_async(OrbContext* ctx, const OrbRenderConfig* cfg);
/* Frames in flight: 1 (default) to 8.  With n > 1, consecutive orb_render_async calls cycle through n sets of frame
 * buffers and n streams, so the kernels of frame i+1 run beside the tail of frame i's persistent traversal kernels
 * (the reference renders strictly one frame at a time: owlLaunch2D blocks, viewer.hpp:565).  ORB_BUF_ACCUM and
 * ORB_BUF_RGBA8 are shared and accumulate in call order; every other buffer exists once per slot.  All calls that
 * follow an orb_render_async -- denoise, combine, downloads, orb_buffer_device_ptr, orb_stream -- refer to THAT
 * frame's slot until the next orb_render_async; orb_sync waits for every frame in flight.  Re-allocates the frame. */
int orb_set_frame_slots(OrbContext* ctx, int32_t n_slots);
/* With several frame slots: resident CTAs per SM the persistent traversal kernels (ray stream, PATH) may take, 0 = all
 * that fit (default).  A full persistent grid keeps the other frames' kernels off the SMs until its last ray ends; on
 * the narrow bands of a tile partition (few rays per lane) a cap of 2 - 4 with 4 - 6 slots is 20 - 35 % faster. */
int orb_set_persistent_ctas_per_sm(OrbContext* ctx, int32_t n);
int32_t orb_frame_slot(OrbContext* ctx);  /* slot of the frame the calls currently refer to */
int32_t orb_frame_slots(OrbContext* ctx);
void* orb_slot_stream(OrbContext* ctx, int32_t slot); /* cudaStream_t of a frame slot (to order a caller's stream behind it) */

/* ---- stand-alone passes over RESIDENT frame buffers (rows [y0,y1), 0,0 = all) ---- */
/* ltcDirectLighingBaseline over the G-buffer (position, normal, albedo, alpha) -> ORB_BUF_LTC;
 * pixels whose normal is (0,0,0) give 0. */
int orb_ltc_gbuffer_pass(OrbContext* ctx, int32_t y0, int32_t y1);
/* a-trous edge-avoiding filter: STO_DIRECT -> STO_DIRECT_DN, STO_NO_VIS -> STO_NO_VIS_DN */
int orb_denoise(OrbContext* ctx, const OrbDenoiseConfig* cfg, int32_t y0, int32_t y1);
/* The two halves of orb_denoise, for tile partitions: a band's owner packs ITS rows into ORB_BUF_DN_GUIDE /
 * ORB_BUF_DN_VAL, the packed rows are exchanged between GPUs (24 B/px instead of re-rendering the filter's
 * 2 * (2^iterations - 1) halo rows), and the filter then runs over the band. */
int orb_denoise_pack(OrbContext* ctx, int32_t y0, int32_t y1);
int orb_denoise_filter(OrbContext* ctx, const OrbDenoiseConfig* cfg, int32_t y0, int32_t y1);
/* ---- tile hand-off of a row-band partition (one context per GPU; the reference has one GPU, viewer.hpp:177) ----
 * Each rank maps its peers' packed denoise buffers once; orb_denoise_band then packs a band and STORES the rows its
 * neighbours' filters read straight into their buffers over NVLink from inside the pack kernel, ordered by epoch
 * flags (no NCCL call, no host synchronisation).  Set-up per (peer, frame slot): the owner calls orb_halo_local,
 * hands the three pointers over -- as they are to a context of the same process, through orb_ipc_export /
 * orb_ipc_open to another process -- and the peer calls orb_halo_connect with them.  Every rank must connect every
 * other rank and call orb_denoise_band exactly once per frame (y0 == y1 for an empty band). */
typedef struct OrbHaloSend {
    int32_t peer_rank; /* rank whose filter reads these rows */
    int32_t y0, y1;    /* rows of THIS band, [y0, y1) */
} OrbHaloSend;
int orb_halo_local(OrbContext* ctx, int32_t slot, void** guide_dev, void** val_dev, void** flags_dev);
int orb_ipc_export(OrbContext* ctx, void* dev_ptr, unsigned char handle[64]);
int orb_ipc_open(OrbContext* ctx, const unsigned char handle[64], void** dev_ptr_out);
int orb_ipc_close(OrbContext* ctx, void* dev_ptr);
int orb_halo_connect(OrbContext* ctx, int32_t my_rank, int32_t peer_rank, int32_t slot, void* guide_dev, void* val_dev,
                     void* flags_dev);
int orb_halo_disconnect(OrbContext* ctx);
int orb_halo_status(OrbContext* ctx, int32_t* status); /* 0 = ok; > 0 = a flag wait timed out (peer lost) */
int orb_denoise_band(OrbContext* ctx, const OrbDenoiseConfig* cfg, int32_t y0, int32_t y1, const OrbHaloSend* sends,
                     int32_t n_sends, const int32_t* recv_ranks, int32_t n_recv);
/* Active (non-light hit) pixels per row of the last rendered frame -> host_counts[height]; rows this context
 * did not render report what its flags hold.  Used to balance row bands by cost. */
int orb_row_active_counts(OrbContext* ctx, int32_t* host_counts);
/* ltc_ratio_estimator.py:8: RATIO_FINAL = LTC * (STO_DIRECT[_DN] / STO_NO_VIS[_DN]) */
/* orb_set_fused_combine(1): the filter's final pass (orb_denoise / orb_denoise_filter / orb_denoise_band) also writes
 * RATIO_FINAL = LTC * (STO_DIRECT_DN / STO_NO_VIS_DN) over the rows it filters -- the same values, bit for bit, as a
 * following orb_ratio_combine(use_denoised = 1), which the caller then leaves out (one launch and 48 B/px of reads). */
int orb_set_fused_combine(OrbContext* ctx, int32_t enable);
int orb_ratio_combine(OrbContext* ctx, int32_t use_denoised, int32_t y0, int32_t y1);

/* ---- buffers ---- replaces owlBufferGetPointer + cudaMemcpy, viewer.hpp:695-713 */
size_t orb_buffer_bytes(OrbContext* ctx, int32_t buffer_id);
void* orb_buffer_device_ptr(OrbContext* ctx, int32_t buffer_id);
/* Buffer `buffer_id` of frame slot `slot` (ORB_BUF_ACCUM / ORB_BUF_RGBA8 are one buffer for every slot). */
void* orb_slot_buffer_device_ptr(OrbContext* ctx, int32_t slot, int32_t buffer_id);
/* One-sided frame assembly of a row-band partition (the reference has one GPU, viewer.hpp:177): rows [y0, y1) of the
 * ACTIVE slot's buffer are copied into the same rows of a peer's buffer -- peer_buffer = that peer's
 * orb_slot_buffer_device_ptr for the same slot and id, as it is (a context of this process) or through orb_ipc_export /
 * orb_ipc_open (another process) -- by the copy engines over NVLink, on a stream of its own behind everything the
 * slot has enqueued: no kernel, no rendezvous between ranks.  last != 0 closes the frame's batch of copies: the slot's
 * stream waits for them, so orb_sync() covers them and later frames do not overwrite rows still being read.  After
 * every rank's copies (a cross-rank barrier behind orb_sync) every rank's buffer holds the whole frame.  Pointers
 * obtained before an orb_resize / orb_set_frame_slots of the peer are stale. */
int orb_copy_rows_to_peer(OrbContext* ctx, int32_t buffer_id, void* peer_buffer, int32_t y0, int32_t y1, int32_t last);
int orb_buffer_download(OrbContext* ctx, int32_t buffer_id, void* host_dst, size_t bytes);
int orb_buffer_download_rows(OrbContext* ctx, int32_t buffer_id, void* host_dst, int32_t y0, int32_t y1);
/* Pipelined read-back for an interactive loop (viewer.hpp:695-713 reads the frame back every frame): returns at
 * once; the rows are snapshotted on the frame stream into one of two device slots per buffer and copied to
 * host_dst on a copy stream while later frames render.  host_dst should be page-locked (a pageable destination
 * makes the copy synchronous) and must stay valid until orb_download_wait, which blocks until every
 * outstanding asynchronous download of this context has landed.  At most two downloads per buffer in flight
 * keep distinct slots; a third waits for the first on the device. */
int orb_buffer_download_rows_async(OrbContext* ctx, int32_t buffer_id, void* host_dst, int32_t y0, int32_t y1);
int orb_download_wait(OrbContext* ctx);
int orb_buffer_upload(OrbContext* ctx, int32_t buffer_id, const void* host_src, size_t bytes);
int orb_get_counters(OrbContext* ctx, OrbCounters* out);

/* ---- host-buffer entry point of BASELINE config 1 (ltc_ratio_estimator.py:4-9):
 * out = ltc * (sto_direct / sto_no_vis) element-wise on n floats; copies in, runs the kernel,
 * copies out. */
int orb_ratio_estimator_host(OrbContext* ctx, const float* ltc, const float* sto_direct, const float* sto_no_vis,
                             float* out, size_t n);

/* ---- ray casting over the committed scene (owl::traceRay semantics, tmin=0, tmax=1e30):
 * n rays, origin/direction as n*3 floats; outputs prim (global id, -1 miss), t, u, v.
 * brute_force: 0 = single-ray BVH2 traversal, 1 = every triangle (hit-id parity checks), 2 = the ray-stream
 * kernel the RATIO / PATH frames use, instrumented: OrbCounters.nodes_visited /
 * tris_tested. */
int orb_trace_rays_host(OrbContext* ctx, const float* origin, const float* direction, size_t n, int32_t brute_force,
                        int32_t* prim, float* t, float* u, float* v);

/* ---- allocation-free ray casting.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops).
 * Same rays -> closest hits as orb_trace_rays_host(brute_force = 2) (the ray-stream kernel the RATIO / PATH
 * frames use), but origin / direction (n*3 floats) and prim / t / u / v (n each) are DEVICE pointers the caller
 * owns, and the kernel's scratch (ray stream, hit records, claim cursor) lives in one file-scope device array
 * of 256 + 48 * ORB_NO_ALLOC_MAX_RAYS bytes (3 MiB) that every GPU loading the library holds until exit.
 * Calls cudaMalloc / cudaFree / no allocating routine; n > ORB_NO_ALLOC_MAX_RAYS is refused with
 * ORB_ERR_INVALID.  Enqueues three kernels on orb_stream() and returns (orb_sync() or the caller's own
 * stream ordering before reading the outputs).  Every call on a device shares the scratch array: calls that
 * overlap on the GPU (two contexts, two streams) race -- serialise them. */
#define ORB_NO_ALLOC_MAX_RAYS 65536u
int orb_trace_rays_no_alloc(OrbContext* ctx, const float* d_origin, const float* d_direction, uint32_t n,
                            int32_t* d_prim, float* d_t, float* d_u, float* d_v);

/* ---- known-answer entry points: run ONE device function in a one-thread kernel ---- */
int orb_unit_rng(OrbContext* ctx, int frame_id, uint32_t px, uint32_t py, uint32_t W, int n, uint32_t* state0,
                 uint32_t* raw, float* flt);
int orb_unit_clip(OrbContext* ctx, const float* abc9, int quad, int* n_out, float* form_factor);
int orb_unit_integrate_edge(OrbContext* ctx, const float* a, const float* b, float* out);
int orb_unit_fetch_ltc(OrbContext* ctx, float alpha, float theta, float* m9_amp);
int orb_unit_mat_inverse(OrbContext* ctx, const float* m9, float* out9);
int orb_unit_onb(OrbContext* ctx, const float* n, float* rows9);
int orb_unit_sample_direction(OrbContext* ctx, const float* wo, float u1, float u2, const float* base, float alpha,
                              float* wi_pdf4);
int orb_unit_evaluate(OrbContext* ctx, const float* wi, const float* wo, const float* base, float alpha,
                      float* out3);
int orb_unit_pdf(OrbContext* ctx, const float* wi, const float* wo, const float* base, float alpha, float* out);
int orb_unit_make_rgba(OrbContext* ctx, const float* c3, uint32_t* out);
int orb_unit_sample_tri(OrbContext* ctx, const float* v9, float u1, float u2, float* out3);
int orb_unit_pdf_a2w(OrbContext* ctx, float pdf, float dist2, float cos_theta, float* out);
/* ltcDirectLighingBaseline for one shading point against the context's lights and LUT */
int orb_unit_ltc_pixel(OrbContext* ctx, const float* cam3, const float* p3, const float* n3, const float* diffuse3,
                       float alpha, float* out3);

#ifdef __cplusplus
}
#endif
#endif /* ORB200_H */

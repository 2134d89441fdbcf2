// orb200.cu -- liborb200.so: frame kernels (sm_100a) and the C-ABI declared in include/orb200.h.
//
// One frame of the reference's raygen megakernel (cuda_src/deviceCode.cu:59-175) becomes a
// short wavefront of kernels over buffers that stay resident in HBM:
//
//   k_primary        camera ray -> BVH closest hit -> closest-hit shading -> 16-B hit record,
//                    six float4 G-buffers, per-pixel flags; non-light hit pixels are appended to
//                    the ACTIVE-PIXEL QUEUE by warp ballot (one atomic per warp)
//   k_ltc            queue -> LTC area-light sum (lights + LUT staged in shared memory)
//   k_shadow_gen -> k_trace_stream_spec -> k_shadow_resolve   queue -> n closest-hit shadow rays per pixel as a
//                    RAY STREAM traced by persistent warps that replace finished rays -> sto_direct / sto_no_vis
//   k_path_persistent  queue -> all bounces of a PATH frame, a lane owns a path (k_path_bounce: literal mode)
//   k_accumulate     colour + running sum -> accum, RGBA8, NaN count
//   k_denoise_*      edge-avoiding a-trous on the packed (direct, no_vis) image
//   k_ratio_combine  ltc * (direct / no_vis)
//   k_refit_*        (bvh_refit.cuh) triangle records and BVH boxes after geometry moved, topology kept
//
// Queue kernels are persistent (grid = SMs x resident CTAs, grid-stride over a DEVICE-side
// count), so a frame needs no host synchronisation.  No kernel is a dense contraction: there
// is no tensor-core work (BASELINE.json north_star).  There is no CPU fallback.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/orb200.h"
#include "bvh_build.cuh"
#include "bvh_sah.cuh"
#include "bvh_refit.cuh"
#include "bvhq.cuh"
#include "pixel.cuh"
#include "halo.cuh"

using namespace orb;

// ---------------------------------------------------------------------------------------------
// error handling
// ---------------------------------------------------------------------------------------------
static thread_local std::string g_err;
static int fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}
#define CK(x)                                                                                      \
    do {                                                                                           \
        cudaError_t e_ = (x);                                                                      \
        if (e_ != cudaSuccess)                                                                     \
            return fail(ORB_ERR_CUDA, std::string(#x) + ": " + cudaGetErrorString(e_));            \
    } while (0)
#define NEED(c, msg) \
    do {             \
        if (!(c)) return fail(ORB_ERR_INVALID, msg); \
    } while (0)

// ---------------------------------------------------------------------------------------------
// device-side frame description
// ---------------------------------------------------------------------------------------------
#define FLAG_HIT 1u
#define FLAG_LIGHT 2u

struct DevCounters {
    unsigned queue_n;       // active (non-light hit) pixels of this frame
    unsigned path_n[2];     // live paths, ping-pong
    unsigned nan_pixels;
    unsigned long long rays_shadow, rays_bounce;
    unsigned long long nodes_visited, tris_tested;  // filled by the instrumented trace only
    unsigned hit_ties, ltc_near_decision;           // filled by a checked render only (k_check_frame)
    // ---- everything above is cleared at the start of a frame; the totals below are not ----
    unsigned long long tot_shadow, tot_bounce, tot_active;
};
#define DEVCOUNTERS_FRAME_BYTES offsetof(DevCounters, tot_shadow)

struct FrameView {
    int W, H, y0, y1;
    float4 *position, *normal, *uv, *albedo, *alpha, *material_id;
    float4 *ltc, *sto_direct, *sto_no_vis, *accum, *hit, *colour;
    uint32_t* rgba8;
    uint32_t* flags;
    uint32_t* queue;
    DevCounters* cnt;
};

#define TRACE_BLOCK 128

// Dynamic shared memory of the traversal kernels: the per-lane stack columns (scene.cuh HybridStack / StridedStack).
extern __shared__ int s_dyn[];

// SD (template parameter of every traversal kernel) selects the stack:
//   SD > 0   HybridStack: SD entries per lane in shared memory (SD * 8 bytes per thread), the rest local
//   SD == 0  thread-local memory only (no shared memory)
//   SD == -1 StridedStack: full-height shared-memory columns, for hierarchies deeper than ORB_LOCAL_STACK_MAX
template <int SD> struct StackOf { typedef HybridStack<SD> type; };
template <> struct StackOf<-1> { typedef StridedStack type; };
template <int SD>
__device__ __forceinline__ void stack_init(HybridStack<SD>& st, int) {
    st.s = reinterpret_cast<unsigned long long*>(s_dyn) + threadIdx.x;
    st.stride = (int)blockDim.x;
}
__device__ __forceinline__ void stack_init(StridedStack& st, int height) {
    st.r = s_dyn + threadIdx.x;
    st.t = reinterpret_cast<float*>(s_dyn + height * blockDim.x) + threadIdx.x;
    st.stride = (int)blockDim.x;
}

struct SmemTracer {  // single-ray closest hit through full-height shared-memory columns (cold paths: literal / fused PATH)
    const SceneView& s;
    int* stk;
    float* stk_t;
    unsigned* rays;
    __device__ Hit operator()(float3 o, float3 d) const {
        ++*rays;
        return trace_closest(s, o, d, stk, stk_t, TRACE_BLOCK);
    }
};

// ---------------------------------------------------------------------------------------------
// k_primary
// ---------------------------------------------------------------------------------------------
// A warp covers an 8x4 pixel tile (coherent rays, 128-byte row segments on the float4 stores);
// a 128-thread block covers 16x8.  QN: inner nodes are the 32-byte quantised ones (bvhq.cuh).
template <int SD, bool QN>
__global__ void __launch_bounds__(TRACE_BLOCK)
k_primary(SceneView s, FrameView f, Camera cam, int frame_id, int renderer, int height, const QNode* __restrict__ qn, QGrid qg) {
    typename StackOf<SD>::type st;
    stack_init(st, height);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int px = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
    const int py = f.y0 + blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
    const bool inside = px < f.W && py < f.y1;
    bool active = false;
    size_t ofs = 0;
    if (inside) {
        ofs = (size_t)py * f.W + px;
        float3 o, d;
        primary_ray(cam, frame_id, px, py, f.W, f.H, o, d);
        const Hit h = QN ? trace_closest_q(s, qn, qg, o, d, st) : trace_closest(s, o, d, st);
        const Surface si = shade_hit(s, h.prim, h.u, h.v);
        st_stream(f.hit + ofs, make_float4(__int_as_float(h.prim), h.u, h.v, h.t));
        st_stream(f.position + ofs, make_float4(si.p.x, si.p.y, si.p.z, 1.f));
        st_stream(f.normal + ofs, make_float4(si.n.x, si.n.y, si.n.z, 1.f));
        st_stream(f.albedo + ofs, make_float4(si.diffuse.x, si.diffuse.y, si.diffuse.z, 1.f));
        st_stream(f.alpha + ofs, make_float4(si.alpha, 0.f, 0.f, 1.f));
        st_stream(f.uv + ofs, make_float4(si.u, si.v, si.diffuse.z, 1.f));
        st_stream(f.material_id + ofs, make_float4((float)si.material_id, 0.f, 0.f, 1.f));
        f.flags[ofs] = (si.hit ? FLAG_HIT : 0u) | (si.is_light ? FLAG_LIGHT : 0u);
        // everything rayGen's mode switch decides without further rays (deviceCode.cu:91-124,146-149)
        float3 c = v3(0.f);
        bool staged = false;  // colour comes from a later stage
        if (!si.hit) c = si.n;
        else if (renderer == ORB_MASK) c = v3(1.f);
        else if (renderer == ORB_POSITION) c = si.p;
        else if (renderer == ORB_DIFFUSE) c = si.diffuse;
        else if (renderer == ORB_ALPHA) c = v3(si.alpha);
        else if (renderer == ORB_NORMALS) c = si.n;
        else if (renderer == ORB_SHADE_NORMALS) c = v3(0.f);  // n_shad is never written
        else if (renderer == ORB_MATERIAL_ID) c = v3((float)si.material_id);
        else if (renderer == ORB_LTC_BASELINE || renderer == ORB_RATIO || renderer == ORB_PATH) {
            if (si.is_light) {
                c = si.emit;
                if (renderer == ORB_RATIO) {
                    const float4 e = make_float4(c.x, c.y, c.z, 1.f);
                    st_stream(f.ltc + ofs, e);
                    st_stream(f.sto_direct + ofs, e);
                    st_stream(f.sto_no_vis + ofs, e);
                }
            } else {
                staged = true;
            }
        } else {
            c = v3(1.f, 0.f, 0.f);
        }
        if (!staged) st_stream(f.colour + ofs, make_float4(c.x, c.y, c.z, 1.f));
        active = staged;
    }
    // warp-ballot compaction into the active-pixel queue
    const unsigned m = __ballot_sync(0xffffffffu, active);
    if (m) {
        unsigned base = 0;
        if (lane == 0) base = atomicAdd(&f.cnt->queue_n, (unsigned)__popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        ORB_DEVICE_ASSERT(!active || base + __popc(m & ((1u << lane) - 1u)) < (unsigned)((f.y1 - f.y0) * f.W));
        if (active) f.queue[base + __popc(m & ((1u << lane) - 1u))] = (uint32_t)ofs;
    }
}

// ---------------------------------------------------------------------------------------------
// k_ltc  (FROM_QUEUE: in-frame over the active queue; else stand-alone over a saved G-buffer)
// ---------------------------------------------------------------------------------------------
#define LTC_BLOCK 128
#define LTC_SMEM_LIGHTS 128        // 9.5 KB of lights in shared memory; more are read from global
#define LTC_SMEM_LUT_TEXELS 3 * 64  // the reference's 8x8 x 3 tables (3 KB); larger LUTs stay global

template <bool FROM_QUEUE>
__global__ void __launch_bounds__(LTC_BLOCK)
k_ltc(SceneView s, FrameView f, float3 cam_pos, float4* __restrict__ out, float4* __restrict__ out2) {
    __shared__ TriLight sh_lights[LTC_SMEM_LIGHTS];
    __shared__ float4 sh_lut[LTC_SMEM_LUT_TEXELS];
    const bool lights_sm = s.n_lights <= LTC_SMEM_LIGHTS;
    const int lut_texels = 3 * s.lut.w * s.lut.h;
    const bool lut_sm = lut_texels <= LTC_SMEM_LUT_TEXELS;
    if (lights_sm) {
        const float* src = reinterpret_cast<const float*>(s.lights);
        float* dst = reinterpret_cast<float*>(sh_lights);
        for (int i = threadIdx.x; i < s.n_lights * 19; i += LTC_BLOCK) dst[i] = src[i];
    }
    if (lut_sm)
        for (int i = threadIdx.x; i < lut_texels; i += LTC_BLOCK) sh_lut[i] = s.lut.data[i];
    __syncthreads();
    LtcLut lut = s.lut;
    if (lut_sm) lut.data = sh_lut;
    const TriLight* lights = lights_sm ? sh_lights : s.lights;
    const size_t first = (size_t)f.y0 * f.W;
    const unsigned n = FROM_QUEUE ? f.cnt->queue_n : (unsigned)((size_t)(f.y1 - f.y0) * f.W);
    for (unsigned i = blockIdx.x * LTC_BLOCK + threadIdx.x; i < n; i += gridDim.x * LTC_BLOCK) {
        const size_t ofs = FROM_QUEUE ? (size_t)f.queue[i] : first + i;
        const float4 nn = ld_stream(f.normal + ofs);
        float3 c = v3(0.f);
        if (FROM_QUEUE || !(nn.x == 0.f && nn.y == 0.f && nn.z == 0.f)) {
            const float4 pp = ld_stream(f.position + ofs);
            const float4 dd = ld_stream(f.albedo + ofs);
            const float4 aa = ld_stream(f.alpha + ofs);
            c = ltc_pixel(lut, lights, s.n_lights, cam_pos, v3(pp), v3(nn), v3(dd), aa.x);
        }
        const float4 o = make_float4(c.x, c.y, c.z, 1.f);
        st_stream(out + ofs, o);
        if (out2) st_stream(out2 + ofs, o);
    }
}

// ---------------------------------------------------------------------------------------------
// Ray streams: generate -> k_trace_stream_spec -> resolve.
//
// ncu on a one-thread-per-pixel shadow kernel (profiles/r1a): incoherent shadow rays leave 8 of 32 lanes
// active on average, because a warp runs until its LONGEST ray ends (2.3 lanes in the triangle block).
// The stream form stores rays (32 B) and hit records (16 B) in HBM and traverses them with
// persistent warps that REPLACE finished rays: a lane whose ray ended writes its hit and, once
// enough lanes are idle, the warp claims new rays from a global cursor (in chunks, one atomic per
// 128 rays) while the other lanes keep their traversal state.  The extra HBM traffic (48 B/ray)
// is ~3 % of the kernel's time at the measured copy rate.
// ---------------------------------------------------------------------------------------------
#define STREAM_CHUNK 128u

// (Tried and removed: longest-first scheduling by the ray's chord through the scene bounds, two
// bins filled by warp-aggregated atomics.  It cost 0.07 ms of coherence on the shadow stage and
// gained nothing: the per-launch floor was not late long rays but the chunk granularity, fixed by
// the guided chunk sizes below.  Also removed after measurement, round 1: leaf batching without
// speculation, a compressed 8-wide BVH, bounded pops, 64-bit stack entries -- DESIGN.md section 3.)
// Ray `pos` of the stream; its hit record goes to hits[slot] (o.w carries the slot, so a generator may order the
// stream as it likes without the consumers of the hits knowing).
__device__ __forceinline__ void stream_put_ray(float4* __restrict__ rays, unsigned pos, float3 o, float3 d, unsigned slot) {
    st_stream(rays + 2 * (size_t)pos, make_float4(o.x, o.y, o.z, __uint_as_float(slot)));
    st_stream(rays + 2 * (size_t)pos + 1, make_float4(d.x, d.y, d.z, 0.f));
}
__device__ __forceinline__ void stream_put_ray(float4* __restrict__ rays, unsigned id, float3 o, float3 d) {
    stream_put_ray(rays, id, o, d, id);
}

// A warp's claim on the shared work list [0, n): `need` idle lanes (ballot `want`) each get the next unclaimed
// index; the warp's private range [w_next, w_end) is refilled from the global cursor by ONE atomic per chunk.
// Guided self-scheduling: chunks shrink (128 -> 32) as the list runs dry, so a small launch or the end of a
// large one is spread over all resident warps instead of being served 4 rounds deep by the few warps that
// grabbed the last full chunks (ncu, PATH bounces 2-7: 0.4 ms each for a few 10k rays before this).
// Returns this lane's index (only meaningful where its `want` bit is set); may be >= n (list exhausted).
__device__ __forceinline__ unsigned warp_claim(unsigned want, unsigned lane, unsigned n, unsigned* __restrict__ cursor,
                                               unsigned& w_next, unsigned& w_end) {
    const unsigned FULL = 0xffffffffu;
    const unsigned need = __popc(want);
    const unsigned left = w_end - w_next;
    unsigned base = w_next;
    unsigned chunk = STREAM_CHUNK;
    if (left < need) {  // uniform: claim another chunk, finish the old range first
        if (lane == 0) {
            const unsigned seen = *reinterpret_cast<volatile unsigned*>(cursor);
            const unsigned rem = n > seen ? n - seen : 0u;
            const unsigned warps = gridDim.x * (blockDim.x / 32);
            chunk = min(STREAM_CHUNK, max(32u, (rem / (2u * warps)) & ~31u));
            base = atomicAdd(cursor, chunk);
        }
        base = __shfl_sync(FULL, base, 0);
        chunk = __shfl_sync(FULL, chunk, 0);
    }
    const unsigned rank = __popc(want & ((1u << lane) - 1u));
    const unsigned idx = left >= need ? w_next + rank : (rank < left ? w_next + rank : base + (rank - left));
    if (left >= need) {
        w_next += need;
    } else {
        w_next = base + (need - left);
        w_end = base + chunk;
    }
    return idx;
}

// Speculative while-while traversal (after Aila & Laine): a lane that reaches a leaf POSTPONES it (one
// slot) and keeps walking inner nodes; the warp leaves the node loop when fewer than `node_min` lanes can
// still advance, and then every lane tests its postponed leaf together.  Postponing is legal because the
// result does not depend on the visiting order; it only delays the pruning by best.t a little.  It keeps
// lanes busy in BOTH phases (profiles/r1c: 19.5 of 32 lanes in the node step, 9.8 in the triangle test).
// QN: the inner nodes are the 32-byte quantised ones of bvhq.cuh (`qn`, grid `qg`; ORB_TRACE_VARIANT=3) -- one
// 256-bit load per node instead of two; same hits (the encoding is conservative).
// SD: where the traversal stack lives (StackOf above).
#ifndef ORB_TRACE_MINB
#define ORB_TRACE_MINB 9  // resident CTAs per SM the register allocation of the stream kernel aims at (56 registers)
#endif
template <int SD, bool COUNT = false, bool QN = false>
__global__ void __launch_bounds__(TRACE_BLOCK, ORB_TRACE_MINB)
k_trace_stream_spec(SceneView s, const float4* __restrict__ rays, float4* __restrict__ hits,
                    const unsigned* __restrict__ n_items, unsigned mult, unsigned* __restrict__ cursor, int refill_min,
                    int node_min, int height, unsigned long long* __restrict__ stats = nullptr,
                    const QNode* __restrict__ qn = nullptr, QGrid qg = QGrid()) {
    unsigned c_nodes = 0, c_tris = 0;  // COUNT: node records fetched / triangle records tested by this lane
    typename StackOf<SD>::type st;
    stack_init(st, height);
    const unsigned n = *n_items * mult;
    const unsigned lane = threadIdx.x & 31u;
    const unsigned FULL = 0xffffffffu;
    unsigned w_next = 0, w_end = 0;  // the warp's claimed range of ray indices (uniform across lanes)
    TravState t;
    t.cur = ORB_TRAVERSAL_DONE;
    t.sp = 0;
    int leaf = 0;  // postponed leaf reference (always negative), 0 = none
    unsigned ray = FULL;
    bool exhausted = false;
    for (;;) {
        const bool idle = t.cur == ORB_TRAVERSAL_DONE && leaf == 0;
        if (idle && ray != FULL) {
            ORB_DEVICE_ASSERT(ray < n);  // the slot a generator wrote into the ray record
            st_stream(hits + ray, make_float4(__int_as_float(t.best.prim), t.best.u, t.best.v, t.best.t));
            ray = FULL;
        }
        const unsigned want = __ballot_sync(FULL, idle && !exhausted);
        const unsigned busy = __ballot_sync(FULL, !idle);
        if (want && ((int)__popc(want) >= refill_min || busy == 0u)) {
            const unsigned idx = warp_claim(want, lane, n, cursor, w_next, w_end);
            if (idle && !exhausted) {
                if (idx < n) {
                    const float4 ro = ld_stream(rays + 2 * (size_t)idx), rd = ld_stream(rays + 2 * (size_t)idx + 1);
                    ray = __float_as_uint(ro.w);  // where this ray's hit record goes
                    if (QN) trav_begin_q(t, s, qg, v3(ro), v3(rd));
                    else trav_begin(t, s, v3(ro), v3(rd));
                } else {
                    exhausted = true;
                }
            }
        }
        if (__ballot_sync(FULL, t.cur != ORB_TRAVERSAL_DONE || leaf != 0) == 0u) break;
        // node phase.  A lane is walking an inner node, holds a leaf, or is done; leaf references are the
        // negative values other than the DONE sentinel.
#pragma unroll 1
        for (int it = 0; it < 64; ++it) {
            if (COUNT && t.cur >= 0) ++c_nodes;
            if (t.cur >= 0) {
                if (QN) trav_node_step_qn(t, qn, st);
                else trav_node_step(t, s, st);
            }
            if (leaf == 0 && t.cur < 0 && t.cur != ORB_TRAVERSAL_DONE) {  // postpone this leaf, keep walking
                leaf = t.cur;
                t.cur = stack_pop(st, t.sp, t.best);
            }
            if ((int)__popc(__ballot_sync(FULL, t.cur >= 0)) < node_min) break;
        }
        // leaf phase
        if (leaf != 0) {
            trav_leaf_test(t, s, leaf, COUNT ? &c_tris : nullptr);
            leaf = 0;
            if (t.cur < 0 && t.cur != ORB_TRAVERSAL_DONE) {  // a second leaf was waiting: it becomes the postponed one
                leaf = t.cur;
                t.cur = stack_pop(st, t.sp, t.best);
            }
        }
    }
    if (COUNT) {
        for (int o = 16; o; o >>= 1) {
            c_nodes += __shfl_xor_sync(FULL, c_nodes, o);
            c_tris += __shfl_xor_sync(FULL, c_tris, o);
        }
        if (lane == 0) {
            atomicAdd(stats, (unsigned long long)c_nodes);
            atomicAdd(stats + 1, (unsigned long long)c_tris);
        }
    }
}

__global__ void k_pack_rays(const float* __restrict__ o, const float* __restrict__ d, unsigned n, float4* __restrict__ rays) {
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) stream_put_ray(rays, i, v3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), v3(d[3 * i], d[3 * i + 1], d[3 * i + 2]));
}
__global__ void k_unpack_hits(const float4* __restrict__ hits, unsigned n, int* __restrict__ prim, float* __restrict__ t,
                              float* __restrict__ u, float* __restrict__ v) {
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 h = hits[i];
    prim[i] = __float_as_int(h.x);
    u[i] = h.y;
    v[i] = h.z;
    t[i] = h.w;
}

// one thread per (active pixel, sample): the ray of ratio_based_shadow's light sample
__global__ void __launch_bounds__(256)
k_shadow_gen(SceneView s, FrameView f, int frame_id, int spp, float4* __restrict__ rays) {
    const unsigned n = f.cnt->queue_n * (unsigned)spp;
    for (unsigned r = blockIdx.x * blockDim.x + threadIdx.x; r < n; r += gridDim.x * blockDim.x) {
        const unsigned i = r / (unsigned)spp, k = r - i * (unsigned)spp;
        const size_t ofs = f.queue[i];
        const int px = (int)(ofs % (size_t)f.W), py = (int)(ofs / (size_t)f.W);
        const float4 pp = __ldg(f.position + ofs), nn = __ldg(f.normal + ofs);
        Rng rng = rng_seed(frame_id, (uint32_t)px, (uint32_t)py, (uint32_t)f.W);
        rng_skip(rng, 2 + 5 * (int)k);  // two jitter draws, five per earlier sample
        const ShadowSample q = shadow_ray_setup(s, v3(pp), v3(nn), rng);
        stream_put_ray(rays, r, q.origin, q.L);
    }
}

// (Measured and removed, round 2: ordering the stream by (chunk of the queue, selected light) so that a warp holds rays
// with nearby origins AND a common target -- trace stage 1.74 ms against 1.61 ms for every chunk size from 256 to
// 16 k pixels, plus 0.09 ms of sorting (profiles/r2f_ab_sort.jsonl).  The four samples of a pixel sit in adjacent
// lanes here and start from the SAME point; most node visits of a shadow ray lie next to its origin, so that
// sharing is worth more than a common target.  Round 1's frame-wide binning by light lost 16 % for the same reason.)
// one thread per active pixel: the n samples in order (sum, / n, grey mean) as rayGen does
__global__ void __launch_bounds__(128)
k_shadow_resolve(SceneView s, FrameView f, float3 cam_pos, int frame_id, int spp, int mode,
                 const float4* __restrict__ hits) {
    const unsigned n = f.cnt->queue_n;
    const float nf = (float)spp;
    if (blockIdx.x == 0 && threadIdx.x == 0 && s.n_lights > 0)
        atomicAdd(&f.cnt->rays_shadow, (unsigned long long)n * (unsigned long long)spp);
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const size_t ofs = f.queue[i];
        const int px = (int)(ofs % (size_t)f.W), py = (int)(ofs / (size_t)f.W);
        const float4 pp = ld_stream(f.position + ofs), nn = ld_stream(f.normal + ofs);
        const float4 dd = ld_stream(f.albedo + ofs), aa = ld_stream(f.alpha + ofs);
        Rng rng = rng_seed(frame_id, (uint32_t)px, (uint32_t)py, (uint32_t)f.W);
        rng_next(rng);
        rng_next(rng);
        float3 sd = v3(0.f), sn = v3(0.f);
        for (int k = 0; k < spp; k++) {
            const ShadowSample q = shadow_ray_setup(s, v3(pp), v3(nn), rng);
            Hit h;
            h.prim = -1;
            h.t = INFINITY;
            h.u = h.v = 0.f;
            if (q.valid) {
                const float4 hr = ld_stream(hits + (size_t)i * spp + k);
                h.prim = __float_as_int(hr.x);
                h.u = hr.y;
                h.v = hr.z;
                h.t = hr.w;
            }
            float3 d, nv;
            shadow_ray_resolve(s, q, h, cam_pos, v3(pp), v3(nn), v3(dd), aa.x, mode, d, nv);
            sd += d;
            sn += nv;
        }
        sd = sd / nf;
        sn = sn / nf;
        const float gd = (sd.x + sd.y + sd.z) / 3.f, gn = (sn.x + sn.y + sn.z) / 3.f;
        st_stream(f.sto_direct + ofs, make_float4(gd, gd, gd, 1.f));
        st_stream(f.sto_no_vis + ofs, make_float4(gn, gn, gn, 1.f));
    }
}

// Same result, one thread per (active pixel, sample) for SPP = 2, 4 or 8: the SPP lanes of a pixel sit side
// by side, so the dependent gathers of a sample (light record, hit record, the hit triangle's normals and
// mesh record) of four times as many samples are in flight; lane 0 of the group then adds the samples in
// rayGen's order (k = 0, 1, ...), which keeps the sums bit-identical to k_shadow_resolve.
template <int SPP>
__global__ void __launch_bounds__(256)
k_shadow_resolve_lanes(SceneView s, FrameView f, float3 cam_pos, int frame_id, int mode, const float4* __restrict__ hits) {
    const unsigned n = f.cnt->queue_n;
    if (blockIdx.x == 0 && threadIdx.x == 0 && s.n_lights > 0)
        atomicAdd(&f.cnt->rays_shadow, (unsigned long long)n * (unsigned long long)SPP);
    const unsigned total = n * (unsigned)SPP;
    const unsigned rounded = (total + 31u) & ~31u;  // whole warps stay in the loop: the shuffles need every lane
    for (unsigned r = blockIdx.x * blockDim.x + threadIdx.x; r < rounded; r += gridDim.x * blockDim.x) {
        const bool live = r < total;
        const unsigned i = r / (unsigned)SPP, k = r % (unsigned)SPP;
        float3 d = v3(0.f), nv = v3(0.f);
        size_t ofs = 0;
        if (live) {
            ofs = f.queue[i];
            const int px = (int)(ofs % (size_t)f.W), py = (int)(ofs / (size_t)f.W);
            const float4 pp = __ldg(f.position + ofs), nn = __ldg(f.normal + ofs);
            const float4 dd = __ldg(f.albedo + ofs), aa = __ldg(f.alpha + ofs);
            Rng rng = rng_seed(frame_id, (uint32_t)px, (uint32_t)py, (uint32_t)f.W);
            rng_skip(rng, 2 + 5 * (int)k);
            const ShadowSample q = shadow_ray_setup(s, v3(pp), v3(nn), rng);
            Hit h;
            h.prim = -1;
            h.t = INFINITY;
            h.u = h.v = 0.f;
            if (q.valid) {
                const float4 hr = ld_stream(hits + r);
                h.prim = __float_as_int(hr.x);
                h.u = hr.y;
                h.v = hr.z;
                h.t = hr.w;
            }
            shadow_ray_resolve(s, q, h, cam_pos, v3(pp), v3(nn), v3(dd), aa.x, mode, d, nv);
        }
        float3 sd = d, sn = nv;
#pragma unroll
        for (int t = 1; t < SPP; t++) {
            sd.x += __shfl_down_sync(0xffffffffu, d.x, t);
            sd.y += __shfl_down_sync(0xffffffffu, d.y, t);
            sd.z += __shfl_down_sync(0xffffffffu, d.z, t);
            sn.x += __shfl_down_sync(0xffffffffu, nv.x, t);
            sn.y += __shfl_down_sync(0xffffffffu, nv.y, t);
            sn.z += __shfl_down_sync(0xffffffffu, nv.z, t);
        }
        if (live && k == 0) {
            sd = sd / (float)SPP;
            sn = sn / (float)SPP;
            const float gd = (sd.x + sd.y + sd.z) / 3.f, gn = (sn.x + sn.y + sn.z) / 3.f;
            st_stream(f.sto_direct + ofs, make_float4(gd, gd, gd, 1.f));
            st_stream(f.sto_no_vis + ofs, make_float4(gn, gn, gn, 1.f));
        }
    }
}

// ---------------------------------------------------------------------------------------------
// k_path_bounce: one iteration of the bounce loop for every live path
// ---------------------------------------------------------------------------------------------
// Path state SoA, 5 float4 per path: (p, alpha) (n, pixel) (diffuse, -) (V, -) (tp, -)
struct PathBuffers {
    float4* s[5];
};

template <bool FIRST>
__global__ void __launch_bounds__(TRACE_BLOCK)
k_path_bounce(SceneView s, FrameView f, float3 cam_pos, int frame_id, int bounce, int mode, int last, PathBuffers in,
              PathBuffers out, int in_slot, int height) {
    int* stk = s_dyn + threadIdx.x;
    float* stk_t = reinterpret_cast<float*>(s_dyn + height * TRACE_BLOCK) + threadIdx.x;
    unsigned rays = 0;
    SmemTracer tr{s, stk, stk_t, &rays};
    const unsigned n = FIRST ? f.cnt->queue_n : f.cnt->path_n[in_slot];
    const int lane = threadIdx.x & 31;
    // round the trip count up so whole warps reach the ballot
    const unsigned n_round = (n + 31u) & ~31u;
    for (unsigned i = blockIdx.x * TRACE_BLOCK + threadIdx.x; i < n_round; i += gridDim.x * TRACE_BLOCK) {
        bool survive = false;
        PathState st;
        size_t ofs = 0;
        if (i < n) {
            if (FIRST) {
                ofs = f.queue[i];
                const float4 pp = ld_stream(f.position + ofs), nn = ld_stream(f.normal + ofs);
                const float4 dd = ld_stream(f.albedo + ofs), aa = ld_stream(f.alpha + ofs);
                st.p = v3(pp);
                st.n = v3(nn);
                st.diffuse = v3(dd);
                st.alpha = aa.x;
                st.V = normalize(cam_pos - st.p);  // si.wo
                st.tp = v3(1.f);
            } else {
                const float4 a = in.s[0][i], b = in.s[1][i], c = in.s[2][i], d = in.s[3][i], e = in.s[4][i];
                st.p = v3(a);
                st.alpha = a.w;
                st.n = v3(b);
                ofs = (size_t)__float_as_uint(b.w);
                st.diffuse = v3(c);
                st.V = v3(d);
                st.tp = v3(e);
            }
            const int px = (int)(ofs % (size_t)f.W), py = (int)(ofs / (size_t)f.W);
            Rng rng = rng_seed(frame_id, (uint32_t)px, (uint32_t)py, (uint32_t)f.W);
            rng_skip(rng, 2 + 5 * bounce);
            float3 colour;
            bool traced;
            const int r = path_bounce(s, tr, st, rng, mode, last != 0, colour, traced);
            if (r == ORB_PATH_DONE) st_stream(f.colour + ofs, make_float4(colour.x, colour.y, colour.z, 1.f));
            else survive = true;
        }
        const unsigned m = __ballot_sync(0xffffffffu, survive);
        if (m) {
            unsigned base = 0;
            if (lane == 0) base = atomicAdd(&f.cnt->path_n[in_slot ^ 1], (unsigned)__popc(m));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (survive) {
                const unsigned o = base + __popc(m & ((1u << lane) - 1u));
                out.s[0][o] = make_float4(st.p.x, st.p.y, st.p.z, st.alpha);
                out.s[1][o] = make_float4(st.n.x, st.n.y, st.n.z, __uint_as_float((unsigned)ofs));
                out.s[2][o] = make_float4(st.diffuse.x, st.diffuse.y, st.diffuse.z, 0.f);
                out.s[3][o] = make_float4(st.V.x, st.V.y, st.V.z, 0.f);
                out.s[4][o] = make_float4(st.tp.x, st.tp.y, st.tp.z, 0.f);
            }
        }
    }
    for (int o = 16; o > 0; o >>= 1) rays += __shfl_xor_sync(0xffffffffu, rays, o);
    if (lane == 0 && rays) atomicAdd(&f.cnt->rays_bounce, (unsigned long long)rays);
}

// ---- ray-stream form of one bounce: k_path_gen -> k_trace_stream -> k_path_shade ----------------
template <bool FIRST>
__device__ __forceinline__ void path_load_state(const FrameView& f, float3 cam_pos, const PathBuffers& in, unsigned i,
                                                PathState& st, size_t& ofs) {
    if (FIRST) {
        ofs = f.queue[i];
        const float4 pp = ld_stream(f.position + ofs), nn = ld_stream(f.normal + ofs);
        const float4 dd = ld_stream(f.albedo + ofs), aa = ld_stream(f.alpha + ofs);
        st.p = v3(pp);
        st.n = v3(nn);
        st.diffuse = v3(dd);
        st.alpha = aa.x;
        st.V = normalize(cam_pos - st.p);  // si.wo
        st.tp = v3(1.f);
    } else {
        const float4 a = in.s[0][i], b = in.s[1][i], c = in.s[2][i], d = in.s[3][i], e = in.s[4][i];
        st.p = v3(a);
        st.alpha = a.w;
        st.n = v3(b);
        ofs = (size_t)__float_as_uint(b.w);
        st.diffuse = v3(c);
        st.V = v3(d);
        st.tp = v3(e);
    }
}

template <bool FIRST>
__global__ void __launch_bounds__(256)
k_path_gen(FrameView f, float3 cam_pos, int frame_id, int bounce, PathBuffers in, int in_slot, float4* __restrict__ rays) {
    const unsigned n = FIRST ? f.cnt->queue_n : f.cnt->path_n[in_slot];
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        PathState st;
        size_t ofs;
        path_load_state<FIRST>(f, cam_pos, in, i, st, ofs);
        const int px = (int)(ofs % (size_t)f.W), py = (int)(ofs / (size_t)f.W);
        Rng rng = rng_seed(frame_id, (uint32_t)px, (uint32_t)py, (uint32_t)f.W);
        rng_skip(rng, 2 + 5 * bounce);
        const PathSample q = path_ray_setup(st, rng);
        stream_put_ray(rays, i, q.origin, q.dir);
    }
}

template <bool FIRST>
__global__ void __launch_bounds__(128)
k_path_shade(SceneView s, FrameView f, float3 cam_pos, int frame_id, int bounce, int last, PathBuffers in, PathBuffers out,
             int in_slot, const float4* __restrict__ hits) {
    const unsigned n = FIRST ? f.cnt->queue_n : f.cnt->path_n[in_slot];
    const int lane = threadIdx.x & 31;
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&f.cnt->rays_bounce, (unsigned long long)n);
    const unsigned n_round = (n + 31u) & ~31u;  // whole warps reach the ballot
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += gridDim.x * blockDim.x) {
        bool survive = false;
        PathState st;
        size_t ofs = 0;
        if (i < n) {
            path_load_state<FIRST>(f, cam_pos, in, i, st, ofs);
            const int px = (int)(ofs % (size_t)f.W), py = (int)(ofs / (size_t)f.W);
            Rng rng = rng_seed(frame_id, (uint32_t)px, (uint32_t)py, (uint32_t)f.W);
            rng_skip(rng, 2 + 5 * bounce);
            const PathSample q = path_ray_setup(st, rng);  // same arithmetic as k_path_gen
            const float4 hr = ld_stream(hits + i);
            Hit h;
            h.prim = __float_as_int(hr.x);
            h.u = hr.y;
            h.v = hr.z;
            h.t = hr.w;
            float3 colour;
            if (path_resolve(s, st, q, h, last != 0, colour) == ORB_PATH_DONE)
                st_stream(f.colour + ofs, make_float4(colour.x, colour.y, colour.z, 1.f));
            else
                survive = true;
        }
        const unsigned m = __ballot_sync(0xffffffffu, survive);
        if (m) {
            unsigned base = 0;
            if (lane == 0) base = atomicAdd(&f.cnt->path_n[in_slot ^ 1], (unsigned)__popc(m));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (survive) {
                const unsigned o = base + __popc(m & ((1u << lane) - 1u));
                out.s[0][o] = make_float4(st.p.x, st.p.y, st.p.z, st.alpha);
                out.s[1][o] = make_float4(st.n.x, st.n.y, st.n.z, __uint_as_float((unsigned)ofs));
                out.s[2][o] = make_float4(st.diffuse.x, st.diffuse.y, st.diffuse.z, 0.f);
                out.s[3][o] = make_float4(st.V.x, st.V.y, st.V.z, 0.f);
                out.s[4][o] = make_float4(st.tp.x, st.tp.y, st.tp.z, 0.f);
            }
        }
    }
}

// ---- persistent form of the whole bounce loop ----------------------------------------------------
// The per-bounce wavefront above pays, for EVERY bounce, three launches, an 80-byte round trip of
// the path state through HBM and -- worst -- the tail of the slowest ray: a few bounce rays skim the
// terrain for a thousand node visits, so each traversal launch has a floor of ~0.18 ms no matter
// how few paths are left (ncu launch list, profiles/r1j: bounces 4-7 take 0.18-0.19 ms each for
// < 10 % of the rays).  Here a lane owns a PATH instead of a ray: when its ray ends it shades
// (path_resolve) and starts the next bounce at once, and when its path ends it claims a new pixel
// from the active-pixel queue.  One launch per frame, one tail per frame, no path state in HBM.
// Shading is batched like the refill: lanes whose ray has ended wait until `refill_min` of them
// have (or nobody can advance), so the divergent shading code runs with a fair number of lanes.
#ifndef ORB_PATH_MINB
// resident CTAs per SM the register allocation of the path kernel aims at.  Measured (profiles/r2l_ab_pmb.jsonl), PATH frame /
// 135-row band: 5 CTAs (95 registers, no spills) 1.899 / 0.812 ms, 6 (80) 1.842 / 0.842, 7 (72, 142 B spilled) 1.759 / 0.833,
// 8 (64, 222 B spilled) 1.740 / 0.863: the kernel is latency-bound, 28 warps beat 20 despite the spills.
#define ORB_PATH_MINB 7
#endif
template <int SD, bool COUNT = false, bool QN = false>
__global__ void __launch_bounds__(TRACE_BLOCK, ORB_PATH_MINB)
k_path_persistent(SceneView s, FrameView f, float3 cam_pos, int frame_id, int depth, unsigned* __restrict__ cursor,
                  int refill_min, int node_min, int height, const QNode* __restrict__ qn, QGrid qg) {
    unsigned c_nodes = 0, c_tris = 0;  // COUNT (ORB_RENDER_COUNT_TRAVERSAL): node / triangle records fetched by this lane
    typename StackOf<SD>::type stk;
    stack_init(stk, height);
    const unsigned n = f.cnt->queue_n;
    const unsigned lane = threadIdx.x & 31u;
    const unsigned FULL = 0xffffffffu;
    unsigned w_next = 0, w_end = 0;
    TravState t;
    t.cur = ORB_TRAVERSAL_DONE;
    t.sp = 0;
    int leaf = 0;
    PathState st;
    // what the shading of a bounce needs from the moment its ray was made (path_ray_setup): the two local
    // directions and the pdf -- 7 registers carried across the traversal instead of re-deriving the basis, the five
    // draws and the BSDF sample a second time per bounce (orb200.cu:912-918 against :962-967)
    float3 q_wi = v3(0.f), q_wo = v3(0.f);
    float q_pdf = 0.f;
    size_t ofs = 0;
    int bounce = 0;
    bool have_path = false, exhausted = false;
    unsigned rays = 0;
    PathBuffers none;
    for (;;) {
        const bool ended = t.cur == ORB_TRAVERSAL_DONE && leaf == 0;  // ray finished, or no path at all
        const unsigned n_ended = __popc(__ballot_sync(FULL, ended && (have_path || !exhausted)));
        const unsigned busy = __ballot_sync(FULL, !ended);
        if (n_ended && ((int)n_ended >= refill_min || busy == 0u)) {
            if (ended && have_path) {  // shade this bounce
                PathSample q;
                q.wi_local = q_wi;
                q.wo_local = q_wo;
                q.bsdf_pdf = q_pdf;
                float3 colour;
                if (path_resolve(s, st, q, t.best, bounce == depth - 1, colour) == ORB_PATH_DONE) {
                    st_stream(f.colour + ofs, make_float4(colour.x, colour.y, colour.z, 1.f));
                    have_path = false;
                } else {
                    ++bounce;
                }
            }
            // lanes without a path claim the next active pixels (chunks of the queue, one atomic per chunk)
            const bool claim = ended && !have_path && !exhausted;
            const unsigned want = __ballot_sync(FULL, claim);
            if (want) {
                const unsigned idx = warp_claim(want, lane, n, cursor, w_next, w_end);
                if (claim) {
                    if (idx < n) {
                        path_load_state<true>(f, cam_pos, none, idx, st, ofs);
                        bounce = 0;
                        have_path = true;
                    } else {
                        exhausted = true;
                    }
                }
            }
            if (ended && have_path) {  // next ray of this path
                const int px = (int)(ofs % (size_t)f.W), py = (int)(ofs / (size_t)f.W);
                Rng rng = rng_seed(frame_id, (uint32_t)px, (uint32_t)py, (uint32_t)f.W);
                rng_skip(rng, 2 + 5 * bounce);
                const PathSample q = path_ray_setup(st, rng);
                q_wi = q.wi_local;
                q_wo = q.wo_local;
                q_pdf = q.bsdf_pdf;
                if (QN) trav_begin_q(t, s, qg, q.origin, q.dir);
                else trav_begin(t, s, q.origin, q.dir);
                ++rays;
            }
        }
        if (__ballot_sync(FULL, have_path) == 0u) break;
#pragma unroll 1
        for (int it = 0; it < 64; ++it) {  // node phase (speculative: one postponed leaf per lane)
            if (COUNT && t.cur >= 0) ++c_nodes;
            if (t.cur >= 0) {
                if (QN) trav_node_step_qn(t, qn, stk);
                else trav_node_step(t, s, stk);
            }
            if (leaf == 0 && t.cur < 0 && t.cur != ORB_TRAVERSAL_DONE) {
                leaf = t.cur;
                t.cur = stack_pop(stk, t.sp, t.best);
            }
            if ((int)__popc(__ballot_sync(FULL, t.cur >= 0)) < node_min) break;
        }
        if (leaf != 0) {  // leaf phase
            trav_leaf_test(t, s, leaf, COUNT ? &c_tris : nullptr);
            leaf = 0;
            if (t.cur < 0 && t.cur != ORB_TRAVERSAL_DONE) {
                leaf = t.cur;
                t.cur = stack_pop(stk, t.sp, t.best);
            }
        }
    }
    for (int o = 16; o > 0; o >>= 1) rays += __shfl_xor_sync(FULL, rays, o);
    if (lane == 0 && rays) atomicAdd(&f.cnt->rays_bounce, (unsigned long long)rays);
    if (COUNT) {
        for (int o = 16; o; o >>= 1) {
            c_nodes += __shfl_xor_sync(FULL, c_nodes, o);
            c_tris += __shfl_xor_sync(FULL, c_tris, o);
        }
        if (lane == 0) {
            atomicAdd(&f.cnt->nodes_visited, (unsigned long long)c_nodes);
            atomicAdd(&f.cnt->tris_tested, (unsigned long long)c_tris);
        }
    }
}

__global__ void k_reset_slot(DevCounters* c, int slot) { c->path_n[slot] = 0; }

// ---------------------------------------------------------------------------------------------
// k_check_frame (ORB_RENDER_CHECKED): what north_star wants "counted and reported"
//   hit_ties           pixels whose primary hit lies within 1e-5 (barycentric) of an edge of its triangle: the ray
//                      passes through the seam between two triangles within the rounding of the ray/triangle test, and
//                      another test or traversal order (OptiX: not in the repository) may report the neighbour's id
//   ltc_near_decision  active pixels whose LTC evaluation is within 1e-4 of a sign decision (ltc_decision_margin)
// One thread per pixel over the frame's hit records and G-buffer; runs after k_primary, outside any timing.
// ---------------------------------------------------------------------------------------------
#define ORB_EDGE_TIE_EPS 1e-5f
__global__ void __launch_bounds__(TRACE_BLOCK)
k_check_frame(SceneView s, FrameView f, Camera cam, int with_ltc) {
    const size_t first = (size_t)f.y0 * f.W, n = (size_t)(f.y1 - f.y0) * f.W;
    unsigned ties = 0, near = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const size_t ofs = first + i;
        const float4 h = f.hit[ofs];
        if (__float_as_int(h.x) < 0) continue;
        const bool tie = fminf(fminf(h.y, h.z), 1.f - h.y - h.z) <= ORB_EDGE_TIE_EPS;
        ties += tie ? 1u : 0u;
        if (with_ltc && (f.flags[ofs] & FLAG_HIT) && !(f.flags[ofs] & FLAG_LIGHT)) {
            const float4 pp = f.position[ofs], nn = f.normal[ofs], aa = f.alpha[ofs];
            const float3 wo = normalize(cam.pos - v3(pp));
            const Mat3 to_local = onb_rows(v3(nn));
            LtcFrame fr;
            ltc_setup(fr, s.lut, v3(pp), v3(0.f), aa.x, to_local, wo);
            const float wl_z = normalize(apply(to_local, wo)).z;
            near += ltc_decision_margin(fr, wl_z, s.lights, s.n_lights) <= 1e-4f ? 1u : 0u;
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        ties += __shfl_xor_sync(0xffffffffu, ties, o);
        near += __shfl_xor_sync(0xffffffffu, near, o);
    }
    if ((threadIdx.x & 31) == 0) {
        if (ties) atomicAdd(&f.cnt->hit_ties, ties);
        if (near) atomicAdd(&f.cnt->ltc_near_decision, near);
    }
}

// ---------------------------------------------------------------------------------------------
// k_accumulate: deviceCode.cu:158-161,172-174
// ---------------------------------------------------------------------------------------------
__global__ void k_accumulate(FrameView f, const float4* __restrict__ src, int accum_id, int count_rays) {
    if (count_rays && blockIdx.x == 0 && threadIdx.x == 0) {  // every earlier kernel of the frame has finished
        f.cnt->tot_shadow += f.cnt->rays_shadow;
        f.cnt->tot_bounce += f.cnt->rays_bounce;
        f.cnt->tot_active += f.cnt->queue_n;
    }
    const size_t first = (size_t)f.y0 * f.W, n = (size_t)(f.y1 - f.y0) * f.W;
    const float scale = 1.f / ((float)accum_id + 1.f);
    unsigned nan = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const size_t ofs = first + i;
        float3 c = v3(0.f);
        if (f.flags[ofs] & FLAG_HIT) c = v3(ld_stream(src + ofs));
        if (accum_id > 0) c = c + v3(__ldcs(f.accum + ofs));  // read-modify-write: no .nc load here
        __stcs(f.accum + ofs, make_float4(c.x, c.y, c.z, 1.f));
        nan += (c.x != c.x || c.y != c.y || c.z != c.z) ? 1u : 0u;
        f.rgba8[ofs] = make_rgba(scale * c);
    }
    for (int o = 16; o > 0; o >>= 1) nan += __shfl_xor_sync(0xffffffffu, nan, o);
    if ((threadIdx.x & 31) == 0 && nan) atomicAdd(&f.cnt->nan_pixels, nan);
}

// ---------------------------------------------------------------------------------------------
// denoise: edge-avoiding a-trous wavelet filter (own definition -- the reference's
// cuda_include/denoise/bilateral.cuh is an empty file).  CPU statement of the same
// filter: oracle/restate.cpp ora_denoise().
//   value  v(q) = (sto_direct.x, sto_no_vis.x)                       packed float2
//   guide  g(q) = (n.xyz, depth = |p - cam|), depth = -1e30 for pixels with materialID == 0, n == 0
//                 (lights, misses) or a non-finite guide: passed through unfiltered, never used as a tap
//   weight w(q) = B3(dx) B3(dy) * max(0, n_c . n_q)^sigma_n * exp(-|d_c - d_q| / (sigma_p * step))
//                 (* exp(-|v_c - v_q| / sigma_c) per image when sigma_c > 0); NaN taps are skipped
// ---------------------------------------------------------------------------------------------
#define DN_SKIP_DEPTH (-1e30f)  // guide.w of a pixel that is neither filtered nor used as a tap
// HaloSendList (halo.cuh): the rows of this band that other ranks' filters read are stored a second time, straight
// into those ranks' buffers over NVLink (peer-mapped pointers) -- the tile hand-off is part of the pack kernel.
__global__ void k_denoise_pack(FrameView f, float3 cam_pos, float4* __restrict__ guide, float2* __restrict__ val,
                               HaloSendList sends) {
    const size_t first = (size_t)f.y0 * f.W, n = (size_t)(f.y1 - f.y0) * f.W;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const size_t ofs = first + i;
        const float4 nn = ld_stream(f.normal + ofs), pp = ld_stream(f.position + ofs);
        const float4 mm = ld_stream(f.material_id + ofs);
        const float4 a = ld_stream(f.sto_direct + ofs), b = ld_stream(f.sto_no_vis + ofs);
        float depth = length(v3(pp) - cam_pos);
        // lights, misses, and pixels whose guide is not finite (every weight would be NaN: the tap-skipping
        // definition passes them through and never uses them as taps either).  The sentinel is so far from
        // any depth that exp(-|dz| / (sigma_p step)) is exactly 0: the filter needs no per-tap test.
        const bool skip = (nn.x == 0.f && nn.y == 0.f && nn.z == 0.f) || mm.x == 0.f ||
                          !(fabsf(nn.x) + fabsf(nn.y) + fabsf(nn.z) < 1e18f) || !(depth < 1e18f);
        const float4 g = make_float4(nn.x, nn.y, nn.z, skip ? DN_SKIP_DEPTH : depth);
        const float2 v = skip ? make_float2(0.f, 0.f) : make_float2(a.x, b.x);  // 0 * weight 0 stays 0 (emit may be inf)
        guide[ofs] = g;
        val[ofs] = v;
        if (sends.n) {
            const int y = (int)(ofs / (size_t)f.W);
#pragma unroll
            for (int k = 0; k < ORB_HALO_MAX_SENDS; k++)
                if (k < sends.n && y >= sends.s[k].y0 && y < sends.s[k].y1) {
                    sends.s[k].guide[ofs] = g;
                    sends.s[k].val[ofs] = v;
                }
        }
    }
}

// MUFU lg2 / ex2 without the denormal fix-up code of __log2f / exp2f (a weight below 2^-126 is 0 either way)
__device__ __forceinline__ float fast_lg2(float x) {
    float r;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float fast_ex2(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
// log2 of the B3 tap weights (3/8, 1/4, 1/16): the tap's B3(dx) B3(dy) enters the weight's exponent as a constant
__device__ __forceinline__ constexpr double b3_lg2_1(int k) { return k == 0 ? -1.4150374992788437 : ((k == 1 || k == -1) ? -2.0 : -4.0); }
__device__ __forceinline__ constexpr float b3_lg2(int i, int j) { return (float)(b3_lg2_1(i) + b3_lg2_1(j)); }

// The 25 taps of one pixel.  INTERIOR: the whole block is further than 2*step from every frame edge, no
// bounds tests.  COLOUR: sigma_c > 0.  STEP > 0: the tap distance as a compile-time constant (the column
// offsets become immediates of the loads); 0: run-time `step`.  Skipped taps (guide sentinel: weight exactly
// 0; NaN value) enter the sums as exact zeros, so the result is the one of the tap-skipping definition
// (oracle/restate.cpp ora_denoise()).  The pass is issue-bound (25 taps x ~23 instructions against 32 B of
// HBM traffic per pixel).
template <bool COLOUR, bool INTERIOR, int STEP>
__device__ __forceinline__ void atrous_taps(const float4* __restrict__ guide, const float2* __restrict__ vin, int W, int H,
                                            int x, int y, int step_rt, float4 gc, float2 vc, float sigma_n,
                                            float inv_sigma_p, float inv_sigma_c, float& sd, float& sn, float& wd,
                                            float& wn) {
    const int step = STEP > 0 ? STEP : step_rt;
#pragma unroll
    for (int j = -2; j <= 2; j++) {
        const int yy = y + j * step;
        if (!INTERIOR && (yy < 0 || yy >= H)) continue;
        const int row = yy * W + x;  // 32-bit pixel offsets: orb_resize refuses frames of 2^31 pixels or more
        const float4* grow = guide + row;
        const float2* vrow = vin + row;
#pragma unroll
        for (int i = -2; i <= 2; i++) {
            const int dx = i * step;
            if (!INTERIOR && (x + dx < 0 || x + dx >= W)) continue;
            const float4 g = __ldg(grow + dx);
            const float2 v = __ldg(vrow + dx);
            // B3 B3 * nd^sigma_n * exp(-|dz| / (sigma_p step)) as ONE ex2 of a sum of logs
            const float nd = fmaxf(0.f, gc.x * g.x + gc.y * g.y + gc.z * g.z);
            const float ln = sigma_n > 0.f ? fmaf(sigma_n, fast_lg2(nd), b3_lg2(i, j)) : b3_lg2(i, j);
            const float w = fast_ex2(fmaf(-fabsf(gc.w - g.w), inv_sigma_p, ln));
            float w0 = w, w1 = w;
            bool ok0 = v.x == v.x, ok1 = v.y == v.y;
            if (COLOUR) {
                w0 *= fast_ex2(-fabsf(vc.x - v.x) * inv_sigma_c);
                w1 *= fast_ex2(-fabsf(vc.y - v.y) * inv_sigma_c);
                ok0 = ok0 && w0 == w0;  // NaN centre value
                ok1 = ok1 && w1 == w1;
            }
            w0 = ok0 ? w0 : 0.f;
            w1 = ok1 ? w1 : 0.f;
            sd += w0 * (ok0 ? v.x : 0.f);
            wd += w0;
            sn += w1 * (ok1 ? v.y : 0.f);
            wn += w1;
        }
    }
}

template <bool FINAL, bool COLOUR, int STEP>
__global__ void __launch_bounds__(256)
k_denoise_atrous(FrameView f, const float4* __restrict__ guide, const float2* __restrict__ vin, float2* __restrict__ vout,
                 float4* __restrict__ out_d, float4* __restrict__ out_n, int step, float sigma_n, float inv_sigma_p,
                 float inv_sigma_c) {
    const int bx = blockIdx.x * 32, by = f.y0 + blockIdx.y * 8;
    const int x = bx + (threadIdx.x & 31);
    const int y = by + (threadIdx.x >> 5);
    if (x >= f.W || y >= f.y1) return;
    const int ofs = y * f.W + x;
    const float4 gc = __ldg(guide + ofs);
    const float2 vc = __ldg(vin + ofs);
    float2 r = vc;
    if (gc.w >= 0.f) {
        float sd = 0.f, sn = 0.f, wd = 0.f, wn = 0.f;
        const int reach = 2 * (STEP > 0 ? STEP : step);
        const bool interior = bx >= reach && bx + 31 + reach < f.W && by >= reach && by + 7 + reach < f.H;
        if (interior)
            atrous_taps<COLOUR, true, STEP>(guide, vin, f.W, f.H, x, y, step, gc, vc, sigma_n, inv_sigma_p, inv_sigma_c, sd,
                                            sn, wd, wn);
        else
            atrous_taps<COLOUR, false, STEP>(guide, vin, f.W, f.H, x, y, step, gc, vc, sigma_n, inv_sigma_p, inv_sigma_c, sd,
                                             sn, wd, wn);
        if (wd > 0.f) r.x = sd / wd;
        if (wn > 0.f) r.y = sn / wn;
    }
    if (FINAL) {
        if (gc.w >= 0.f) {
            out_d[ofs] = make_float4(r.x, r.x, r.x, 1.f);
            out_n[ofs] = make_float4(r.y, r.y, r.y, 1.f);
        } else {  // lights / misses: the unfiltered float4 (emit on light pixels, deviceCode.cu:121-123)
            out_d[ofs] = f.sto_direct[ofs];
            out_n[ofs] = f.sto_no_vis[ofs];
        }
    } else {
        vout[ofs] = r;
    }
}

// Four pixels per thread, `step` rows apart: pixel k's tap rows are y + (k - 2 .. k + 2) step, so the four share
// eight rows of taps -- 40 tap loads for 100 tap evaluations instead of 100.  The one-pixel kernel sits at the L1
// data pipe's roof as much as at the issue roof (6 wavefronts and 23 instructions per tap and warp, 4 schedulers
// on one pipe); this form needs 2.4 wavefronts per evaluation.  A NaN value is skipped as a tap of ITS image only:
// the test runs once per LOADED tap (value -> 0, mask -> 0), not once per evaluation, and an evaluation is four
// multiply-adds (two sums, two masked weight sums) -- the same products and sums in the same order as
// atrous_taps, so the two kernels agree bit for bit (test_atrous_four_pixels_per_thread_equals_one_pixel_kernel).
// sigma_n > 0 and no colour term (the host launches k_denoise_atrous otherwise).  ratio_out != nullptr (FINAL): the
// pass also writes RATIO_FINAL = ltc * (direct / no_vis), the k_ratio_combine of the denoised pair.
// Rows of a CTA: step <= 8: 32 consecutive rows (8 / step groups of 4 step rows); step >= 16: the rows
// by + ty + k step of a 4 step-row span, step / 8 CTAs per span.
template <bool FINAL, int STEP, bool INTERIOR>
__device__ __forceinline__ void atrous4_body(const FrameView& f, const float4* __restrict__ guide, const float2* __restrict__ vin,
                                             float2* __restrict__ vout, float4* __restrict__ out_d, float4* __restrict__ out_n,
                                             const float4* __restrict__ ltc, float4* __restrict__ ratio_out, int s, int x,
                                             int base, float sigma_n, float inv_sigma_p) {
    const int W = f.W, H = f.H;
    float4 gc[4];
    float2 vc[4];
    bool any = false;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int y = base + k * s;
        gc[k] = make_float4(0.f, 0.f, 0.f, DN_SKIP_DEPTH);
        vc[k] = make_float2(0.f, 0.f);
        if (y < f.y1) {
            gc[k] = __ldg(guide + y * W + x);
            vc[k] = __ldg(vin + y * W + x);
        }
        any = any || gc[k].w >= 0.f;
    }
    float sd[4] = {0.f, 0.f, 0.f, 0.f}, sn[4] = {0.f, 0.f, 0.f, 0.f}, wd[4] = {0.f, 0.f, 0.f, 0.f}, wn[4] = {0.f, 0.f, 0.f, 0.f};
    if (any) {
#pragma unroll
        for (int m = 0; m < 8; m++) {
            const int yy = base + (m - 2) * s;
            const bool rowok = INTERIOR || (yy >= 0 && yy < H);
            const float4* grow = guide + yy * W + x;
            const float2* vrow = vin + yy * W + x;
#pragma unroll
            for (int i = -2; i <= 2; i++) {
                const int dx = i * s;
                // a tap outside the frame: the skip sentinel and a zero value enter every sum as exact zeros
                float4 g = make_float4(0.f, 0.f, 0.f, DN_SKIP_DEPTH);
                float2 v = make_float2(0.f, 0.f);
                if (INTERIOR || (rowok && x + dx >= 0 && x + dx < W)) {
                    ORB_DEVICE_ASSERT(yy >= 0 && yy < H && x + dx >= 0 && x + dx < W);
                    g = __ldg(grow + dx);
                    v = __ldg(vrow + dx);
                }
                const float m0 = v.x == v.x ? 1.f : 0.f, m1 = v.y == v.y ? 1.f : 0.f;
                v.x = v.x == v.x ? v.x : 0.f;
                v.y = v.y == v.y ? v.y : 0.f;
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const int j = m - k - 2;
                    if (j < -2 || j > 2) continue;
                    const float nd = fmaxf(0.f, gc[k].x * g.x + gc[k].y * g.y + gc[k].z * g.z);
                    const float w = fast_ex2(fmaf(-fabsf(gc[k].w - g.w), inv_sigma_p, fmaf(sigma_n, fast_lg2(nd), b3_lg2(i, j))));
                    sd[k] = fmaf(w, v.x, sd[k]);
                    sn[k] = fmaf(w, v.y, sn[k]);
                    wd[k] = fmaf(w, m0, wd[k]);
                    wn[k] = fmaf(w, m1, wn[k]);
                }
            }
        }
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int y = base + k * s;
        if (y >= f.y1) continue;
        const int ofs = y * W + x;
        float2 r = vc[k];
        const bool act = gc[k].w >= 0.f;
        if (act) {
            if (wd[k] > 0.f) r.x = sd[k] / wd[k];
            if (wn[k] > 0.f) r.y = sn[k] / wn[k];
        }
        if (FINAL) {
            float4 od, on;
            if (act) {
                od = make_float4(r.x, r.x, r.x, 1.f);
                on = make_float4(r.y, r.y, r.y, 1.f);
            } else {  // lights / misses: the unfiltered float4 (emit on light pixels, deviceCode.cu:121-123)
                od = f.sto_direct[ofs];
                on = f.sto_no_vis[ofs];
            }
            out_d[ofs] = od;
            out_n[ofs] = on;
            if (ratio_out) {
                const float4 l = ld_stream(ltc + ofs);
                st_stream(ratio_out + ofs, make_float4(l.x * (od.x / on.x), l.y * (od.y / on.y), l.z * (od.z / on.z), 1.f));
            }
        } else {
            vout[ofs] = r;
        }
    }
}

#ifndef ORB_ATROUS4_MINB
#define ORB_ATROUS4_MINB 3
#endif
template <bool FINAL, int STEP>
__global__ void __launch_bounds__(256, ORB_ATROUS4_MINB)
k_denoise_atrous4(FrameView f, const float4* __restrict__ guide, const float2* __restrict__ vin, float2* __restrict__ vout,
                  float4* __restrict__ out_d, float4* __restrict__ out_n, const float4* __restrict__ ltc,
                  float4* __restrict__ ratio_out, int step_rt, float sigma_n, float inv_sigma_p) {
    const int s = STEP > 0 ? STEP : step_rt;
    const int ty = threadIdx.x >> 5;
    const int bx = blockIdx.x * 32;
    const int x = bx + (threadIdx.x & 31);
    int by, base, last;  // the CTA's first row, this thread's first row, the CTA's last row
    if (s <= 8) {
        by = f.y0 + blockIdx.y * 32;
        base = by + (ty / s) * 4 * s + (ty % s);
        last = by + 31;
    } else {
        const int per_span = s >= 8 ? s >> 3 : 1;
        by = f.y0 + ((int)blockIdx.y / per_span) * 4 * s + ((int)blockIdx.y % per_span) * 8;
        base = by + ty;
        last = by + 7 + 3 * s;
    }
    if (x >= f.W || base >= f.y1) return;
    const int reach = 2 * s;
    const bool interior = bx >= reach && bx + 31 + reach < f.W && by >= reach && last + reach < f.H;
    if (interior)
        atrous4_body<FINAL, STEP, true>(f, guide, vin, vout, out_d, out_n, ltc, ratio_out, s, x, base, sigma_n, inv_sigma_p);
    else
        atrous4_body<FINAL, STEP, false>(f, guide, vin, vout, out_d, out_n, ltc, ratio_out, s, x, base, sigma_n, inv_sigma_p);
}
static dim3 atrous4_grid(int W, int rows, int step) {
    if (step <= 8) return dim3((W + 31) / 32, (rows + 31) / 32);
    return dim3((W + 31) / 32, ((rows + 4 * step - 1) / (4 * step)) * (step / 8));
}
typedef void (*Atrous4Kernel)(FrameView, const float4*, const float2*, float2*, float4*, float4*, const float4*, float4*, int,
                              float, float);
template <bool FINAL>
static Atrous4Kernel atrous4_kernel(int step) {
    switch (step) {
        case 1: return k_denoise_atrous4<FINAL, 1>;
        case 2: return k_denoise_atrous4<FINAL, 2>;
        case 4: return k_denoise_atrous4<FINAL, 4>;
        case 8: return k_denoise_atrous4<FINAL, 8>;
        case 16: return k_denoise_atrous4<FINAL, 16>;
        default: return k_denoise_atrous4<FINAL, 0>;
    }
}

typedef void (*AtrousKernel)(FrameView, const float4*, const float2*, float2*, float4*, float4*, int, float, float, float);
template <bool FINAL>
static AtrousKernel atrous_kernel(bool colour, int step) {
    if (colour) return k_denoise_atrous<FINAL, true, 0>;
    switch (step) {
        case 1: return k_denoise_atrous<FINAL, false, 1>;
        case 2: return k_denoise_atrous<FINAL, false, 2>;
        case 4: return k_denoise_atrous<FINAL, false, 4>;
        case 8: return k_denoise_atrous<FINAL, false, 8>;
        case 16: return k_denoise_atrous<FINAL, false, 16>;
        default: return k_denoise_atrous<FINAL, false, 0>;
    }
}

// ---------------------------------------------------------------------------------------------
// ratio combine: ltc_ratio_estimator.py:8  colors = ltc * (stoDirect / stoNoVis)
// ---------------------------------------------------------------------------------------------
__global__ void k_ratio_combine(const float4* __restrict__ ltc, const float4* __restrict__ sd,
                                const float4* __restrict__ sn, float4* __restrict__ out, size_t first, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float4 a = ld_stream(ltc + first + i), b = ld_stream(sd + first + i), c = ld_stream(sn + first + i);
        st_stream(out + first + i, make_float4(a.x * (b.x / c.x), a.y * (b.y / c.y), a.z * (b.z / c.z), 1.f));
    }
}

// flat float arrays (the python RGB layout): n4 float4 groups + scalar tail
__global__ void k_ratio_combine_flat(const float* __restrict__ ltc, const float* __restrict__ sd,
                                     const float* __restrict__ sn, float* __restrict__ out, size_t n) {
    const size_t n4 = n / 4;
    const float4 *a4 = reinterpret_cast<const float4*>(ltc), *b4 = reinterpret_cast<const float4*>(sd),
                 *c4 = reinterpret_cast<const float4*>(sn);
    float4* o4 = reinterpret_cast<float4*>(out);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 a = ld_stream(a4 + i), b = ld_stream(b4 + i), c = ld_stream(c4 + i);
        st_stream(o4 + i, make_float4(a.x * (b.x / c.x), a.y * (b.y / c.y), a.z * (b.z / c.z), a.w * (b.w / c.w)));
    }
    if (blockIdx.x == 0)
        for (size_t i = 4 * n4 + threadIdx.x; i < n; i += blockDim.x) out[i] = ltc[i] * (sd[i] / sn[i]);
}

// ---------------------------------------------------------------------------------------------
// ray casting entry point (tests, tools) and unit kernels
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TRACE_BLOCK)
k_trace_rays(SceneView s, const float* __restrict__ org, const float* __restrict__ dir, size_t n, int brute,
             int* __restrict__ prim, float* __restrict__ t, float* __restrict__ u, float* __restrict__ v, int height,
             DevCounters* cnt) {
    int* stk = s_dyn + threadIdx.x;
    float* stk_t = reinterpret_cast<float*>(s_dyn + height * TRACE_BLOCK) + threadIdx.x;
    const size_t i = (size_t)blockIdx.x * TRACE_BLOCK + threadIdx.x;
    if (i >= n) return;
    const float3 o = v3(org[3 * i], org[3 * i + 1], org[3 * i + 2]), d = v3(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]);
    unsigned nn = 0, nt = 0;
    const Hit h = brute ? trace_brute(s, o, d) : trace_closest(s, o, d, stk, stk_t, TRACE_BLOCK, &nn, &nt);
    prim[i] = h.prim;
    t[i] = h.t;
    u[i] = h.u;
    v[i] = h.v;
    if (cnt) {
        atomicAdd(&cnt->nodes_visited, (unsigned long long)nn);
        atomicAdd(&cnt->tris_tested, (unsigned long long)nt);
    }
}

enum { U_RNG, U_CLIP, U_EDGE, U_FETCH, U_MINV, U_ONB, U_SAMPLE, U_EVAL, U_PDF, U_RGBA, U_SAMPLE_TRI, U_A2W, U_LTC_PIXEL };

__global__ void k_unit(int op, SceneView s, const float* in, float* out) {
    if (threadIdx.x || blockIdx.x) return;
    switch (op) {
        case U_RNG: {  // in: frame, px, py, W, n   out: state0, raw[n], flt[n] (as bits / floats)
            Rng r = rng_seed((int)in[0], (uint32_t)in[1], (uint32_t)in[2], (uint32_t)in[3]);
            const int n = (int)in[4];
            out[0] = __uint_as_float(r.s);
            for (int i = 0; i < n; i++) out[1 + i] = __uint_as_float(rng_next(r));
            for (int i = 0; i < n; i++) out[1 + n + i] = rng_nextf(r);
        } break;
        case U_CLIP: {  // in: a b c quad  out: n_out, form factor
            int n_out = 0;
            const float ff = clipped_form_factor(v3(in[0], in[1], in[2]), v3(in[3], in[4], in[5]), v3(in[6], in[7], in[8]),
                                                 in[9] != 0.f, n_out);
            out[0] = (float)n_out;
            out[1] = ff;
        } break;
        case U_EDGE:
            out[0] = edge_integral(v3(in[0], in[1], in[2]), v3(in[3], in[4], in[5]));
            break;
        case U_FETCH: {
            Mat3 m;
            float amp;
            s.lut.fetch(in[0], in[1], m, amp);
            out[0] = m.r0.x; out[1] = m.r0.y; out[2] = m.r0.z;
            out[3] = m.r1.x; out[4] = m.r1.y; out[5] = m.r1.z;
            out[6] = m.r2.x; out[7] = m.r2.y; out[8] = m.r2.z;
            out[9] = amp;
        } break;
        case U_MINV: {
            Mat3 m;
            m.r0 = v3(in[0], in[1], in[2]); m.r1 = v3(in[3], in[4], in[5]); m.r2 = v3(in[6], in[7], in[8]);
            const Mat3 r = mat3_inverse(m);
            out[0] = r.r0.x; out[1] = r.r0.y; out[2] = r.r0.z;
            out[3] = r.r1.x; out[4] = r.r1.y; out[5] = r.r1.z;
            out[6] = r.r2.x; out[7] = r.r2.y; out[8] = r.r2.z;
        } break;
        case U_ONB: {
            const Mat3 r = onb_rows(v3(in[0], in[1], in[2]));
            out[0] = r.r0.x; out[1] = r.r0.y; out[2] = r.r0.z;
            out[3] = r.r1.x; out[4] = r.r1.y; out[5] = r.r1.z;
            out[6] = r.r2.x; out[7] = r.r2.y; out[8] = r.r2.z;
        } break;
        case U_SAMPLE: {  // in: wo, u1, u2, base, alpha
            float pdf = 0.f;
            const float3 wi = bsdf_sample(v3(in[0], in[1], in[2]), in[3], in[4], pdf, v3(in[5], in[6], in[7]), in[8]);
            out[0] = wi.x; out[1] = wi.y; out[2] = wi.z; out[3] = pdf;
        } break;
        case U_EVAL: {  // in: wi, wo, base, alpha
            const float3 r = bsdf_evaluate(v3(in[0], in[1], in[2]), v3(in[3], in[4], in[5]), v3(in[6], in[7], in[8]), in[9]);
            out[0] = r.x; out[1] = r.y; out[2] = r.z;
        } break;
        case U_PDF:
            out[0] = bsdf_pdf(v3(in[0], in[1], in[2]), v3(in[3], in[4], in[5]), v3(in[6], in[7], in[8]), in[9]);
            break;
        case U_RGBA:
            out[0] = __uint_as_float(make_rgba(v3(in[0], in[1], in[2])));
            break;
        case U_SAMPLE_TRI: {
            const float3 r = sample_point_on_triangle(v3(in[0], in[1], in[2]), v3(in[3], in[4], in[5]),
                                                      v3(in[6], in[7], in[8]), in[9], in[10]);
            out[0] = r.x; out[1] = r.y; out[2] = r.z;
        } break;
        case U_A2W:
            out[0] = pdf_area_to_solid_angle(in[0], in[1], in[2]);
            break;
        case U_LTC_PIXEL: {  // in: cam(3) p(3) n(3) diffuse(3) alpha
            const float3 c = ltc_pixel(s.lut, s.lights, s.n_lights, v3(in[0], in[1], in[2]), v3(in[3], in[4], in[5]),
                                       v3(in[6], in[7], in[8]), v3(in[9], in[10], in[11]), in[12]);
            out[0] = c.x; out[1] = c.y; out[2] = c.z;
        } break;
    }
}

// ---------------------------------------------------------------------------------------------
// host side: context
// ---------------------------------------------------------------------------------------------
struct HostMesh {
    int n_tris;
    MeshRec rec;  // tex filled at commit
    std::vector<uint8_t> tex;
};

// Everything ONE FRAME IN FLIGHT owns.  A context has 1 (default) to 4 of these (orb_set_frame_slots): with several,
// orb_render_async cycles through them, each on its own stream, so the kernels of frame i+1 fill the SMs that
// the persistent traversal kernel of frame i leaves idle while its last rays finish (the fixed tail of every stage,
// profiles/r1k_stage_floors.json: what limits a narrow band of a tile partition) and the issue slots it leaves
// empty while it waits on node fetches.  Frames are independent except for the running sum: `accum` and RGBA8 are
// SHARED between the slots and frame i+1's k_accumulate is ordered behind everything frame i's slot has enqueued.
struct DlSlot { void* dev = nullptr; size_t cap = 0; cudaEvent_t ready = nullptr, done = nullptr; };
struct FrameSlot {
    cudaStream_t stream = nullptr;
    bool own_stream = true;
    void* buf[ORB_BUF_COUNT] = {nullptr};  // ORB_BUF_ACCUM / ORB_BUF_RGBA8: the same pointers in every slot
    float4* colour = nullptr;
    uint32_t* flags = nullptr;
    uint32_t* queue = nullptr;
    DevCounters* cnt = nullptr;
    float4* path_state[2][5] = {{nullptr}};
    float4* stream_rays = nullptr;   // 2 float4 per ray
    float4* stream_hits = nullptr;   // 1 float4 per ray
    size_t stream_cap = 0;
    unsigned* stream_cursor = nullptr;
    float2* dn_val_b = nullptr;  // ping-pong partner of ORB_BUF_DN_VAL
    // orb_buffer_download_rows_async: the frame stream snapshots the rows into a device slot (two per buffer), the
    // copy stream moves the slot to the host while the next frame renders.
    DlSlot dl[ORB_BUF_COUNT][2];
    int dl_next[ORB_BUF_COUNT] = {};
    cudaEvent_t pending_dl[8] = {};  // in-place read-backs of this slot's buffers the slot's next frame has to wait for
    int n_pending_dl = 0;
    cudaEvent_t ev_order = nullptr;  // "everything enqueued on this slot so far" for the other slot's k_accumulate
    cudaEvent_t ev_peer_ready = nullptr, ev_peer_done = nullptr;  // orb_copy_rows_to_peer: the frame is finished / its copies have left
    bool peer_batch_open = false;
    int last_spp = 0;
    bool rendered = false;           // a frame has been enqueued on this slot since the last resize
    uint32_t halo_epoch = 0;         // frames whose packed denoise rows this slot has pushed to its peers (halo.cuh)
};
#define ORB_MAX_SLOTS 8
#define ORB_MAX_PENDING_DL 8

// The context IS its active slot (the FrameSlot base: c->stream, c->buf, ... are the active frame's); the other
// slot is parked in `parked` and swapped in by select_slot().
struct OrbContext : FrameSlot {
    int device = 0;
    int n_sm = 148;
    int n_slots = 1, cur_slot = 0;
    FrameSlot parked[ORB_MAX_SLOTS];
    // scene staging (host)
    std::vector<float> h_pos, h_nrm, h_uv;
    std::vector<int> h_prim_mesh;
    std::vector<HostMesh> meshes;
    std::vector<OrbTriLight> h_lights;
    std::vector<float> h_lut;
    int lut_w = 0, lut_h = 0, lut_quant = 0;
    // scene (device)
    float *d_pos = nullptr, *d_nrm = nullptr, *d_uv = nullptr;
    int* d_prim_mesh = nullptr;
    MeshRec* d_meshes = nullptr;
    std::vector<void*> d_tex;
    TriLight* d_lights = nullptr;
    float4* d_lut = nullptr;
    BvhBuildResult bvh;
    SceneView view;
    bool committed = false;
    QNode* qnodes = nullptr;  // the hierarchy's inner nodes as 32-byte records on a 16-bit grid (bvhq.cuh)
    QGrid qgrid;
    bool geometry_dirty = false;  // orb_scene_update_mesh moved corners that orb_scene_refit has not seen yet
    int height = 1;  // stack entries per thread
    // frame
    int W = 0, H = 0;
    int refill_min = 12;
    int node_min = 16;
    // traversal kernels: 1 = 64-byte fp32 nodes (default), 3 = the 32-byte quantised nodes of bvhq.cuh (ORB_TRACE_VARIANT).
    // Measured with the hybrid stack (profiles/r2c_ab_trace.jsonl): trace stage 1.626 ms against 1.726 ms, primary 0.412 against
    // 0.457 -- the decode (12 LOP3 / PRMT per node) and 4 more registers cost more than the second 256-bit load.
    int trace_variant = 1;
    // > 0: shrink the stream kernel's grid until every lane expects this many rays.  Measured on a 135-row band:
    // 0.428 / 0.430 / 0.453 / 0.606 / 0.961 ms for 0 / 4 / 8 / 16 / 32 -- parallelism beats shorter tails.
    int min_rays_per_lane = 0;
    // persistent kernels: resident CTAs per SM the traversal kernels may take, 0 = all that fit.  With two frames
    // in flight a full persistent grid would keep the other frame's CTAs out until it exits (ORB_TRACE_CTAS_PER_SM).
    int trace_ctas_per_sm = 0;
    cudaStream_t copy_stream = nullptr;
    cudaStream_t peer_stream = nullptr;  // orb_copy_rows_to_peer: copy-engine transfers of finished bands into the peers' frames
    int dl_snapshot_kernel = 1;
    int dl_inplace = 0;  // several frame slots: 1 = read per-slot buffers back in place instead of through a copy-engine snapshot (ORB_DL_INPLACE)
    int resolve_lanes = 1;  // RATIO resolve: one thread per (pixel, sample) for spp 2 / 4 / 8
    int atrous4 = 1;         // a-trous passes as k_denoise_atrous4 (four pixels per thread); 0 = one pixel per thread (ORB_ATROUS4)
    int fused_combine = 0;   // the final a-trous pass also writes RATIO_FINAL (orb_set_fused_combine)
    int path_shade_min = 20;  // k_path_persistent: lanes with an ended ray before the warp shades / refills
    int path_node_min = 8;
    int path_persistent = 1;  // PATH intended mode: one persistent kernel for all bounces (0: per-bounce ray streams)
    int local_stack = 1;  // 0: full-height shared-memory stacks everywhere (round 1: 2.96 vs 2.51 ms for the 1080p shadow stage)
    int l2_persist_mb = 0;  // > 0: pin the BVH nodes in a persisting-L2 window of this many MB (ORB_L2_PERSIST_MB)
    // threads per CTA of the persistent traversal kernels (32 / 64 / 128, ORB_STREAM_BLOCK).  A CTA keeps its registers and
    // shared memory until its LAST warp has finished: with frames in flight, small CTAs hand the SM to the other
    // frame's kernels warp by warp while a traversal kernel's last rays run
    int stream_block = 128;
    int stack_smem = 12;  // HybridStack: stack entries per lane held in shared memory (0, 8, 12 or 16; ORB_STACK_SMEM)
    float* unit_in = nullptr;
    float* unit_out = nullptr;
    Camera cam;
    bool have_cam = false;
    OrbCounters counters;
    // per-stage timing (orb_set_profiling)
    bool profiling = false;
    struct StageSpan { int stage; int launches; cudaEvent_t e0, e1; int slot; };
    std::vector<StageSpan> spans;          // recorded since the last orb_get_stage_times
    std::vector<cudaEvent_t> free_events;  // recycled
    HaloLinks halo;                        // peer-mapped denoise buffers of the other ranks of a tile partition (halo.cuh)
};

static void select_slot(OrbContext* c, int k) {
    if (k == c->cur_slot) return;
    c->parked[c->cur_slot] = *static_cast<FrameSlot*>(c);
    *static_cast<FrameSlot*>(c) = c->parked[k];
    c->cur_slot = k;
}

// RAII span: records an event pair around the launches of one stage when profiling is on
struct StageTimer {
    OrbContext* c;
    OrbContext::StageSpan sp;
    bool on;
    static cudaEvent_t get(OrbContext* c) {
        cudaEvent_t e = nullptr;
        if (!c->free_events.empty()) { e = c->free_events.back(); c->free_events.pop_back(); }
        else cudaEventCreate(&e);
        return e;
    }
    cudaStream_t stream;
    StageTimer(OrbContext* ctx, int stage, int launches, cudaStream_t on_stream = nullptr, bool enable = true)
        : c(ctx), on(ctx->profiling && enable), stream(on_stream ? on_stream : ctx->stream) {
        if (!on) return;
        sp.stage = stage; sp.launches = launches; sp.slot = ctx->cur_slot;
        sp.e0 = get(c); sp.e1 = get(c);
        cudaEventRecord(sp.e0, stream);
    }
    ~StageTimer() {
        if (!on) return;
        cudaEventRecord(sp.e1, stream);
        c->spans.push_back(sp);
    }
};

static size_t buffer_elem_bytes(int id) { return id == ORB_BUF_RGBA8 ? 4 : (id == ORB_BUF_DN_VAL ? 8 : 16); }

static void free_scene(OrbContext* c) {
    cudaFree(c->d_pos); cudaFree(c->d_nrm); cudaFree(c->d_uv); cudaFree(c->d_prim_mesh); cudaFree(c->d_meshes);
    for (void* p : c->d_tex) cudaFree(p);
    c->d_tex.clear();
    cudaFree(c->bvh.tri); cudaFree(c->bvh.tri_b); cudaFree(c->bvh.nodes);
    cudaFree(c->qnodes);
    c->qnodes = nullptr;
    c->d_pos = c->d_nrm = c->d_uv = nullptr;
    c->d_prim_mesh = nullptr;
    c->d_meshes = nullptr;
    c->bvh = BvhBuildResult();
    c->committed = false;
    c->geometry_dirty = false;
}

static size_t buffer_elem_bytes(int id);

static void free_slot_frame(FrameSlot& f, bool owns_shared) {
    for (int i = 0; i < ORB_BUF_COUNT; i++) {
        if (owns_shared || (i != ORB_BUF_ACCUM && i != ORB_BUF_RGBA8)) cudaFree(f.buf[i]);
        f.buf[i] = nullptr;
    }
    cudaFree(f.colour); cudaFree(f.flags); cudaFree(f.queue);
    for (int a = 0; a < 2; a++) for (int b = 0; b < 5; b++) { cudaFree(f.path_state[a][b]); f.path_state[a][b] = nullptr; }
    cudaFree(f.dn_val_b);
    cudaFree(f.stream_rays); cudaFree(f.stream_hits);
    f.stream_rays = f.stream_hits = nullptr;
    f.stream_cap = 0;
    f.colour = nullptr; f.flags = nullptr; f.queue = nullptr; f.dn_val_b = nullptr;
    f.rendered = false;
}

static void free_frame(OrbContext* c) {
    select_slot(c, 0);
    free_slot_frame(*c, true);
    for (int k = 1; k < ORB_MAX_SLOTS; k++) free_slot_frame(c->parked[k], false);
    c->W = c->H = 0;
}

// stream and counters of a slot (everything of a slot that does not depend on the frame size)
static int init_slot(FrameSlot& f) {
    if (!f.stream) {
        CK(cudaStreamCreateWithFlags(&f.stream, cudaStreamNonBlocking));
        f.own_stream = true;
    }
    if (!f.cnt) {
        CK(cudaMalloc(&f.cnt, sizeof(DevCounters)));
        CK(cudaMemset(f.cnt, 0, sizeof(DevCounters)));
    }
    return ORB_OK;
}

static void destroy_slot(FrameSlot& f) {
    if (f.stream) cudaStreamSynchronize(f.stream);
    cudaFree(f.stream_cursor);
    cudaFree(f.cnt);
    f.stream_cursor = nullptr;
    f.cnt = nullptr;
    for (auto& per_buf : f.dl)
        for (auto& sl : per_buf) {
            cudaFree(sl.dev);
            if (sl.ready) cudaEventDestroy(sl.ready);
            if (sl.done) cudaEventDestroy(sl.done);
            sl = DlSlot();
        }
    if (f.ev_order) cudaEventDestroy(f.ev_order);
    f.ev_order = nullptr;
    if (f.ev_peer_ready) cudaEventDestroy(f.ev_peer_ready);
    if (f.ev_peer_done) cudaEventDestroy(f.ev_peer_done);
    f.ev_peer_ready = f.ev_peer_done = nullptr;
    f.peer_batch_open = false;
    if (f.own_stream && f.stream) cudaStreamDestroy(f.stream);
    f.stream = nullptr;
}

// frame buffers of one slot; slot 0 owns the shared accum / RGBA8
static int alloc_slot_frame(FrameSlot& f, const FrameSlot* shared_from, int W, int H) {
    const size_t n = (size_t)W * H;
    for (int i = 0; i < ORB_BUF_COUNT; i++) {
        if (shared_from && (i == ORB_BUF_ACCUM || i == ORB_BUF_RGBA8)) {
            f.buf[i] = shared_from->buf[i];
            continue;
        }
        CK(cudaMalloc(&f.buf[i], n * buffer_elem_bytes(i)));
        CK(cudaMemset(f.buf[i], 0, n * buffer_elem_bytes(i)));
    }
    CK(cudaMalloc(&f.colour, n * 16)); CK(cudaMemset(f.colour, 0, n * 16));
    CK(cudaMalloc(&f.flags, n * 4)); CK(cudaMemset(f.flags, 0, n * 4));
    CK(cudaMalloc(&f.queue, n * 4));
    return ORB_OK;
}

static void refresh_view(OrbContext* c) {
    SceneView& v = c->view;
    v.tri = c->bvh.tri;
    v.tri_b = c->bvh.tri_b;
    v.nodes = c->bvh.nodes;
    v.root_ref = c->bvh.root_ref;
    v.n_tris = (int)c->h_prim_mesh.size();
    v.n_nodes = c->bvh.n_nodes;
    v.pos = c->d_pos;
    v.nrm = c->d_nrm;
    v.uv = c->d_uv;
    v.prim_mesh = c->d_prim_mesh;
    v.meshes = c->d_meshes;
    v.lights = c->d_lights;
    v.n_lights = (int)c->h_lights.size();
    v.lut.data = c->d_lut;
    v.lut.w = c->lut_w;
    v.lut.h = c->lut_h;
    v.lut.quantise = c->lut_quant;
    if (!c->committed) v.n_tris = 0;
}

static FrameView frame_view(OrbContext* c, int y0, int y1) {
    FrameView f;
    f.W = c->W; f.H = c->H; f.y0 = y0; f.y1 = y1;
    f.position = (float4*)c->buf[ORB_BUF_POSITION]; f.normal = (float4*)c->buf[ORB_BUF_NORMAL];
    f.uv = (float4*)c->buf[ORB_BUF_UV]; f.albedo = (float4*)c->buf[ORB_BUF_ALBEDO];
    f.alpha = (float4*)c->buf[ORB_BUF_ALPHA]; f.material_id = (float4*)c->buf[ORB_BUF_MATERIAL_ID];
    f.ltc = (float4*)c->buf[ORB_BUF_LTC]; f.sto_direct = (float4*)c->buf[ORB_BUF_STO_DIRECT];
    f.sto_no_vis = (float4*)c->buf[ORB_BUF_STO_NO_VIS]; f.accum = (float4*)c->buf[ORB_BUF_ACCUM];
    f.hit = (float4*)c->buf[ORB_BUF_HIT]; f.colour = c->colour; f.rgba8 = (uint32_t*)c->buf[ORB_BUF_RGBA8];
    f.flags = c->flags; f.queue = c->queue; f.cnt = c->cnt;
    return f;
}

static bool rows_ok(OrbContext* c, int& y0, int& y1) {
    if (y0 == 0 && y1 == 0) y1 = c->H;
    return y0 >= 0 && y1 <= c->H && y0 < y1;
}

static size_t trace_smem(OrbContext* c) { return (size_t)c->height * TRACE_BLOCK * 8; }

// The attribute belongs to the (kernel, device) pair, not to a context: with two committed contexts on a
// device, the one with the deeper BVH must keep its allowance, so the limit only ever grows.
template <class K>
static cudaError_t allow_smem(K kernel, size_t bytes) {
    static std::mutex mu;
    static std::map<std::pair<const void*, int>, size_t> granted;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lock(mu);
    size_t& g = granted[std::make_pair((const void*)kernel, dev)];
    if (bytes <= g) return cudaSuccess;
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e == cudaSuccess) g = bytes;
    return e;
}

// persistent grid: resident CTAs per SM x SM count (`cap` > 0 limits the CTAs per SM: room for another frame's kernels)
template <class K>
static int persistent_grid(OrbContext* c, K kernel, int block, size_t smem, int cap = 0) {
    int per_sm = 1;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, block, smem) != cudaSuccess || per_sm < 1) per_sm = 1;
    if (cap > 0) per_sm = std::min(per_sm, cap);
    return per_sm * c->n_sm;
}

// ---- which instantiation of a traversal kernel a context launches --------------------------------------------
// stack form (template parameter SD of the kernels) and the dynamic shared memory it needs
static int stack_mode(OrbContext* c) { return (c->local_stack && c->height <= ORB_LOCAL_STACK_MAX) ? c->stack_smem : -1; }
static size_t stack_bytes(OrbContext* c, int sd, int block = TRACE_BLOCK) { return sd < 0 ? (size_t)c->height * block * 8 : (size_t)sd * 8 * block; }
// the 64-byte fp32 nodes (default) or the 32-byte quantised ones (ORB_TRACE_VARIANT=3)
static bool use_qnodes(OrbContext* c) { return c->trace_variant == 3 && c->qnodes != nullptr; }

typedef void (*StreamKernel)(SceneView, const float4*, float4*, const unsigned*, unsigned, unsigned*, int, int, int,
                             unsigned long long*, const QNode*, QGrid);
typedef void (*PrimaryKernel)(SceneView, FrameView, Camera, int, int, int, const QNode*, QGrid);
typedef void (*PathKernel)(SceneView, FrameView, float3, int, int, unsigned*, int, int, int, const QNode*, QGrid);
template <int SD> static StreamKernel stream_kernel_sd(bool count, bool qn) {
    return qn ? (count ? k_trace_stream_spec<SD, true, true> : k_trace_stream_spec<SD, false, true>)
              : (count ? k_trace_stream_spec<SD, true, false> : k_trace_stream_spec<SD, false, false>);
}
template <int SD> static PathKernel path_kernel_sd(bool count, bool qn) {
    return qn ? (count ? k_path_persistent<SD, true, true> : k_path_persistent<SD, false, true>)
              : (count ? k_path_persistent<SD, true, false> : k_path_persistent<SD, false, false>);
}
template <int SD> static PrimaryKernel primary_kernel_sd(bool qn) { return qn ? k_primary<SD, true> : k_primary<SD, false>; }
#define ORB_SD_SWITCH(sd, F, ...)                \
    switch (sd) {                                \
        case 0: return F<0>(__VA_ARGS__);        \
        case 8: return F<8>(__VA_ARGS__);        \
        case 12: return F<12>(__VA_ARGS__);      \
        case 16: return F<16>(__VA_ARGS__);      \
        default: return F<-1>(__VA_ARGS__);      \
    }
static StreamKernel stream_kernel(int sd, bool count, bool qn) { ORB_SD_SWITCH(sd, stream_kernel_sd, count, qn) }
static PathKernel path_kernel(int sd, bool count, bool qn) { ORB_SD_SWITCH(sd, path_kernel_sd, count, qn) }
static PrimaryKernel primary_kernel(int sd, bool qn) { ORB_SD_SWITCH(sd, primary_kernel_sd, qn) }

__global__ void k_quantise_nodes(const BvhNode* __restrict__ nodes, int n, QGrid g, QNode* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = q_encode(nodes[i], g);
}

// (re-)encode the inner nodes as the 32-byte records the traversal kernels read; after a build and after a refit.
static int build_qnodes(OrbContext* c) {
    if (c->bvh.n_nodes <= 0 || c->bvh.root_ref < 0) return ORB_OK;
    BvhNode root;
    CK(cudaMemcpyAsync(&root, c->bvh.nodes + c->bvh.root_ref, sizeof root, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    c->qgrid = qgrid_from_root(root);
    if (!c->qnodes) CK(cudaMalloc(&c->qnodes, sizeof(QNode) * (size_t)c->bvh.n_nodes));
    k_quantise_nodes<<<(c->bvh.n_nodes + 255) / 256, 256, 0, c->stream>>>(c->bvh.nodes, c->bvh.n_nodes, c->qgrid, c->qnodes);
    CK(cudaGetLastError());
    c->counters.kernel_launches++;
    return ORB_OK;
}

// ORB_L2_PERSIST_MB: an access-policy window over the node array on every frame stream -- node lines are kept in
// the persisting part of L2 while the ray stream and the frame buffers pass through the rest.
static int apply_l2_window(OrbContext* c) {
    if (c->l2_persist_mb <= 0 || c->bvh.n_nodes <= 0) return ORB_OK;
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, c->device));
    const bool qn = use_qnodes(c);
    const void* base = qn ? (const void*)c->qnodes : (const void*)c->bvh.nodes;
    size_t bytes = (size_t)c->bvh.n_nodes * (qn ? sizeof(QNode) : sizeof(BvhNode));
    size_t want = std::min<size_t>((size_t)c->l2_persist_mb << 20, (size_t)prop.persistingL2CacheMaxSize);
    if (want == 0) return ORB_OK;
    CK(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want));
    bytes = std::min<size_t>(bytes, (size_t)prop.accessPolicyMaxWindowSize);
    cudaStreamAttrValue a;
    memset(&a, 0, sizeof a);
    a.accessPolicyWindow.base_ptr = const_cast<void*>(base);
    a.accessPolicyWindow.num_bytes = bytes;
    a.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)want / (double)bytes);
    a.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    a.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    for (int k = 0; k < c->n_slots; k++) {
        cudaStream_t st = k == c->cur_slot ? c->stream : c->parked[k].stream;
        if (st) CK(cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &a));
    }
    fprintf(stderr, "orb200: L2 window %.1f MB over the nodes (persisting L2 %.1f MB, hit ratio %.2f)\n", bytes / 1048576.0,
            want / 1048576.0, a.accessPolicyWindow.hitRatio);
    return ORB_OK;
}

extern "C" {

int orb_api_version(void) { return ORB_API_VERSION; }
const char* orb_last_error(void) { return g_err.c_str(); }

int orb_create(int device, OrbContext** out) {
    NEED(out, "orb_create: out is NULL");
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
        cudaGetLastError();
        return fail(ORB_ERR_NO_DEVICE, "orb_create: no CUDA device (this library has no CPU path)");
    }
    NEED(device >= 0 && device < n, "orb_create: bad device index");
    CK(cudaSetDevice(device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    OrbContext* c = new OrbContext();
    c->device = device;
    c->n_sm = prop.multiProcessorCount;
    {
        int rc = init_slot(*c);
        if (rc) return rc;
    }
    CK(cudaMalloc(&c->unit_in, 64 * sizeof(float)));
    CK(cudaMalloc(&c->unit_out, 4096 * sizeof(float)));
    memset(&c->counters, 0, sizeof(c->counters));
    memset(&c->view, 0, sizeof(c->view));
    if (const char* e = getenv("ORB_REFILL_MIN")) c->refill_min = std::max(1, std::min(32, atoi(e)));  // tuning knobs
    if (const char* e = getenv("ORB_LOCAL_STACK")) c->local_stack = atoi(e);
    if (const char* e = getenv("ORB_STACK_SMEM")) { const int v = atoi(e); if (v == 0 || v == 8 || v == 12 || v == 16) c->stack_smem = v; }
    if (const char* e = getenv("ORB_L2_PERSIST_MB")) c->l2_persist_mb = std::max(0, atoi(e));
    if (const char* e = getenv("ORB_STREAM_BLOCK")) { const int v = atoi(e); if (v == 32 || v == 64 || v == 128) c->stream_block = v; }
    if (const char* e = getenv("ORB_TRACE_CTAS_PER_SM")) c->trace_ctas_per_sm = std::max(0, atoi(e));
    if (const char* e = getenv("ORB_NODE_MIN")) c->node_min = std::max(1, std::min(32, atoi(e)));
    if (const char* e = getenv("ORB_TRACE_VARIANT")) c->trace_variant = atoi(e);
    if (const char* e = getenv("ORB_PATH_PERSISTENT")) c->path_persistent = atoi(e);
    if (const char* e = getenv("ORB_RESOLVE_LANES")) c->resolve_lanes = atoi(e);
    if (const char* e = getenv("ORB_MIN_RAYS_PER_LANE")) c->min_rays_per_lane = atoi(e);
    if (const char* e = getenv("ORB_DL_SNAPSHOT_KERNEL")) c->dl_snapshot_kernel = atoi(e);
    if (const char* e = getenv("ORB_DL_INPLACE")) c->dl_inplace = atoi(e);
    if (const char* e = getenv("ORB_ATROUS4")) c->atrous4 = atoi(e);
    if (const char* e = getenv("ORB_PATH_SHADE_MIN")) c->path_shade_min = std::max(1, std::min(32, atoi(e)));
    if (const char* e = getenv("ORB_PATH_NODE_MIN")) c->path_node_min = std::max(1, std::min(32, atoi(e)));
    *out = c;
    return ORB_OK;
}

static void dl_trace_report();
void orb_destroy(OrbContext* c) {
    if (!c) return;
    dl_trace_report();
    cudaSetDevice(c->device);
    select_slot(c, 0);
    cudaStreamSynchronize(c->stream);
    for (int k = 1; k < ORB_MAX_SLOTS; k++)
        if (c->parked[k].stream) cudaStreamSynchronize(c->parked[k].stream);
    if (c->copy_stream) { cudaStreamSynchronize(c->copy_stream); cudaStreamDestroy(c->copy_stream); }
    if (c->peer_stream) { cudaStreamSynchronize(c->peer_stream); cudaStreamDestroy(c->peer_stream); }
    halo_disconnect(c->halo);
    cudaFree(c->halo.flags);
    free_scene(c);
    free_frame(c);
    for (auto& sp : c->spans) { cudaEventDestroy(sp.e0); cudaEventDestroy(sp.e1); }
    for (auto e : c->free_events) cudaEventDestroy(e);
    cudaFree(c->d_lights); cudaFree(c->d_lut); cudaFree(c->unit_in); cudaFree(c->unit_out);
    destroy_slot(*c);
    for (int k = 1; k < ORB_MAX_SLOTS; k++) destroy_slot(c->parked[k]);
    delete c;
}

// Frames in flight (1 to 4).  With n > 1, consecutive orb_render_async calls cycle through n sets of frame buffers
// and n streams; `accum` / RGBA8 stay shared and accumulate in call order.  Re-allocates the frame buffers.
int orb_set_frame_slots(OrbContext* c, int32_t n) {
    NEED(c && n >= 1 && n <= ORB_MAX_SLOTS, "orb_set_frame_slots: 1 to 8 slots");
    CK(cudaSetDevice(c->device));
    int rc = orb_sync(c);
    if (rc) return rc;
    NEED(c->halo.n_peers == 0, "orb_set_frame_slots: disconnect the halo peers first (their buffers are re-allocated)");
    const int W = c->W, H = c->H;
    free_frame(c);  // back on slot 0
    for (int k = 1; k < ORB_MAX_SLOTS; k++) {
        if (k < n) {
            rc = init_slot(c->parked[k]);
            if (rc) return rc;
        } else {
            destroy_slot(c->parked[k]);
        }
    }
    c->n_slots = n;
    return W > 0 ? orb_resize(c, W, H) : ORB_OK;
}

int32_t orb_frame_slot(OrbContext* c) { return c ? c->cur_slot : -1; }
void* orb_slot_stream(OrbContext* c, int32_t k) {
    if (!c || k < 0 || k >= c->n_slots) return nullptr;
    return (void*)(k == c->cur_slot ? c->stream : c->parked[k].stream);
}
int32_t orb_frame_slots(OrbContext* c) { return c ? c->n_slots : 0; }

void* orb_stream(OrbContext* c) { return c ? (void*)c->stream : nullptr; }

int orb_set_stream(OrbContext* c, void* stream) {
    NEED(c, "orb_set_stream: ctx is NULL");
    CK(cudaSetDevice(c->device));
    CK(cudaStreamSynchronize(c->stream));
    if (c->own_stream) cudaStreamDestroy(c->stream);
    c->stream = (cudaStream_t)stream;
    c->own_stream = false;
    return ORB_OK;
}

int orb_sync(OrbContext* c) {  // every frame in flight
    NEED(c, "orb_sync: ctx is NULL");
    CK(cudaSetDevice(c->device));
    CK(cudaStreamSynchronize(c->stream));
    for (int k = 0; k < c->n_slots; k++)
        if (k != c->cur_slot && c->parked[k].stream) CK(cudaStreamSynchronize(c->parked[k].stream));
    return ORB_OK;
}

int orb_set_profiling(OrbContext* c, int enable) {
    NEED(c, "orb_set_profiling: ctx is NULL");
    c->profiling = enable != 0;
    return ORB_OK;
}

// The recorded spans themselves (a timeline of the frames in flight), times in ms relative to the first span's start;
// does not clear them (orb_get_stage_times does).  Returns the number of spans through *n (at most `max` are written).
int orb_get_stage_spans(OrbContext* c, OrbStageSpan* out, int32_t max, int32_t* n) {
    NEED(c && n && (out || max == 0), "orb_get_stage_spans: NULL argument");
    {
        int rc = orb_sync(c);
        if (rc) return rc;
    }
    *n = (int32_t)c->spans.size();
    for (int i = 0; i < (int)c->spans.size() && i < max; i++) {
        const auto& sp = c->spans[i];
        float a = 0.f, b = 0.f;
        CK(cudaEventElapsedTime(&a, c->spans[0].e0, sp.e0));
        CK(cudaEventElapsedTime(&b, c->spans[0].e0, sp.e1));
        out[i].stage = sp.stage;
        out[i].slot = sp.slot;
        out[i].t0_ms = a;
        out[i].t1_ms = b;
    }
    return ORB_OK;
}

int orb_get_stage_times(OrbContext* c, OrbStageTimes* out) {
    NEED(c && out, "orb_get_stage_times: NULL argument");
    {
        int rc = orb_sync(c);
        if (rc) return rc;
    }
    memset(out, 0, sizeof(*out));
    for (auto& sp : c->spans) {
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, sp.e0, sp.e1));
        out->ms[sp.stage] += (double)ms;
        out->launches[sp.stage] += (uint64_t)sp.launches;
        c->free_events.push_back(sp.e0);
        c->free_events.push_back(sp.e1);
    }
    c->spans.clear();
    return ORB_OK;
}

// ---- scene -----------------------------------------------------------------------------------
int orb_scene_clear(OrbContext* c) {
    NEED(c, "orb_scene_clear: ctx is NULL");
    CK(cudaSetDevice(c->device));
    CK(cudaStreamSynchronize(c->stream));
    free_scene(c);
    c->h_pos.clear(); c->h_nrm.clear(); c->h_uv.clear(); c->h_prim_mesh.clear(); c->meshes.clear();
    refresh_view(c);
    return ORB_OK;
}

int orb_scene_add_mesh(OrbContext* c, const OrbMeshDesc* m) {
    NEED(c && m, "orb_scene_add_mesh: NULL argument");
    NEED(!c->committed, "orb_scene_add_mesh: scene already committed (orb_scene_clear first)");
    NEED(m->n_triangles >= 0 && m->n_vertices >= 0, "orb_scene_add_mesh: negative count");
    NEED(m->n_triangles == 0 || (m->vertex && m->normal && m->texcoord && m->index), "orb_scene_add_mesh: NULL array");
    NEED(c->h_prim_mesh.size() + (size_t)m->n_triangles < (1u << 28), "orb_scene_add_mesh: too many triangles");
    const int mesh_id = (int)c->meshes.size();
    for (int t = 0; t < m->n_triangles; t++)
        for (int k = 0; k < 3; k++) {
            const int vi = m->index[3 * t + k];
            NEED(vi >= 0 && vi < m->n_vertices, "orb_scene_add_mesh: index out of range");
        }
    // un-index: 3 corners per global primitive (gathered only at the closest hit)
    const size_t base = c->h_prim_mesh.size();
    c->h_pos.resize(9 * (base + m->n_triangles));
    c->h_nrm.resize(9 * (base + m->n_triangles));
    c->h_uv.resize(6 * (base + m->n_triangles));
    for (int t = 0; t < m->n_triangles; t++) {
        for (int k = 0; k < 3; k++) {
            const int vi = m->index[3 * t + k];
            for (int a = 0; a < 3; a++) {
                c->h_pos[9 * (base + t) + 3 * k + a] = m->vertex[3 * (size_t)vi + a];
                c->h_nrm[9 * (base + t) + 3 * k + a] = m->normal[3 * (size_t)vi + a];
            }
            c->h_uv[6 * (base + t) + 2 * k] = m->texcoord[2 * (size_t)vi];
            c->h_uv[6 * (base + t) + 2 * k + 1] = m->texcoord[2 * (size_t)vi + 1];
        }
        c->h_prim_mesh.push_back(mesh_id);
    }
    HostMesh hm;
    hm.n_tris = m->n_triangles;
    memset(&hm.rec, 0, sizeof(hm.rec));
    hm.rec.diffuse = v3(m->diffuse[0], m->diffuse[1], m->diffuse[2]);
    hm.rec.alpha = m->alpha;
    hm.rec.emit = v3(m->emit[0], m->emit[1], m->emit[2]);
    hm.rec.is_light = m->is_light != 0;
    hm.rec.material_id = m->material_id;
    if (m->diffuse_tex_rgba8) {
        NEED(m->tex_w > 0 && m->tex_h > 0, "orb_scene_add_mesh: bad texture size");
        hm.rec.tex_w = m->tex_w;
        hm.rec.tex_h = m->tex_h;
        hm.tex.assign(m->diffuse_tex_rgba8, m->diffuse_tex_rgba8 + 4 * (size_t)m->tex_w * m->tex_h);
    }
    c->meshes.push_back(std::move(hm));
    return ORB_OK;
}

int orb_scene_set_lights(OrbContext* c, const OrbTriLight* lights, int32_t n) {
    NEED(c && n >= 0 && (n == 0 || lights), "orb_scene_set_lights: bad argument");
    CK(cudaSetDevice(c->device));
    CK(cudaStreamSynchronize(c->stream));
    c->h_lights.assign(lights, lights + n);
    cudaFree(c->d_lights);
    c->d_lights = nullptr;
    if (n) {
        CK(cudaMalloc(&c->d_lights, sizeof(TriLight) * n));
        CK(cudaMemcpy(c->d_lights, lights, sizeof(TriLight) * n, cudaMemcpyHostToDevice));
    }
    refresh_view(c);
    return ORB_OK;
}

int orb_scene_set_ltc_lut(OrbContext* c, const float* m1, const float* m2, const float* m3, int32_t w, int32_t h,
                          int32_t quantise) {
    NEED(c && m1 && m2 && m3 && w > 0 && h > 0, "orb_scene_set_ltc_lut: bad argument");
    CK(cudaSetDevice(c->device));
    CK(cudaStreamSynchronize(c->stream));
    const size_t n = 4 * (size_t)w * h;
    c->h_lut.resize(3 * n);
    memcpy(c->h_lut.data(), m1, n * 4);
    memcpy(c->h_lut.data() + n, m2, n * 4);
    memcpy(c->h_lut.data() + 2 * n, m3, n * 4);
    cudaFree(c->d_lut);
    CK(cudaMalloc(&c->d_lut, 3 * n * 4));
    CK(cudaMemcpy(c->d_lut, c->h_lut.data(), 3 * n * 4, cudaMemcpyHostToDevice));
    c->lut_w = w; c->lut_h = h; c->lut_quant = quantise;
    refresh_view(c);
    return ORB_OK;
}

int orb_scene_commit(OrbContext* c, const OrbBvhOptions* opt) {
    NEED(c, "orb_scene_commit: ctx is NULL");
    NEED(!c->committed, "orb_scene_commit: already committed");
    CK(cudaSetDevice(c->device));
    const size_t n = c->h_prim_mesh.size();
    if (n) {
        CK(cudaMalloc(&c->d_pos, 36 * n)); CK(cudaMalloc(&c->d_nrm, 36 * n)); CK(cudaMalloc(&c->d_uv, 24 * n));
        CK(cudaMalloc(&c->d_prim_mesh, 4 * n));
        CK(cudaMemcpyAsync(c->d_pos, c->h_pos.data(), 36 * n, cudaMemcpyHostToDevice, c->stream));
        CK(cudaMemcpyAsync(c->d_nrm, c->h_nrm.data(), 36 * n, cudaMemcpyHostToDevice, c->stream));
        CK(cudaMemcpyAsync(c->d_uv, c->h_uv.data(), 24 * n, cudaMemcpyHostToDevice, c->stream));
        CK(cudaMemcpyAsync(c->d_prim_mesh, c->h_prim_mesh.data(), 4 * n, cudaMemcpyHostToDevice, c->stream));
    }
    std::vector<MeshRec> recs;
    for (auto& hm : c->meshes) {
        MeshRec r = hm.rec;
        r.tex = nullptr;
        if (!hm.tex.empty()) {
            void* p = nullptr;
            CK(cudaMalloc(&p, hm.tex.size()));
            CK(cudaMemcpyAsync(p, hm.tex.data(), hm.tex.size(), cudaMemcpyHostToDevice, c->stream));
            c->d_tex.push_back(p);
            r.tex = (const uchar4*)p;
        }
        recs.push_back(r);
    }
    if (!recs.empty()) {
        CK(cudaMalloc(&c->d_meshes, sizeof(MeshRec) * recs.size()));
        CK(cudaMemcpyAsync(c->d_meshes, recs.data(), sizeof(MeshRec) * recs.size(), cudaMemcpyHostToDevice, c->stream));
    }
    CK(cudaStreamSynchronize(c->stream));
    int max_leaf = opt ? opt->max_leaf_tris : 0;
    if (const char* e = getenv("ORB_MAX_LEAF")) max_leaf = atoi(e);  // tuning knob
    int builder = opt ? opt->builder : ORB_BVH_SAH;
    if (const char* e = getenv("ORB_BVH_BUILDER")) builder = atoi(e);
    NEED(builder == ORB_BVH_SAH || builder == ORB_BVH_LBVH, "orb_scene_commit: unknown BVH builder");
    cudaEvent_t eb0, eb1;
    CK(cudaEventCreate(&eb0));
    CK(cudaEventCreate(&eb1));
    CK(cudaEventRecord(eb0, c->stream));
    if (builder == ORB_BVH_LBVH) CK(build_bvh(c->d_pos, (int)n, max_leaf, c->stream, c->bvh));
    else CK(build_bvh_sah(c->d_pos, (int)n, max_leaf, c->stream, c->bvh));
    {
        int rcq = build_qnodes(c);
        if (rcq) return rcq;
    }
    CK(cudaEventRecord(eb1, c->stream));
    CK(cudaEventSynchronize(eb1));
    float build_ms = 0.f;
    CK(cudaEventElapsedTime(&build_ms, eb0, eb1));
    cudaEventDestroy(eb0);
    cudaEventDestroy(eb1);
    c->counters.bvh_build_us = (uint64_t)(build_ms * 1000.f);
    c->counters.kernel_launches += c->bvh.launches;
    c->counters.bvh_nodes = (uint64_t)c->bvh.n_nodes;
    c->counters.bvh_triangles = (uint64_t)n;
    c->height = c->bvh.height + 2;
    if (trace_smem(c) > 200 * 1024)
        return fail(ORB_ERR_INVALID, "orb_scene_commit: BVH too deep for the shared-memory traversal stack");
    CK(allow_smem(k_path_bounce<true>, trace_smem(c)));
    CK(allow_smem(k_path_bounce<false>, trace_smem(c)));
    CK(allow_smem(k_trace_rays, trace_smem(c)));
    c->committed = true;
    {
        int rcw = apply_l2_window(c);
        if (rcw) return rcw;
    }
    // host staging is no longer needed
    std::vector<float>().swap(c->h_pos);
    std::vector<float>().swap(c->h_nrm);
    std::vector<float>().swap(c->h_uv);
    refresh_view(c);
    return ORB_OK;
}

// ---- animated geometry: new corners for a mesh of the committed scene, then a refit ----------
int orb_scene_update_mesh(OrbContext* c, int32_t mesh_index, const float* vertex, const float* normal, int32_t n_vertices,
                          const int32_t* index, int32_t n_triangles) {
    NEED(c, "orb_scene_update_mesh: ctx is NULL");
    NEED(c->committed, "orb_scene_update_mesh: scene not committed");
    NEED(mesh_index >= 0 && (size_t)mesh_index < c->meshes.size(), "orb_scene_update_mesh: no such mesh");
    NEED(n_triangles == c->meshes[mesh_index].n_tris,
         "orb_scene_update_mesh: triangle count differs from the mesh that was added (topology is fixed; rebuild with "
         "orb_scene_clear + orb_scene_commit)");
    if (n_triangles == 0) return ORB_OK;
    NEED(vertex && normal && index && n_vertices > 0, "orb_scene_update_mesh: NULL array");
    size_t base = 0;
    for (int m = 0; m < mesh_index; m++) base += (size_t)c->meshes[m].n_tris;
    std::vector<float> pos(9 * (size_t)n_triangles), nrm(9 * (size_t)n_triangles);
    for (int t = 0; t < n_triangles; t++)
        for (int k = 0; k < 3; k++) {
            const int vi = index[3 * t + k];
            NEED(vi >= 0 && vi < n_vertices, "orb_scene_update_mesh: index out of range");
            for (int a = 0; a < 3; a++) {
                pos[9 * (size_t)t + 3 * k + a] = vertex[3 * (size_t)vi + a];
                nrm[9 * (size_t)t + 3 * k + a] = normal[3 * (size_t)vi + a];
            }
        }
    CK(cudaSetDevice(c->device));
    CK(cudaStreamSynchronize(c->stream));  // frames in flight still read the old corners
    c->geometry_dirty = true;
    CK(cudaMemcpy(c->d_pos + 9 * base, pos.data(), 4 * pos.size(), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(c->d_nrm + 9 * base, nrm.data(), 4 * nrm.size(), cudaMemcpyHostToDevice));
    return ORB_OK;
}

int orb_scene_refit(OrbContext* c) {
    NEED(c, "orb_scene_refit: ctx is NULL");
    NEED(c->committed, "orb_scene_refit: scene not committed");
    CK(cudaSetDevice(c->device));
    struct EventPair {  // destroyed on every return path
        cudaEvent_t a = nullptr, b = nullptr;
        ~EventPair() {
            if (a) cudaEventDestroy(a);
            if (b) cudaEventDestroy(b);
        }
    } ev;
    CK(cudaEventCreate(&ev.a));
    CK(cudaEventCreate(&ev.b));
    cudaEvent_t e0 = ev.a, e1 = ev.b;
    CK(cudaEventRecord(e0, c->stream));
    const int before = c->bvh.launches;
    CK(refit_bvh(c->d_pos, (int)c->h_prim_mesh.size(), c->stream, c->bvh));
    {
        int rcq = build_qnodes(c);
        if (rcq) return rcq;
    }
    CK(cudaEventRecord(e1, c->stream));
    CK(cudaEventSynchronize(e1));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    c->counters.bvh_refit_us = (uint64_t)(ms * 1000.f);
    c->counters.kernel_launches += (uint64_t)(c->bvh.launches - before);
    c->geometry_dirty = false;
    return ORB_OK;
}

// ---- camera / frame --------------------------------------------------------------------------
int orb_set_camera(OrbContext* c, const float pos[3], const float d00[3], const float du[3], const float dv[3]) {
    NEED(c && pos && d00 && du && dv, "orb_set_camera: NULL argument");
    c->cam.pos = v3(pos[0], pos[1], pos[2]);
    c->cam.d00 = v3(d00[0], d00[1], d00[2]);
    c->cam.du = v3(du[0], du[1], du[2]);
    c->cam.dv = v3(dv[0], dv[1], dv[2]);
    c->have_cam = true;
    return ORB_OK;
}

int orb_resize(OrbContext* c, int32_t W, int32_t H) {
    NEED(c && W > 0 && H > 0, "orb_resize: bad size");
    NEED((size_t)W * (size_t)H < ((size_t)1 << 31), "orb_resize: frames are limited to 2^31 - 1 pixels (32-bit pixel offsets)");
    CK(cudaSetDevice(c->device));
    {
        int rc = orb_sync(c);
        if (rc) return rc;
    }
    NEED(c->halo.n_peers == 0, "orb_resize: disconnect the halo peers first (their buffers are re-allocated)");
    free_frame(c);  // leaves slot 0 active
    {
        int rc = alloc_slot_frame(*c, nullptr, W, H);
        if (rc) return rc;
        for (int k = 1; k < c->n_slots; k++) {
            rc = alloc_slot_frame(c->parked[k], c, W, H);
            if (rc) return rc;
        }
    }
    c->W = W; c->H = H;
    return ORB_OK;
}

// ---- launch ----------------------------------------------------------------------------------
static int ensure_stream(OrbContext* c, size_t n_rays) {
    if (!c->stream_cursor) CK(cudaMalloc(&c->stream_cursor, 64));
    if (c->stream_cap >= n_rays) return ORB_OK;
    cudaFree(c->stream_rays);
    cudaFree(c->stream_hits);
    c->stream_rays = c->stream_hits = nullptr;
    c->stream_cap = 0;
    // growth with headroom: the bands of a tile partition move by a few rows when they are re-balanced, and a
    // cudaFree / cudaMalloc pair synchronises the device (milliseconds beside peers' mapped memory)
    const size_t cap = std::min(n_rays + n_rays / 4, std::max(n_rays, (size_t)c->W * c->H * 8));
    CK(cudaMalloc(&c->stream_rays, cap * 32));
    CK(cudaMalloc(&c->stream_hits, cap * 16));
    c->stream_cap = cap;
    return ORB_OK;
}

// closest hits of the rays in c->stream_rays -> c->stream_hits; the ray count is *n_items * mult on the device
static int launch_trace_stream(OrbContext* c, const unsigned* n_items, unsigned mult, bool count = false,
                               bool timed = false, size_t n_upper = 0) {
    cudaStream_t st = c->stream;
    CK(cudaMemsetAsync(c->stream_cursor, 0, 4, st));
    StageTimer tm(c, ORB_STAGE_TRACE, 1, nullptr, timed);
    const int sd = stack_mode(c);
    const int block = c->stream_block;
    const size_t smem = stack_bytes(c, sd, block);
    const bool qn = use_qnodes(c);
    StreamKernel k = stream_kernel(sd, count, qn);
    if (smem > 48 * 1024) CK(allow_smem(k, smem));
    int g = persistent_grid(c, k, block, smem, c->n_slots > 1 ? c->trace_ctas_per_sm : 0);
    // a short stream (a narrow band of a tile partition) on a full persistent grid is mostly tail: every lane
    // gets a handful of rays and idles while the slowest ray of its warp finishes.  Fewer CTAs, more rays each.
    if (c->min_rays_per_lane > 0 && n_upper > 0)
        g = std::max(c->n_sm, std::min(g, (int)((n_upper + (size_t)block * c->min_rays_per_lane - 1) /
                                                 ((size_t)block * c->min_rays_per_lane))));
    k<<<g, block, smem, st>>>(c->view, c->stream_rays, c->stream_hits, n_items, mult, c->stream_cursor, c->refill_min,
                                    c->node_min, c->height, count ? &c->cnt->nodes_visited : nullptr,
                                    qn ? c->qnodes : (const QNode*)nullptr, qn ? c->qgrid : QGrid());
    c->counters.kernel_launches++;
    return ORB_OK;
}

static int ensure_path_state(OrbContext* c) {
    if (c->path_state[0][0]) return ORB_OK;
    const size_t n = (size_t)c->W * c->H;
    for (int a = 0; a < 2; a++)
        for (int b = 0; b < 5; b++) CK(cudaMalloc(&c->path_state[a][b], n * 16));
    return ORB_OK;
}

// Two frame slots: `accum` and RGBA8 are shared, so this frame's k_accumulate (on the active slot's stream) is ordered
// behind EVERYTHING the previous frame's slot has enqueued so far -- its own k_accumulate, and whatever read the shared
// buffers after it (snapshots of orb_buffer_download_rows_async, a caller's copies on that slot's stream).
static int order_behind_other_slot(OrbContext* c, int prev_slot, bool prev_rendered) {
    if (c->n_slots < 2 || prev_slot == c->cur_slot || !prev_rendered) return ORB_OK;
    FrameSlot& p = c->parked[prev_slot];
    if (!p.ev_order) CK(cudaEventCreateWithFlags(&p.ev_order, cudaEventDisableTiming));
    CK(cudaEventRecord(p.ev_order, p.stream));
    CK(cudaStreamWaitEvent(c->stream, p.ev_order, 0));
    return ORB_OK;
}

int orb_render_async(OrbContext* c, const OrbRenderConfig* cfg) {
    NEED(c && cfg, "orb_render: NULL argument");
    NEED(c->committed, "orb_render: scene not committed");
    NEED(!c->geometry_dirty, "orb_render: orb_scene_update_mesh without orb_scene_refit (the BVH is stale)");
    NEED(c->W > 0, "orb_render: orb_resize first");
    NEED(c->have_cam, "orb_render: orb_set_camera first");
    NEED(c->lut_w > 0 || (cfg->renderer_type != ORB_LTC_BASELINE && cfg->renderer_type != ORB_RATIO),
         "orb_render: orb_scene_set_ltc_lut first");
    int y0 = cfg->y0, y1 = cfg->y1;
    NEED(rows_ok(c, y0, y1), "orb_render: bad row range");
    const int mode = cfg->mode;
    NEED(mode == 0 || mode == 1, "orb_render: mode must be 0 (literal) or 1 (intended)");
    const int renderer = cfg->renderer_type;
    const int spp = mode == 0 ? 4 : (cfg->shadow_spp > 0 ? cfg->shadow_spp : 4);      // deviceCode.cu:128
    const int depth = mode == 0 ? 1 : (cfg->max_depth > 0 ? cfg->max_depth : 1);      // deviceCode.cu:151
    CK(cudaSetDevice(c->device));
    // two frames in flight: this frame takes the other slot (its own buffers and stream); everything the caller
    // enqueues next (denoise, combine, downloads, orb_stream()) refers to it until the next orb_render_async
    const int prev_slot = c->cur_slot;
    const bool prev_rendered = c->rendered;
    if (c->n_slots > 1) select_slot(c, (c->cur_slot + 1) % c->n_slots);
    c->rendered = true;
    cudaStream_t st = c->stream;
    for (int i = 0; i < c->n_pending_dl; i++) CK(cudaStreamWaitEvent(st, c->pending_dl[i], 0));  // in-place read-backs of this slot
    c->n_pending_dl = 0;
    FrameView f = frame_view(c, y0, y1);
    const int frame_id = cfg->accum_id + 10007;  // deviceCode.cu:65
    const size_t smem = trace_smem(c);
    const int count_rays = (cfg->flags & ORB_RENDER_NO_COUNT) ? 0 : 1;
    CK(cudaMemsetAsync(c->cnt, 0, DEVCOUNTERS_FRAME_BYTES, st));
    {
        StageTimer tm(c, ORB_STAGE_PRIMARY, 1);
        dim3 grid((c->W + 15) / 16, (y1 - y0 + 7) / 8);
        const int sd = stack_mode(c);
        const size_t sm = stack_bytes(c, sd);
        const bool qn = use_qnodes(c);
        PrimaryKernel k = primary_kernel(sd, qn);
        if (sm > 48 * 1024) CK(allow_smem(k, sm));
        k<<<grid, TRACE_BLOCK, sm, st>>>(c->view, f, c->cam, frame_id, renderer, c->height, qn ? c->qnodes : (const QNode*)nullptr,
                                         qn ? c->qgrid : QGrid());
        c->counters.kernel_launches++;
        if (count_rays) c->counters.rays_primary += (uint64_t)(y1 - y0) * c->W;
    }
    if (cfg->flags & ORB_RENDER_CHECKED) {  // ties of the primary hits, LTC pixels next to a clip decision -> OrbCounters
        const size_t npx = (size_t)(y1 - y0) * c->W;
        const int with_ltc = (renderer == ORB_LTC_BASELINE || renderer == ORB_RATIO) && c->lut_w > 0;
        k_check_frame<<<(unsigned)std::min<size_t>((npx + TRACE_BLOCK - 1) / TRACE_BLOCK, (size_t)c->n_sm * 16), TRACE_BLOCK, 0, st>>>(
            c->view, f, c->cam, with_ltc);
        c->counters.kernel_launches++;
    }
    const float4* colour_src = c->colour;
    auto launch_ltc = [&](cudaStream_t ls) {
        StageTimer tm(c, ORB_STAGE_LTC, 1, ls);
        const int g = persistent_grid(c, k_ltc<true>, LTC_BLOCK, 0);
        if (renderer == ORB_RATIO) {
            // light pixels already hold emit in `ltc` (k_primary), so `ltc` IS the frame colour
            k_ltc<true><<<g, LTC_BLOCK, 0, ls>>>(c->view, f, c->cam.pos, f.ltc, nullptr);
            colour_src = f.ltc;
        } else {
            k_ltc<true><<<g, LTC_BLOCK, 0, ls>>>(c->view, f, c->cam.pos, f.colour, nullptr);
        }
        c->counters.kernel_launches++;
    };
    if (renderer == ORB_LTC_BASELINE || renderer == ORB_RATIO) launch_ltc(st);
    if (renderer == ORB_RATIO) {
        {
            const size_t npx = (size_t)(y1 - y0) * c->W;
            int rc = ensure_stream(c, npx * (size_t)spp);
            if (rc) return rc;
            const int n_lights = (int)c->h_lights.size();
            if (n_lights > 0) {
                {
                    StageTimer tm(c, ORB_STAGE_SHADOW, 1);
                    const int gg = (int)std::min<size_t>((npx * spp + 255) / 256, (size_t)c->n_sm * 8);
                    k_shadow_gen<<<gg, 256, 0, st>>>(c->view, f, frame_id, spp, c->stream_rays);
                    c->counters.kernel_launches++;
                }
                rc = launch_trace_stream(c, &c->cnt->queue_n, (unsigned)spp, (cfg->flags & ORB_RENDER_COUNT_TRAVERSAL) != 0, true,
                                         npx * (size_t)spp);
                if (rc) return rc;
            }
            StageTimer tm(c, ORB_STAGE_SHADOW, 1);
            const int gr = (int)std::min<size_t>((npx + 127) / 128, (size_t)c->n_sm * 8);
            const int gl = (int)std::min<size_t>((npx * spp + 255) / 256, (size_t)c->n_sm * 8);
            if (spp == 4 && c->resolve_lanes)
                k_shadow_resolve_lanes<4><<<gl, 256, 0, st>>>(c->view, f, c->cam.pos, frame_id, mode, c->stream_hits);
            else if (spp == 2 && c->resolve_lanes)
                k_shadow_resolve_lanes<2><<<gl, 256, 0, st>>>(c->view, f, c->cam.pos, frame_id, mode, c->stream_hits);
            else if (spp == 8 && c->resolve_lanes)
                k_shadow_resolve_lanes<8><<<gl, 256, 0, st>>>(c->view, f, c->cam.pos, frame_id, mode, c->stream_hits);
            else
                k_shadow_resolve<<<gr, 128, 0, st>>>(c->view, f, c->cam.pos, frame_id, spp, mode, c->stream_hits);
            c->counters.kernel_launches++;
        }
        c->last_spp = spp;
    }
    if (renderer == ORB_PATH) {
        int rc = ORB_OK;
        const bool stream = mode == ORB_MODE_INTENDED && !(cfg->flags & ORB_RENDER_FUSED_SHADOW);
        const bool persistent = stream && c->path_persistent && !(cfg->flags & ORB_RENDER_PATH_WAVEFRONT);
        if (persistent) {
            rc = ensure_stream(c, 1);  // the cursor
            if (rc) return rc;
            StageTimer tm(c, ORB_STAGE_PATH, 1);
            CK(cudaMemsetAsync(c->stream_cursor, 0, 4, st));
            const int sd = stack_mode(c);
            const int block = c->stream_block;
            const size_t sm = stack_bytes(c, sd, block);
            const bool qn = use_qnodes(c);
            PathKernel k = path_kernel(sd, (cfg->flags & ORB_RENDER_COUNT_TRAVERSAL) != 0, qn);
            if (sm > 48 * 1024) CK(allow_smem(k, sm));
            const int g = persistent_grid(c, k, block, sm, c->n_slots > 1 ? c->trace_ctas_per_sm : 0);
            k<<<g, block, sm, st>>>(c->view, f, c->cam.pos, frame_id, depth, c->stream_cursor, c->path_shade_min,
                                          c->path_node_min, c->height, qn ? c->qnodes : (const QNode*)nullptr,
                                          qn ? c->qgrid : QGrid());
            c->counters.kernel_launches++;
        } else {
        rc = ensure_path_state(c);
        if (rc) return rc;
        StageTimer tm(c, ORB_STAGE_PATH, stream ? 4 * depth - 1 : 2 * depth - 1);
        PathBuffers pb[2];
        for (int a = 0; a < 2; a++)
            for (int b = 0; b < 5; b++) pb[a].s[b] = c->path_state[a][b];
        const size_t npx = (size_t)(y1 - y0) * c->W;
        if (stream) {
            rc = ensure_stream(c, npx);
            if (rc) return rc;
        }
        const int g1 = persistent_grid(c, k_path_bounce<true>, TRACE_BLOCK, smem);
        const int g2 = persistent_grid(c, k_path_bounce<false>, TRACE_BLOCK, smem);
        const int gg = (int)std::min<size_t>((npx + 255) / 256, (size_t)c->n_sm * 8);
        const int gs = (int)std::min<size_t>((npx + 127) / 128, (size_t)c->n_sm * 8);
        for (int b = 0; b < depth; b++) {
            const int last = b == depth - 1;
            const int in_slot = b & 1;  // bounce b reads path_n[in_slot] (b > 0), writes path_n[in_slot ^ 1]
            if (b > 0) {
                k_reset_slot<<<1, 1, 0, st>>>(c->cnt, in_slot ^ 1);
                c->counters.kernel_launches++;
            }
            if (stream) {  // generate -> persistent traversal with refill -> shade + compaction
                if (b == 0) {
                    k_path_gen<true><<<gg, 256, 0, st>>>(f, c->cam.pos, frame_id, b, pb[0], in_slot, c->stream_rays);
                    rc = launch_trace_stream(c, &c->cnt->queue_n, 1u);
                    if (rc) return rc;
                    k_path_shade<true><<<gs, 128, 0, st>>>(c->view, f, c->cam.pos, frame_id, b, last, pb[0], pb[1], in_slot,
                                                           c->stream_hits);
                } else {
                    k_path_gen<false><<<gg, 256, 0, st>>>(f, c->cam.pos, frame_id, b, pb[in_slot], in_slot, c->stream_rays);
                    rc = launch_trace_stream(c, &c->cnt->path_n[in_slot], 1u);
                    if (rc) return rc;
                    k_path_shade<false><<<gs, 128, 0, st>>>(c->view, f, c->cam.pos, frame_id, b, last, pb[in_slot],
                                                            pb[in_slot ^ 1], in_slot, c->stream_hits);
                }
                c->counters.kernel_launches += 2;
            } else if (b == 0) {
                k_path_bounce<true><<<g1, TRACE_BLOCK, smem, st>>>(c->view, f, c->cam.pos, frame_id, b, mode, last, pb[0],
                                                                    pb[1], in_slot, c->height);
                c->counters.kernel_launches++;
            } else {
                k_path_bounce<false><<<g2, TRACE_BLOCK, smem, st>>>(c->view, f, c->cam.pos, frame_id, b, mode, last,
                                                                     pb[in_slot], pb[in_slot ^ 1], in_slot, c->height);
                c->counters.kernel_launches++;
            }
        }
        }
    }
    {
        int rc = order_behind_other_slot(c, prev_slot, prev_rendered);
        if (rc) return rc;
        StageTimer tm(c, ORB_STAGE_ACCUMULATE, 1);
        const size_t n = (size_t)(y1 - y0) * c->W;
        const int g = (int)std::min<size_t>((n + 255) / 256, (size_t)c->n_sm * 8);
        k_accumulate<<<g, 256, 0, st>>>(f, colour_src, cfg->accum_id, count_rays);
        c->counters.kernel_launches++;
    }
    CK(cudaGetLastError());
    return ORB_OK;
}

int orb_render(OrbContext* c, const OrbRenderConfig* cfg) {
    int rc = orb_render_async(c, cfg);
    if (rc) return rc;
    return orb_sync(c);
}

// ---- stand-alone passes ----------------------------------------------------------------------
int orb_ltc_gbuffer_pass(OrbContext* c, int32_t y0, int32_t y1) {
    NEED(c && c->W > 0, "orb_ltc_gbuffer_pass: orb_resize first");
    NEED(c->have_cam, "orb_ltc_gbuffer_pass: orb_set_camera first");
    NEED(c->lut_w > 0, "orb_ltc_gbuffer_pass: orb_scene_set_ltc_lut first");
    NEED(rows_ok(c, y0, y1), "orb_ltc_gbuffer_pass: bad row range");
    CK(cudaSetDevice(c->device));
    refresh_view(c);
    FrameView f = frame_view(c, y0, y1);
    const int g = persistent_grid(c, k_ltc<false>, LTC_BLOCK, 0);
    {
        StageTimer tm(c, ORB_STAGE_LTC_GBUFFER, 1);
        k_ltc<false><<<g, LTC_BLOCK, 0, c->stream>>>(c->view, f, c->cam.pos, f.ltc, nullptr);
    }
    c->counters.kernel_launches++;
    CK(cudaGetLastError());
    return ORB_OK;
}

// rows the a-trous pass `it` (step 2^it) of an `iters`-pass filter must WRITE beyond a band so that the
// later passes find their taps: sum of 2 * 2^j over j > it; it = -1 gives the rows the packed input must cover
static int denoise_reach(int iters, int it) { return 2 * ((1 << iters) - (1 << (it + 1))); }

int orb_denoise_pack(OrbContext* c, int32_t y0, int32_t y1) {
    NEED(c && c->W > 0, "orb_denoise_pack: orb_resize first");
    NEED(c->have_cam, "orb_denoise_pack: orb_set_camera first");
    NEED(rows_ok(c, y0, y1), "orb_denoise_pack: bad row range");
    CK(cudaSetDevice(c->device));
    FrameView f = frame_view(c, y0, y1);
    const size_t np = (size_t)(y1 - y0) * c->W;
    const int gp = (int)std::min<size_t>((np + 255) / 256, (size_t)c->n_sm * 8);
    {
        StageTimer tm(c, ORB_STAGE_DENOISE_PACK, 1);
        HaloSendList none;
        none.n = 0;
        k_denoise_pack<<<gp, 256, 0, c->stream>>>(f, c->cam.pos, (float4*)c->buf[ORB_BUF_DN_GUIDE], (float2*)c->buf[ORB_BUF_DN_VAL], none);
    }
    c->counters.kernel_launches++;
    CK(cudaGetLastError());
    return ORB_OK;
}

int orb_denoise_filter(OrbContext* c, const OrbDenoiseConfig* cfg, int32_t y0, int32_t y1) {
    NEED(c && c->W > 0, "orb_denoise_filter: orb_resize first");
    NEED(rows_ok(c, y0, y1), "orb_denoise_filter: bad row range");
    OrbDenoiseConfig d = {5, 32.f, 0.25f, 0.f};
    if (cfg) d = *cfg;
    if (d.iterations <= 0) d.iterations = 5;
    NEED(d.iterations <= 8, "orb_denoise: at most 8 iterations");
    NEED(d.sigma_position > 0.f && d.sigma_position <= 1e12f, "orb_denoise: sigma_position must be in (0, 1e12]");
    CK(cudaSetDevice(c->device));
    const size_t n = (size_t)c->W * c->H;
    if (!c->dn_val_b) CK(cudaMalloc(&c->dn_val_b, n * 8));
    const float4* guide = (const float4*)c->buf[ORB_BUF_DN_GUIDE];
    float2* val[2] = {(float2*)c->buf[ORB_BUF_DN_VAL], c->dn_val_b};
    FrameView f = frame_view(c, y0, y1);
    StageTimer tm_dn(c, ORB_STAGE_DENOISE, d.iterations);
    int cur = 0;
    for (int it = 0; it < d.iterations; it++) {
        const int step = 1 << it;
        const float log2e = 1.4426950408889634f;  // the kernel evaluates exp(x) as exp2(x log2 e)
        const float isp = log2e / (d.sigma_position * (float)step);
        const float isc = d.sigma_colour > 0.f ? log2e / d.sigma_colour : 0.f;
        const bool four = c->atrous4 && isc == 0.f && d.sigma_normal > 0.f;
        if (four && it == d.iterations - 1) {
            Atrous4Kernel k = atrous4_kernel<true>(step);
            k<<<atrous4_grid(c->W, y1 - y0, step), 256, 0, c->stream>>>(
                f, guide, val[cur], nullptr, (float4*)c->buf[ORB_BUF_STO_DIRECT_DN], (float4*)c->buf[ORB_BUF_STO_NO_VIS_DN],
                (const float4*)c->buf[ORB_BUF_LTC], c->fused_combine ? (float4*)c->buf[ORB_BUF_RATIO_FINAL] : nullptr, step,
                d.sigma_normal, isp);
        } else if (four) {
            const int r = denoise_reach(d.iterations, it);
            FrameView part = frame_view(c, std::max(0, y0 - r), std::min(c->H, y1 + r));
            Atrous4Kernel k = atrous4_kernel<false>(step);
            k<<<atrous4_grid(c->W, part.y1 - part.y0, step), 256, 0, c->stream>>>(part, guide, val[cur], val[cur ^ 1], nullptr,
                                                                                  nullptr, nullptr, nullptr, step,
                                                                                  d.sigma_normal, isp);
            cur ^= 1;
        } else if (it == d.iterations - 1) {
            dim3 gb((c->W + 31) / 32, (y1 - y0 + 7) / 8);
            AtrousKernel k = atrous_kernel<true>(isc > 0.f, step);
            k<<<gb, 256, 0, c->stream>>>(f, guide, val[cur], nullptr, (float4*)c->buf[ORB_BUF_STO_DIRECT_DN],
                                         (float4*)c->buf[ORB_BUF_STO_NO_VIS_DN], step, d.sigma_normal, isp, isc);
        } else {
            const int r = denoise_reach(d.iterations, it);
            FrameView part = frame_view(c, std::max(0, y0 - r), std::min(c->H, y1 + r));
            dim3 grid((c->W + 31) / 32, (part.y1 - part.y0 + 7) / 8);
            AtrousKernel k = atrous_kernel<false>(isc > 0.f, step);
            k<<<grid, 256, 0, c->stream>>>(part, guide, val[cur], val[cur ^ 1], nullptr, nullptr, step, d.sigma_normal, isp,
                                           isc);
            cur ^= 1;
        }
        c->counters.kernel_launches++;
        if (c->fused_combine && !four && it == d.iterations - 1) {  // same contract without the fused kernel
            const size_t first = (size_t)y0 * c->W, np = (size_t)(y1 - y0) * c->W;
            const int g = (int)std::min<size_t>((np + 255) / 256, (size_t)c->n_sm * 8);
            k_ratio_combine<<<g, 256, 0, c->stream>>>((const float4*)c->buf[ORB_BUF_LTC], (const float4*)c->buf[ORB_BUF_STO_DIRECT_DN],
                                                      (const float4*)c->buf[ORB_BUF_STO_NO_VIS_DN],
                                                      (float4*)c->buf[ORB_BUF_RATIO_FINAL], first, np);
            c->counters.kernel_launches++;
        }
    }
    if (d.iterations > 1 && ((d.iterations - 1) & 1)) {
        // an odd number of ping-pong passes left pass 0's output in ORB_BUF_DN_VAL; nothing to restore:
        // the packed input is rewritten by the next orb_denoise_pack
    }
    CK(cudaGetLastError());
    return ORB_OK;
}

// ---- tile hand-off between the ranks of a row-band partition (halo.cuh) ---------------------------------------
// CUDA loads kernels lazily, and a first-time load can wait for running kernels to finish.  A flag-wait kernel of
// the hand-off runs until ANOTHER stream's kernels have run, so every kernel a frame may launch behind it has to be
// resident before the first wait is enqueued (CUDA programming guide, "Lazy Loading": concurrent kernels).
__global__ void k_snapshot(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t n);
static int preload_frame_kernels() {
    cudaFuncAttributes a;
    const int sds[5] = {-1, 0, 8, 12, 16};
    for (int sd : sds)
        for (int q = 0; q < 2; q++) {
            CK(cudaFuncGetAttributes(&a, primary_kernel(sd, q != 0)));
            for (int cnt = 0; cnt < 2; cnt++) {
                CK(cudaFuncGetAttributes(&a, stream_kernel(sd, cnt != 0, q != 0)));
                CK(cudaFuncGetAttributes(&a, path_kernel(sd, cnt != 0, q != 0)));
            }
        }
    const int steps[6] = {0, 1, 2, 4, 8, 16};
    for (int st : steps) {
        CK(cudaFuncGetAttributes(&a, atrous_kernel<true>(false, st)));
        CK(cudaFuncGetAttributes(&a, atrous_kernel<false>(false, st)));
    }
    for (int st : steps) {
        CK(cudaFuncGetAttributes(&a, atrous4_kernel<true>(st)));
        CK(cudaFuncGetAttributes(&a, atrous4_kernel<false>(st)));
    }
    CK(cudaFuncGetAttributes(&a, atrous_kernel<true>(true, 0)));
    CK(cudaFuncGetAttributes(&a, atrous_kernel<false>(true, 0)));
    CK(cudaFuncGetAttributes(&a, k_ltc<true>));
    CK(cudaFuncGetAttributes(&a, k_ltc<false>));
    CK(cudaFuncGetAttributes(&a, k_shadow_gen));
    CK(cudaFuncGetAttributes(&a, k_shadow_resolve));
    CK(cudaFuncGetAttributes(&a, k_shadow_resolve_lanes<2>));
    CK(cudaFuncGetAttributes(&a, k_shadow_resolve_lanes<4>));
    CK(cudaFuncGetAttributes(&a, k_shadow_resolve_lanes<8>));
    CK(cudaFuncGetAttributes(&a, k_path_bounce<true>));
    CK(cudaFuncGetAttributes(&a, k_path_bounce<false>));
    CK(cudaFuncGetAttributes(&a, k_path_gen<true>));
    CK(cudaFuncGetAttributes(&a, k_path_gen<false>));
    CK(cudaFuncGetAttributes(&a, k_path_shade<true>));
    CK(cudaFuncGetAttributes(&a, k_path_shade<false>));
    CK(cudaFuncGetAttributes(&a, k_reset_slot));
    CK(cudaFuncGetAttributes(&a, k_accumulate));
    CK(cudaFuncGetAttributes(&a, k_denoise_pack));
    CK(cudaFuncGetAttributes(&a, k_ratio_combine));
    CK(cudaFuncGetAttributes(&a, k_snapshot));
    CK(cudaFuncGetAttributes(&a, k_halo_signal));
    CK(cudaFuncGetAttributes(&a, k_halo_wait));
    return ORB_OK;
}

static FrameSlot& slot_ref(OrbContext* c, int k) { return k == c->cur_slot ? *static_cast<FrameSlot*>(c) : c->parked[k]; }

// Device pointers a peer needs to push its rows here: this context's ORB_BUF_DN_GUIDE / ORB_BUF_DN_VAL of `slot` and
// its control block.  Same process: pass them to the peer's orb_halo_connect; other process: through orb_ipc_export.
int orb_halo_local(OrbContext* c, int32_t slot, void** guide, void** val, void** flags) {
    NEED(c && guide && val && flags, "orb_halo_local: NULL argument");
    NEED(c->W > 0, "orb_halo_local: orb_resize first");
    NEED(slot >= 0 && slot < c->n_slots, "orb_halo_local: no such frame slot");
    CK(cudaSetDevice(c->device));
    if (!c->halo.flags) {
        int rc = preload_frame_kernels();
        if (rc) return rc;
        CK(cudaMalloc(&c->halo.flags, ORB_HALO_BLOCK_BYTES));
        CK(cudaMemset(c->halo.flags, 0, ORB_HALO_BLOCK_BYTES));
    }
    FrameSlot& f = slot_ref(c, slot);
    *guide = f.buf[ORB_BUF_DN_GUIDE];
    *val = f.buf[ORB_BUF_DN_VAL];
    *flags = c->halo.flags;
    return ORB_OK;
}

// cudaIpc handles (64 bytes) for the pointers above; the opening side gets a pointer valid in ITS process.
int orb_ipc_export(OrbContext* c, void* dev_ptr, unsigned char handle[64]) {
    NEED(c && dev_ptr && handle, "orb_ipc_export: NULL argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t size");
    CK(cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    CK(cudaIpcGetMemHandle(&h, dev_ptr));
    memcpy(handle, &h, 64);
    return ORB_OK;
}
int orb_ipc_open(OrbContext* c, const unsigned char handle[64], void** out) {
    NEED(c && handle && out, "orb_ipc_open: NULL argument");
    CK(cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    CK(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
    return ORB_OK;
}
int orb_ipc_close(OrbContext* c, void* ptr) {
    NEED(c && ptr, "orb_ipc_close: NULL argument");
    CK(cudaSetDevice(c->device));
    CK(cudaIpcCloseMemHandle(ptr));
    return ORB_OK;
}

// Register peer `peer_rank`'s buffers of `slot` (pointers valid in this process: orb_halo_local of a context in the
// same process, or orb_ipc_open of an exported handle).  my_rank = this context's index in the partition.
int orb_halo_connect(OrbContext* c, int32_t my_rank, int32_t peer_rank, int32_t slot, void* guide, void* val, void* flags) {
    NEED(c && guide && val && flags, "orb_halo_connect: NULL argument");
    NEED(my_rank >= 0 && my_rank < ORB_HALO_MAX_PEERS && peer_rank >= 0 && peer_rank < ORB_HALO_MAX_PEERS && peer_rank != my_rank,
         "orb_halo_connect: ranks must be distinct and below 8");
    NEED(slot >= 0 && slot < c->n_slots, "orb_halo_connect: no such frame slot");
    HaloLinks& h = c->halo;
    NEED(h.flags, "orb_halo_connect: orb_halo_local first (it allocates this context's control block)");
    NEED(h.n_peers == 0 || h.my_rank == my_rank, "orb_halo_connect: my_rank changed");
    h.my_rank = my_rank;
    int i = 0;
    while (i < h.n_peers && h.peer[i].rank != peer_rank) i++;
    if (i == h.n_peers) {
        NEED(h.n_peers < ORB_HALO_MAX_PEERS, "orb_halo_connect: too many peers");
        h.peer[h.n_peers++].rank = peer_rank;
    }
    h.peer[i].guide[slot] = (float4*)guide;
    h.peer[i].val[slot] = (float2*)val;
    h.peer[i].flags = (uint32_t*)flags;
    return ORB_OK;
}

int orb_halo_disconnect(OrbContext* c) {
    NEED(c, "orb_halo_disconnect: ctx is NULL");
    int rc = orb_sync(c);
    if (rc) return rc;
    halo_disconnect(c->halo);
    if (c->halo.flags) CK(cudaMemset(c->halo.flags, 0, ORB_HALO_WORDS * 4));
    for (int k = 0; k < ORB_MAX_SLOTS; k++) slot_ref(c, k).halo_epoch = 0;
    return ORB_OK;
}

// 0 = every wait so far was answered; k > 0 = a wait gave up after its timeout (k - 1 = index in its flag list)
int orb_halo_status(OrbContext* c, int32_t* status) {
    NEED(c && status, "orb_halo_status: NULL argument");
    *status = 0;
    if (!c->halo.flags) return ORB_OK;
    int rc = orb_sync(c);
    if (rc) return rc;
    uint32_t e = 0;
    CK(cudaMemcpy(&e, c->halo.flags + ORB_HALO_ERR, 4, cudaMemcpyDeviceToHost));
    *status = (int32_t)e;
    return ORB_OK;
}

// The denoise of one band of a partitioned frame, hand-off included (asynchronous, on the frame's stream):
//   wait until every peer has consumed what this slot pushed two frames ago -> pack rows [y0, y1) and store the rows
//   listed in `sends` (peer rank, row range: what that rank's filter reads of this band) into the peers' buffers ->
//   publish READY to all peers -> wait for READY from `recv_ranks` (the ranks whose rows this band's filter reads)
//   -> filter -> publish CONSUMED to all peers.
// A rank with an empty band passes y0 == y1 (no kernels but the flags).  All ranks call this once per frame.
int orb_denoise_band(OrbContext* c, const OrbDenoiseConfig* cfg, int32_t y0, int32_t y1, const OrbHaloSend* sends,
                     int32_t n_sends, const int32_t* recv_ranks, int32_t n_recv) {
    NEED(c && c->W > 0, "orb_denoise_band: orb_resize first");
    NEED(c->have_cam, "orb_denoise_band: orb_set_camera first");
    NEED(y0 >= 0 && y1 <= c->H && y0 <= y1, "orb_denoise_band: bad row range");
    NEED(n_sends >= 0 && n_sends <= ORB_HALO_MAX_SENDS && (n_sends == 0 || sends), "orb_denoise_band: at most 4 sends");
    NEED(n_recv >= 0 && n_recv <= ORB_HALO_MAX_PEERS && (n_recv == 0 || recv_ranks), "orb_denoise_band: bad receive list");
    HaloLinks& h = c->halo;
    NEED(h.flags && h.n_peers > 0, "orb_denoise_band: no peers connected (orb_halo_connect)");
    CK(cudaSetDevice(c->device));
    const int slot = c->cur_slot;
    const uint32_t epoch = ++c->halo_epoch;
    const long long timeout = 4000000000ll;  // ~2 s of SM clock
    HaloFlagList to_ready, to_consumed, from_consumed, from_ready;
    to_ready.n = to_consumed.n = from_consumed.n = from_ready.n = 0;
    HaloSendList sl;
    sl.n = 0;
    for (int i = 0; i < h.n_peers; i++) {
        NEED(h.peer[i].guide[slot] && h.peer[i].val[slot], "orb_denoise_band: a peer has no buffers for this frame slot");
        to_ready.p[to_ready.n++] = h.peer[i].flags + ORB_HALO_READY(slot, h.my_rank);
        to_consumed.p[to_consumed.n++] = h.peer[i].flags + ORB_HALO_CONSUMED(slot, h.my_rank);
        from_consumed.p[from_consumed.n++] = h.flags + ORB_HALO_CONSUMED(slot, h.peer[i].rank);
    }
    for (int k = 0; k < n_sends; k++) {
        int i = 0;
        while (i < h.n_peers && h.peer[i].rank != sends[k].peer_rank) i++;
        NEED(i < h.n_peers, "orb_denoise_band: send to a rank that is not connected");
        NEED(sends[k].y0 >= y0 && sends[k].y1 <= y1 && sends[k].y0 <= sends[k].y1, "orb_denoise_band: send rows outside the band");
        if (sends[k].y1 > sends[k].y0) sl.s[sl.n++] = HaloSend{h.peer[i].guide[slot], h.peer[i].val[slot], sends[k].y0, sends[k].y1};
    }
    for (int k = 0; k < n_recv; k++) {
        NEED(recv_ranks[k] >= 0 && recv_ranks[k] < ORB_HALO_MAX_PEERS && recv_ranks[k] != h.my_rank, "orb_denoise_band: bad receive rank");
        from_ready.p[from_ready.n++] = h.flags + ORB_HALO_READY(slot, recv_ranks[k]);
    }
    cudaStream_t st = c->stream;
    {
        StageTimer tm(c, ORB_STAGE_HALO, 2);
        if (epoch > 1) k_halo_wait<<<1, 32, 0, st>>>(from_consumed, epoch - 1, h.flags + ORB_HALO_ERR, timeout);
    }
    if (y1 > y0) {
        FrameView f = frame_view(c, y0, y1);
        const size_t np = (size_t)(y1 - y0) * c->W;
        const int gp = (int)std::min<size_t>((np + 255) / 256, (size_t)c->n_sm * 8);
        StageTimer tm(c, ORB_STAGE_DENOISE_PACK, 1);
        k_denoise_pack<<<gp, 256, 0, st>>>(f, c->cam.pos, (float4*)c->buf[ORB_BUF_DN_GUIDE], (float2*)c->buf[ORB_BUF_DN_VAL], sl);
        c->counters.kernel_launches++;
    }
    {
        StageTimer tm(c, ORB_STAGE_HALO, 2);
        k_halo_signal<<<1, 32, 0, st>>>(to_ready, epoch);
        if (from_ready.n) k_halo_wait<<<1, 32, 0, st>>>(from_ready, epoch, h.flags + ORB_HALO_ERR, timeout);
        c->counters.kernel_launches += 2;
    }
    CK(cudaGetLastError());
    if (y1 > y0) {
        int rc = orb_denoise_filter(c, cfg, y0, y1);
        if (rc) return rc;
    }
    k_halo_signal<<<1, 32, 0, st>>>(to_consumed, epoch);
    c->counters.kernel_launches++;
    CK(cudaGetLastError());
    return ORB_OK;
}

// pack + filter over rows [y0, y1): the packed input covers the band plus the filter's reach, so those rows
// must hold valid G-buffer / sto data (a single-GPU frame, or halo rows rendered by the band's owner)
int orb_denoise(OrbContext* c, const OrbDenoiseConfig* cfg, int32_t y0, int32_t y1) {
    NEED(c && c->W > 0, "orb_denoise: orb_resize first");
    NEED(rows_ok(c, y0, y1), "orb_denoise: bad row range");
    int iters = cfg && cfg->iterations > 0 ? cfg->iterations : 5;
    NEED(iters <= 8, "orb_denoise: at most 8 iterations");
    const int r = denoise_reach(iters, -1);
    int rc = orb_denoise_pack(c, std::max(0, y0 - r), std::min(c->H, y1 + r));
    if (rc) return rc;
    return orb_denoise_filter(c, cfg, y0, y1);
}

// active (non-light hit) pixels of every row of the last frame, for cost-balanced tile partitions
// orb_buffer_download_rows_async: rows -> device slot on the SMs, so the copy engines only carry the host copy
__global__ void k_snapshot(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}

__global__ void k_row_active(FrameView f, int* __restrict__ counts) {
    const int y = blockIdx.x;
    int n = 0;
    for (int x = threadIdx.x; x < f.W; x += blockDim.x) {
        const uint32_t fl = f.flags[(size_t)y * f.W + x];
        n += ((fl & FLAG_HIT) && !(fl & FLAG_LIGHT)) ? 1 : 0;
    }
    for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
    if ((threadIdx.x & 31) == 0 && n) atomicAdd(&counts[y], n);
}

int orb_row_active_counts(OrbContext* c, int32_t* host_counts) {
    NEED(c && c->W > 0 && host_counts, "orb_row_active_counts: bad argument");
    CK(cudaSetDevice(c->device));
    int* d = nullptr;
    CK(cudaMalloc(&d, sizeof(int) * c->H));
    CK(cudaMemsetAsync(d, 0, sizeof(int) * c->H, c->stream));
    k_row_active<<<c->H, 128, 0, c->stream>>>(frame_view(c, 0, c->H), d);
    c->counters.kernel_launches++;
    CK(cudaMemcpyAsync(host_counts, d, sizeof(int) * c->H, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaFree(d));
    return ORB_OK;
}

int orb_set_persistent_ctas_per_sm(OrbContext* c, int32_t n) {
    NEED(c && n >= 0, "orb_set_persistent_ctas_per_sm: bad argument");
    c->trace_ctas_per_sm = n;
    return ORB_OK;
}

int orb_set_fused_combine(OrbContext* c, int32_t enable) {
    NEED(c, "orb_set_fused_combine: ctx is NULL");
    c->fused_combine = enable ? 1 : 0;
    return ORB_OK;
}

int orb_ratio_combine(OrbContext* c, int32_t use_denoised, int32_t y0, int32_t y1) {
    NEED(c && c->W > 0, "orb_ratio_combine: orb_resize first");
    NEED(rows_ok(c, y0, y1), "orb_ratio_combine: bad row range");
    CK(cudaSetDevice(c->device));
    const size_t first = (size_t)y0 * c->W, n = (size_t)(y1 - y0) * c->W;
    const int g = (int)std::min<size_t>((n + 255) / 256, (size_t)c->n_sm * 8);
    StageTimer tm(c, ORB_STAGE_COMBINE, 1);
    k_ratio_combine<<<g, 256, 0, c->stream>>>(
        (const float4*)c->buf[ORB_BUF_LTC],
        (const float4*)c->buf[use_denoised ? ORB_BUF_STO_DIRECT_DN : ORB_BUF_STO_DIRECT],
        (const float4*)c->buf[use_denoised ? ORB_BUF_STO_NO_VIS_DN : ORB_BUF_STO_NO_VIS],
        (float4*)c->buf[ORB_BUF_RATIO_FINAL], first, n);
    c->counters.kernel_launches++;
    CK(cudaGetLastError());
    return ORB_OK;
}

// ---- buffers ---------------------------------------------------------------------------------
size_t orb_buffer_bytes(OrbContext* c, int32_t id) {
    if (!c || id < 0 || id >= ORB_BUF_COUNT) return 0;
    return (size_t)c->W * c->H * buffer_elem_bytes(id);
}

void* orb_buffer_device_ptr(OrbContext* c, int32_t id) {
    if (!c || id < 0 || id >= ORB_BUF_COUNT) return nullptr;
    return c->buf[id];
}

void* orb_slot_buffer_device_ptr(OrbContext* c, int32_t slot, int32_t id) {
    if (!c || id < 0 || id >= ORB_BUF_COUNT || slot < 0 || slot >= c->n_slots) return nullptr;
    return slot_ref(c, slot).buf[id];
}

// One-sided frame assembly of a row-band partition: rows [y0, y1) of the active slot's buffer go into the same rows of
// a PEER's buffer (a pointer valid in this process: another context's orb_slot_buffer_device_ptr, or orb_ipc_open of
// its exported handle) -- cudaMemcpyAsync over NVLink on the context's peer-copy stream, behind everything the slot
// has enqueued.  No kernel and no rendezvous: an NCCL all-gather is a kernel whose CTAs spin until every peer's
// kernel has arrived, and the ranks of a deep frame pipeline are frames apart (bench.py: 0.2 - 0.4 ms per frame at
// N = 2).  `last` closes the frame's batch: the slot's stream then waits for the copies, so the slot's next frame and
// (through the accumulate ordering) the next frame's writes to the shared ACCUM / RGBA8 rows come after them.
int orb_copy_rows_to_peer(OrbContext* c, int32_t id, void* peer_buffer, int32_t y0, int32_t y1, int32_t last) {
    NEED(c && peer_buffer && id >= 0 && id < ORB_BUF_COUNT, "orb_copy_rows_to_peer: bad argument");
    NEED(c->W > 0 && y0 >= 0 && y0 <= y1 && y1 <= c->H, "orb_copy_rows_to_peer: bad row range");
    CK(cudaSetDevice(c->device));
    if (!c->peer_stream) CK(cudaStreamCreateWithFlags(&c->peer_stream, cudaStreamNonBlocking));
    if (!c->ev_peer_ready) {
        CK(cudaEventCreateWithFlags(&c->ev_peer_ready, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&c->ev_peer_done, cudaEventDisableTiming));
    }
    if (!c->peer_batch_open) {
        CK(cudaEventRecord(c->ev_peer_ready, c->stream));
        CK(cudaStreamWaitEvent(c->peer_stream, c->ev_peer_ready, 0));
        c->peer_batch_open = true;
    }
    const size_t row = (size_t)c->W * buffer_elem_bytes(id);
    if (y1 > y0)
        CK(cudaMemcpyAsync((char*)peer_buffer + row * y0, (const char*)c->buf[id] + row * y0, row * (size_t)(y1 - y0),
                           cudaMemcpyDefault, c->peer_stream));
    if (last) {
        CK(cudaEventRecord(c->ev_peer_done, c->peer_stream));
        CK(cudaStreamWaitEvent(c->stream, c->ev_peer_done, 0));
        c->peer_batch_open = false;
    }
    return ORB_OK;
}

int orb_buffer_download(OrbContext* c, int32_t id, void* dst, size_t bytes) {
    NEED(c && dst && id >= 0 && id < ORB_BUF_COUNT, "orb_buffer_download: bad argument");
    NEED(bytes <= orb_buffer_bytes(c, id), "orb_buffer_download: too many bytes");
    CK(cudaSetDevice(c->device));
    CK(cudaMemcpyAsync(dst, c->buf[id], bytes, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return ORB_OK;
}

int orb_buffer_download_rows(OrbContext* c, int32_t id, void* dst, int32_t y0, int32_t y1) {
    NEED(c && dst && id >= 0 && id < ORB_BUF_COUNT, "orb_buffer_download_rows: bad argument");
    NEED(rows_ok(c, y0, y1), "orb_buffer_download_rows: bad row range");
    CK(cudaSetDevice(c->device));
    const size_t row = (size_t)c->W * buffer_elem_bytes(id);
    CK(cudaMemcpyAsync(dst, (const char*)c->buf[id] + row * y0, row * (y1 - y0), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return ORB_OK;
}

// ORB_DL_TRACE=1: host time spent in each CUDA call of orb_buffer_download_rows_async, printed by orb_destroy
static double g_dl_trace[8];
static long g_dl_calls;
static int g_dl_trace_on = -1;
struct DlTrace {
    int k;
    std::chrono::steady_clock::time_point t0;
    explicit DlTrace(int kk) : k(kk), t0(std::chrono::steady_clock::now()) {}
    ~DlTrace() {
        if (g_dl_trace_on > 0) g_dl_trace[k] += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
};
static void dl_trace_report() {
    if (g_dl_trace_on > 0 && g_dl_calls)
        fprintf(stderr, "ORB_DL_TRACE: %ld calls; ms per call: event-wait %.4f  snapshot %.4f  record+wait %.4f  d2h %.4f  record %.4f  realloc %.4f\n",
                g_dl_calls, g_dl_trace[0] / g_dl_calls, g_dl_trace[1] / g_dl_calls, g_dl_trace[2] / g_dl_calls,
                g_dl_trace[3] / g_dl_calls, g_dl_trace[4] / g_dl_calls, g_dl_trace[5] / g_dl_calls);
}

int orb_buffer_download_rows_async(OrbContext* c, int32_t id, void* dst, int32_t y0, int32_t y1) {
    NEED(c && dst && id >= 0 && id < ORB_BUF_COUNT, "orb_buffer_download_rows_async: bad argument");
    if (g_dl_trace_on < 0) g_dl_trace_on = getenv("ORB_DL_TRACE") ? 1 : 0;
    g_dl_calls++;
    NEED(rows_ok(c, y0, y1), "orb_buffer_download_rows_async: bad row range");
    CK(cudaSetDevice(c->device));
    if (!c->copy_stream) CK(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    const size_t row = (size_t)c->W * buffer_elem_bytes(id), bytes = row * (y1 - y0);
    DlSlot& sl = c->dl[id][c->dl_next[id]];
    c->dl_next[id] ^= 1;
    if (!sl.ready) {
        CK(cudaEventCreateWithFlags(&sl.ready, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
    }
    const bool shared = id == ORB_BUF_ACCUM || id == ORB_BUF_RGBA8;
    if (c->n_slots > 1 && !shared && c->dl_inplace) {
        // Two frame slots: this buffer belongs to the frame's slot and is not written again before the NEXT frame on
        // the slot (two frames later), so the copy stream reads it in place -- no snapshot.  (A snapshot KERNEL would
        // queue for SMs behind the other slot's persistent traversal kernel and hold back everything enqueued on this
        // slot after it: measured, PATH e2e 2.97 ms against 1.99 ms device time.)  The slot's next render waits for
        // the copy (pending_dl), which by then has long finished.
        CK(cudaEventRecord(sl.ready, c->stream));
        CK(cudaStreamWaitEvent(c->copy_stream, sl.ready, 0));
        CK(cudaMemcpyAsync(dst, (const char*)c->buf[id] + row * y0, bytes, cudaMemcpyDeviceToHost, c->copy_stream));
        CK(cudaEventRecord(sl.done, c->copy_stream));
        if (c->n_pending_dl < ORB_MAX_PENDING_DL) c->pending_dl[c->n_pending_dl++] = sl.done;
        else CK(cudaStreamWaitEvent(c->stream, sl.done, 0));
        return ORB_OK;
    }
    if (sl.cap < bytes) {
        DlTrace t(5);
        CK(cudaStreamSynchronize(c->copy_stream));
        CK(cudaFree(sl.dev));
        sl.dev = nullptr, sl.cap = 0;
        CK(cudaMalloc(&sl.dev, bytes));
        sl.cap = bytes;
    } else {
        DlTrace t(0);
        CK(cudaStreamWaitEvent(c->stream, sl.done, 0));  // the slot's previous host copy has left it
    }
    // one slot: a kernel snapshots the rows (the copy engines only carry the host copy); two slots, shared accum /
    // RGBA8: a copy-engine snapshot, which needs no SM while the other slot's persistent kernels hold them all
    if (c->dl_snapshot_kernel && c->n_slots == 1 && bytes % 16 == 0 && (row * y0) % 16 == 0) {
        DlTrace t(1);
        const size_t n16 = bytes / 16;
        k_snapshot<<<(unsigned)std::min<size_t>((n16 + 255) / 256, (size_t)c->n_sm * 8), 256, 0, c->stream>>>(
            (const uint4*)((const char*)c->buf[id] + row * y0), (uint4*)sl.dev, n16);
        CK(cudaGetLastError());
    } else {
        DlTrace t(1);
        CK(cudaMemcpyAsync(sl.dev, (const char*)c->buf[id] + row * y0, bytes, cudaMemcpyDeviceToDevice, c->stream));
    }
    {
        DlTrace t(2);
        CK(cudaEventRecord(sl.ready, c->stream));
        CK(cudaStreamWaitEvent(c->copy_stream, sl.ready, 0));
    }
    {
        DlTrace t(3);
        CK(cudaMemcpyAsync(dst, sl.dev, bytes, cudaMemcpyDeviceToHost, c->copy_stream));
    }
    DlTrace t(4);
    CK(cudaEventRecord(sl.done, c->copy_stream));
    return ORB_OK;
}

int orb_download_wait(OrbContext* c) {
    NEED(c, "orb_download_wait: ctx is NULL");
    CK(cudaSetDevice(c->device));
    if (c->copy_stream) CK(cudaStreamSynchronize(c->copy_stream));
    return ORB_OK;
}

int orb_buffer_upload(OrbContext* c, int32_t id, const void* src, size_t bytes) {
    NEED(c && src && id >= 0 && id < ORB_BUF_COUNT, "orb_buffer_upload: bad argument");
    NEED(bytes <= orb_buffer_bytes(c, id), "orb_buffer_upload: too many bytes");
    CK(cudaSetDevice(c->device));
    CK(cudaMemcpyAsync(c->buf[id], src, bytes, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return ORB_OK;
}

int orb_get_counters(OrbContext* c, OrbCounters* out) {
    NEED(c && out, "orb_get_counters: NULL argument");
    CK(cudaSetDevice(c->device));
    DevCounters d;
    CK(cudaMemcpyAsync(&d, c->cnt, sizeof(d), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    *out = c->counters;
    out->rays_shadow = d.rays_shadow;   // of the last frame
    out->rays_bounce = d.rays_bounce;
    out->nan_pixels = d.nan_pixels;
    out->active_pixels = d.queue_n;
    out->nodes_visited = d.nodes_visited;
    out->tris_tested = d.tris_tested;
    out->hit_ties = d.hit_ties;
    out->ltc_near_decision = d.ltc_near_decision;
    out->bvh_height = (uint64_t)c->bvh.height;
    out->rays_shadow_total = d.tot_shadow;
    out->rays_bounce_total = d.tot_bounce;
    out->active_pixels_total = d.tot_active;
    for (int k = 0; k < c->n_slots; k++) {  // the running totals are kept per slot
        if (k == c->cur_slot || !c->parked[k].cnt) continue;
        DevCounters o;
        CK(cudaMemcpyAsync(&o, c->parked[k].cnt, sizeof(o), cudaMemcpyDeviceToHost, c->parked[k].stream));
        CK(cudaStreamSynchronize(c->parked[k].stream));
        out->rays_shadow_total += o.tot_shadow;
        out->rays_bounce_total += o.tot_bounce;
        out->active_pixels_total += o.tot_active;
    }
    return ORB_OK;
}

// ---- host-buffer entry points ----------------------------------------------------------------
int orb_ratio_estimator_host(OrbContext* c, const float* ltc, const float* sd, const float* sn, float* out, size_t n) {
    NEED(c && ltc && sd && sn && out, "orb_ratio_estimator_host: NULL argument");
    if (n == 0) return ORB_OK;
    CK(cudaSetDevice(c->device));
    const size_t np = (n + 3) & ~(size_t)3;  // 16-byte aligned sub-arrays
    float* d = nullptr;
    CK(cudaMalloc(&d, 4 * np * sizeof(float)));
    CK(cudaMemcpyAsync(d, ltc, n * 4, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(d + np, sd, n * 4, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(d + 2 * np, sn, n * 4, cudaMemcpyHostToDevice, c->stream));
    const int g = (int)std::max<size_t>(1, std::min<size_t>((n / 4 + 255) / 256, (size_t)c->n_sm * 8));
    k_ratio_combine_flat<<<g, 256, 0, c->stream>>>(d, d + np, d + 2 * np, d + 3 * np, n);
    c->counters.kernel_launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out, d + 3 * np, n * 4, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaFree(d));
    return ORB_OK;
}

int orb_trace_rays_host(OrbContext* c, const float* origin, const float* direction, size_t n, int32_t brute,
                        int32_t* prim, float* t, float* u, float* v) {
    NEED(c && origin && direction && prim && t && u && v, "orb_trace_rays_host: NULL argument");
    NEED(c->committed, "orb_trace_rays_host: scene not committed");
    NEED(!c->geometry_dirty, "orb_trace_rays_host: orb_scene_update_mesh without orb_scene_refit (the BVH is stale)");
    if (n == 0) return ORB_OK;
    CK(cudaSetDevice(c->device));
    float *d_o = nullptr, *d_d = nullptr, *d_t = nullptr;
    int* d_p = nullptr;
    CK(cudaMalloc(&d_o, 12 * n)); CK(cudaMalloc(&d_d, 12 * n)); CK(cudaMalloc(&d_t, 12 * n)); CK(cudaMalloc(&d_p, 4 * n));
    CK(cudaMemcpyAsync(d_o, origin, 12 * n, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(d_d, direction, 12 * n, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemsetAsync(c->cnt, 0, DEVCOUNTERS_FRAME_BYTES, c->stream));
    const int g = (int)((n + TRACE_BLOCK - 1) / TRACE_BLOCK);
    if (brute == 2) {  // through the ray-stream kernel the RATIO / PATH frames use (instrumented twin)
        NEED(n < ((size_t)1 << 31), "orb_trace_rays_host: too many rays for the stream");
        int rc = ensure_stream(c, n);
        if (rc) return rc;
        const unsigned nn = (unsigned)n;
        CK(cudaMemcpyAsync(&c->cnt->queue_n, &nn, 4, cudaMemcpyHostToDevice, c->stream));
        k_pack_rays<<<(nn + 255) / 256, 256, 0, c->stream>>>(d_o, d_d, nn, c->stream_rays);
        rc = launch_trace_stream(c, &c->cnt->queue_n, 1u, true);
        if (rc) return rc;
        k_unpack_hits<<<(nn + 255) / 256, 256, 0, c->stream>>>(c->stream_hits, nn, d_p, d_t, d_t + n, d_t + 2 * n);
        c->counters.kernel_launches += 2;
    } else {
        k_trace_rays<<<g, TRACE_BLOCK, trace_smem(c), c->stream>>>(c->view, d_o, d_d, n, brute, d_p, d_t, d_t + n, d_t + 2 * n,
                                                                   c->height, c->cnt);
        c->counters.kernel_launches++;
    }
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(prim, d_p, 4 * n, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(t, d_t, 4 * n, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(u, d_t + n, 4 * n, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(v, d_t + 2 * n, 4 * n, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    cudaFree(d_o); cudaFree(d_d); cudaFree(d_t); cudaFree(d_p);
    return ORB_OK;
}

// ---- allocation-free ray casting -------------------------------------------------------------
// orb_trace_rays_no_alloc(): the ray-stream kernel (k_trace_stream_spec) with its scratch -- the packed ray
// stream (32 B / ray), the hit records (16 B / ray), the claim cursor and the device-side ray count -- in ONE
// file-scope device array instead of cudaMalloc'd buffers, for callers that may not allocate (CUDA graph
// capture, serving loops).  The array is a fixed ORB_NO_ALLOC_SCRATCH_BYTES on every GPU that loads the
// library, until exit.  Every call on a device shares it: calls that overlap on the GPU race.
#define NOALLOC_HDR_BYTES 256u  // [0] claim cursor, [16] ray count (both cleared / set by k_noalloc_pack)
#define ORB_NO_ALLOC_SCRATCH_BYTES (NOALLOC_HDR_BYTES + (size_t)ORB_NO_ALLOC_MAX_RAYS * 48u)
static __device__ __align__(256) unsigned char g_noalloc_scratch[ORB_NO_ALLOC_SCRATCH_BYTES];

// `g_noalloc_scratch` in HOST code is only a handle for the symbol, not a device address; the address is asked of
// the runtime once per device (with that device current) and cached.
static int noalloc_scratch(int device, unsigned char** out) {
    static std::mutex mu;
    static unsigned char* base[64] = {};
    NEED(device >= 0 && device < 64, "orb_trace_rays_no_alloc: device index out of range");
    std::lock_guard<std::mutex> lock(mu);
    if (!base[device]) {
        unsigned char* p = nullptr;
        CK(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
        NEED(p != nullptr && ((uintptr_t)p & 255u) == 0, "orb_trace_rays_no_alloc: scratch symbol has no aligned device address");
        base[device] = p;
    }
    *out = base[device];
    return ORB_OK;
}

// rays into the scratch stream; thread 0 resets what must start at zero (the claim cursor) and publishes the count
__global__ void k_noalloc_pack(const float* __restrict__ o, const float* __restrict__ d, unsigned n, unsigned char* __restrict__ scratch) {
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) {
        reinterpret_cast<unsigned*>(scratch)[0] = 0u;
        reinterpret_cast<unsigned*>(scratch)[4] = n;
    }
    float4* rays = reinterpret_cast<float4*>(scratch + NOALLOC_HDR_BYTES);
    if (i < n) stream_put_ray(rays, i, v3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), v3(d[3 * i], d[3 * i + 1], d[3 * i + 2]));
}

int orb_trace_rays_no_alloc(OrbContext* c, const float* d_origin, const float* d_direction, uint32_t n, int32_t* d_prim,
                            float* d_t, float* d_u, float* d_v) {
    NEED(c && d_origin && d_direction && d_prim && d_t && d_u && d_v, "orb_trace_rays_no_alloc: NULL argument");
    NEED(c->committed, "orb_trace_rays_no_alloc: scene not committed");
    NEED(!c->geometry_dirty, "orb_trace_rays_no_alloc: orb_scene_update_mesh without orb_scene_refit (the BVH is stale)");
    if (n > ORB_NO_ALLOC_MAX_RAYS) {
        char msg[160];
        snprintf(msg, sizeof msg, "orb_trace_rays_no_alloc: %u rays exceed the fixed scratch (ORB_NO_ALLOC_MAX_RAYS = %u)", n,
                 (unsigned)ORB_NO_ALLOC_MAX_RAYS);
        return fail(ORB_ERR_INVALID, msg);
    }
    if (n == 0) return ORB_OK;
    CK(cudaSetDevice(c->device));
    unsigned char* scratch = nullptr;
    int rc = noalloc_scratch(c->device, &scratch);
    if (rc) return rc;
    unsigned* cursor = reinterpret_cast<unsigned*>(scratch);
    const unsigned* count = cursor + 4;
    float4* rays = reinterpret_cast<float4*>(scratch + NOALLOC_HDR_BYTES);
    float4* hits = rays + 2 * (size_t)ORB_NO_ALLOC_MAX_RAYS;
    cudaStream_t st = c->stream;
    k_noalloc_pack<<<(n + 255) / 256, 256, 0, st>>>(d_origin, d_direction, n, scratch);
    const int sd = stack_mode(c);
    const int block = c->stream_block;
    const size_t smem = stack_bytes(c, sd, block);
    const bool qn = use_qnodes(c);
    StreamKernel k = stream_kernel(sd, false, qn);
    if (smem > 48 * 1024) CK(allow_smem(k, smem));
    int g = persistent_grid(c, k, block, smem, 0);
    g = std::max(1, std::min(g, (int)((n + (unsigned)block - 1) / (unsigned)block)));
    k<<<g, block, smem, st>>>(c->view, rays, hits, count, 1u, cursor, c->refill_min, c->node_min, c->height, nullptr,
                              qn ? c->qnodes : (const QNode*)nullptr, qn ? c->qgrid : QGrid());
    k_unpack_hits<<<(n + 255) / 256, 256, 0, st>>>(hits, n, d_prim, d_t, d_u, d_v);
    c->counters.kernel_launches += 3;
    CK(cudaGetLastError());
    return ORB_OK;
}

// ---- known-answer entry points ---------------------------------------------------------------
static int run_unit(OrbContext* c, int op, const float* in, int n_in, float* out, int n_out) {
    NEED(c, "orb_unit: ctx is NULL");
    NEED(n_in <= 64 && n_out <= 4096, "orb_unit: too many values");
    CK(cudaSetDevice(c->device));
    refresh_view(c);
    CK(cudaMemcpyAsync(c->unit_in, in, n_in * 4, cudaMemcpyHostToDevice, c->stream));
    k_unit<<<1, 32, 0, c->stream>>>(op, c->view, c->unit_in, c->unit_out);
    c->counters.kernel_launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out, c->unit_out, n_out * 4, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return ORB_OK;
}

int orb_unit_rng(OrbContext* c, int frame_id, uint32_t px, uint32_t py, uint32_t W, int n, uint32_t* state0,
                 uint32_t* raw, float* flt) {
    NEED(n >= 0 && n <= 1000, "orb_unit_rng: n out of range");
    const float in[5] = {(float)frame_id, (float)px, (float)py, (float)W, (float)n};
    std::vector<float> out(1 + 2 * n);
    int rc = run_unit(c, U_RNG, in, 5, out.data(), 1 + 2 * n);
    if (rc) return rc;
    memcpy(state0, &out[0], 4);
    if (n) memcpy(raw, &out[1], 4 * n);
    if (n) memcpy(flt, &out[1 + n], 4 * n);
    return ORB_OK;
}

int orb_unit_clip(OrbContext* c, const float* abc9, int quad, int* n_out, float* form_factor) {
    float in[10];
    memcpy(in, abc9, 36);
    in[9] = quad ? 1.f : 0.f;
    float out[2];
    int rc = run_unit(c, U_CLIP, in, 10, out, 2);
    if (rc) return rc;
    *n_out = (int)out[0];
    *form_factor = out[1];
    return ORB_OK;
}

int orb_unit_integrate_edge(OrbContext* c, const float* a, const float* b, float* out) {
    float in[6];
    memcpy(in, a, 12);
    memcpy(in + 3, b, 12);
    return run_unit(c, U_EDGE, in, 6, out, 1);
}

int orb_unit_fetch_ltc(OrbContext* c, float alpha, float theta, float* m9_amp) {
    NEED(c && c->lut_w > 0, "orb_unit_fetch_ltc: orb_scene_set_ltc_lut first");
    const float in[2] = {alpha, theta};
    return run_unit(c, U_FETCH, in, 2, m9_amp, 10);
}

int orb_unit_mat_inverse(OrbContext* c, const float* m9, float* out9) { return run_unit(c, U_MINV, m9, 9, out9, 9); }
int orb_unit_onb(OrbContext* c, const float* n, float* rows9) { return run_unit(c, U_ONB, n, 3, rows9, 9); }

int orb_unit_sample_direction(OrbContext* c, const float* wo, float u1, float u2, const float* base, float alpha,
                              float* wi_pdf4) {
    const float in[9] = {wo[0], wo[1], wo[2], u1, u2, base[0], base[1], base[2], alpha};
    return run_unit(c, U_SAMPLE, in, 9, wi_pdf4, 4);
}

int orb_unit_evaluate(OrbContext* c, const float* wi, const float* wo, const float* base, float alpha, float* out3) {
    const float in[10] = {wi[0], wi[1], wi[2], wo[0], wo[1], wo[2], base[0], base[1], base[2], alpha};
    return run_unit(c, U_EVAL, in, 10, out3, 3);
}

int orb_unit_pdf(OrbContext* c, const float* wi, const float* wo, const float* base, float alpha, float* out) {
    const float in[10] = {wi[0], wi[1], wi[2], wo[0], wo[1], wo[2], base[0], base[1], base[2], alpha};
    return run_unit(c, U_PDF, in, 10, out, 1);
}

int orb_unit_make_rgba(OrbContext* c, const float* c3, uint32_t* out) {
    float o;
    int rc = run_unit(c, U_RGBA, c3, 3, &o, 1);
    if (rc) return rc;
    memcpy(out, &o, 4);
    return ORB_OK;
}

int orb_unit_sample_tri(OrbContext* c, const float* v9, float u1, float u2, float* out3) {
    float in[11];
    memcpy(in, v9, 36);
    in[9] = u1;
    in[10] = u2;
    return run_unit(c, U_SAMPLE_TRI, in, 11, out3, 3);
}

int orb_unit_pdf_a2w(OrbContext* c, float pdf, float dist2, float cos_theta, float* out) {
    const float in[3] = {pdf, dist2, cos_theta};
    return run_unit(c, U_A2W, in, 3, out, 1);
}

int orb_unit_ltc_pixel(OrbContext* c, const float* cam3, const float* p3, const float* n3, const float* diffuse3,
                       float alpha, float* out3) {
    NEED(c && c->lut_w > 0, "orb_unit_ltc_pixel: orb_scene_set_ltc_lut first");
    const float in[13] = {cam3[0], cam3[1], cam3[2], p3[0], p3[1], p3[2], n3[0], n3[1], n3[2],
                          diffuse3[0], diffuse3[1], diffuse3[2], alpha};
    return run_unit(c, U_LTC_PIXEL, in, 13, out3, 3);
}

}  // extern "C"

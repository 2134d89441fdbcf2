// bvh8.cuh -- compressed 8-wide BVH for the ray-stream kernels ("BVH8-compressed
// builder"), after Ylitie, Karras & Laine, "Efficient Incoherent Ray Traversal on GPUs Through Compressed
// Wide BVHs" (HPG 2017): the published algorithm restated, nothing taken from the reference (which leaves
// the acceleration structure to OptiX: include/viewer.hpp:506-507,516-517).
//
// Why: the BVH2 stream kernel is bound by its own node fetches (ncu profiles/r1m: L1 data pipe 86 % busy,
// 36 nodes x 64 B per shadow ray).  An 80-byte node here decides EIGHT subtrees: child boxes are 8-bit
// offsets on a per-node power-of-two grid, children sit in slots ordered by octant so that `slot ^ octant`
// is a front-to-back order without sorting, and a stack entry is a (node group, hit mask) pair.
//
// Result parity: quantised boxes CONTAIN the BVH2 boxes they were made from (rounded outwards, checked at
// build time), every box whose entry distance is <= the current closest t is opened, and the hit rule is
// hit_consider() (closest t, ties to the lowest primitive id) -- so the hit is the brute-force hit, bit
// for bit, as with the BVH2 (tests: test_hostcheck.py on the CPU, test_gpu_frames.py on the GPU).
//
// HBM layout:
//   nodes8  80 B per node, five 128-bit loads:
//             n0 = (origin.xyz, {ex, ey, ez, imask})      quantum 2^(e-127) per axis; imask: slots with inner nodes
//             n1 = (child_base, tri_base, meta[0..3], meta[4..7])
//             n2 = (qlo.x[0..7], qlo.y[0..7])  n3 = (qlo.z[0..7], qhi.x[0..7])  n4 = (qhi.y[0..7], qhi.z[0..7])
//           meta of an inner child in slot s: 001|11sss (24 + s); of a leaf of k <= 3 triangles at offset o
//           (< 24) of the node's triangle run: unary(k)|ooooo; empty slot: 0.
//   tri8    the 48-byte triangle records of scene.cuh, in node order (a node's leaf triangles are consecutive).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "scene.cuh"

namespace orb {

struct Bvh8Node {
    float px, py, pz;
    uint8_t ex, ey, ez, imask;
    uint32_t child_base, tri_base;
    uint8_t meta[8];
    uint8_t qlox[8], qloy[8], qloz[8], qhix[8], qhiy[8], qhiz[8];
};
static_assert(sizeof(Bvh8Node) == 80, "Bvh8Node layout");

#define ORB_BVH8_STACK 32  // one entry per level at most; deeper trees fall back to the BVH2 kernels

// ---------------------------------------------------------------------------------------------
// host: collapse a BVH2 (leaves of <= 3 triangles) into wide nodes
// ---------------------------------------------------------------------------------------------
struct Bvh8Host {
    std::vector<Bvh8Node> nodes;
    std::vector<int> tri_order;  // position in tri8 -> position in the BVH2's triangle array
    int height = 0;
};

struct Bvh8Child {
    int ref;
    float lo[3], hi[3];
};

inline void bvh2_children(const BvhNode& n, Bvh8Child& a, Bvh8Child& b) {
    a.ref = n.c0;
    a.lo[0] = n.a.x; a.lo[1] = n.a.y; a.lo[2] = n.a.z;
    a.hi[0] = n.a.w; a.hi[1] = n.b.x; a.hi[2] = n.b.y;
    b.ref = n.c1;
    b.lo[0] = n.b.z; b.lo[1] = n.b.w; b.lo[2] = n.c.x;
    b.hi[0] = n.c.y; b.hi[1] = n.c.z; b.hi[2] = n.c.w;
}

// root_lo / root_hi: bounds of the whole scene, used only when the BVH2 is a single leaf.
// Returns false if a leaf holds more than 3 triangles or a node's leaves more than 24.
inline bool bvh8_collapse(const BvhNode* n2, int root_ref, const float* root_lo, const float* root_hi, Bvh8Host& out) {
    out.nodes.clear();
    out.tri_order.clear();
    out.height = 0;
    struct Work { int node8, ref2, depth; };
    std::vector<Work> queue;
    out.nodes.emplace_back();
    queue.push_back({0, root_ref, 1});
    for (size_t head = 0; head < queue.size(); head++) {
        const Work w = queue[head];
        out.height = w.depth > out.height ? w.depth : out.height;
        Bvh8Child ch[8];
        int n = 0;
        if (w.ref2 < 0) {
            ch[0].ref = w.ref2;
            for (int a = 0; a < 3; a++) { ch[0].lo[a] = root_lo[a]; ch[0].hi[a] = root_hi[a]; }
            n = 1;
        } else {
            bvh2_children(n2[w.ref2], ch[0], ch[1]);
            n = 2;
            while (n < 8) {  // open the inner child with the largest surface area
                int best = -1;
                float best_area = -1.f;
                for (int i = 0; i < n; i++) {
                    if (ch[i].ref < 0) continue;
                    const float dx = ch[i].hi[0] - ch[i].lo[0], dy = ch[i].hi[1] - ch[i].lo[1], dz = ch[i].hi[2] - ch[i].lo[2];
                    const float area = dx * dy + dy * dz + dz * dx;
                    if (area > best_area) { best_area = area; best = i; }
                }
                if (best < 0) break;
                Bvh8Child a, b;
                bvh2_children(n2[ch[best].ref], a, b);
                ch[best] = a;
                ch[n++] = b;
            }
        }
        // node box and quantisation grid
        float lo[3], hi[3];
        for (int a = 0; a < 3; a++) {
            lo[a] = ch[0].lo[a]; hi[a] = ch[0].hi[a];
            for (int i = 1; i < n; i++) { lo[a] = std::fmin(lo[a], ch[i].lo[a]); hi[a] = std::fmax(hi[a], ch[i].hi[a]); }
        }
        // grid: origin a little below the box, quantum 2^e with 255 quanta reaching a little above it, so that
        // every child plane can sit `margin` quanta outside the BVH2 box it bounds
        int e[3];
        float org[3];
        for (int a = 0; a < 3; a++) {
            const double ext = (double)hi[a] - (double)lo[a];
            int ea = ext > 0.0 ? (int)std::ceil(std::log2(ext / 255.0)) : -126;
            ea = ea < -100 ? -100 : ea;
            for (;; ea++) {
                if (ea > 100) return false;  // the traversal scales the quantum by 2^15
                const double g = std::ldexp(1.0, ea);
                org[a] = (float)((double)lo[a] - g / 16.0);
                if ((double)org[a] <= (double)lo[a] - g / 32.0 && (double)org[a] + 255.0 * g >= (double)hi[a] + g / 32.0) break;
            }
            e[a] = ea;
        }
        // slots: child c in slot s scores dot(centroid_c - centre, sign(s)); x = bit 2, y = bit 1, z = bit 0
        int child_in[8];
        for (int s = 0; s < 8; s++) child_in[s] = -1;
        double score[8][8];
        for (int i = 0; i < n; i++)
            for (int s = 0; s < 8; s++) {
                double v = 0.0;
                for (int a = 0; a < 3; a++) {
                    const double c = 0.5 * ((double)ch[i].lo[a] + ch[i].hi[a]) - 0.5 * ((double)lo[a] + hi[a]);
                    v += ((s >> (2 - a)) & 1) ? c : -c;
                }
                score[i][s] = v;
            }
        for (int k = 0; k < n; k++) {  // greedy: best remaining (child, slot) pair
            int bi = -1, bs = -1;
            double bv = -1e300;
            for (int i = 0; i < n; i++) {
                bool used = false;
                for (int s = 0; s < 8; s++) used = used || child_in[s] == i;
                if (used) continue;
                for (int s = 0; s < 8; s++)
                    if (child_in[s] < 0 && score[i][s] > bv) { bv = score[i][s]; bi = i; bs = s; }
            }
            child_in[bs] = bi;
        }
        for (bool improved = true; improved;) {  // pairwise exchanges (also into empty slots) until none helps
            improved = false;
            for (int s1 = 0; s1 < 8; s1++)
                for (int s2 = s1 + 1; s2 < 8; s2++) {
                    const int a = child_in[s1], b = child_in[s2];
                    const double now = (a >= 0 ? score[a][s1] : 0.0) + (b >= 0 ? score[b][s2] : 0.0);
                    const double swp = (a >= 0 ? score[a][s2] : 0.0) + (b >= 0 ? score[b][s1] : 0.0);
                    if (swp > now + 1e-12 * (std::fabs(now) + std::fabs(swp))) {
                        child_in[s1] = b; child_in[s2] = a; improved = true;
                    }
                }
        }
        Bvh8Node nd;
        nd.px = org[0]; nd.py = org[1]; nd.pz = org[2];
        nd.ex = (uint8_t)(e[0] + 127); nd.ey = (uint8_t)(e[1] + 127); nd.ez = (uint8_t)(e[2] + 127);
        nd.imask = 0;
        nd.child_base = (uint32_t)out.nodes.size();
        nd.tri_base = (uint32_t)out.tri_order.size();
        int tri_off = 0;
        for (int s = 0; s < 8; s++) {
            const int i = child_in[s];
            uint8_t* q[6] = {nd.qlox, nd.qloy, nd.qloz, nd.qhix, nd.qhiy, nd.qhiz};
            if (i < 0) {
                nd.meta[s] = 0;
                for (int a = 0; a < 3; a++) { q[a][s] = 255; q[3 + a][s] = 0; }
                continue;
            }
            for (int a = 0; a < 3; a++) {
                const double p = (double)org[a], g = std::ldexp(1.0, e[a]);
                const double margin = 1.0 / 64.0;  // outwards, in quanta: the traversal's plane rounding (w_byte_m)
                int ql = (int)std::floor(((double)ch[i].lo[a] - p) / g - margin);
                int qh = (int)std::ceil(((double)ch[i].hi[a] - p) / g + margin);
                ql = ql < 0 ? 0 : (ql > 255 ? 255 : ql);
                qh = qh < 0 ? 0 : (qh > 255 ? 255 : qh);
                while (ql > 0 && p + ql * g > (double)ch[i].lo[a] - margin * g) ql--;
                while (qh < 255 && p + qh * g < (double)ch[i].hi[a] + margin * g) qh++;
                if (p + ql * g > (double)ch[i].lo[a] - margin * g || p + qh * g < (double)ch[i].hi[a] + margin * g)
                    return false;  // not a superset with margin: cannot happen with the grid chosen above
                q[a][s] = (uint8_t)ql;
                q[3 + a][s] = (uint8_t)qh;
            }
            if (ch[i].ref >= 0) {
                nd.imask |= (uint8_t)(1u << s);
                nd.meta[s] = (uint8_t)((1u << 5) | (24u + (unsigned)s));
                queue.push_back({(int)out.nodes.size(), ch[i].ref, w.depth + 1});
                out.nodes.emplace_back();
            } else {
                const int lr = ~ch[i].ref;
                const int first = lr >> 3, count = (lr & 7) + 1;
                if (count > 3 || tri_off + count > 24) return false;
                nd.meta[s] = (uint8_t)((((1u << count) - 1u) << 5) | (unsigned)tri_off);
                for (int k = 0; k < count; k++) out.tri_order.push_back(first + k);
                tri_off += count;
            }
        }
        out.nodes[w.node8] = nd;
    }
    return true;
}

// ---------------------------------------------------------------------------------------------
// traversal (host and device)
// ---------------------------------------------------------------------------------------------
struct Bvh8View {
    const float4* nodes;  // 5 float4 per node
    const float4* tri;    // 3 float4 per triangle, node order
};

struct WideState {
    float3 o, d, inv;
    Hit best;
    uint32_t ng_base, ng_hits;  // node group: first child of the parent, (hit bits 31..24) | imask
    uint32_t octinv4;
    int sp;
    bool done;
};

#ifdef __CUDA_ARCH__
// prmt.b32 itself: selector nibbles 8..15 replicate the SIGN of byte (nibble & 7), which __byte_perm does not promise
__device__ __forceinline__ uint32_t w_prmt(uint32_t a, uint32_t b, uint32_t sel) {
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
    return r;
}
#endif
ORB_HD uint32_t w_sext_bytes(uint32_t x) {  // every byte -> 0xff if its top bit is set, else 0
#ifdef __CUDA_ARCH__
    return w_prmt(x, 0u, 0xba98u);
#else
    uint32_t r = 0;
    for (int j = 0; j < 4; j++)
        if ((x >> (8 * j + 7)) & 1u) r |= 0xffu << (8 * j);
    return r;
#endif
}
ORB_HD int w_bfind(uint32_t x) {
#ifdef __CUDA_ARCH__
    return 31 - __clz((int)x);
#else
    return 31 - __builtin_clz(x);
#endif
}
ORB_HD int w_popc(uint32_t x) {
#ifdef __CUDA_ARCH__
    return __popc(x);
#else
    return __builtin_popcount(x);
#endif
}
// byte j of w as the float 1 + q * 2^-15 (0x3F80qq00): one PRMT, no int-to-float conversion (that pipe runs
// at a quarter of the FMA rate and 48 conversions per node made it the bottleneck).  A plane distance is
// then fma(m, A, o - A) with A = quantum * 2^15 / d; rounding o - A moves a plane by at most 2^-9 quantum,
// which the builder's outward margin of 2^-6 quantum covers.
ORB_HD float w_byte_m(uint32_t w, int j) {
#ifdef __CUDA_ARCH__
    return __uint_as_float(w_prmt(w, 0x3f800000u, 0x7604u | ((unsigned)j << 4)));
#else
    uint32_t u = 0x3f800000u | (((w >> (8 * j)) & 0xffu) << 8);
    float f;
    memcpy(&f, &u, 4);
    return f;
#endif
}
ORB_HD uint32_t w_ubyte(uint32_t w, int j) {
#ifdef __CUDA_ARCH__
    return w_prmt(w, 0u, 0x4440u | (unsigned)j);
#else
    return (w >> (8 * j)) & 0xffu;
#endif
}
ORB_HD uint32_t w_bits(float f) {
#ifdef __CUDA_ARCH__
    return __float_as_uint(f);
#else
    uint32_t u;
    memcpy(&u, &f, 4);
    return u;
#endif
}
ORB_HD float w_float(uint32_t u) {
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f;
    memcpy(&f, &u, 4);
    return f;
#endif
}

ORB_HD void wide_begin(WideState& t, int n_tris, float3 o, float3 d) {
    t.o = o;
    t.d = d;
    t.inv = v3(fabsf(d.x) < 1e-30f ? copysignf(1e30f, d.x) : 1.f / d.x, fabsf(d.y) < 1e-30f ? copysignf(1e30f, d.y) : 1.f / d.y,
               fabsf(d.z) < 1e-30f ? copysignf(1e30f, d.z) : 1.f / d.z);
    t.best.t = INFINITY;
    t.best.u = t.best.v = 0.f;
    t.best.prim = -1;
    t.sp = 0;
    const uint32_t oct = (t.inv.x < 0.f ? 4u : 0u) | (t.inv.y < 0.f ? 2u : 0u) | (t.inv.z < 0.f ? 1u : 0u);
    t.octinv4 = (7u - oct) * 0x01010101u;
    t.ng_base = 0u;
    t.ng_hits = 0x80000000u;  // "the root was hit"
    const float chk = (o.x + o.y + o.z) + (d.x + d.y + d.z);  // NaN rays miss at once (see trav_begin)
    t.done = n_tris == 0 || chk != chk;
}

// One step of a live ray: open the nearest unopened node of the current group, test the triangles of its
// hit leaves, then take the next group from the stack if this one is used up.
template <class StackPtr>
ORB_HD void wide_step(WideState& t, const Bvh8View& s, StackPtr stk, int stride, unsigned* n_nodes = nullptr,
                      unsigned* n_tris = nullptr) {
    // ---- choose the child: highest hit bit = nearest in this ray's octant order ----
    const uint32_t hits = t.ng_hits;
    const int bit = w_bfind(hits);
    const uint32_t rest = hits & ~(1u << bit);
    if (rest > 0x00ffffffu) {  // siblings still to open
        stk[2 * t.sp * stride] = t.ng_base;
        stk[(2 * t.sp + 1) * stride] = rest;
        ++t.sp;
    }
    const uint32_t slot = (uint32_t)(bit - 24) ^ (t.octinv4 & 7u);
    const uint32_t node = t.ng_base + (uint32_t)w_popc(hits & ~(0xffffffffu << slot));  // inner siblings in lower slots
    const float4* np = s.nodes + 5 * (size_t)node;
#ifdef __CUDA_ARCH__
    const float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
#else
    const float4 n0 = np[0], n1 = np[1], n2 = np[2], n3 = np[3], n4 = np[4];
#endif
    if (n_nodes) ++*n_nodes;
    const uint32_t ebits = w_bits(n0.w);
    // A = quantum * 2^15 / d per axis (see w_byte_m), o = (origin - ray origin) / d - A
    const float ax = w_float(((ebits & 0xffu) + 15u) << 23) * t.inv.x;
    const float ay = w_float((((ebits >> 8) & 0xffu) + 15u) << 23) * t.inv.y;
    const float az = w_float((((ebits >> 16) & 0xffu) + 15u) << 23) * t.inv.z;
    const float ox = (n0.x - t.o.x) * t.inv.x - ax, oy = (n0.y - t.o.y) * t.inv.y - ay, oz = (n0.z - t.o.z) * t.inv.z - az;
    const float lim = t.best.prim < 0 ? ORB_TMAX : t.best.t;
    uint32_t hitmask = 0u;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int h = 0; h < 2; h++) {
        const uint32_t meta4 = w_bits(h ? n1.w : n1.z);
        const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
        const uint32_t inner_mask4 = w_sext_bytes(is_inner4 << 3);
        const uint32_t bit_index4 = (meta4 ^ (t.octinv4 & inner_mask4)) & 0x1f1f1f1fu;
        const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
        const uint32_t lox = w_bits(h ? n2.y : n2.x), loy = w_bits(h ? n2.w : n2.z), loz = w_bits(h ? n3.y : n3.x);
        const uint32_t hix = w_bits(h ? n3.w : n3.z), hiy = w_bits(h ? n4.y : n4.x), hiz = w_bits(h ? n4.w : n4.z);
        const uint32_t nx = t.inv.x < 0.f ? hix : lox, fx = t.inv.x < 0.f ? lox : hix;  // near / far planes
        const uint32_t ny = t.inv.y < 0.f ? hiy : loy, fy = t.inv.y < 0.f ? loy : hiy;
        const uint32_t nz = t.inv.z < 0.f ? hiz : loz, fz = t.inv.z < 0.f ? loz : hiz;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int j = 0; j < 4; j++) {
            const float t0 = fmaxf(fmaxf(xfma(w_byte_m(nx, j), ax, ox), xfma(w_byte_m(ny, j), ay, oy)),
                                   fmaxf(xfma(w_byte_m(nz, j), az, oz), 0.f));
            const float t1 = fminf(fminf(xfma(w_byte_m(fx, j), ax, ox), xfma(w_byte_m(fy, j), ay, oy)),
                                   fminf(xfma(w_byte_m(fz, j), az, oz), lim));
            if (t0 <= t1) hitmask |= w_ubyte(child_bits4, j) << w_ubyte(bit_index4, j);
        }
    }
    t.ng_base = w_bits(n1.x);
    t.ng_hits = (hitmask & 0xff000000u) | (ebits >> 24);
    // ---- triangles of the leaves that were hit ----
    uint32_t tg = hitmask & 0x00ffffffu;
    const uint32_t tri_base = w_bits(n1.y);
    while (tg) {
        const int k = w_bfind(tg);
        tg &= ~(1u << k);
        const float4* tp = s.tri + 3 * (size_t)(tri_base + (uint32_t)k);
#ifdef __CUDA_ARCH__
        const float4 t0 = __ldg(tp), t1 = __ldg(tp + 1), t2 = __ldg(tp + 2);
#else
        const float4 t0 = tp[0], t1 = tp[1], t2 = tp[2];
#endif
        if (n_tris) ++*n_tris;
        float tt, u, v;
        if (tri_test(t0, t1, t2, t.o, t.d, 0.f, ORB_TMAX, tt, u, v)) hit_consider(t.best, tt, u, v, __float_as_int_hd(t0.w));
    }
    // ---- next group ----
    if (t.ng_hits <= 0x00ffffffu) {
        if (t.sp > 0) {
            --t.sp;
            t.ng_base = stk[2 * t.sp * stride];
            t.ng_hits = stk[(2 * t.sp + 1) * stride];
        } else {
            t.done = true;
        }
    }
}

template <class StackPtr>
ORB_HD Hit trace_closest_wide(const Bvh8View& s, int n_tris, float3 o, float3 d, StackPtr stk, int stride,
                              unsigned* n_nodes = nullptr, unsigned* n_tris_tested = nullptr) {
    WideState t;
    wide_begin(t, n_tris, o, d);
    while (!t.done) wide_step(t, s, stk, stride, n_nodes, n_tris_tested);
    return t.best;
}

}  // namespace orb

// bvhq.cuh -- 32-byte BVH2 nodes: both child boxes on a 16-bit grid over the scene box, so ONE 256-bit
// load fetches a node (the fp32 node of scene.cuh takes two).  ncu on the ray-stream kernel
// (profiles/r1o_ncu_summary.md) shows the L1 data pipe 86 % busy with node fetches in which every lane
// reads its own node; this halves those wavefronts and the node footprint (64 -> 32 MB at 1 M triangles).
//
// STATUS (round 1): the format, its conservative encoder and the traversal step are verified on the CPU
// against brute force (tests/test_hostcheck.py: hits bit-identical, node visits +0.1 % at 40 k triangles) and run
// in the GPU ray-stream kernel behind ORB_TRACE_VARIANT=3 (k_trace_stream_spec<.., QN = true>: one LDG.256 and
// 12 more LOP3 per node).  Measured on the bench frame: trace stage 1.669 ms against 1.691 ms, 35.98 node
// fetches per ray against 35.86 (profiles/r1q_qnodes_bench.json); hit-parity and refit tests pass under it
// (tools/ab_qnodes.sh; tests/test_gpu_frames.py test_quantised_node_stream_kernel).  Not the default yet: the
// whole GPU suite has not run under it.
//
// Same traversal as scene.cuh otherwise (owl::traceRay -> optixTrace replacement; call sites
// cuda_src/deviceCode.cu:76, cuda_include/ratio/ratio.cuh:50, cuda_include/path/path.cuh:219).
//
// Grid: plane(q) = origin + q * step per axis, q in [0, 65535]; origin / step from the root's two child
// boxes (every descendant box lies inside them) with four quanta of margin.
// Encoding is conservative: lo -> floor - 1, hi -> ceil + 1 (computed in double).  The extra quantum pays
// for the decode, which never converts an integer: the 16-bit value is OR-ed into the mantissa of 2^23
// (one LOP3 / PRMT), f = 8388608 + q, and the ray carries
//     inv' = step * inv,   noi' = (origin - o) * inv - 8388608 * inv'
// so that the slab distance is the same single FMA as for fp32 nodes, f * inv' + noi' -- box_test() is reused
// unchanged.  noi' is rounded at the magnitude 2^23 * step * |inv|, i.e. to half a quantum of plane position;
// all other roundings are below 2^-24 of the scene extent.  Hence decoded planes are never inside the true box,
// and the closest hit (closest t, ties to the lowest primitive id) is the one every other hierarchy finds.
#pragma once
#include <stdint.h>

#include "scene.cuh"

namespace orb {

struct QNode {
    uint32_t w[6];  // u16 pairs (low half first): c0.lo.x c0.lo.y | c0.lo.z c0.hi.x | c0.hi.y c0.hi.z | same for c1
    int c0, c1;     // child references as in BvhNode
};
static_assert(sizeof(QNode) == 32, "QNode layout");

struct QGrid {
    float3 origin, step;
};

// From the root node of the fp32 hierarchy (both child boxes, already padded by the builder).
ORB_HD QGrid qgrid_from_root(const BvhNode& r) {
    const float lo[3] = {fminf(r.a.x, r.b.z), fminf(r.a.y, r.b.w), fminf(r.a.z, r.c.x)};
    const float hi[3] = {fmaxf(r.a.w, r.c.y), fmaxf(r.b.x, r.c.z), fmaxf(r.b.y, r.c.w)};
    float st[3], og[3];
    for (int a = 0; a < 3; a++) {
        const float ext = fmaxf(hi[a] - lo[a], 1e-20f);
        st[a] = ext / 65527.f;  // 65535 - 8: four quanta of margin at either end
        og[a] = lo[a] - 4.f * st[a];
    }
    QGrid g;
    g.origin = v3(og[0], og[1], og[2]);
    g.step = v3(st[0], st[1], st[2]);
    return g;
}

ORB_HD uint32_t q_lo(float x, float origin, float step) {
    const double q = floor(((double)x - (double)origin) / (double)step) - 1.0;
    return (uint32_t)(q < 0.0 ? 0.0 : (q > 65535.0 ? 65535.0 : q));
}
ORB_HD uint32_t q_hi(float x, float origin, float step) {
    const double q = ceil(((double)x - (double)origin) / (double)step) + 1.0;
    return (uint32_t)(q < 0.0 ? 0.0 : (q > 65535.0 ? 65535.0 : q));
}

ORB_HD QNode q_encode(const BvhNode& n, const QGrid& g) {
    QNode o;
    const uint32_t l0x = q_lo(n.a.x, g.origin.x, g.step.x), l0y = q_lo(n.a.y, g.origin.y, g.step.y),
                   l0z = q_lo(n.a.z, g.origin.z, g.step.z);
    const uint32_t h0x = q_hi(n.a.w, g.origin.x, g.step.x), h0y = q_hi(n.b.x, g.origin.y, g.step.y),
                   h0z = q_hi(n.b.y, g.origin.z, g.step.z);
    const uint32_t l1x = q_lo(n.b.z, g.origin.x, g.step.x), l1y = q_lo(n.b.w, g.origin.y, g.step.y),
                   l1z = q_lo(n.c.x, g.origin.z, g.step.z);
    const uint32_t h1x = q_hi(n.c.y, g.origin.x, g.step.x), h1y = q_hi(n.c.z, g.origin.y, g.step.y),
                   h1z = q_hi(n.c.w, g.origin.z, g.step.z);
    o.w[0] = l0x | (l0y << 16);
    o.w[1] = l0z | (h0x << 16);
    o.w[2] = h0y | (h0z << 16);
    o.w[3] = l1x | (l1y << 16);
    o.w[4] = l1z | (h1x << 16);
    o.w[5] = h1y | (h1z << 16);
    o.c0 = n.c0;
    o.c1 = n.c1;
    return o;
}

// 8388608 + q as a float, without an integer conversion
ORB_HD float q_f_lo(uint32_t w) { return __int_as_float_hd((int)(0x4B000000u | (w & 0xffffu))); }
ORB_HD float q_f_hi(uint32_t w) { return __int_as_float_hd((int)(0x4B000000u | (w >> 16))); }

// trav_begin() for quantised nodes: t.inv / t.noi become the grid-space constants box_test() consumes.
ORB_HD void trav_begin_q(TravState& t, const SceneView& s, const QGrid& g, float3 o, float3 d) {
    trav_begin(t, s, o, d);
    // trav_begin() substitutes +-1e30 for the reciprocal of a zero component; 2^23 * step * 1e30 overflows once
    // the grid step exceeds ~40 (scene extent ~2.6e6), the slab distances become inf - inf and every node is
    // rejected.  Clamp the reciprocal so the grid-space constants stay finite: the slab distances of such an
    // axis remain far beyond any hit distance with the right sign, which is all the test needs.
    const float lx = fminf(1e30f, 1e37f / (8388608.f * g.step.x)), ly = fminf(1e30f, 1e37f / (8388608.f * g.step.y)),
                lz = fminf(1e30f, 1e37f / (8388608.f * g.step.z));
    const float3 inv = v3(fmaxf(fminf(t.inv.x, lx), -lx), fmaxf(fminf(t.inv.y, ly), -ly), fmaxf(fminf(t.inv.z, lz), -lz));
    t.inv = v3(g.step.x * inv.x, g.step.y * inv.y, g.step.z * inv.z);
    t.noi = v3(xfma(-8388608.f, t.inv.x, (g.origin.x - o.x) * inv.x), xfma(-8388608.f, t.inv.y, (g.origin.y - o.y) * inv.y),
               xfma(-8388608.f, t.inv.z, (g.origin.z - o.z) * inv.z));
}

// One inner node of the quantised hierarchy; stack and order exactly as trav_node_step().
template <class Stack>
ORB_HD void trav_node_step_qn(TravState& t, const QNode* __restrict__ nodes, Stack& st) {
    const QNode* np = nodes + t.cur;
    uint32_t w0, w1, w2, w3, w4, w5;
    int c0, c1;
#ifdef __CUDA_ARCH__
    asm("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r"(w0), "=r"(w1), "=r"(w2), "=r"(w3), "=r"(w4), "=r"(w5), "=r"(c0), "=r"(c1)
        : "l"(np));
#else
    w0 = np->w[0], w1 = np->w[1], w2 = np->w[2], w3 = np->w[3], w4 = np->w[4], w5 = np->w[5];
    c0 = np->c0, c1 = np->c1;
#endif
    const float lim = t.best.prim < 0 ? ORB_TMAX : t.best.t;
    float tn0, tn1;
    const bool h0 = box_test(q_f_lo(w0), q_f_hi(w0), q_f_lo(w1), q_f_hi(w1), q_f_lo(w2), q_f_hi(w2), t.noi, t.inv, lim, tn0);
    const bool h1 = box_test(q_f_lo(w3), q_f_hi(w3), q_f_lo(w4), q_f_hi(w4), q_f_lo(w5), q_f_hi(w5), t.noi, t.inv, lim, tn1);
    trav_descend(t, st, h0, h1, tn0, tn1, c0, c1);
}

template <class Stack>
ORB_HD Hit trace_closest_q(const SceneView& s, const QNode* nodes, const QGrid& g, float3 o, float3 d, Stack& st,
                           unsigned* n_nodes = nullptr, unsigned* n_tris = nullptr) {
    TravState t;
    trav_begin_q(t, s, g, o, d);
    while (t.cur != ORB_TRAVERSAL_DONE) {
        while (t.cur >= 0) {
            if (n_nodes) ++*n_nodes;
            trav_node_step_qn(t, nodes, st);
        }
        if (t.cur == ORB_TRAVERSAL_DONE) break;
        trav_leaf_step(t, s, st, n_tris);
    }
    return t.best;
}
ORB_HD Hit trace_closest_q(const SceneView& s, const QNode* nodes, const QGrid& g, float3 o, float3 d, int* stk,
                           float* stk_t, int stride, unsigned* n_nodes = nullptr, unsigned* n_tris = nullptr) {
    StridedStack st{stk, stk_t, stride};
    return trace_closest_q(s, nodes, g, o, d, st, n_nodes, n_tris);
}

}  // namespace orb

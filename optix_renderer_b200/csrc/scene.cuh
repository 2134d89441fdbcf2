// scene.cuh -- scene data as the kernels see it in HBM, the ray/triangle test, BVH traversal and
// the closest-hit / miss shading record.
//
// Replaces, for the per-pixel hot path:
//   owl::traceRay -> optixTrace over OWL's IAS -> per-mesh GAS        (call sites
//       cuda_src/deviceCode.cu:76, cuda_include/ratio/ratio.cuh:50, cuda_include/path/path.cuh:219;
//       build include/viewer.hpp:506-507,516-517)                      -> trace_closest()
//   __closesthit__triangleMeshCH, cuda_include/hit_miss.cuh:14-50      -> shade_hit()
//   __miss__miss, cuda_include/hit_miss.cuh:52-63                      -> Surface defaults
//   barycentricInterpolate, cuda_include/utils.cuh:10-29
//
// HBM layout (all arrays 16-byte aligned, built once per scene, replicated per GPU):
//   tri, tri_b   per triangle in BVH-leaf order: 32 bytes (v0.xyz, global prim id | e1.xyz, e2.x) + 8 bytes
//            (e2.y, e2.z), with e1 = v1 - v0, e2 = v2 - v0 rounded in fp32 -- the operands the checker's
//            Moeller-Trumbore forms at test time.  One 256-bit and one 64-bit load per lane.
//   nodes    64-byte BVH2 nodes, each holding BOTH child boxes and both child references, so
//            one node fetch (two 256-bit loads) decides two subtrees; the traversal kernels read the
//            32-byte quantised copy of them (bvhq.cuh), the builders and the refit these.
//   pos/nrm/uv  per-global-primitive corner attributes, un-indexed like the reference's meshes
//            (src/Model.cpp:186-201): 9 + 9 + 6 floats, gathered only at the closest hit.
//   prim_mesh, meshes   global prim -> mesh record (TriangleMeshData without the pointers).
// Global primitive id = triangles of earlier meshes + local primitive index, meshes in scene
// order (surfaces, then lights; src/scene.cpp:5-13) = (instance, primitive) of the reference.
#pragma once
#include "ltc.cuh"
#include "vec.cuh"

namespace orb {

struct BvhNode {
    float4 a;  // c0.lo.x c0.lo.y c0.lo.z c0.hi.x
    float4 b;  // c0.hi.y c0.hi.z c1.lo.x c1.lo.y
    float4 c;  // c1.lo.z c1.hi.x c1.hi.y c1.hi.z
    int c0, c1;  // child reference: >= 0 inner node index, < 0 leaf = ~((first << 3) | (count-1))
    int pad0, pad1;
};
static_assert(sizeof(BvhNode) == 64, "BvhNode layout");

ORB_HD int leaf_ref(int first, int count) { return ~((first << 3) | (count - 1)); }

struct MeshRec {  // TriangleMeshData, cuda_include/common.cuh:130-151, minus pointers
    float3 diffuse;
    float alpha;
    float3 emit;
    int is_light;
    uint32_t material_id;
    int tex_w, tex_h;
    int pad;
    const uchar4* tex;  // RGBA8 diffuse texture or nullptr (nearest / clamp, viewer.hpp:441-447)
};

struct SceneView {
    const float4* tri;    // 2 x float4 per triangle: (v0.xyz, prim id) (e1.xyz, e2.x)
    const float2* tri_b;  // (e2.y, e2.z)
    const BvhNode* nodes;
    int root_ref;  // child-style reference of the root; 0 triangles -> n_tris == 0
    int n_tris;
    int n_nodes;   // inner nodes (ORB_CHECKS: bound of every node reference)
    const float* pos;
    const float* nrm;
    const float* uv;
    const int* prim_mesh;
    const MeshRec* meshes;
    const TriLight* lights;
    int n_lights;
    LtcLut lut;
};

struct Hit {
    float t, u, v;
    int prim;  // global primitive id, -1 = miss
};

#define ORB_TMAX 1e30f  // owl::RayT default tmax; tmin = 0

// Triangle record k as the three operands of tri_test(): (v0, prim id), (e1, -), (e2, -).
ORB_HD void tri_record(float* a8, float* b2, const float* p9, int prim) {  // writer side: builders, refit, CPU checks
    a8[0] = p9[0], a8[1] = p9[1], a8[2] = p9[2], a8[3] = __int_as_float_hd(prim);
    a8[4] = xsub(p9[3], p9[0]), a8[5] = xsub(p9[4], p9[1]), a8[6] = xsub(p9[5], p9[2]);
    a8[7] = xsub(p9[6], p9[0]);
    b2[0] = xsub(p9[7], p9[1]), b2[1] = xsub(p9[8], p9[2]);
}
ORB_HD void tri_fetch(const SceneView& s, int k, float4& t0, float4& t1, float4& t2) {
    ORB_DEVICE_ASSERT(k >= 0 && k < s.n_tris);
#ifdef __CUDA_ARCH__
    float e2x;
#if defined(ORB_TRI_FETCH) && ORB_TRI_FETCH == 1  // A/B: two 128-bit loads instead of one 256-bit load
    t0 = __ldg(s.tri + 2 * (size_t)k);
    t1 = __ldg(s.tri + 2 * (size_t)k + 1);
    e2x = t1.w;
#else
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(t0.x), "=f"(t0.y), "=f"(t0.z), "=f"(t0.w), "=f"(t1.x), "=f"(t1.y), "=f"(t1.z), "=f"(e2x)
        : "l"(s.tri + 2 * (size_t)k));
#endif
    const float2 b = __ldg(s.tri_b + k);
#else
    const float4 a0 = s.tri[2 * (size_t)k], a1 = s.tri[2 * (size_t)k + 1];
    const float2 b = s.tri_b[k];
    const float e2x = a1.w;
    t0 = a0;
    t1 = a1;
#endif
    t1.w = 0.f;
    t2 = make_float4(e2x, b.x, b.y, 0.f);
}

// Moeller-Trumbore exactly as oracle/trace_cpu.h tri_test(): every rounding step is explicit.
ORB_HD bool tri_test(float4 t0, float4 t1, float4 t2, float3 o, float3 d, float tmin, float tmax, float& t, float& u,
                     float& v) {
    const float px = xfma(d.y, t2.z, -xmul(d.z, t2.y));
    const float py = xfma(d.z, t2.x, -xmul(d.x, t2.z));
    const float pz = xfma(d.x, t2.y, -xmul(d.y, t2.x));
    const float det = xfma(t1.z, pz, xfma(t1.y, py, xmul(t1.x, px)));
    if (det == 0.f) return false;
    const float inv = xdiv(1.f, det);
    const float sx = xsub(o.x, t0.x), sy = xsub(o.y, t0.y), sz = xsub(o.z, t0.z);
    const float uu = xmul(xfma(sz, pz, xfma(sy, py, xmul(sx, px))), inv);
    if (!(uu >= 0.f && uu <= 1.f)) return false;
    const float qx = xfma(sy, t1.z, -xmul(sz, t1.y));
    const float qy = xfma(sz, t1.x, -xmul(sx, t1.z));
    const float qz = xfma(sx, t1.y, -xmul(sy, t1.x));
    const float vv = xmul(xfma(d.z, qz, xfma(d.y, qy, xmul(d.x, qx))), inv);
    if (!(vv >= 0.f && xadd(uu, vv) <= 1.f)) return false;
    const float tt = xmul(xfma(t2.z, qz, xfma(t2.y, qy, xmul(t2.x, qx))), inv);
    if (!(tt > tmin && tt < tmax)) return false;
    t = tt;
    u = uu;
    v = vv;
    return true;
}

// closest t wins; equal t goes to the lowest global primitive id (deterministic in any order)
ORB_HD void hit_consider(Hit& best, float t, float u, float v, int prim) {
    if (t < best.t || (t == best.t && prim < best.prim)) {
        best.t = t;
        best.u = u;
        best.v = v;
        best.prim = prim;
    }
}

// Conservative slab test: entry distance (widened by a few ulp) or a negative value on a miss.
// Boxes are already padded by the builder, so a box is never rejected for a triangle the test
// above would accept at t <= tbest.
// Plane distances as ONE fma each, plane * inv + (-o * inv): half the arithmetic of (plane - o) * inv.
// Its extra error is about one ulp of the origin's coordinate in space, far inside the builder's
// padding of 1e-5 x scene extent.
ORB_HD bool box_test(float lox, float loy, float loz, float hix, float hiy, float hiz, float3 noi, float3 inv,
                     float tbest, float& tnear) {
    const float ax = xfma(lox, inv.x, noi.x), bx = xfma(hix, inv.x, noi.x);
    const float ay = xfma(loy, inv.y, noi.y), by = xfma(hiy, inv.y, noi.y);
    const float az = xfma(loz, inv.z, noi.z), bz = xfma(hiz, inv.z, noi.z);
    float t0 = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.f));
    float t1 = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), tbest));
    t0 = t0 - fabsf(t0) * 1e-6f;
    t1 = t1 + fabsf(t1) * 1e-6f;
    tnear = t0;
    return t0 <= t1;
}

// Closest-hit traversal, "while-while" form (Aila & Laine, HPG 2009): a lane walks INNER nodes
// until it holds a leaf (or is finished) and only then does the warp run the triangle tests, so
// the expensive Moeller-Trumbore block executes with most lanes active instead of the two or
// three that happened to reach a leaf in the same iteration (ncu, profiles/r1a: 2.3 active lanes
// in the leaf block of the if-if form).  The result does not depend on the visiting order (closest t,
// ties to the lowest primitive id), so any BVH and any order agree bit for bit.
#define ORB_TRAVERSAL_DONE ((int)0x80000000)  // never a leaf reference: first + count < 2^28

// ---- traversal stacks: entry = (child reference, entry distance of its box) --------------------------------
// StridedStack: two arrays indexed [i * stride] -- plain arrays on the CPU, thread-local arrays (stride 1), or
// per-lane columns of shared memory (entry e of thread t at [e * BLOCK + t]: any mix of stack depths inside a
// warp is bank-conflict free).
struct StridedStack {
    int* r;
    float* t;
    int stride;
    ORB_HD void put(int i, int ref, float tn) const {
        r[i * stride] = ref;
        t[i * stride] = tn;
    }
    ORB_HD float dist(int i) const { return t[i * stride]; }
    ORB_HD int ref(int i) const { return r[i * stride]; }
};

// HybridStack: the D entries nearest the root in SHARED memory (one 64-bit word per entry, columns as above),
// deeper ones in thread-local memory.  Why (ncu --set full of the round-1 kernel, profiles/r2b): the ray-stream
// kernel is bound by the L1 data pipe -- 402 M wavefronts per 5 M-ray launch = 82 % of one wavefront per cycle
// and SM -- because every lane's access is its own wavefront: two per node (64-byte node = 2 x LDG.256), three
// per triangle (3 x LDG.128), and the thread-local stack costs one per DISTINCT stack depth in the warp for each
// of its four accesses per push / pop (incoherent rays: nearly one per lane).  The local stack also takes ~4 KB
// of L1 per warp (147 KB per SM at 36 warps) from the node data.  A shared-memory column is ONE wavefront per
// warp and instruction whatever the depths, and leaves L1 to the nodes.  (A full-height shared stack was
// measured slower in round 1: 27 KB per CTA cut the resident warps.  D entries cost 8 D bytes per thread.)
#define ORB_LOCAL_STACK_MAX 64
template <int D>
struct HybridStack {
    unsigned long long* s;  // this thread's column: entry e at s[e * stride]
    int stride;             // threads per CTA (columns side by side); 1 for a private array
    unsigned long long l[ORB_LOCAL_STACK_MAX - D];
    ORB_HD static unsigned long long pack(int ref, float tn) {
        return ((unsigned long long)(unsigned)__float_as_int_hd(tn) << 32) | (unsigned long long)(unsigned)ref;
    }
    ORB_HD void put(int i, int ref, float tn) {
        ORB_DEVICE_ASSERT(i >= 0 && i < ORB_LOCAL_STACK_MAX);
        if (i < D) s[i * stride] = pack(ref, tn);
        else l[i - D] = pack(ref, tn);
    }
    ORB_HD unsigned long long get(int i) const {
        ORB_DEVICE_ASSERT(i >= 0 && i < ORB_LOCAL_STACK_MAX);
        return i < D ? s[i * stride] : l[i - D];
    }
};

// pop: skip subtrees that start beyond the current closest hit
template <class Stack>
ORB_HD int stack_pop(const Stack& st, int& sp, const Hit& best) {
    for (;;) {
        if (sp == 0) return ORB_TRAVERSAL_DONE;
        --sp;
        if (best.prim < 0 || !(st.dist(sp) > best.t)) return st.ref(sp);
    }
}
template <int D>
ORB_HD int stack_pop(const HybridStack<D>& st, int& sp, const Hit& best) {
    for (;;) {
        if (sp == 0) return ORB_TRAVERSAL_DONE;
        --sp;
        const unsigned long long e = st.get(sp);  // one 64-bit load: distance and reference together
        if (best.prim < 0 || !(__int_as_float_hd((int)(unsigned)(e >> 32)) > best.t)) return (int)(unsigned)(e & 0xffffffffull);
    }
}

// Traversal state of one ray; a "macro step" walks inner nodes until a leaf is reached, tests
// that leaf, and pops.  trace_closest() runs macro steps to completion; the ray-stream kernel
// (orb200.cu k_trace_stream_spec) interleaves them with refills so finished lanes get new rays.
struct TravState {
    float3 o, d, inv, noi;  // noi = -o * inv
    Hit best;
    int cur, sp;
};

ORB_HD void trav_begin(TravState& t, const SceneView& s, float3 o, float3 d) {
    t.o = o;
    t.d = d;
    // a zero (or denormal) component would make plane * inv + noi an inf - inf; 1e30 keeps every slab
    // distance finite with the right sign for coordinates up to 1e7
    t.inv = v3(fabsf(d.x) < 1e-30f ? copysignf(1e30f, d.x) : 1.f / d.x, fabsf(d.y) < 1e-30f ? copysignf(1e30f, d.y) : 1.f / d.y,
               fabsf(d.z) < 1e-30f ? copysignf(1e30f, d.z) : 1.f / d.z);
    t.noi = v3(-(o.x * t.inv.x), -(o.y * t.inv.y), -(o.z * t.inv.z));
    t.best.t = INFINITY;
    t.best.u = t.best.v = 0.f;
    t.best.prim = -1;
    t.sp = 0;
    // A ray with a NaN component can never pass tri_test (every comparison there is false), but the
    // slab test would ACCEPT every box for it (fminf/fmaxf drop NaNs) and walk the whole tree:
    // the literal path tracer produces such rays (NaN sources Q16/Q17).  Miss at once.
    const float chk = (o.x + o.y + o.z) + (d.x + d.y + d.z);
    t.cur = (s.n_tris == 0 || chk != chk) ? ORB_TRAVERSAL_DONE : s.root_ref;
}

// After both child boxes of a node were tested: near child next, far child pushed, or pop.
template <class Stack>
ORB_HD void trav_descend(TravState& t, Stack& st, bool h0, bool h1, float tn0, float tn1, int c0, int c1) {
    if (h0 && h1) {
        const bool swap = tn1 < tn0;
        st.put(t.sp, swap ? c0 : c1, swap ? tn0 : tn1);
        ++t.sp;
        t.cur = swap ? c1 : c0;
    } else if (h0) {
        t.cur = c0;
    } else if (h1) {
        t.cur = c1;
    } else {
        t.cur = stack_pop(st, t.sp, t.best);
    }
}

// One inner node (t.cur >= 0): both child boxes of a 64-byte node.
template <class Stack>
ORB_HD void trav_node_step(TravState& t, const SceneView& s, Stack& st) {
    ORB_DEVICE_ASSERT(t.cur >= 0 && t.cur < s.n_nodes);
    const BvhNode* np = s.nodes + t.cur;
#ifdef __CUDA_ARCH__
    // two 256-bit loads (sm_100: LDG.E.256) instead of four 128-bit ones: every lane reads its own node and
    // every load instruction of a lane is its own L1 wavefront
    float4 na, nb, nc;
    int c0, c1, pad0, pad1;
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(na.x), "=f"(na.y), "=f"(na.z), "=f"(na.w), "=f"(nb.x), "=f"(nb.y), "=f"(nb.z), "=f"(nb.w)
        : "l"(np));
    asm("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(nc.x), "=f"(nc.y), "=f"(nc.z), "=f"(nc.w), "=r"(c0), "=r"(c1), "=r"(pad0), "=r"(pad1)
        : "l"(reinterpret_cast<const char*>(np) + 32));
    (void)pad0;
    (void)pad1;
#else
    const float4 na = np->a, nb = np->b, nc = np->c;
    const int c0 = np->c0, c1 = np->c1;
#endif
    const float lim = t.best.prim < 0 ? ORB_TMAX : t.best.t;
    float tn0, tn1;
    const bool h0 = box_test(na.x, na.y, na.z, na.w, nb.x, nb.y, t.noi, t.inv, lim, tn0);
    const bool h1 = box_test(nb.z, nb.w, nc.x, nc.y, nc.z, nc.w, t.noi, t.inv, lim, tn1);
    trav_descend(t, st, h0, h1, tn0, tn1, c0, c1);
}

// The 1..8 triangles of one leaf reference against the ray; updates t.best only.  A triangle record is one
// 32-byte word of `tri` (v0.xyz, prim id | e1.xyz, e2.x) and one 8-byte word of `tri_b` (e2.y, e2.z): two loads
// per lane instead of the three 16-byte ones of a 48-byte record, 40 bytes of cache footprint instead of 48.
ORB_HD void trav_leaf_test(TravState& t, const SceneView& s, int leaf, unsigned* n_tris = nullptr) {
    const int lr = ~leaf;
    const int first = lr >> 3, count = (lr & 7) + 1;
    for (int i = 0; i < count; i++) {
        float4 t0, t1, t2;
        tri_fetch(s, first + i, t0, t1, t2);
        if (n_tris) ++*n_tris;
        float tt, u, v;
        if (tri_test(t0, t1, t2, t.o, t.d, 0.f, ORB_TMAX, tt, u, v)) hit_consider(t.best, tt, u, v, __float_as_int_hd(t0.w));
    }
}

// One leaf (t.cur < 0, not DONE): its triangles, then pop.
template <class Stack>
ORB_HD void trav_leaf_step(TravState& t, const SceneView& s, Stack& st, unsigned* n_tris = nullptr) {
    trav_leaf_test(t, s, t.cur, n_tris);
    t.cur = stack_pop(st, t.sp, t.best);
}

template <class Stack>
ORB_HD Hit trace_closest(const SceneView& s, float3 o, float3 d, Stack& st, unsigned* n_nodes = nullptr,
                         unsigned* n_tris = nullptr) {
    TravState t;
    trav_begin(t, s, o, d);
    while (t.cur != ORB_TRAVERSAL_DONE) {
        while (t.cur >= 0) {
            if (n_nodes) ++*n_nodes;
            trav_node_step(t, s, st);
        }
        if (t.cur == ORB_TRAVERSAL_DONE) break;
        trav_leaf_step(t, s, st, n_tris);
    }
    return t.best;
}
ORB_HD Hit trace_closest(const SceneView& s, float3 o, float3 d, int* stk, float* stk_t, int stride,
                         unsigned* n_nodes = nullptr, unsigned* n_tris = nullptr) {
    StridedStack st{stk, stk_t, stride};
    return trace_closest(s, o, d, st, n_nodes, n_tris);
}

// Brute force over every triangle record: the GPU-side reference for hit-id parity tests.
ORB_HD Hit trace_brute(const SceneView& s, float3 o, float3 d) {
    Hit best;
    best.t = INFINITY;
    best.u = best.v = 0.f;
    best.prim = -1;
    for (int i = 0; i < s.n_tris; i++) {
        float4 t0, t1, t2;
        tri_fetch(s, i, t0, t1, t2);
        float t, u, v;
        if (tri_test(t0, t1, t2, o, d, 0.f, ORB_TMAX, t, u, v)) hit_consider(best, t, u, v, __float_as_int_hd(t0.w));
    }
    return best;
}

// ---- closest-hit / miss record ---------------------------------------------------------------
// SurfaceInteraction (cuda_include/common.cuh:163-183) reduced to the fields the path reads.
struct Surface {
    bool hit, is_light;
    uint32_t material_id;
    float3 p, n, diffuse, emit;
    float u, v;  // si.uv
    float alpha, area;
};

ORB_HD Surface surface_miss() {  // __miss__miss + SurfaceInteraction defaults, const_color = 0 (viewer.hpp:530)
    Surface s;
    s.hit = false;
    s.is_light = false;
    s.material_id = 0;
    s.p = s.n = s.diffuse = s.emit = v3(0.f);
    s.u = s.v = 0.f;
    s.alpha = 0.f;
    s.area = 0.f;
    return s;
}

// (1-u-v)*a + u*b + v*c, each step rounded (barycentricInterpolate, utils.cuh:10-29)
ORB_HD float bary1(float w, float u, float v, float a, float b, float c) {
    return xadd(xadd(xmul(w, a), xmul(u, b)), xmul(v, c));
}

ORB_HD Surface shade_hit(const SceneView& s, int prim, float u, float v) {
    if (prim < 0) return surface_miss();
    Surface r;
    const MeshRec& m = s.meshes[s.prim_mesh[prim]];
    r.hit = true;
    r.material_id = m.material_id;
    const float* P = s.pos + 9 * (size_t)prim;
    const float* N = s.nrm + 9 * (size_t)prim;
    const float* T = s.uv + 6 * (size_t)prim;
    const float3 v1 = v3(P[0], P[1], P[2]), v2 = v3(P[3], P[4], P[5]), v3_ = v3(P[6], P[7], P[8]);
    r.area = 0.5f * length(cross(v1 - v2, v3_ - v2));
    const float w = xsub(xsub(1.f, u), v);
    r.p = v3(bary1(w, u, v, v1.x, v2.x, v3_.x), bary1(w, u, v, v1.y, v2.y, v3_.y), bary1(w, u, v, v1.z, v2.z, v3_.z));
    r.u = fabsf(fmodf(bary1(w, u, v, T[0], T[2], T[4]), 1.f));
    r.v = fabsf(fmodf(bary1(w, u, v, T[1], T[3], T[5]), 1.f));
    r.n = xnormalize(v3(bary1(w, u, v, N[0], N[3], N[6]), bary1(w, u, v, N[1], N[4], N[7]), bary1(w, u, v, N[2], N[5], N[8])));
    r.diffuse = m.diffuse;
    if (m.tex) {  // point sampling, normalised coordinates, clamp; texel read as unorm8 -> float
        int i = (int)floorf(r.u * (float)m.tex_w), j = (int)floorf(r.v * (float)m.tex_h);
        i = imin(imax(i, 0), m.tex_w - 1);
        j = imin(imax(j, 0), m.tex_h - 1);
        const uchar4 c = m.tex[(size_t)j * m.tex_w + i];
        r.diffuse = v3((float)c.x * (1.f / 255.f), (float)c.y * (1.f / 255.f), (float)c.z * (1.f / 255.f));
    }
    r.alpha = owl_max(0.01f, owl_min(m.alpha, 1.f));  // owl::clamp, hit_miss.cuh:45-46
    r.emit = m.emit;
    r.is_light = m.is_light != 0;
    return r;
}

}  // namespace orb

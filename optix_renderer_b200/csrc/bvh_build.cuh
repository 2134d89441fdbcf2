// bvh_build.cuh -- GPU LBVH builder (device-only; included by orb200.cu).
//
// Replaces owlTrianglesGeomGroupCreate + owlGroupBuildAccel per mesh and the instance group over
// them (include/viewer.hpp:506-507,516-517 -> optixAccelBuild inside un-vendored OWL/OptiX):
// all meshes are flattened into one triangle soup addressed by global primitive id and ONE
// hierarchy is built over it.
//
// Pipeline (all on the context's stream, no host round trip until the final height read-back):
//   1. k_prim_bounds     per-triangle box + scene box (warp-reduced float atomics)
//   2. k_morton          63-bit Morton code of the box centre (21 bits per axis)
//   3. cub radix sort    (code, primitive) pairs -- stable, so equal codes keep primitive order
//                        and every GPU builds the identical tree (tile partition)
//   4. k_hierarchy       Karras 2012: one thread per internal node finds its range and split
//   5. k_refit           bottom-up boxes + subtree height, second arrival at a node proceeds
//   6. k_emit            64-byte traversal nodes holding both child boxes; subtrees of
//                        <= max_leaf triangles become leaves (contiguous runs of `tri`)
//   7. k_tri_records     (v0, e1, e2, global id) in leaf order
#pragma once
#include <cub/device/device_radix_sort.cuh>

#include "scene.cuh"

namespace orb {

__device__ __forceinline__ unsigned f2ord(float f) {
    const unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float ord2f(unsigned u) {
    const unsigned b = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#ifdef __CUDA_ARCH__
    return __uint_as_float(b);
#else
    float f;
    memcpy(&f, &b, 4);
    return f;
#endif
}

// scene_box: 6 ordered-uint words: lo.xyz (min), hi.xyz (max)
__global__ void k_prim_bounds(const float* __restrict__ pos, int n, float4* __restrict__ plo, float4* __restrict__ phi,
                              unsigned* __restrict__ scene_box) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    if (i < n) {
        const float* p = pos + 9 * (size_t)i;
#pragma unroll
        for (int k = 0; k < 3; k++)
#pragma unroll
            for (int a = 0; a < 3; a++) {
                lo[a] = fminf(lo[a], p[3 * k + a]);
                hi[a] = fmaxf(hi[a], p[3 * k + a]);
            }
        plo[i] = make_float4(lo[0], lo[1], lo[2], 0.f);
        phi[i] = make_float4(hi[0], hi[1], hi[2], 0.f);
    }
#pragma unroll
    for (int a = 0; a < 3; a++)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
        }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; a++) {
            if (lo[a] <= hi[a]) {
                atomicMin(scene_box + a, f2ord(lo[a]));
                atomicMax(scene_box + 3 + a, f2ord(hi[a]));
            }
        }
    }
}

__device__ __forceinline__ unsigned long long spread21(unsigned v) {  // 21 bits -> every third bit
    unsigned long long x = v & 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

__global__ void k_morton(const float4* __restrict__ plo, const float4* __restrict__ phi, int n,
                         const unsigned* __restrict__ scene_box, unsigned long long* __restrict__ keys,
                         unsigned* __restrict__ vals) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 lo = plo[i], hi = phi[i];
    const float c[3] = {0.5f * (lo.x + hi.x), 0.5f * (lo.y + hi.y), 0.5f * (lo.z + hi.z)};
    unsigned q[3];
#pragma unroll
    for (int a = 0; a < 3; a++) {
        const float slo = ord2f(scene_box[a]), shi = ord2f(scene_box[3 + a]);
        const float ext = shi - slo;
        float t = ext > 0.f ? (c[a] - slo) / ext : 0.f;
        t = fminf(fmaxf(t, 0.f), 1.f);
        q[a] = min((unsigned)(t * 2097152.f), 2097151u);
    }
    keys[i] = (spread21(q[0]) << 2) | (spread21(q[1]) << 1) | spread21(q[2]);
    vals[i] = (unsigned)i;
}

// common-prefix length of sorted keys i and j; ties broken by position (Karras 2012, sec. 4)
__device__ __forceinline__ int lbvh_delta(const unsigned long long* __restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const unsigned long long a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz((unsigned)i ^ (unsigned)j);
    return __clzll((long long)(a ^ b));
}

// child encoding inside the builder: >= 0 internal node, < 0 sorted leaf position ~k
__global__ void k_hierarchy(const unsigned long long* __restrict__ keys, int n, int2* __restrict__ child,
                            int2* __restrict__ range, int* __restrict__ node_parent, int* __restrict__ leaf_parent) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = (lbvh_delta(keys, n, i, i + 1) - lbvh_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = lbvh_delta(keys, n, i, i - d);
    int lmax = 2;
    while (lbvh_delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (lbvh_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = lbvh_delta(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
        if (lbvh_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
        if (t <= 1) break;
    }
    const int gamma = i + s * d + min(d, 0);
    const int first = min(i, j), last = max(i, j);
    int left, right;
    if (first == gamma) {
        left = ~gamma;
        leaf_parent[gamma] = i;
    } else {
        left = gamma;
        node_parent[gamma] = i;
    }
    if (last == gamma + 1) {
        right = ~(gamma + 1);
        leaf_parent[gamma + 1] = i;
    } else {
        right = gamma + 1;
        node_parent[gamma + 1] = i;
    }
    child[i] = make_int2(left, right);
    range[i] = make_int2(first, last);
    if (i == 0) node_parent[0] = -1;
}

__device__ __forceinline__ void box_of(int ref, const float4* nlo, const float4* nhi, const float4* plo,
                                       const float4* phi, const unsigned* sorted, float4& lo, float4& hi) {
    if (ref >= 0) {
        lo = __ldcg(nlo + ref);  // written by another thread of this launch: bypass L1
        hi = __ldcg(nhi + ref);
    } else {
        const unsigned g = sorted[~ref];
        lo = plo[g];
        hi = phi[g];
    }
}

__global__ void k_refit(int n, const int2* __restrict__ child, const int* __restrict__ node_parent,
                        const int* __restrict__ leaf_parent, const float4* __restrict__ plo,
                        const float4* __restrict__ phi, const unsigned* __restrict__ sorted, float4* nlo, float4* nhi,
                        int* height, unsigned* flag) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    int node = leaf_parent[k];
    while (node >= 0) {
        __threadfence();
        if (atomicAdd(flag + node, 1u) == 0u) return;  // first arrival: the sibling finishes this node
        __threadfence();
        const int2 c = child[node];
        float4 l0, h0, l1, h1;
        box_of(c.x, nlo, nhi, plo, phi, sorted, l0, h0);
        box_of(c.y, nlo, nhi, plo, phi, sorted, l1, h1);
        const int hl = c.x >= 0 ? __ldcg(height + c.x) : 0, hr = c.y >= 0 ? __ldcg(height + c.y) : 0;
        nlo[node] = make_float4(fminf(l0.x, l1.x), fminf(l0.y, l1.y), fminf(l0.z, l1.z), 0.f);
        nhi[node] = make_float4(fmaxf(h0.x, h1.x), fmaxf(h0.y, h1.y), fmaxf(h0.z, h1.z), 0.f);
        height[node] = 1 + max(hl, hr);
        node = node_parent[node];
    }
}

__global__ void k_emit(int n, int max_leaf, float pad, const int2* __restrict__ child, const int2* __restrict__ range,
                       const float4* __restrict__ plo, const float4* __restrict__ phi,
                       const unsigned* __restrict__ sorted, const float4* __restrict__ nlo,
                       const float4* __restrict__ nhi, BvhNode* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int2 r = range[i];
    BvhNode nd;
    if (r.y - r.x + 1 <= max_leaf) {  // collapsed into a leaf of its parent: never referenced
        nd.a = nd.b = nd.c = make_float4(0.f, 0.f, 0.f, 0.f);
        nd.c0 = nd.c1 = nd.pad0 = nd.pad1 = 0;
        out[i] = nd;
        return;
    }
    const int2 c = child[i];
    int ref[2];
    float4 lo[2], hi[2];
    const int cc[2] = {c.x, c.y};
#pragma unroll
    for (int k = 0; k < 2; k++) {
        box_of(cc[k], nlo, nhi, plo, phi, sorted, lo[k], hi[k]);
        if (cc[k] < 0) {
            ref[k] = leaf_ref(~cc[k], 1);
        } else {
            const int2 cr = range[cc[k]];
            const int cnt = cr.y - cr.x + 1;
            ref[k] = cnt <= max_leaf ? leaf_ref(cr.x, cnt) : cc[k];
        }
    }
    nd.a = make_float4(lo[0].x - pad, lo[0].y - pad, lo[0].z - pad, hi[0].x + pad);
    nd.b = make_float4(hi[0].y + pad, hi[0].z + pad, lo[1].x - pad, lo[1].y - pad);
    nd.c = make_float4(lo[1].z - pad, hi[1].x + pad, hi[1].y + pad, hi[1].z + pad);
    nd.c0 = ref[0];
    nd.c1 = ref[1];
    nd.pad0 = nd.pad1 = 0;
    out[i] = nd;
}

__global__ void k_tri_records(const float* __restrict__ pos, const unsigned* __restrict__ sorted, int n,
                              float4* __restrict__ tri, float2* __restrict__ tri_b) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const unsigned g = sorted[k];
    float a[8], b[2];
    tri_record(a, b, pos + 9 * (size_t)g, (int)g);
    tri[2 * (size_t)k + 0] = make_float4(a[0], a[1], a[2], a[3]);
    tri[2 * (size_t)k + 1] = make_float4(a[4], a[5], a[6], a[7]);
    tri_b[k] = make_float2(b[0], b[1]);
}

struct BvhBuildResult {
    float4* tri = nullptr;    // 32 B per triangle (scene.cuh tri_record)
    float2* tri_b = nullptr;  // 8 B per triangle
    BvhNode* nodes = nullptr;
    int root_ref = 0;
    int height = 0;
    int n_nodes = 0;
    int launches = 0;
};

#define ORB_CU(x)                      \
    do {                               \
        cudaError_t e_ = (x);          \
        if (e_ != cudaSuccess) return e_; \
    } while (0)

// Scratch arrays of a build / refit: freed when the function returns, on the error paths of ORB_CU too.
struct DevFreeOnExit {
    void** v[40];
    int n = 0;
    void add(void* pp) { if (n < 40) v[n++] = static_cast<void**>(pp); }
    ~DevFreeOnExit() {
        for (int i = 0; i < n; i++) {
            cudaFree(*v[i]);
            *v[i] = nullptr;
        }
    }
};

// d_pos: 9 floats per global primitive, already on the device.  Synchronises the stream once.
inline cudaError_t build_bvh(const float* d_pos, int n, int max_leaf, cudaStream_t st, BvhBuildResult& out) {
    out = BvhBuildResult();
    if (n <= 0) return cudaSuccess;
    max_leaf = max_leaf <= 0 ? 4 : (max_leaf > 8 ? 8 : max_leaf);
    const int B = 256, G = (n + B - 1) / B;
    float4 *plo = nullptr, *phi = nullptr, *nlo = nullptr, *nhi = nullptr;
    unsigned *scene_box = nullptr, *vals = nullptr, *vals_s = nullptr, *flag = nullptr;
    unsigned long long *keys = nullptr, *keys_s = nullptr;
    int2 *child = nullptr, *range = nullptr;
    int *node_parent = nullptr, *leaf_parent = nullptr, *height = nullptr;
    void* cub_tmp = nullptr;
    DevFreeOnExit scratch;
    for (void* pp : {(void*)&plo, (void*)&phi, (void*)&nlo, (void*)&nhi, (void*)&scene_box, (void*)&vals, (void*)&vals_s,
                     (void*)&flag, (void*)&keys, (void*)&keys_s, (void*)&child, (void*)&range, (void*)&node_parent,
                     (void*)&leaf_parent, (void*)&height, (void*)&cub_tmp})
        scratch.add(pp);
    const size_t ni = (size_t)(n > 1 ? n - 1 : 1);
    ORB_CU(cudaMalloc(&plo, sizeof(float4) * n));
    ORB_CU(cudaMalloc(&phi, sizeof(float4) * n));
    ORB_CU(cudaMalloc(&scene_box, 6 * sizeof(unsigned)));
    ORB_CU(cudaMalloc(&keys, 8 * (size_t)n));
    ORB_CU(cudaMalloc(&keys_s, 8 * (size_t)n));
    ORB_CU(cudaMalloc(&vals, 4 * (size_t)n));
    ORB_CU(cudaMalloc(&vals_s, 4 * (size_t)n));
    ORB_CU(cudaMalloc(&out.tri, sizeof(float4) * 2 * (size_t)n));
    ORB_CU(cudaMalloc(&out.tri_b, sizeof(float2) * (size_t)n));
    const unsigned box_init[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
    ORB_CU(cudaMemcpyAsync(scene_box, box_init, sizeof(box_init), cudaMemcpyHostToDevice, st));
    k_prim_bounds<<<G, B, 0, st>>>(d_pos, n, plo, phi, scene_box);
    k_morton<<<G, B, 0, st>>>(plo, phi, n, scene_box, keys, vals);
    size_t tmp_bytes = 0;
    ORB_CU(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys, keys_s, vals, vals_s, n, 0, 63, st));
    ORB_CU(cudaMalloc(&cub_tmp, tmp_bytes ? tmp_bytes : 16));
    ORB_CU(cub::DeviceRadixSort::SortPairs(cub_tmp, tmp_bytes, keys, keys_s, vals, vals_s, n, 0, 63, st));
    k_tri_records<<<G, B, 0, st>>>(d_pos, vals_s, n, out.tri, out.tri_b);
    out.launches = 4;
    unsigned box_h[6];
    ORB_CU(cudaMemcpyAsync(box_h, scene_box, sizeof(box_h), cudaMemcpyDeviceToHost, st));
    if (n <= max_leaf) {
        out.root_ref = leaf_ref(0, n);
        out.height = 0;
        ORB_CU(cudaStreamSynchronize(st));
    } else {
        ORB_CU(cudaMalloc(&child, sizeof(int2) * ni));
        ORB_CU(cudaMalloc(&range, sizeof(int2) * ni));
        ORB_CU(cudaMalloc(&node_parent, sizeof(int) * ni));
        ORB_CU(cudaMalloc(&leaf_parent, sizeof(int) * n));
        ORB_CU(cudaMalloc(&height, sizeof(int) * ni));
        ORB_CU(cudaMalloc(&flag, sizeof(unsigned) * ni));
        ORB_CU(cudaMalloc(&nlo, sizeof(float4) * ni));
        ORB_CU(cudaMalloc(&nhi, sizeof(float4) * ni));
        ORB_CU(cudaMalloc(&out.nodes, sizeof(BvhNode) * ni));
        ORB_CU(cudaMemsetAsync(flag, 0, sizeof(unsigned) * ni, st));
        ORB_CU(cudaMemsetAsync(height, 0, sizeof(int) * ni, st));
        ORB_CU(cudaStreamSynchronize(st));  // scene box on the host (for the padding)
        float ext = 0.f;
        for (int a = 0; a < 3; a++) ext = fmaxf(ext, ord2f(box_h[3 + a]) - ord2f(box_h[a]));
        const float pad = ext * 1e-5f;
        const int Gi = (n - 1 + B - 1) / B;
        k_hierarchy<<<Gi, B, 0, st>>>(keys_s, n, child, range, node_parent, leaf_parent);
        k_refit<<<G, B, 0, st>>>(n, child, node_parent, leaf_parent, plo, phi, vals_s, nlo, nhi, height, flag);
        k_emit<<<Gi, B, 0, st>>>(n, max_leaf, pad, child, range, plo, phi, vals_s, nlo, nhi, out.nodes);
        out.launches += 3;
        ORB_CU(cudaMemcpyAsync(&out.height, height, sizeof(int), cudaMemcpyDeviceToHost, st));
        ORB_CU(cudaStreamSynchronize(st));
        out.root_ref = 0;
        out.n_nodes = n - 1;
    }
    ORB_CU(cudaGetLastError());
    return cudaSuccess;
}

}  // namespace orb

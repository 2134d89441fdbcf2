// bvh_refit.cuh -- refit of a committed BVH after its triangles moved (animated lights / meshes,
// SURVEY.md 8(f) rank 4).  Replaces the reference's only way to move geometry, a full
// owlGroupBuildAccel of the GAS and the IAS (include/viewer.hpp:506-507,516-517).
//
// The hierarchy keeps its topology and its leaf order; what is recomputed on the device:
//   1. k_refit_tris     the 48-byte triangle records (v0, e1, e2, global id), in place, from `pos`
//   2. k_refit_scene    the scene box (ordered-uint atomics) -> the builder's padding 1e-5 x extent
//   3. k_refit_parents  parent index of every inner node (the traversal nodes do not store it)
//   4. k_refit_nodes    bottom-up: a thread starts at every node whose two children are leaves,
//                       writes that node's (padded) child boxes and its own unpadded box, and climbs;
//                       at a node with two inner children the second arrival proceeds (atomic
//                       counter, fence, .cg loads), as in Karras 2012.
// Works for either builder's output (binned SAH, Morton LBVH): both emit the same BvhNode array
// in which a slot with c0 == c1 == 0 is dead (a subtree that was collapsed into a leaf; the
// root is never a child, so no live node refers to node 0 twice).
// Results after a refit are exact in the same sense as after a build: boxes are conservative,
// closest t wins, ties to the lowest primitive id -- any valid hierarchy gives the same hits.
#pragma once
#include "bvh_build.cuh"

namespace orb {

__global__ void k_refit_tris(const float* __restrict__ pos, int n, float4* __restrict__ tri, float2* __restrict__ tri_b,
                             unsigned* __restrict__ scene_box) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    if (k < n) {
        const int g = __float_as_int(tri[2 * (size_t)k].w);
        const float* p = pos + 9 * (size_t)g;
        float a[8], b[2];
        tri_record(a, b, p, g);
        tri[2 * (size_t)k + 0] = make_float4(a[0], a[1], a[2], a[3]);
        tri[2 * (size_t)k + 1] = make_float4(a[4], a[5], a[6], a[7]);
        tri_b[k] = make_float2(b[0], b[1]);
#pragma unroll
        for (int a = 0; a < 3; a++) {
            lo[a] = fminf(fminf(p[a], p[3 + a]), p[6 + a]);
            hi[a] = fmaxf(fmaxf(p[a], p[3 + a]), p[6 + a]);
        }
    }
#pragma unroll
    for (int a = 0; a < 3; a++) {
        for (int o = 16; o; o >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
        }
        if ((threadIdx.x & 31) == 0 && lo[a] <= hi[a]) {
            atomicMin(scene_box + a, f2ord(lo[a]));
            atomicMax(scene_box + 3 + a, f2ord(hi[a]));
        }
    }
}

__device__ __forceinline__ bool refit_dead(const BvhNode* nodes, int i, int& c0, int& c1) {
    c0 = nodes[i].c0;
    c1 = nodes[i].c1;
    return c0 == 0 && c1 == 0;
}

__global__ void k_refit_parents(int n_nodes, const BvhNode* __restrict__ nodes, int* __restrict__ parent) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    int c0, c1;
    if (refit_dead(nodes, i, c0, c1)) return;
    if (c0 >= 0) parent[c0] = i;
    if (c1 >= 0) parent[c1] = i;
}

// unpadded box of one child: a leaf from the corner positions of its triangles, an inner node from sub_lo/hi
__device__ __forceinline__ void refit_child_box(int ref, const float4* __restrict__ tri, const float* __restrict__ pos,
                                                const float4* sub_lo, const float4* sub_hi, float3& lo, float3& hi) {
    if (ref >= 0) {
        const float4 l = __ldcg(sub_lo + ref), h = __ldcg(sub_hi + ref);
        lo = v3(l.x, l.y, l.z);
        hi = v3(h.x, h.y, h.z);
        return;
    }
    const int r = ~ref, first = r >> 3, count = (r & 7) + 1;
    lo = v3(INFINITY);
    hi = v3(-INFINITY);
    for (int k = first; k < first + count; k++) {
        const int g = __float_as_int(tri[2 * (size_t)k].w);
        const float* p = pos + 9 * (size_t)g;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            lo.x = fminf(lo.x, p[3 * c]);
            lo.y = fminf(lo.y, p[3 * c + 1]);
            lo.z = fminf(lo.z, p[3 * c + 2]);
            hi.x = fmaxf(hi.x, p[3 * c]);
            hi.y = fmaxf(hi.y, p[3 * c + 1]);
            hi.z = fmaxf(hi.z, p[3 * c + 2]);
        }
    }
}

__global__ void k_refit_nodes(int n_nodes, BvhNode* __restrict__ nodes, const int* __restrict__ parent,
                              unsigned* __restrict__ arrivals, float4* sub_lo, float4* sub_hi,
                              const float4* __restrict__ tri, const float* __restrict__ pos,
                              const unsigned* __restrict__ scene_box) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    int c0, c1;
    if (refit_dead(nodes, i, c0, c1)) return;
    if (c0 >= 0 || c1 >= 0) return;  // a thread starts where both children are leaves
    float ext = 0.f;
#pragma unroll
    for (int a = 0; a < 3; a++) ext = fmaxf(ext, ord2f(scene_box[3 + a]) - ord2f(scene_box[a]));
    const float pad = ext * 1e-5f;  // the builders' padding (bvh_sah.cuh, bvh_build.cuh)
    int node = i;
    for (;;) {
        float3 lo0, hi0, lo1, hi1;
        refit_child_box(c0, tri, pos, sub_lo, sub_hi, lo0, hi0);
        refit_child_box(c1, tri, pos, sub_lo, sub_hi, lo1, hi1);
        nodes[node].a = make_float4(lo0.x - pad, lo0.y - pad, lo0.z - pad, hi0.x + pad);
        nodes[node].b = make_float4(hi0.y + pad, hi0.z + pad, lo1.x - pad, lo1.y - pad);
        nodes[node].c = make_float4(lo1.z - pad, hi1.x + pad, hi1.y + pad, hi1.z + pad);
        sub_lo[node] = make_float4(fminf(lo0.x, lo1.x), fminf(lo0.y, lo1.y), fminf(lo0.z, lo1.z), 0.f);
        sub_hi[node] = make_float4(fmaxf(hi0.x, hi1.x), fmaxf(hi0.y, hi1.y), fmaxf(hi0.z, hi1.z), 0.f);
        const int p = parent[node];
        if (p < 0) return;  // the root
        __threadfence();
        c0 = nodes[p].c0;
        c1 = nodes[p].c1;
        const unsigned needed = (c0 >= 0) + (c1 >= 0);
        if (atomicAdd(arrivals + p, 1u) + 1u < needed) return;  // the sibling subtree is not finished: its thread goes on
        __threadfence();
        node = p;
    }
}

// d_pos: 9 floats per global primitive (already updated).  All work is enqueued on `st`; the scratch
// arrays are freed after a stream synchronise.  `launches` is incremented by the kernels launched.
inline cudaError_t refit_bvh(const float* d_pos, int n, cudaStream_t st, BvhBuildResult& bvh) {
    if (n <= 0 || !bvh.tri) return cudaSuccess;
    const int B = 256, G = (n + B - 1) / B;
    unsigned* scene_box = nullptr;
    int* parent = nullptr;
    unsigned* arrivals = nullptr;
    float4 *sub_lo = nullptr, *sub_hi = nullptr;
    DevFreeOnExit scratch;
    for (void* pp : {(void*)&scene_box, (void*)&parent, (void*)&arrivals, (void*)&sub_lo, (void*)&sub_hi}) scratch.add(pp);
    ORB_CU(cudaMalloc(&scene_box, 6 * sizeof(unsigned)));
    ORB_CU(cudaMemsetAsync(scene_box, 0xff, 3 * sizeof(unsigned), st));  // mins start at the largest key
    ORB_CU(cudaMemsetAsync(scene_box + 3, 0, 3 * sizeof(unsigned), st));
    k_refit_tris<<<G, B, 0, st>>>(d_pos, n, bvh.tri, bvh.tri_b, scene_box);
    bvh.launches++;
    const int nn = bvh.n_nodes;
    if (nn > 0 && bvh.root_ref >= 0) {
        const int Gn = (nn + B - 1) / B;
        ORB_CU(cudaMalloc(&parent, sizeof(int) * (size_t)nn));
        ORB_CU(cudaMalloc(&arrivals, sizeof(unsigned) * (size_t)nn));
        ORB_CU(cudaMalloc(&sub_lo, sizeof(float4) * (size_t)nn));
        ORB_CU(cudaMalloc(&sub_hi, sizeof(float4) * (size_t)nn));
        ORB_CU(cudaMemsetAsync(parent, 0xff, sizeof(int) * (size_t)nn, st));
        ORB_CU(cudaMemsetAsync(arrivals, 0, sizeof(unsigned) * (size_t)nn, st));
        k_refit_parents<<<Gn, B, 0, st>>>(nn, bvh.nodes, parent);
        k_refit_nodes<<<Gn, B, 0, st>>>(nn, bvh.nodes, parent, arrivals, sub_lo, sub_hi, bvh.tri, d_pos, scene_box);
        bvh.launches += 2;
    }
    ORB_CU(cudaStreamSynchronize(st));
    ORB_CU(cudaGetLastError());
    return cudaSuccess;
}

}  // namespace orb

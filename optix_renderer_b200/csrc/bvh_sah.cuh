// bvh_sah.cuh -- GPU binned-SAH builder (device-only; included by orb200.cu).
//
// Same role and same output format as bvh_build.cuh (the LBVH builder): it replaces the GAS/IAS
// builds of include/viewer.hpp:506-507,516-517.  On the bench scene a binned surface-area-
// heuristic tree needs ~20 % fewer node visits per ray than the Morton-code tree (shadow rays
// 48 -> 38, camera rays 28 -> 22, bounce rays 40 -> 31 inner nodes per ray),
// and traversal is where the frame's time goes, so this is the default builder.
//
// Breadth-first, one level of the tree per round, everything on the device:
//   state     slot i of the working order holds primitive prim_id[i] and the node prim_node[i]
//             it currently belongs to; slots of one node are contiguous (sorted by node index)
//   k_sah_bin     every primitive of a splittable node adds its box to 3 x 16 centroid bins of
//                 that node (block-local histogram in shared memory for the block's first node,
//                 global atomics for the rest)
//   k_sah_split   one thread per node: sweeps the bins of the three axes, picks the split with
//                 the lowest  SA(L) n(L) + SA(R) n(R), allocates the two children and gives them
//                 the exact union boxes of their bins; a node whose centroids all fall into one
//                 bin is halved by position instead
//   k_sah_assign  every primitive moves to the child its bin belongs to
//   cub sort      stable radix sort of (node, primitive) pairs -> slots contiguous per node again
//   k_sah_first   first slot of every node
// until a level creates no children.  Leaves are nodes of <= max_leaf primitives.  The order in
// which children are numbered depends on atomic arrival and may differ from run to run; the tree's
// SHAPE and every query result do not (closest t, ties to the lowest primitive id).
#pragma once
#include "bvh_build.cuh"

namespace orb {

#define SAH_BINS 16

struct SahBin {  // boxes in the ordered-uint encoding of f2ord()
    unsigned cnt;
    unsigned lo[3], hi[3];
    unsigned pad;
};
static_assert(sizeof(SahBin) == 32, "SahBin layout");

struct SahNodes {
    float4* lo;   // box min (w unused)
    float4* hi;   // box max
    int* count;   // primitives below
    int* first;   // first slot (valid after k_sah_first of the current round)
    int* left;    // left child (right = left + 1), -1 = leaf, -3 = not decided yet
    int* split;   // axis * SAH_BINS + bin (primitives with bin <= this go left), -2 = halve by position
    int* slot;    // index of the node's bin block in this round, -1 = none (leaf)
};

__device__ __forceinline__ int sah_bin_of(float c, float lo, float hi) {
    const float ext = hi - lo;
    if (!(ext > 0.f)) return 0;
    const int b = (int)((c - lo) / ext * (float)SAH_BINS);
    return min(max(b, 0), SAH_BINS - 1);
}

__global__ void k_sah_clear(SahBin* __restrict__ bins, int n_bins) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_bins) return;
    SahBin b;
    b.cnt = 0;
    b.lo[0] = b.lo[1] = b.lo[2] = 0xffffffffu;
    b.hi[0] = b.hi[1] = b.hi[2] = 0u;
    b.pad = 0;
    bins[i] = b;
}

__device__ __forceinline__ void sah_bin_add(SahBin* b, const float4& lo, const float4& hi) {
    atomicAdd(&b->cnt, 1u);
    atomicMin(&b->lo[0], f2ord(lo.x));
    atomicMin(&b->lo[1], f2ord(lo.y));
    atomicMin(&b->lo[2], f2ord(lo.z));
    atomicMax(&b->hi[0], f2ord(hi.x));
    atomicMax(&b->hi[1], f2ord(hi.y));
    atomicMax(&b->hi[2], f2ord(hi.z));
}

#define SAH_BLOCK 256
__global__ void __launch_bounds__(SAH_BLOCK)
k_sah_bin(int n, int level_begin, const unsigned* __restrict__ prim_node, const unsigned* __restrict__ prim_id,
          const float4* __restrict__ plo, const float4* __restrict__ phi, SahNodes nd, SahBin* __restrict__ bins) {
    __shared__ SahBin sh[3 * SAH_BINS];
    __shared__ int sh_used;
    const int i = blockIdx.x * SAH_BLOCK + threadIdx.x;
    const int node0 = (int)prim_node[min(blockIdx.x * SAH_BLOCK, n - 1)];
    if (threadIdx.x < 3 * SAH_BINS) {
        SahBin b;
        b.cnt = 0;
        b.lo[0] = b.lo[1] = b.lo[2] = 0xffffffffu;
        b.hi[0] = b.hi[1] = b.hi[2] = 0u;
        b.pad = 0;
        sh[threadIdx.x] = b;
    }
    if (threadIdx.x == 0) sh_used = 0;
    __syncthreads();
    if (i < n) {
        const int node = (int)prim_node[i];
        const int slot = node >= level_begin ? nd.slot[node] : -1;
        if (slot >= 0) {
            const unsigned g = prim_id[i];
            const float4 lo = plo[g], hi = phi[g];
            const float4 nlo = nd.lo[node], nhi = nd.hi[node];
            const int bx = sah_bin_of(0.5f * (lo.x + hi.x), nlo.x, nhi.x);
            const int by = sah_bin_of(0.5f * (lo.y + hi.y), nlo.y, nhi.y);
            const int bz = sah_bin_of(0.5f * (lo.z + hi.z), nlo.z, nhi.z);
            SahBin* base = node == node0 ? sh : bins + (size_t)slot * 3 * SAH_BINS;
            if (node == node0) sh_used = 1;
            sah_bin_add(base + bx, lo, hi);
            sah_bin_add(base + SAH_BINS + by, lo, hi);
            sah_bin_add(base + 2 * SAH_BINS + bz, lo, hi);
        }
    }
    __syncthreads();
    if (sh_used && threadIdx.x < 3 * SAH_BINS && sh[threadIdx.x].cnt) {
        SahBin* dst = bins + (size_t)nd.slot[node0] * 3 * SAH_BINS + threadIdx.x;
        const SahBin& s = sh[threadIdx.x];
        atomicAdd(&dst->cnt, s.cnt);
#pragma unroll
        for (int a = 0; a < 3; a++) {
            atomicMin(&dst->lo[a], s.lo[a]);
            atomicMax(&dst->hi[a], s.hi[a]);
        }
    }
}

struct SahBox {
    float lo[3], hi[3];
};
__device__ __forceinline__ void sah_box_empty(SahBox& b) {
    b.lo[0] = b.lo[1] = b.lo[2] = INFINITY;
    b.hi[0] = b.hi[1] = b.hi[2] = -INFINITY;
}
__device__ __forceinline__ void sah_box_add(SahBox& b, const SahBin& s) {
#pragma unroll
    for (int a = 0; a < 3; a++) {
        b.lo[a] = fminf(b.lo[a], ord2f(s.lo[a]));
        b.hi[a] = fmaxf(b.hi[a], ord2f(s.hi[a]));
    }
}
__device__ __forceinline__ float sah_box_area(const SahBox& b) {
    const float x = b.hi[0] - b.lo[0], y = b.hi[1] - b.lo[1], z = b.hi[2] - b.lo[2];
    return x * y + y * z + z * x;
}

// counters[0] = nodes allocated so far, counters[1] = bin blocks handed out for the NEXT round
__global__ void k_sah_split(int level_begin, int level_end, int max_leaf, SahNodes nd, const SahBin* __restrict__ bins,
                            int* __restrict__ counters) {
    const int k = level_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= level_end) return;
    const int count = nd.count[k];
    const int slot = nd.slot[k];
    if (slot < 0) {  // leaf
        nd.left[k] = -1;
        return;
    }
    const SahBin* B = bins + (size_t)slot * 3 * SAH_BINS;
    float best = INFINITY;
    int best_axis = -1, best_bin = -1;
    for (int axis = 0; axis < 3; axis++) {
        const SahBin* A = B + axis * SAH_BINS;
        float r_area[SAH_BINS];
        unsigned r_cnt[SAH_BINS];
        SahBox acc;
        sah_box_empty(acc);
        unsigned c = 0;
        for (int i = SAH_BINS - 1; i > 0; i--) {
            const SahBin s = A[i];
            if (s.cnt) sah_box_add(acc, s);
            c += s.cnt;
            r_area[i] = c ? sah_box_area(acc) : 0.f;
            r_cnt[i] = c;
        }
        sah_box_empty(acc);
        c = 0;
        for (int i = 0; i < SAH_BINS - 1; i++) {
            const SahBin s = A[i];
            if (s.cnt) sah_box_add(acc, s);
            c += s.cnt;
            if (c == 0 || r_cnt[i + 1] == 0) continue;
            const float cost = sah_box_area(acc) * (float)c + r_area[i + 1] * (float)r_cnt[i + 1];
            if (cost < best) {
                best = cost;
                best_axis = axis;
                best_bin = i;
            }
        }
    }
    const int left = atomicAdd(&counters[0], 2);
    int cl, cr;
    SahBox bl, br;
    if (best_axis >= 0) {
        const SahBin* A = B + best_axis * SAH_BINS;
        sah_box_empty(bl);
        sah_box_empty(br);
        unsigned nl = 0, nr = 0;
        for (int i = 0; i < SAH_BINS; i++) {
            const SahBin s = A[i];
            if (!s.cnt) continue;
            if (i <= best_bin) { sah_box_add(bl, s); nl += s.cnt; }
            else { sah_box_add(br, s); nr += s.cnt; }
        }
        cl = (int)nl;
        cr = (int)nr;
        nd.split[k] = best_axis * SAH_BINS + best_bin;
    } else {  // every centroid in one bin on every axis: halve by position, children keep the parent's box
        const float4 lo = nd.lo[k], hi = nd.hi[k];
        bl.lo[0] = lo.x; bl.lo[1] = lo.y; bl.lo[2] = lo.z;
        bl.hi[0] = hi.x; bl.hi[1] = hi.y; bl.hi[2] = hi.z;
        br = bl;
        cl = count / 2;
        cr = count - cl;
        nd.split[k] = -2;
    }
    nd.left[k] = left;
    nd.lo[left] = make_float4(bl.lo[0], bl.lo[1], bl.lo[2], 0.f);
    nd.hi[left] = make_float4(bl.hi[0], bl.hi[1], bl.hi[2], 0.f);
    nd.lo[left + 1] = make_float4(br.lo[0], br.lo[1], br.lo[2], 0.f);
    nd.hi[left + 1] = make_float4(br.hi[0], br.hi[1], br.hi[2], 0.f);
    nd.count[left] = cl;
    nd.count[left + 1] = cr;
    nd.left[left] = nd.left[left + 1] = -3;
    nd.slot[left] = cl > max_leaf ? atomicAdd(&counters[1], 1) : -1;
    nd.slot[left + 1] = cr > max_leaf ? atomicAdd(&counters[1], 1) : -1;
}

__global__ void k_sah_assign(int n, int level_begin, unsigned* __restrict__ prim_node, const unsigned* __restrict__ prim_id,
                             const float4* __restrict__ plo, const float4* __restrict__ phi, SahNodes nd) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int node = (int)prim_node[i];
    if (node < level_begin) return;
    const int left = nd.left[node];
    if (left < 0) return;  // leaf
    const int sp = nd.split[node];
    bool go_left;
    if (sp == -2) {
        go_left = i - nd.first[node] < nd.count[node] / 2;
    } else {
        const int axis = sp / SAH_BINS, bin = sp % SAH_BINS;
        const unsigned g = prim_id[i];
        const float4 lo = plo[g], hi = phi[g];
        const float4 nlo = nd.lo[node], nhi = nd.hi[node];
        const float c = axis == 0 ? 0.5f * (lo.x + hi.x) : (axis == 1 ? 0.5f * (lo.y + hi.y) : 0.5f * (lo.z + hi.z));
        const float a0 = axis == 0 ? nlo.x : (axis == 1 ? nlo.y : nlo.z), a1 = axis == 0 ? nhi.x : (axis == 1 ? nhi.y : nhi.z);
        go_left = sah_bin_of(c, a0, a1) <= bin;
    }
    prim_node[i] = (unsigned)(go_left ? left : left + 1);
}

__global__ void k_sah_first(int n, const unsigned* __restrict__ prim_node, SahNodes nd) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (i == 0 || prim_node[i] != prim_node[i - 1]) nd.first[prim_node[i]] = i;
}

__global__ void k_sah_emit(int n_nodes, float pad, SahNodes nd, BvhNode* __restrict__ out) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_nodes) return;
    BvhNode o;
    const int left = nd.left[k];
    if (left < 0) {  // leaves are referenced by (first, count), never as nodes
        o.a = o.b = o.c = make_float4(0.f, 0.f, 0.f, 0.f);
        o.c0 = o.c1 = o.pad0 = o.pad1 = 0;
        out[k] = o;
        return;
    }
    int ref[2];
    float4 lo[2], hi[2];
#pragma unroll
    for (int j = 0; j < 2; j++) {
        const int c = left + j;
        lo[j] = nd.lo[c];
        hi[j] = nd.hi[c];
        ref[j] = nd.left[c] < 0 ? leaf_ref(nd.first[c], nd.count[c]) : c;
    }
    o.a = make_float4(lo[0].x - pad, lo[0].y - pad, lo[0].z - pad, hi[0].x + pad);
    o.b = make_float4(hi[0].y + pad, hi[0].z + pad, lo[1].x - pad, lo[1].y - pad);
    o.c = make_float4(lo[1].z - pad, hi[1].x + pad, hi[1].y + pad, hi[1].z + pad);
    o.c0 = ref[0];
    o.c1 = ref[1];
    o.pad0 = o.pad1 = 0;
    out[k] = o;
}

__global__ void k_sah_init(int n, unsigned* __restrict__ prim_node, SahNodes nd, const unsigned* __restrict__ scene_box,
                           int max_leaf, int* __restrict__ counters) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) prim_node[i] = 0u;
    if (i == 0) {
        nd.lo[0] = make_float4(ord2f(scene_box[0]), ord2f(scene_box[1]), ord2f(scene_box[2]), 0.f);
        nd.hi[0] = make_float4(ord2f(scene_box[3]), ord2f(scene_box[4]), ord2f(scene_box[5]), 0.f);
        nd.count[0] = n;
        nd.first[0] = 0;
        nd.left[0] = -3;
        nd.slot[0] = n > max_leaf ? 0 : -1;
        counters[0] = 1;
        counters[1] = 0;
    }
}

// d_pos: 9 floats per global primitive, already on the device.  One small read-back per level.
inline cudaError_t build_bvh_sah(const float* d_pos, int n, int max_leaf, cudaStream_t st, BvhBuildResult& out) {
    out = BvhBuildResult();
    if (n <= 0) return cudaSuccess;
    max_leaf = max_leaf <= 0 ? 4 : (max_leaf > 8 ? 8 : max_leaf);
    const int B = SAH_BLOCK, G = (n + B - 1) / B;
    const size_t max_nodes = 2 * (size_t)n + 2;
    const size_t max_slots = (size_t)n / (size_t)(max_leaf + 1) + 2;  // nodes that can still be split
    float4 *plo = nullptr, *phi = nullptr;
    unsigned *scene_box = nullptr, *node_a = nullptr, *node_b = nullptr, *id_a = nullptr, *id_b = nullptr;
    unsigned long long *keys = nullptr, *keys_s = nullptr;
    SahNodes nd = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    SahBin* bins = nullptr;
    int* counters = nullptr;
    void* cub_tmp = nullptr;
    DevFreeOnExit scratch;
    for (void* pp : {(void*)&plo, (void*)&phi, (void*)&scene_box, (void*)&node_a, (void*)&node_b, (void*)&id_a, (void*)&id_b,
                     (void*)&keys, (void*)&keys_s, (void*)&nd.lo, (void*)&nd.hi, (void*)&nd.count, (void*)&nd.first,
                     (void*)&nd.left, (void*)&nd.split, (void*)&nd.slot, (void*)&bins, (void*)&counters, (void*)&cub_tmp})
        scratch.add(pp);
    ORB_CU(cudaMalloc(&plo, sizeof(float4) * n));
    ORB_CU(cudaMalloc(&phi, sizeof(float4) * n));
    ORB_CU(cudaMalloc(&scene_box, 6 * sizeof(unsigned)));
    ORB_CU(cudaMalloc(&keys, 8 * (size_t)n));
    ORB_CU(cudaMalloc(&keys_s, 8 * (size_t)n));
    ORB_CU(cudaMalloc(&node_a, 4 * (size_t)n));
    ORB_CU(cudaMalloc(&node_b, 4 * (size_t)n));
    ORB_CU(cudaMalloc(&id_a, 4 * (size_t)n));
    ORB_CU(cudaMalloc(&id_b, 4 * (size_t)n));
    ORB_CU(cudaMalloc(&out.tri, sizeof(float4) * 2 * (size_t)n));
    ORB_CU(cudaMalloc(&out.tri_b, sizeof(float2) * (size_t)n));
    const unsigned box_init[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
    ORB_CU(cudaMemcpyAsync(scene_box, box_init, sizeof(box_init), cudaMemcpyHostToDevice, st));
    k_prim_bounds<<<G, B, 0, st>>>(d_pos, n, plo, phi, scene_box);
    // start from the Morton order: neighbouring slots are neighbouring primitives (block-local bins hit)
    k_morton<<<G, B, 0, st>>>(plo, phi, n, scene_box, keys, id_a);
    size_t tmp_bytes = 0, tmp2 = 0;
    ORB_CU(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys, keys_s, id_a, id_b, n, 0, 63, st));
    ORB_CU(cub::DeviceRadixSort::SortPairs(nullptr, tmp2, node_a, node_b, id_a, id_b, n, 0, 32, st));
    tmp_bytes = tmp_bytes > tmp2 ? tmp_bytes : tmp2;
    ORB_CU(cudaMalloc(&cub_tmp, tmp_bytes ? tmp_bytes : 16));
    ORB_CU(cub::DeviceRadixSort::SortPairs(cub_tmp, tmp_bytes, keys, keys_s, id_a, id_b, n, 0, 63, st));
    std::swap(id_a, id_b);  // id_a: primitives in Morton order
    out.launches = 3;
    unsigned box_h[6];
    ORB_CU(cudaMemcpyAsync(box_h, scene_box, sizeof(box_h), cudaMemcpyDeviceToHost, st));
    if (n <= max_leaf) {
        k_tri_records<<<G, B, 0, st>>>(d_pos, id_a, n, out.tri, out.tri_b);
        out.launches++;
        out.root_ref = leaf_ref(0, n);
        out.height = 0;
        ORB_CU(cudaStreamSynchronize(st));
    } else {
        ORB_CU(cudaMalloc(&nd.lo, sizeof(float4) * max_nodes));
        ORB_CU(cudaMalloc(&nd.hi, sizeof(float4) * max_nodes));
        ORB_CU(cudaMalloc(&nd.count, sizeof(int) * max_nodes));
        ORB_CU(cudaMalloc(&nd.first, sizeof(int) * max_nodes));
        ORB_CU(cudaMalloc(&nd.left, sizeof(int) * max_nodes));
        ORB_CU(cudaMalloc(&nd.split, sizeof(int) * max_nodes));
        ORB_CU(cudaMalloc(&nd.slot, sizeof(int) * max_nodes));
        ORB_CU(cudaMalloc(&bins, sizeof(SahBin) * 3 * SAH_BINS * max_slots));
        ORB_CU(cudaMalloc(&counters, 2 * sizeof(int)));
        k_sah_init<<<G, B, 0, st>>>(n, node_a, nd, scene_box, max_leaf, counters);
        out.launches++;
        int key_bits = 1;
        while ((1ull << key_bits) < max_nodes) key_bits++;
        int level_begin = 0, level_end = 1, slots = 1, depth = 0;
        while (level_begin < level_end) {
            const int n_level = level_end - level_begin;
            if (slots > 0) {
                const int nb = slots * 3 * SAH_BINS;
                k_sah_clear<<<(nb + 255) / 256, 256, 0, st>>>(bins, nb);
                k_sah_bin<<<G, B, 0, st>>>(n, level_begin, node_a, id_a, plo, phi, nd, bins);
                out.launches += 2;
            }
            ORB_CU(cudaMemsetAsync(counters + 1, 0, sizeof(int), st));
            k_sah_split<<<(n_level + 127) / 128, 128, 0, st>>>(level_begin, level_end, max_leaf, nd, bins, counters);
            out.launches++;
            int h_cnt[2];
            ORB_CU(cudaMemcpyAsync(h_cnt, counters, sizeof(h_cnt), cudaMemcpyDeviceToHost, st));
            ORB_CU(cudaStreamSynchronize(st));
            if (h_cnt[0] > level_end) {  // children were created: move the primitives down and regroup
                k_sah_assign<<<G, B, 0, st>>>(n, level_begin, node_a, id_a, plo, phi, nd);
                ORB_CU(cub::DeviceRadixSort::SortPairs(cub_tmp, tmp_bytes, node_a, node_b, id_a, id_b, n, 0, key_bits, st));
                std::swap(node_a, node_b);
                std::swap(id_a, id_b);
                k_sah_first<<<G, B, 0, st>>>(n, node_a, nd);
                out.launches += 3;
                depth++;
            }
            level_begin = level_end;
            level_end = h_cnt[0];
            slots = h_cnt[1];
            if ((size_t)slots > max_slots) return cudaErrorUnknown;  // cannot happen: each takes > max_leaf primitives
        }
        const int n_nodes = level_end;
        float ext = 0.f;
        for (int a = 0; a < 3; a++) ext = fmaxf(ext, ord2f(box_h[3 + a]) - ord2f(box_h[a]));
        const float pad = ext * 1e-5f;
        ORB_CU(cudaMalloc(&out.nodes, sizeof(BvhNode) * (size_t)n_nodes));
        k_sah_emit<<<(n_nodes + 255) / 256, 256, 0, st>>>(n_nodes, pad, nd, out.nodes);
        k_tri_records<<<G, B, 0, st>>>(d_pos, id_a, n, out.tri, out.tri_b);
        out.launches += 2;
        ORB_CU(cudaStreamSynchronize(st));
        out.root_ref = 0;
        out.height = depth;
        out.n_nodes = n_nodes;
    }
    ORB_CU(cudaGetLastError());
    return cudaSuccess;
}

}  // namespace orb

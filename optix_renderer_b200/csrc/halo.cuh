// halo.cuh -- the tile hand-off of a row-band partition, fused into the denoise pack kernel.
//
// The reference renders on ONE GPU (owlContextCreate(nullptr, 1), include/viewer.hpp:177); there is nothing to
// restate.  With the frame split into row bands (one per GPU) the a-trous filter of a band reads
// 2 * (2^iterations - 1) packed rows of its neighbours (guide float4 + value float2, 24 B/px).  Round 1 exchanged
// them with NCCL point-to-point calls between pack and filter: 0.08 - 0.11 ms of every 0.95 ms frame at N = 8, all
// of it on the critical path.  Here every rank maps its peers' ORB_BUF_DN_GUIDE / ORB_BUF_DN_VAL (cudaIpc handles,
// exchanged once) and k_denoise_pack STORES the rows a neighbour will read straight into that neighbour's buffers
// over NVLink while it packs its own -- compute and transfer are one kernel.  Ordering is by epoch flags in a small
// control block per rank, written with system-scope stores and awaited by one-thread kernels on the frame's stream:
//
//   ready[slot][src]     written by rank `src` into MY block after its pack of this slot's frame: its rows are here
//   consumed[slot][dst]  written by rank `dst` into MY block after its filter: my rows in ITS buffers may be
//                        overwritten by the next frame that uses this slot
//
// Every rank signals BOTH flags to ALL connected peers for EVERY frame (also when it has no rows for them), so a
// band plan that changes between frames (cost re-balancing) never waits on a stale flag; epochs are per slot and
// advance in lockstep because every rank renders every frame.  A wait gives up after ~2 s, records the failure in
// the control block (orb_halo_status) and lets the stream continue: a lost peer must not hang the GPU.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace orb {

#define ORB_HALO_MAX_PEERS 8
#define ORB_HALO_MAX_SLOTS 8
#define ORB_HALO_MAX_SENDS 4  // row ranges one pack launch pushes (a band narrower than the filter's reach has > 2)

// control block layout (uint32 words): [0] error flag, then ready[slot][rank], then consumed[slot][rank]
#define ORB_HALO_ERR 0
#define ORB_HALO_READY(slot, rank) (8 + (slot) * ORB_HALO_MAX_PEERS + (rank))
#define ORB_HALO_CONSUMED(slot, rank) (8 + (ORB_HALO_MAX_SLOTS + (slot)) * ORB_HALO_MAX_PEERS + (rank))
#define ORB_HALO_WORDS (8 + 2 * ORB_HALO_MAX_SLOTS * ORB_HALO_MAX_PEERS)
#define ORB_HALO_BLOCK_BYTES (4u << 20)  // its own allocation of >= 2 MB: an IPC handle then maps it at offset 0

struct HaloPeer {
    int rank = -1;
    float4* guide[ORB_HALO_MAX_SLOTS] = {nullptr};  // the peer's ORB_BUF_DN_GUIDE, mapped here
    float2* val[ORB_HALO_MAX_SLOTS] = {nullptr};    // the peer's ORB_BUF_DN_VAL
    uint32_t* flags = nullptr;                                // the peer's control block
};

struct HaloLinks {
    int my_rank = 0, n_peers = 0;
    HaloPeer peer[ORB_HALO_MAX_PEERS];
    uint32_t* flags = nullptr;  // MY control block (device memory, ORB_HALO_BLOCK_BYTES)
};

// what one pack launch pushes: rows [y0, y1) of this frame into one peer's buffers
struct HaloSend {
    float4* guide;
    float2* val;
    int y0, y1;
};
struct HaloSendList {
    HaloSend s[ORB_HALO_MAX_SENDS];
    int n;
};
struct HaloFlagList {  // flag words to write (remote) or to wait for (local)
    uint32_t* p[ORB_HALO_MAX_PEERS];
    int n;
};

inline void halo_disconnect(HaloLinks& h) {
    for (int i = 0; i < h.n_peers; i++) h.peer[i] = HaloPeer();
    h.n_peers = 0;
}

// one thread: publish `epoch` in every listed (peer-mapped) flag word.  The fence orders the pack kernel's peer
// stores -- complete at the kernel boundary before this one -- ahead of the flag for an observer on another GPU.
__global__ void k_halo_signal(HaloFlagList f, uint32_t epoch) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    __threadfence_system();
    for (int i = 0; i < f.n; i++) *reinterpret_cast<volatile uint32_t*>(f.p[i]) = epoch;
    __threadfence_system();
}

// one thread: spin until every listed LOCAL flag word has reached `epoch` (epochs only grow; wrap after 2^32 frames)
__global__ void k_halo_wait(HaloFlagList f, uint32_t epoch, uint32_t* err, long long timeout_cycles) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const long long t0 = clock64();
    for (int i = 0; i < f.n; i++) {
        volatile uint32_t* p = reinterpret_cast<volatile uint32_t*>(f.p[i]);
        while ((int32_t)(*p - epoch) < 0) {
            if (clock64() - t0 > timeout_cycles) {
                *err = 1u + (uint32_t)i;  // which wait gave up; the stream goes on (no hang), the host reports it
                return;
            }
            __nanosleep(200);
        }
    }
    __threadfence_system();  // acquire: the rows behind the flag are visible to the kernels that follow
}

}  // namespace orb

// rng.cuh -- per-pixel random stream.
//
// Restates cuda_include/lcg_random.cuh:11-62: seed = murmur3 fmix32 of (pixel index, frame id)
// mixed from hash 0, then a 32-bit LCG (a = 1664525, c = 1013904223) whose state is returned
// and scaled by 2^-32 after an int->float conversion (round-to-nearest, so 1.0f is reachable,
// SURVEY.md Q21).  Draw ORDER is the contract between the wavefront stages:
//   draw 0,1              pixel jitter x, y                     cuda_src/deviceCode.cu:68
//   draw 2+5i .. 6+5i     shadow sample i / bounce i: rand1.x rand1.y rand2.x rand2.y select
//                                                               ratio.cuh:29-33, path.cuh:165-169
// Every stage re-derives its stream position from (pixel, frame) with skip(), so no RNG state
// travels through HBM for the G-buffer / LTC / shadow passes.
#pragma once
#include "vec.cuh"

namespace orb {

struct Rng {
    uint32_t s;
};

ORB_HD uint32_t rotl32(uint32_t v, int r) { return (v << r) | (v >> (32 - r)); }

ORB_HD uint32_t murmur_mix(uint32_t h, uint32_t k) {
    k *= 0xcc9e2d51u;
    k = rotl32(k, 15);
    k *= 0x1b873593u;
    h ^= k;
    return rotl32(h, 13) * 5u + 0xe6546b64u;
}

ORB_HD uint32_t murmur_fmix(uint32_t h) {
    h ^= h >> 16;
    h *= 0x85ebca6bu;
    h ^= h >> 13;
    h *= 0xc2b2ae35u;
    h ^= h >> 16;
    return h;
}

// get_rng(frame_id, pixel, dims): the caller passes frame_id = accumId + 10007.
ORB_HD Rng rng_seed(int frame_id, uint32_t px, uint32_t py, uint32_t width) {
    Rng r;
    r.s = murmur_fmix(murmur_mix(murmur_mix(0u, px + py * width), (uint32_t)frame_id));
    return r;
}

ORB_HD uint32_t rng_next(Rng& r) {
    r.s = r.s * 1664525u + 1013904223u;
    return r.s;
}

// ldexp((float)u, -32): the int->float conversion rounds to nearest even, the scale is exact.
ORB_HD float rng_nextf(Rng& r) { return (float)rng_next(r) * 2.3283064365386963e-10f; }

// Advance by n draws (n is small: 2 + 5*sample).
ORB_HD void rng_skip(Rng& r, int n) {
    uint32_t a = 1u, c = 0u;  // composition of n affine steps
    for (int i = 0; i < n; i++) {
        c = c * 1664525u + 1013904223u;
        a *= 1664525u;
    }
    r.s = r.s * a + c;
}

}  // namespace orb

// vec.cuh -- fp32 3-vector helpers for the sm_100a kernels.
//
// The reference does all of its per-pixel arithmetic on owl::common::vec3f (un-vendored OWL,
// SURVEY.md A.2): plain fp32 component-wise operators, dot = x*x' + y*y' + z*z',
// normalize(v) = v * rsqrt(dot(v,v)).  These helpers restate that algebra on CUDA's float3.
// Kernels are compiled WITHOUT --use_fast_math (IEEE div/sqrt, fmad contraction on), the
// flags OWL's embed_ptx is assumed to use.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <string.h>

// ORB_CHECKS build (tools/build_checked.sh -> liborb200_chk.so): device-side assertions on every index the kernels
// form from claimed work items, stack depths, node / triangle references and tap coordinates.  compute-sanitizer is
// not available on this pool (profiles/r2m_sanitizer_closed.txt); the GPU suite run against the checked library is
// the bounds evidence instead.  A failed assertion prints file:line and turns the next CUDA call into
// cudaErrorAssert, so the test that triggered it fails.
#if defined(ORB_CHECKS) && defined(__CUDA_ARCH__)
#include <assert.h>
#define ORB_DEVICE_ASSERT(c) assert(c)
#else
#define ORB_DEVICE_ASSERT(c) ((void)0)
#endif

#define ORB_PI_D 3.141592653589793238  // include/common.h:4 -- a DOUBLE literal in the reference
#define ORB_PI_F ((float)ORB_PI_D)
#define ORB_EPS 1e-5f  // cuda_include/frostbite.cuh:8

#ifdef __CUDACC__
#define ORB_HD __host__ __device__ __forceinline__
#define ORB_D __device__ __forceinline__
#else  // plain C++ build of the same headers: tests/hostcheck only
#define ORB_HD inline
#define ORB_D inline
#endif

namespace orb {

ORB_HD float3 v3(float x, float y, float z) { return make_float3(x, y, z); }
ORB_HD float3 v3(float s) { return make_float3(s, s, s); }
ORB_HD float3 v3(const float4& f) { return make_float3(f.x, f.y, f.z); }
ORB_HD float3 operator+(float3 a, float3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
ORB_HD float3 operator-(float3 a, float3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
ORB_HD float3 operator*(float3 a, float3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
ORB_HD float3 operator/(float3 a, float3 b) { return v3(a.x / b.x, a.y / b.y, a.z / b.z); }
ORB_HD float3 operator*(float3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
ORB_HD float3 operator*(float s, float3 a) { return v3(s * a.x, s * a.y, s * a.z); }
ORB_HD float3 operator/(float3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
ORB_HD float3 operator-(float3 a) { return v3(-a.x, -a.y, -a.z); }
ORB_HD float3& operator+=(float3& a, float3 b) { a = a + b; return a; }
ORB_HD float3& operator*=(float3& a, float3 b) { a = a * b; return a; }
ORB_HD float dot(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
ORB_HD float3 cross(float3 a, float3 b) {
    return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
// rsqrt: one MUFU.RSQ on the device (<= 2 ulp); the CPU build used by tests/hostcheck divides.
ORB_HD float orb_rsqrt(float x) {
#ifdef __CUDA_ARCH__
    float r;  // bare MUFU.RSQ: without .ftz the compiler wraps it in a denormal-scaling sequence (4 instructions)
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
#else
    return 1.f / sqrtf(x);
#endif
}
// quotient to 2 ulp (MUFU.RCP + FMUL) where the reference's IEEE division is not needed for parity:
// |b| must be in [2^-126, 2^126] (all uses: bounded denominators)
ORB_HD float orb_fdiv(float a, float b) {
#ifdef __CUDA_ARCH__
    return __fdividef(a, b);
#else
    return a / b;
#endif
}
ORB_HD float3 normalize(float3 a) { return a * orb_rsqrt(dot(a, a)); }
ORB_HD float length(float3 a) { return sqrtf(dot(a, a)); }
ORB_HD int __float_as_int_hd(float f) {
#ifdef __CUDA_ARCH__
    return __float_as_int(f);
#else
    int i;
    memcpy(&i, &f, 4);
    return i;
#endif
}
ORB_HD float __int_as_float_hd(int i) {
#ifdef __CUDA_ARCH__
    return __int_as_float(i);
#else
    float f;
    memcpy(&f, &i, 4);
    return f;
#endif
}
ORB_HD int imin(int a, int b) { return a < b ? a : b; }
ORB_HD int imax(int a, int b) { return a > b ? a : b; }

// ---- "exact" fp32 arithmetic ------------------------------------------------------------------
// Single IEEE-rounded operations that the compiler may NOT contract into FMAs.  Everything that
// DEFINES a ray (camera direction, hit point, interpolated normal, light sample, shadow
// direction) and the ray/triangle test itself are written with these, so primitive ids and
// barycentrics are bit-identical to the CPU checker (oracle/trace_cpu.h, compiled with
// -ffp-contract=off) and do not depend on nvcc's contraction choices.  The CPU build of these
// headers (tests/hostcheck) is compiled with -ffp-contract=off, where plain operators are exact.
#ifdef __CUDA_ARCH__
ORB_HD float xmul(float a, float b) { return __fmul_rn(a, b); }
ORB_HD float xadd(float a, float b) { return __fadd_rn(a, b); }
ORB_HD float xsub(float a, float b) { return __fsub_rn(a, b); }
ORB_HD float xfma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
ORB_HD float xdiv(float a, float b) { return __fdiv_rn(a, b); }
ORB_HD float xsqrt(float a) { return __fsqrt_rn(a); }
#else
ORB_HD float xmul(float a, float b) { return a * b; }
ORB_HD float xadd(float a, float b) { return a + b; }
ORB_HD float xsub(float a, float b) { return a - b; }
ORB_HD float xfma(float a, float b, float c) { return __builtin_fmaf(a, b, c); }
ORB_HD float xdiv(float a, float b) { return a / b; }
ORB_HD float xsqrt(float a) { return sqrtf(a); }
#endif
ORB_HD float3 xadd(float3 a, float3 b) { return v3(xadd(a.x, b.x), xadd(a.y, b.y), xadd(a.z, b.z)); }
ORB_HD float3 xsub(float3 a, float3 b) { return v3(xsub(a.x, b.x), xsub(a.y, b.y), xsub(a.z, b.z)); }
ORB_HD float3 xmul(float s, float3 a) { return v3(xmul(s, a.x), xmul(s, a.y), xmul(s, a.z)); }
ORB_HD float3 xmul(float3 a, float s) { return v3(xmul(a.x, s), xmul(a.y, s), xmul(a.z, s)); }
ORB_HD float xdot(float3 a, float3 b) { return xadd(xadd(xmul(a.x, b.x), xmul(a.y, b.y)), xmul(a.z, b.z)); }
// owl normalize as the checker's shim spells it: a * (1 / sqrt(dot(a,a)))
ORB_HD float3 xnormalize(float3 a) { return xmul(a, xdiv(1.f, xsqrt(xdot(a, a)))); }
ORB_HD float xlength(float3 a) { return xsqrt(xdot(a, a)); }

// owl::common::max / min (assumed `a<b?b:a`): NaN in `a` propagates.
ORB_HD float owl_max(float a, float b) { return a < b ? b : a; }
ORB_HD float owl_min(float a, float b) { return a < b ? a : b; }

// 3x3 matrix as three ROW vectors, like the reference's `vec3f mat[3]`.
struct Mat3 {
    float3 r0, r1, r2;
};
// apply_mat, cuda_include/utils.cuh:70-74
ORB_HD float3 apply(const Mat3& m, float3 v) { return v3(dot(m.r0, v), dot(m.r1, v), dot(m.r2, v)); }
// matrixTranspose, cuda_include/utils.cuh:142-155
ORB_HD Mat3 transpose(const Mat3& m) {
    Mat3 t;
    t.r0 = v3(m.r0.x, m.r1.x, m.r2.x);
    t.r1 = v3(m.r0.y, m.r1.y, m.r2.y);
    t.r2 = v3(m.r0.z, m.r1.z, m.r2.z);
    return t;
}

// 128-bit streaming accesses for frame buffers that are touched once per pass
// (keeps the scene / BVH working set in L1 and L2).
#ifdef __CUDACC__
ORB_D float4 ld_stream(const float4* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p));
    return r;
}
ORB_D void st_stream(float4* p, float4 v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
                 "f"(v.w)
                 : "memory");
}

#endif  // __CUDACC__

}  // namespace orb

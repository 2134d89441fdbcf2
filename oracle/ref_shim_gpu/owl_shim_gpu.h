// owl_shim_gpu.h -- TEST INFRASTRUCTURE (oracle/_ref/libref_gpu_*.so only; never linked into the product).
//
// The device-side twin of ../ref_shim/owl_shim.h: the parts of owl-project/owl (@4416b979, un-vendored) and of
// the OptiX 7 device API that the reference's device translation unit touches, as __host__ __device__ code,
// so that cuda_src/deviceCode.cu + cuda_include/** compile UNMODIFIED with nvcc for sm_100a and run as the
// literal one-thread-per-pixel megakernel on the B200 (SURVEY.md 8(d) "GPU baseline to beat").
// Same assumptions about OWL semantics as the host shim.  Differences from the host shim:
//   * no draw-order fix-up: nvcc device code evaluates vec2f(lcg_randomf(rng), lcg_randomf(rng)) left to right;
//   * tex2D<float4> is the real texture unit (cudaTextureObject_t made by the driver), not an emulation;
//   * the per-thread hooks (launch index, PRD, program data, hit attributes) live in a __shared__ slot per thread.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <texture_fetch_functions.h>
#include <vector_functions.h>
#include <vector_types.h>

#define SHD __host__ __device__ inline

typedef unsigned long long OptixTraversableHandle;

// P3 parameters; referenced only by the patched copies of the intended build.
__device__ int shim_shadow_spp = 4;  // one TU only (ref_gpu_driver.cu)
__device__ int shim_max_depth = 1;

#define OWL_TERMINAL_BLUE ""
#define OWL_TERMINAL_LIGHT_BLUE ""
#define OWL_TERMINAL_DEFAULT ""

#define OPTIX_RAYGEN_PROGRAM(name) __device__ void __raygen__##name
#define OPTIX_CLOSEST_HIT_PROGRAM(name) __device__ void __closesthit__##name
#define OPTIX_MISS_PROGRAM(name) __device__ void __miss__##name

namespace owl {
namespace common {

// Default constructors are EMPTY, as OWL's vec_t() {} is: a __constant__ LaunchParams needs that, and it is what
// the reference's real build runs with (the host shim zero-fills instead, for determinism).
struct vec2i {
    int x, y;
    SHD vec2i() {}
    SHD vec2i(int a, int b) : x(a), y(b) {}
};
struct vec3i {
    int x, y, z;
    SHD vec3i() {}
    SHD vec3i(int a, int b, int c) : x(a), y(b), z(c) {}
};

struct vec2f {
    union { float x; float u; };
    union { float y; float v; };
    SHD vec2f() {}
    SHD vec2f(float a) : x(a), y(a) {}
    template <class A, class B> SHD vec2f(A a, B b) : x((float)a), y((float)b) {}
    SHD explicit vec2f(const vec2i& i) : x((float)i.x), y((float)i.y) {}
};
typedef vec2f vec2;

struct vec3f {
    float x, y, z;
    SHD vec3f() {}
    template <class A> SHD vec3f(A a) : x((float)a), y((float)a), z((float)a) {}
    template <class A, class B, class C> SHD vec3f(A a, B b, C c) : x((float)a), y((float)b), z((float)c) {}
    SHD float& operator[](int i) { return (&x)[i]; }
    SHD const float& operator[](int i) const { return (&x)[i]; }
};

struct vec4f {
    float x, y, z, w;
    SHD vec4f() {}
    SHD vec4f(const float4& f) : x(f.x), y(f.y), z(f.z), w(f.w) {}
};

SHD float abs(float a) { return fabsf(a); }
SHD float sqrt(float a) { return sqrtf(a); }
SHD float cos(float a) { return cosf(a); }
SHD float sin(float a) { return sinf(a); }
SHD float max(float a, float b) { return a < b ? b : a; }
SHD float min(float a, float b) { return a < b ? a : b; }
SHD float clamp(float f, float lo, float hi) { return max(lo, min(f, hi)); }

#define OWLSHIM_VOP(op)                                                                                      \
    SHD vec3f operator op(const vec3f& a, const vec3f& b) { return vec3f(a.x op b.x, a.y op b.y, a.z op b.z); } \
    SHD vec3f operator op(const vec3f& a, float b) { return vec3f(a.x op b, a.y op b, a.z op b); }              \
    SHD vec3f operator op(float a, const vec3f& b) { return vec3f(a op b.x, a op b.y, a op b.z); }              \
    SHD vec2f operator op(const vec2f& a, const vec2f& b) { return vec2f(a.x op b.x, a.y op b.y); }             \
    SHD vec2f operator op(const vec2f& a, float b) { return vec2f(a.x op b, a.y op b); }                        \
    SHD vec2f operator op(float a, const vec2f& b) { return vec2f(a op b.x, a op b.y); }
OWLSHIM_VOP(+)
OWLSHIM_VOP(-)
OWLSHIM_VOP(*)
OWLSHIM_VOP(/)
#undef OWLSHIM_VOP

SHD vec3f operator-(const vec3f& a) { return vec3f(-a.x, -a.y, -a.z); }
SHD vec2f operator-(const vec2f& a) { return vec2f(-a.x, -a.y); }
SHD vec3f& operator+=(vec3f& a, const vec3f& b) { a = a + b; return a; }
SHD vec3f& operator-=(vec3f& a, const vec3f& b) { a = a - b; return a; }
SHD vec3f& operator*=(vec3f& a, const vec3f& b) { a = a * b; return a; }
SHD vec3f& operator*=(vec3f& a, float b) { a = a * b; return a; }
SHD vec3f& operator/=(vec3f& a, float b) { a = a / b; return a; }

SHD float dot(const vec3f& a, const vec3f& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
SHD vec3f cross(const vec3f& a, const vec3f& b) {
    return vec3f(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
SHD float length(const vec3f& a) { return sqrtf(dot(a, a)); }
// as in the host shim (owl's own is v * rsqrt(dot(v,v)); <= 1 ulp apart, SURVEY.md 8(c) caveat 5)
SHD vec3f normalize(const vec3f& a) { return a * (1.f / sqrtf(dot(a, a))); }

}  // namespace common
using namespace common;

template <int RayType, int NumRayTypes>
struct RayT {
    common::vec3f origin, direction;
    float tmin = 0.f, tmax = 1e30f, time = 0.f;
};

SHD uint32_t make_rgba(const common::vec3f& c) {
    const float f[3] = {c.x * 256.f, c.y * 256.f, c.z * 256.f};
    uint32_t q[3];
    for (int k = 0; k < 3; k++) q[k] = (f[k] != f[k]) ? 0u : (f[k] >= 255.f ? 255u : (f[k] <= 0.f ? 0u : (uint32_t)(int)f[k]));
    return q[0] | (q[1] << 8) | (q[2] << 16) | (0xffu << 24);
}

// ---- per-thread hooks; the slot is written by oracle/ref_shim_gpu/ref_gpu_driver.cu ----------------
struct ShimCtx {
    const void* programData;
    void* prd;
    float2 bary;
    unsigned int prim;
    int traces, firstPrim;
    float firstU, firstV;
};
#define SHIM_BLOCK_X 16
#define SHIM_BLOCK_Y 8
__device__ __forceinline__ ShimCtx& shim_ctx() {
    __shared__ ShimCtx slots[SHIM_BLOCK_X * SHIM_BLOCK_Y];
    return slots[threadIdx.y * SHIM_BLOCK_X + threadIdx.x];
}
__device__ void shim_trace(const common::vec3f& org, const common::vec3f& dir, void* prd);

template <class T> __device__ inline const T& getProgramData() { return *(const T*)shim_ctx().programData; }
template <class T> __device__ inline T& getPRD() { return *(T*)shim_ctx().prd; }
__device__ inline common::vec2i getLaunchIndex() {
    return common::vec2i((int)(blockIdx.x * blockDim.x + threadIdx.x), (int)(blockIdx.y * blockDim.y + threadIdx.y));
}
template <class Ray, class PRD>
__device__ inline void traceRay(OptixTraversableHandle, const Ray& ray, PRD& prd) {
    shim_trace(ray.origin, ray.direction, (void*)&prd);
}
}  // namespace owl

// common.cuh:59-62 declares RadianceRay only in device passes; nvcc's host pass of this TU still parses the
// __device__ functions that use it (a typedef may be repeated)
typedef owl::RayT<0, 2> RadianceRay;

__device__ inline float2 optixGetTriangleBarycentrics() { return owl::shim_ctx().bary; }
__device__ inline unsigned int optixGetPrimitiveIndex() { return owl::shim_ctx().prim; }

// restate.cpp -- TEST INFRASTRUCTURE ONLY (oracle/liboracle.so).  Never linked into or loaded by
// the product; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
// reference legs may load it.
//
// A plain-C++ CPU restatement of the reference's per-pixel hot path, one function per reference
// function, each citing the file:line (relative to /root/reference) it follows.  It implements
// oracle/oracle_api.h, literal and intended (patch set P1-P3) selected at run
// time.  PINNING: the reference ships no tests or golden vectors, so this
// restatement is pinned against the reference's OWN device source compiled unmodified as host
// C++ (oracle/_ref/libref_literal.so, libref_intended.so; tests/test_oracle.py compares the two
// bit for bit) and against the known-answer vectors generated from it (tests/golden/).
// What stays UNPINNED, because the code lives in absent third-party layers (owl-project/owl
// @4416b979, OptiX 7, the CUDA texture unit): the ray/triangle test and traversal order, and
// the texture filter; both are the documented stand-ins of oracle/trace_cpu.h.  The denoise
// filter has no reference implementation at all (cuda_include/denoise/bilateral.cuh is empty):
// ora_denoise() below is the DEFINITION the CUDA kernel is checked against ("parity unpinned").
//
// Arithmetic notes: fp32 with the double sub-expressions the reference's double `PI` literal and
// unsuffixed constants cause (include/common.h:4); evaluation order as written there; compiled
// with -ffp-contract=off so nothing is fused.  RNG draws are taken left to right (nvcc device
// order).
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <atomic>
#include <thread>
#include <vector>

#include "oracle_api.h"
#include "trace_cpu.h"

namespace {

const double kPi = 3.141592653589793238;  // include/common.h:4 (double)
const float kEps = float(1e-5);           // cuda_include/frostbite.cuh:8

struct V3 {
    float x, y, z;
};
inline V3 mk(float x, float y, float z) { return V3{x, y, z}; }
inline V3 mk(const float* p) { return V3{p[0], p[1], p[2]}; }
inline V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator*(V3 a, V3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V3 operator/(V3 a, V3 b) { return mk(a.x / b.x, a.y / b.y, a.z / b.z); }
inline V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(float s, V3 a) { return mk(s * a.x, s * a.y, s * a.z); }
inline V3 operator/(V3 a, float s) { return mk(a.x / s, a.y / s, a.z / s); }
inline V3 operator-(V3 a) { return mk(-a.x, -a.y, -a.z); }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 cross(V3 a, V3 b) { return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline float length(V3 a) { return sqrtf(dot(a, a)); }
// owl::normalize = v * rsqrt(dot(v,v)); 1/sqrt on the CPU like the _ref shim
inline V3 normalize(V3 a) { return a * (1.f / sqrtf(dot(a, a))); }
inline float owl_max(float a, float b) { return a < b ? b : a; }  // owl::common::max (A.2 item 4)
inline float owl_min(float a, float b) { return a < b ? a : b; }

struct M3 {
    V3 r[3];
};

// ---- cuda_include/lcg_random.cuh -------------------------------------------------------------
uint32_t murmur_mix(uint32_t hash, uint32_t k) {  // :11-28
    k *= 0xcc9e2d51u;
    k = (k << 15) | (k >> 17);
    k *= 0x1b873593u;
    hash ^= k;
    return ((hash << 13) | (hash >> 19)) * 5u + 0xe6546b64u;
}
uint32_t murmur_fin(uint32_t h) {  // :30-39
    h ^= h >> 16;
    h *= 0x85ebca6bu;
    h ^= h >> 13;
    h *= 0xc2b2ae35u;
    h ^= h >> 16;
    return h;
}
struct Rng {
    uint32_t s;
};
uint32_t rng_u32(Rng& r) {  // lcg_random :41-47
    r.s = r.s * 1664525u + 1013904223u;
    return r.s;
}
float rng_f(Rng& r) { return ldexpf((float)rng_u32(r), -32); }  // lcg_randomf :49-52
Rng rng_make(int frame, uint32_t px, uint32_t py, uint32_t W) {  // get_rng :54-62
    Rng r;
    r.s = murmur_mix(0u, px + py * W);
    r.s = murmur_mix(r.s, (uint32_t)frame);
    r.s = murmur_fin(r.s);
    return r;
}

// ---- cuda_include/utils.cuh --------------------------------------------------------------------
V3 apply_mat(const M3& m, V3 v) { return mk(dot(m.r[0], v), dot(m.r[1], v), dot(m.r[2], v)); }  // :70-74

M3 transpose(const M3& m) {  // matrixTranspose :142-155
    M3 t;
    t.r[0] = mk(m.r[0].x, m.r[1].x, m.r[2].x);
    t.r[1] = mk(m.r[0].y, m.r[1].y, m.r[2].y);
    t.r[2] = mk(m.r[0].z, m.r[1].z, m.r[2].z);
    return t;
}

// matrixInverse :77-138 -- Gauss-Jordan with full pivoting (pbrt) on the 3x3
M3 inverse(const M3& m) {
    float a[3][3];
    for (int i = 0; i < 3; i++) {
        a[i][0] = m.r[i].x;
        a[i][1] = m.r[i].y;
        a[i][2] = m.r[i].z;
    }
    int indxc[3], indxr[3], ipiv[3] = {0, 0, 0};
    for (int i = 0; i < 3; i++) {
        int irow = 0, icol = 0;
        float big = 0.f;
        for (int j = 0; j < 3; j++) {
            if (ipiv[j] == 1) continue;
            for (int k = 0; k < 3; k++)
                if (ipiv[k] == 0 && fabsf(a[j][k]) >= big) {
                    big = fabsf(a[j][k]);
                    irow = j;
                    icol = k;
                }
        }
        ++ipiv[icol];
        if (irow != icol)
            for (int k = 0; k < 3; k++) {
                const float t = a[irow][k];
                a[irow][k] = a[icol][k];
                a[icol][k] = t;
            }
        indxr[i] = irow;
        indxc[i] = icol;
        const float pivinv = 1.f / a[icol][icol];
        a[icol][icol] = 1.f;
        for (int j = 0; j < 3; j++) a[icol][j] *= pivinv;
        for (int j = 0; j < 3; j++) {
            if (j == icol) continue;
            const float save = a[j][icol];
            a[j][icol] = 0.f;
            for (int k = 0; k < 3; k++) a[j][k] -= a[icol][k] * save;
        }
    }
    for (int j = 2; j >= 0; j--)
        if (indxr[j] != indxc[j])
            for (int k = 0; k < 3; k++) {
                const float t = a[k][indxr[j]];
                a[k][indxr[j]] = a[k][indxc[j]];
                a[k][indxc[j]] = t;
            }
    M3 o;
    for (int i = 0; i < 3; i++) o.r[i] = mk(a[i][0], a[i][1], a[i][2]);
    return o;
}

// orthonormalBasis :168-190 -- rows (c1, c2, n); `1. / (1. + n.z)` and `1. - n.x*n.x*a` are
// double expressions rounded to float on assignment / construction.
void onb(V3 n, M3& mat, M3& inv) {
    V3 c1, c2;
    if (n.z < -0.999999f) {
        c1 = mk(0.f, -1.f, 0.f);
        c2 = mk(-1.f, 0.f, 0.f);
    } else {
        const float a = (float)(1. / (1. + (double)n.z));
        const float b = -n.x * n.y * a;
        c1 = normalize(mk((float)(1. - (double)(n.x * n.x * a)), b, -n.x));
        c2 = normalize(mk(b, (float)(1. - (double)(n.y * n.y * a)), -n.y));
    }
    mat.r[0] = c1;
    mat.r[1] = c2;
    mat.r[2] = n;
    inv = transpose(mat);
}

V3 sample_point_on_triangle(V3 v1, V3 v2, V3 v3, float u1, float u2) {  // :194-199
    const float su1 = sqrtf(u1);
    return (1 - su1) * v1 + su1 * ((1 - u2) * v2 + u2 * v3);
}

float balance_heuristic(int nf, float fPdf, int ng, float gPdf) {  // :207-209
    return (nf * fPdf) / (nf * fPdf + ng * gPdf);
}

V3 check_positive(V3 c) { return mk(owl_max(c.x, 0.f), owl_max(c.y, 0.f), owl_max(c.z, 0.f)); }  // :219-226

}  // namespace

// ================================================================================================
// scene state (one scene at a time, like the _ref libraries)
// ================================================================================================
namespace {

struct Light {  // TriLight, cuda_include/common.cuh:64-72 (19 floats)
    V3 v1, v2, v3, cg, normal, emit;
    float area;
};

struct Mesh {  // TriangleMeshData, cuda_include/common.cuh:130-151
    std::vector<float> vtx, nrm, uv;
    std::vector<int32_t> idx;
    std::vector<uint8_t> tex;
    int texW = 0, texH = 0;
    V3 diffuse, emit;
    float alpha = 0.f;
    bool isLight = false;
    uint32_t materialID = 0;
};

struct Scene {
    std::vector<Mesh> meshes;
    std::vector<int> primMesh, primLocal;
    ocpu::TriSoup soup;
    bool use_bvh = true;
    std::vector<Light> lights;
    std::vector<float> lut[3];
    int lutW = 0, lutH = 0, lutQuant = 0;
    V3 camPos, d00, du, dv;
    int mode = 0, shadow_spp = 4, max_depth = 1;
} S;

// SurfaceInteraction, cuda_include/common.cuh:163-183 (fields the path reads), with its defaults
struct Surf {
    bool hit = false;
    V3 p = {0, 0, 0};
    float u = 0.f, v = 0.f;
    V3 wo = {0, 0, 0}, wo_local = {0, 0, 0};
    uint32_t materialID = 0;
    V3 n_geom = {0, 0, 0}, n_shad = {0, 0, 0}, diffuse = {0, 0, 0};
    float alpha = 0.f;
    V3 emit = {0, 0, 0};
    bool isLight = false;
    float area = 0.f;
    M3 to_local, to_world;
};

// ---- cuda_include/ltc/ltc_utils.cuh ----------------------------------------------------------
// fetchLtcMat :10-23.  `theta * 0.99f / (0.5 * PI)` is float*float -> double division -> float.
void fetch_ltc(float alpha, float theta, M3& m, float& amplitude) {
    theta = (float)((double)(theta * 0.99f) / (0.5 * kPi));
    float r[3][4];
    for (int i = 0; i < 3; i++)
        ocpu::tex2d_linear_clamp_f4(S.lut[i].data(), S.lutW, S.lutH, theta, alpha, S.lutQuant, r[i]);
    for (int i = 0; i < 3; i++) m.r[i] = mk(r[i][0], r[i][1], r[i][2]);
    amplitude = r[2][3];
}

// integrateEdgeVec / integrateEdge :26-44.  The x <= 0 branch is evaluated in double:
// 0.5 * (1.0f / sqrt(max(1.0f - x*x, 1e-7))) - v, with max(float,double) -> double.
float integrate_edge(V3 v1, V3 v2) {
    const float x = dot(v1, v2);
    const float y = fabsf(x);
    const float a = 0.8543985f + (0.4965155f + 0.0145206f * y) * y;
    const float b = 3.4175940f + (4.1616724f + y) * y;
    const float v = a / b;
    float ts;
    if (x > 0.0f) {
        ts = v;
    } else {
        const double m = fmax((double)(1.0f - x * x), 1e-7);
        ts = (float)(0.5 * (1.0 / sqrt(m)) - (double)v);
    }
    return (cross(v1, v2) * ts).z;
}

// ---- cuda_include/ltc/polygon_utils.cuh ------------------------------------------------------
V3 iz0(V3 lhs, V3 rhs) {  // :7-22 -- intersection with z = 0, re-normalised
    const float f = lhs.z / (lhs.z - rhs.z);
    const float cx = -f * lhs.x + lhs.x, cy = -f * lhs.y + lhs.y;
    const float tx = f * rhs.x + cx, ty = f * rhs.y + cy;
    return normalize(mk(tx, ty, 0.f));
}

// clipPolygon :33-120 -- the 25-case table keyed by vertex_count | (above-horizon bits << 3)
int clip_polygon(int vertex_count, V3 v[5]) {
    if (vertex_count == 0) return 0;
    int mask = vertex_count;
    for (int i = 0; i < 4; i++)
        if (v[i].z > 0.0f && (i < 3 || i < vertex_count)) mask |= 1 << (i + 3);
    int vc = 0;
    switch (mask) {
        case 3: vc = 0; break;
        case 59: vc = 3; v[3] = v[0]; break;
        case 11: vc = 3; v[1] = iz0(v[0], v[1]); v[2] = iz0(v[2], v[0]); v[3] = v[0]; break;
        case 19: vc = 3; v[0] = iz0(v[0], v[1]); v[2] = iz0(v[1], v[2]); v[3] = v[0]; break;
        case 35: vc = 3; v[0] = iz0(v[2], v[0]); v[1] = iz0(v[1], v[2]); v[3] = v[0]; break;
        case 27: vc = 4; v[3] = iz0(v[2], v[0]); v[2] = iz0(v[1], v[2]); v[4] = v[0]; break;
        case 51: vc = 4; v[3] = iz0(v[2], v[0]); v[0] = iz0(v[0], v[1]); v[4] = v[0]; break;
        case 43: vc = 4; v[3] = v[2]; v[2] = iz0(v[1], v[2]); v[1] = iz0(v[0], v[1]); v[4] = v[0]; break;
        case 4: vc = 0; break;
        case 124: vc = 4; v[4] = v[0]; break;
        case 12: vc = 3; v[1] = iz0(v[0], v[1]); v[2] = iz0(v[3], v[0]); v[3] = v[0]; break;
        case 20: vc = 3; v[0] = iz0(v[0], v[1]); v[2] = iz0(v[1], v[2]); v[3] = v[0]; break;
        case 36: vc = 3; v[0] = iz0(v[2], v[3]); v[1] = iz0(v[1], v[2]); v[3] = v[0]; break;
        case 68: vc = 3; v[1] = iz0(v[3], v[0]); v[0] = v[3]; v[2] = iz0(v[2], v[3]); break;
        case 28: vc = 4; v[2] = iz0(v[1], v[2]); v[3] = iz0(v[3], v[0]); v[4] = v[0]; break;
        case 52: vc = 4; v[0] = iz0(v[0], v[1]); v[3] = iz0(v[2], v[3]); v[4] = v[0]; break;
        case 100: vc = 4; v[0] = iz0(v[3], v[0]); v[1] = iz0(v[1], v[2]); v[4] = v[0]; break;
        case 76: vc = 4; v[1] = iz0(v[0], v[1]); v[2] = iz0(v[2], v[3]); v[4] = v[0]; break;
        case 60: vc = 5; v[4] = iz0(v[3], v[0]); v[3] = iz0(v[2], v[3]); break;
        case 116: vc = 5; v[4] = iz0(v[3], v[0]); v[0] = iz0(v[0], v[1]); break;
        case 108: vc = 5; v[4] = v[0]; v[0] = iz0(v[0], v[1]); v[1] = iz0(v[1], v[2]); break;
        case 92: vc = 5; v[4] = v[3]; v[3] = iz0(v[2], v[3]); v[2] = iz0(v[1], v[2]); break;
        default: vc = 0; break;
    }
    return vc;
}

// closed-polygon edge sum used by integrateOverPolygon :82-95,103-125 (abs of the sum; 0 unless
// n is one of the listed counts)
float polygon_sum(const V3* v, int n, int n_max) {
    if (n < 3 || n > n_max) return 0.f;
    float s = integrate_edge(v[0], v[1]);
    for (int i = 1; i < n; i++) s += integrate_edge(v[i], v[(i + 1) % n]);
    return fabsf(s);
}

// integrateOverPolygon :47-127
V3 integrate_over_polygon(const Surf& si, const M3& ltc_inv, float amplitude, const M3& iso, const Light& L) {
    V3 a = normalize(L.v1 - si.p), b = normalize(L.v2 - si.p), c = normalize(L.v3 - si.p);
    const V3 cg = normalize(a + b + c);
    if (dot(-cg, L.normal) < 0.f) return mk(0.f, 0.f, 0.f);
    a = normalize(apply_mat(si.to_local, a));
    b = normalize(apply_mat(si.to_local, b));
    c = normalize(apply_mat(si.to_local, c));
    a = normalize(apply_mat(iso, a));
    b = normalize(apply_mat(iso, b));
    c = normalize(apply_mat(iso, c));
    V3 dv[5] = {a, b, c, a, a};
    const int dn = clip_polygon(3, dv);
    const float diffuse_shading = polygon_sum(dv, dn, 4);  // a diffuse count of 5 cannot occur from 3
    const V3 ta = normalize(apply_mat(ltc_inv, a)), tb = normalize(apply_mat(ltc_inv, b)),
             tc = normalize(apply_mat(ltc_inv, c));
    V3 lv[5] = {ta, tb, tc, ta, ta};
    const int ln = clip_polygon(dn, lv);  // vertex_count = the diffuse clip's result
    const float ggx_shading = polygon_sum(lv, ln, 5);
    return (si.diffuse * L.emit * diffuse_shading) + (amplitude * L.emit * ggx_shading);
}

// ltcDirectLighingBaseline, cuda_src/deviceCode.cu:24-56
V3 ltc_direct(const Surf& si) {
    const V3 wo_local = normalize(apply_mat(si.to_local, si.wo));
    if (wo_local.z < 0.f) return mk(0.f, 0.f, 0.f);
    const float theta = acosf(wo_local.z);  // sphericalTheta, utils.cuh:202-204
    M3 ltc_mat;
    float amplitude = 1.f;
    fetch_ltc(si.alpha, theta, ltc_mat, amplitude);
    const M3 ltc_inv = inverse(ltc_mat);
    M3 iso;
    iso.r[0] = normalize(mk(wo_local.x, wo_local.y, 0.f));
    iso.r[2] = mk(0.f, 0.f, 1.f);
    iso.r[1] = normalize(cross(iso.r[2], iso.r[0]));
    V3 color = mk(0.f, 0.f, 0.f);
    for (size_t l = 0; l < S.lights.size(); l++)
        color = color + integrate_over_polygon(si, ltc_inv, amplitude, iso, S.lights[l]);
    return color;
}

// Classification aid for the parity tests (not a reference function): how close the pixel's LTC evaluation comes
// to one of its SIGN DECISIONS -- wo_local.z < 0 (deviceCode.cu:27), the back-face test (ltc_utils.cuh:58-60), the
// above-horizon bit of every vertex fed to the two clips (polygon_utils.cuh:27-31).  A pixel whose margin is below
// ~1e-4 takes, or sits next to, another branch of the 25-case table when a direction differs in the last ulps
// (libdevice vs glibc, fma contraction, the texture unit's weights): a "case flip" pixel in north_star's terms.
float ltc_decision_margin(const Surf& si) {
    const V3 wo_local = normalize(apply_mat(si.to_local, si.wo));
    float m = fabsf(wo_local.z);
    if (wo_local.z < 0.f) return m;
    const float theta = acosf(wo_local.z);
    M3 ltc_mat;
    float amplitude = 1.f;
    fetch_ltc(si.alpha, theta, ltc_mat, amplitude);
    const M3 ltc_inv = inverse(ltc_mat);
    M3 iso;
    iso.r[0] = normalize(mk(wo_local.x, wo_local.y, 0.f));
    iso.r[2] = mk(0.f, 0.f, 1.f);
    iso.r[1] = normalize(cross(iso.r[2], iso.r[0]));
    for (size_t l = 0; l < S.lights.size(); l++) {
        const Light& L = S.lights[l];
        V3 a = normalize(L.v1 - si.p), b = normalize(L.v2 - si.p), c = normalize(L.v3 - si.p);
        const float bf = dot(-normalize(a + b + c), L.normal);
        m = fminf(m, fabsf(bf));
        if (bf < 0.f) continue;
        a = normalize(apply_mat(iso, normalize(apply_mat(si.to_local, a))));
        b = normalize(apply_mat(iso, normalize(apply_mat(si.to_local, b))));
        c = normalize(apply_mat(iso, normalize(apply_mat(si.to_local, c))));
        m = fminf(m, fminf(fabsf(a.z), fminf(fabsf(b.z), fabsf(c.z))));
        if (!(a.z > 0.f || b.z > 0.f || c.z > 0.f)) continue;  // the diffuse clip returned 0: the second clip does nothing
        const V3 ta = normalize(apply_mat(ltc_inv, a)), tb = normalize(apply_mat(ltc_inv, b)),
                 tc = normalize(apply_mat(ltc_inv, c));
        m = fminf(m, fminf(fabsf(ta.z), fminf(fabsf(tb.z), fabsf(tc.z))));
    }
    return m;
}

}  // namespace

// ================================================================================================
// cuda_include/frostbite.cuh (the reachable subset) and cuda_include/material.cuh
// ================================================================================================
namespace {

inline float cos_theta(V3 w) { return w.z; }                                    // frostbite.cuh:19
inline float cos_theta2(V3 w) { return w.z * w.z; }                             // :20
inline float sin_theta2(V3 w) { return fmaxf(0.0f, 1.0f - cos_theta2(w)); }     // :21
inline float tan_theta2(V3 w) { return sin_theta2(w) / cos_theta2(w); }         // :24
inline bool same_hemisphere(V3 w, V3 wp) { return w.z * wp.z > 0.0f; }          // :26-28

// clampDot :14-16 (zero = false at every live call site)
float clamp_dot(V3 a, V3 b, bool zero) { return fmaxf(dot(a, b), zero ? 0.f : kEps); }

V3 fr_schlick(float cosThetaI, V3 f0) {  // Fr_Schlick :36-41
    const float a = fmaxf(0.0f, 1.0f - cosThetaI);
    const float a2 = a * a;
    const float a5 = a2 * a2 * a;
    return f0 + (mk(1.f, 1.f, 1.f) - f0) * a5;
}

float d_ggx(V3 wh, float alpha) {  // D_GGX :43-47
    const float alpha2 = alpha * alpha;
    const float a = 1.0f + cos_theta2(wh) * (alpha2 - 1.0f);
    return alpha2 / ((float)kPi * a * a);
}

float g1_smith_ggx(V3 w, float alpha) {  // G1_Smith_GGX :49-56
    const float t2 = tan_theta2(w);
    if ((double)t2 > 1e5) return 0.0f;
    const float alpha2 = alpha * alpha;
    const float lambda = (-1.0f + sqrtf(alpha2 * t2 + 1.0f)) / 2.0f;
    return 1.0f / (1.0f + lambda);
}

// G2_SmithHeightCorrelated_GGX :62-72 -- returns 0 unless BOTH tan^2 <= 1e-5
float g2_height_correlated(V3 wi, V3 wo, float alpha) {
    const float tO = tan_theta2(wo), tI = tan_theta2(wi);
    if ((double)tO > 1e-5 || (double)tI > 1e-5) return 0.f;
    const float alpha2 = alpha * alpha;
    const float lo = (-1.0f + sqrtf(alpha2 * tO + 1.0f)) / 2.0f;
    const float li = (-1.0f + sqrtf(alpha2 * tI + 1.0f)) / 2.0f;
    return 1.0f / (1.0f + lo + li);
}

V3 diffuse_lambert(V3 wi, V3 wo, V3 colour) {  // diffuse_Lambert :80-86
    if (!same_hemisphere(wi, wo)) return mk(0.f, 0.f, 0.f);
    return colour / (float)kPi;
}

V3 microfacet_reflection_ggx(V3 wi, V3 wo, V3 f0, float eta, float alpha) {  // :88-113
    if (!same_hemisphere(wi, wo) || cos_theta(wi) == 0.0f || cos_theta(wo) == 0.0f) return mk(0.f, 0.f, 0.f);
    V3 wh = wi + wo;
    if (wh.x == 0.0f && wh.y == 0.0f && wh.z == 0.0f) return mk(0.f, 0.f, 0.f);
    wh = normalize(wh);
    V3 F;
    if (eta < 1.0f) {
        const float c = dot(wi, wh);
        F = c * c > 0.0f ? fr_schlick(fabsf(c), f0) : mk(1.f, 1.f, 1.f);
    } else {
        F = fr_schlick(fabsf(dot(wh, wo)), f0);
    }
    const float G = g2_height_correlated(wi, wo, alpha);
    const float D = d_ggx(wh, alpha);
    return F * G * D / (4.0f * fabsf(cos_theta(wi)) * fabsf(cos_theta(wo)));
}

V3 sample_cosine_hemisphere(float u1, float u2) {  // sampleCosineHemisphere :160-165
    const float c = sqrtf(fmaxf(0.0f, 1.0f - u1));
    const float s = sqrtf(u1);
    const float phi = 2.0f * (float)kPi * u2;
    return mk(s * cosf(phi), s * sinf(phi), c);
}

float pdf_cosine_hemisphere(V3 wi, V3 wo) {  // :167-169
    return same_hemisphere(wi, wo) ? cos_theta(wi) / (float)kPi : 0.0f;
}

float pdf_ggx_vndf_reflection(V3 wi, V3 wo, float alpha) {  // pdfGGX_VNDF_reflection :234-243
    if (!same_hemisphere(wi, wo)) return 0.0f;
    const V3 wh = normalize(wi + wo);
    const float pdf_h = g1_smith_ggx(wo, alpha) * d_ggx(wh, alpha) * fabsf(dot(wh, wo)) / fabsf(cos_theta(wo));
    const float dwh_dwi = 1.0f / (4.0f * dot(wi, wh));
    return pdf_h * dwh_dwi;
}

// material.cuh evaluate :7-22 -- 0.5 Lambert + 0.5 GGX with f0 = base colour, eta = 0 and
// alpha^2 handed down as the roughness (D_GGX squares it again)
V3 evaluate(V3 wi, V3 wo, V3 base, float alpha) {
    const float alpha2 = alpha * alpha;
    const V3 d = diffuse_lambert(wi, wo, base);
    const V3 s = microfacet_reflection_ggx(wi, wo, base, 0.f, alpha2);
    return 0.5f * d + 0.5f * s;
}

// computeLobeProbabilities :26-41
void lobe_probabilities(V3 base, float& pDiffuse, float& pSpecular) {
    pDiffuse = fmaxf(base.x, fmaxf(base.y, base.z)) * 0.5f;
    pSpecular = fmaxf(base.x, fmaxf(base.y, base.z));
    const float norm = 1.0f / (pDiffuse + pSpecular);
    pDiffuse *= norm;
    pSpecular *= norm;
}

float remap(float value, float low1, float high1, float low2, float high2) {  // :44-47 (helper_math clamp)
    const float r = low2 + (value - low1) * (high2 - low2) / (high1 - low1);
    return fmaxf(low2, fminf(r, high2));
}

// sample_direction :50-91.  pDiffuse is forced to 1 (always the cosine lobe, Q17) but the pdf
// still adds pSpecular * pdfVNDF.  A draw u1 == 1.0f (possible, Q21) falls to the specular
// branch of the reference; with pDiffuse = 1 that branch is the `else` below.
V3 sample_direction(V3 wo, float u1, float u2, float* pdf_out, V3 base, float alpha);

}  // namespace

namespace {

V3 sample_ggx_vndf(V3 wo, float alpha, float u1, float u2) {  // sampleGGX_VNDF, frostbite.cuh:208-232
    const V3 h = normalize(mk(alpha * wo.x, alpha * wo.y, wo.z));
    const float l2 = h.x * h.x + h.y * h.y;
    const V3 b1 = l2 > 0.0f ? mk(-h.y, h.x, 0.0f) * (1.0f / sqrtf(l2)) : mk(1.0f, 0.0f, 0.0f);
    const V3 b2 = cross(h, b1);
    const float r = sqrtf(u1);
    const float phi = 2.0f * (float)kPi * u2;
    const float t1 = r * cosf(phi);
    float t2 = r * sinf(phi);
    const float s = 0.5f * (1.0f + h.z);
    t2 = (1.0f - s) * sqrtf(1.0f - t1 * t1) + s * t2;
    const V3 wh = t1 * b1 + t2 * b2 + sqrtf(fmaxf(0.0f, 1.0f - t1 * t1 - t2 * t2)) * h;
    return normalize(mk(alpha * wh.x, alpha * wh.y, fmaxf(0.0f, wh.z)));
}

V3 sample_direction(V3 wo, float u1, float u2, float* pdf_out, V3 base, float alpha) {
    float pDiffuse, pSpecular;
    lobe_probabilities(base, pDiffuse, pSpecular);
    pDiffuse = 1.f;  // material.cuh:57
    V3 wi;
    if (u1 < pDiffuse) {
        u1 = remap(u1, 0.0f, pDiffuse - kEps, 0.0f, 1 - kEps);
        wi = copysignf(1.f, cos_theta(wo)) * sample_cosine_hemisphere(u1, u2);
        wi = normalize(wi);
    } else {
        u1 = remap(u1, pDiffuse, pDiffuse + pSpecular - kEps, 0.0f, 1 - kEps);
        const float sg = (float)copysign(1., (double)cos_theta(wo));
        const V3 wo_upper = sg * wo;
        const V3 wh = sg * sample_ggx_vndf(wo_upper, alpha, u1, u2);
        if (dot(wo, wh) < 0.0f) return mk(0.f, 0.f, 0.f);
        wi = 2.0f * wh * dot(wh, wo) - wh;  // as written (material.cuh:78)
        if (!same_hemisphere(wi, wo)) return mk(0.f, 0.f, 0.f);
    }
    if (pdf_out) *pdf_out = pDiffuse * pdf_cosine_hemisphere(wi, wo) + pSpecular * pdf_ggx_vndf_reflection(wi, wo, alpha);
    return wi;
}

float material_pdf(V3 wi, V3 wo, V3 base, float alpha) {  // pdf, material.cuh:94-104
    float pDiffuse, pSpecular;
    lobe_probabilities(base, pDiffuse, pSpecular);
    return pDiffuse * pdf_cosine_hemisphere(wi, wo) + pSpecular * pdf_ggx_vndf_reflection(wi, wo, alpha);
}

// ---- cuda_include/path/path.cuh helpers ------------------------------------------------------
float light_pdf_area(int idx) { return 1 / (S.lights[idx].area * (int)S.lights.size()); }  // sampleLightPdf :17-22

float pdf_a2w(float pdf, float dist2, float cosTheta) {  // pdfA2W :26-33
    const float a = fabsf(cosTheta);
    if ((double)a < 1e-8) return 0.0f;
    return pdf * dist2 / a;
}

// ---- owl::traceRay + the two hit programs ----------------------------------------------------
// traceRay (cuda_src/deviceCode.cu:76, ratio/ratio.cuh:50, path/path.cuh:219) -> the stand-in
// closest hit of trace_cpu.h, then __closesthit__triangleMeshCH (cuda_include/hit_miss.cuh:14-50)
// or __miss__miss (:52-63) on `si` (which keeps whatever it held in untouched fields).
struct TraceStats {
    long long rays = 0;
    int first_prim = -1;
    float first_u = 0.f, first_v = 0.f;
    int traces = 0;
};

inline V3 bary3(const float* a, const float* b, const float* c, float u, float v) {  // utils.cuh:10-19
    return (1.f - u - v) * mk(a) + u * mk(b) + v * mk(c);
}

void trace(V3 org, V3 dir, Surf& si, TraceStats& st) {
    const float o[3] = {org.x, org.y, org.z}, d[3] = {dir.x, dir.y, dir.z};
    const ocpu::Hit h = S.use_bvh ? S.soup.closest(o, d, 0.f, 1e30f) : S.soup.brute(o, d, 0.f, 1e30f);
    st.rays++;
    if (st.traces++ == 0) {
        st.first_prim = h.prim;
        st.first_u = h.u;
        st.first_v = h.v;
    }
    if (h.prim < 0) {  // miss: const_color = 0 (include/viewer.hpp:530)
        si.hit = false;
        si.isLight = false;
        si.materialID = 0;
        si.n_geom = mk(0.f, 0.f, 0.f);
        si.diffuse = mk(0.f, 0.f, 0.f);
        return;
    }
    const Mesh& m = S.meshes[S.primMesh[h.prim]];
    const int32_t* ix = &m.idx[3 * (size_t)S.primLocal[h.prim]];
    const float *p0 = &m.vtx[3 * (size_t)ix[0]], *p1 = &m.vtx[3 * (size_t)ix[1]], *p2 = &m.vtx[3 * (size_t)ix[2]];
    si.hit = true;
    si.materialID = m.materialID;
    si.area = 0.5f * length(cross(mk(p0) - mk(p1), mk(p2) - mk(p1)));
    si.p = bary3(p0, p1, p2, h.u, h.v);
    const float *t0 = &m.uv[2 * (size_t)ix[0]], *t1 = &m.uv[2 * (size_t)ix[1]], *t2 = &m.uv[2 * (size_t)ix[2]];
    const float w = 1.f - h.u - h.v;
    const float tu = w * t0[0] + h.u * t1[0] + h.v * t2[0], tv = w * t0[1] + h.u * t1[1] + h.v * t2[1];
    si.u = fabsf(fmodf(tu, 1.));
    si.v = fabsf(fmodf(tv, 1.));
    si.n_geom = normalize(bary3(&m.nrm[3 * (size_t)ix[0]], &m.nrm[3 * (size_t)ix[1]], &m.nrm[3 * (size_t)ix[2]], h.u, h.v));
    si.diffuse = m.diffuse;
    if (!m.tex.empty()) {
        float c[4];
        ocpu::tex2d_nearest_clamp_rgba8(m.tex.data(), m.texW, m.texH, si.u, si.v, c);
        si.diffuse = mk(c[0], c[1], c[2]);
    }
    si.alpha = owl_max(0.01f, owl_min(m.alpha, 1.f));  // owl::clamp(alpha, 0.01, 1)
    si.emit = m.emit;
    si.isLight = m.isLight;
}

// ---- ratio_based_shadow, cuda_include/ratio/ratio.cuh:11-126 with max_ray_depth = 1 -----------
void ratio_based_shadow(const Surf& si, Rng& rng, TraceStats& st, V3& shadowed, V3& unshadowed) {
    V3 color = mk(0.f, 0.f, 0.f), sto = mk(0.f, 0.f, 0.f);
    const V3 tp = mk(1.f, 1.f, 1.f);
    if (si.hit) {
        if (si.isLight) {
            color = color + tp * si.emit;
            sto = si.emit * tp;
        } else {
            const float r1x = rng_f(rng), r1y = rng_f(rng);
            rng_f(rng);  // rand2: drawn, unused
            rng_f(rng);
            const int n = (int)S.lights.size();
            const int idx = (int)(rng_f(rng) * (n - 1));
            const float lightPdf = light_pdf_area(idx);
            const Light& lt = S.lights[idx];
            const V3 newPos = sample_point_on_triangle(lt.v1, lt.v2, lt.v3, r1x, r1y);  // sampleLight, path.cuh:7-14
            const V3 L = normalize(newPos - si.p);
            float dist = length(newPos - si.p);
            dist = dist * dist;
            Surf light_si;
            trace(si.p + si.n_geom * float(1e-3), L, light_si, st);
            const float lightPdfW = pdf_a2w(lightPdf, dist, dot(-L, light_si.n_geom));
            const float brdfPdf = 0.f;
            V3 brdf = mk(0.f, 0.f, 0.f);  // committed source, ratio.cuh:58
            if (S.mode == 1) {            // patch P1: the evaluate() the comment on that line stands for
                const V3 wi_local = normalize(apply_mat(si.to_local, L));
                brdf = evaluate(wi_local, si.wo_local, si.diffuse, si.alpha);
            }
            const float misW = balance_heuristic(1, lightPdfW, 1, brdfPdf);
            if (light_si.isLight) {
                color = color + misW * light_si.emit * tp * brdf * clamp_dot(si.n_geom, L, false) / lightPdfW;
                color = check_positive(color);
            }
            sto = misW * lt.emit * tp * brdf * clamp_dot(si.n_geom, L, false) / lightPdfW;
            sto = check_positive(sto);
        }
    }
    shadowed = check_positive(color);
    unshadowed = check_positive(sto);
}

// ---- estimatePathTracing, cuda_include/path/path.cuh:145-258 ----------------------------------
V3 estimate_path_tracing(const Surf& si, Rng& rng, TraceStats& st, int max_depth) {
    V3 color = mk(0.f, 0.f, 0.f);
    Surf cur = si;
    V3 tp = mk(1.f, 1.f, 1.f);
    V3 V = cur.wo;
    for (int depth = 0; depth < max_depth; depth++) {
        if (!cur.hit) break;
        if (cur.isLight) {
            color = color + tp * cur.emit;
            break;
        }
        onb(cur.n_geom, cur.to_world, cur.to_local);  // outputs swapped w.r.t. rayGen (Q16)
        const V3 wo_local = normalize(apply_mat(cur.to_local, V));
        rng_f(rng);  // rand1: drawn, unused
        rng_f(rng);
        const float r2x = rng_f(rng), r2y = rng_f(rng);
        rng_f(rng);  // light selection draw, unused
        const float cur_alpha = cur.alpha;
        float bsdfPdf = 0.f;
        const V3 wi_local = sample_direction(wo_local, r2x, r2y, &bsdfPdf, cur.diffuse, cur_alpha);
        const V3 org = cur.p + kEps * cur.n_geom;
        const V3 dir = normalize(apply_mat(cur.to_world, wi_local));
        if (S.mode == 0) {  // committed debug stub :216-217
            color = wi_local;
            continue;
        }
        Surf nxt;
        nxt.materialID = 0;  // brdf_si{0}
        trace(org, dir, nxt, st);
        if (!nxt.hit) break;
        const V3 brdf = evaluate(wi_local, wo_local, cur.diffuse, cur_alpha);
        if (nxt.isLight) {
            const float lightPdf = 1 / (nxt.area * (int)S.lights.size());
            float dist = length(nxt.p - cur.p);
            dist *= dist;
            const float lightPdfW = pdf_a2w(lightPdf, dist, cos_theta(wi_local));
            const float misW = balance_heuristic(1, bsdfPdf, 1, lightPdfW);
            color = color + misW * nxt.emit * tp / bsdfPdf;
            break;
        }
        tp = tp * (cos_theta(wi_local) * brdf / bsdfPdf);
        V = normalize(cur.p - nxt.p);
        cur = nxt;
        cur.alpha = 0.f;
        cur.wo = V;
    }
    return mk(owl_max(color.x, kEps), owl_max(color.y, kEps), owl_max(color.z, kEps));
}

uint32_t make_rgba(V3 c) {  // owl::make_rgba; NaN -> 0 like cvt.rzi
    auto q = [](float f) -> uint32_t {
        if (f != f) return 0u;
        const float s = f * 256.f;
        if (s >= 2147483648.f) return 255u;
        if (s <= -2147483648.f) return 0u;
        const int i = (int)s;
        return (uint32_t)(i < 0 ? 0 : (i > 255 ? 255 : i));
    };
    return q(c.x) | (q(c.y) << 8) | (q(c.z) << 16) | (0xffu << 24);
}

}  // namespace

// ================================================================================================
// rayGen, cuda_src/deviceCode.cu:59-175 -- one pixel
// ================================================================================================
namespace {

inline void put4(float* buf, size_t o, float x, float y, float z, float w) {
    if (!buf) return;
    buf[4 * o] = x;
    buf[4 * o + 1] = y;
    buf[4 * o + 2] = z;
    buf[4 * o + 3] = w;
}

void ray_gen(int x, int y, int W, int H, int rendererType, int accumId, const OraBuffers* b, TraceStats& st) {
    const size_t o = (size_t)x + (size_t)W * y;
    Rng rng = rng_make(accumId + 10007, (uint32_t)x, (uint32_t)y, (uint32_t)W);
    const float jx = rng_f(rng), jy = rng_f(rng);  // left to right
    const float su = ((float)x + jx) / (float)W, sv = ((float)y + jy) / (float)H;
    const V3 org = S.camPos;
    const V3 dir = normalize(S.d00 + su * S.du + sv * S.dv);
    Surf si;
    trace(org, dir, si, st);
    si.wo = normalize(org - si.p);
    onb(si.n_geom, si.to_local, si.to_world);
    si.wo_local = normalize(apply_mat(si.to_local, si.wo));

    V3 color = mk(0.f, 0.f, 0.f);
    if (!si.hit) color = si.n_geom;
    else if (rendererType == 5) color = mk(1.f, 1.f, 1.f);                      // MASK
    else if (rendererType == 4) color = si.p;                                   // POSITION
    else if (rendererType == 0) color = si.diffuse;                             // DIFFUSE
    else if (rendererType == 1) color = mk(si.alpha, si.alpha, si.alpha);       // ALPHA
    else if (rendererType == 2) color = si.n_geom;                              // NORMALS
    else if (rendererType == 3) color = si.n_shad;                              // SHADE_NORMALS (never written)
    else if (rendererType == 6) color = mk((float)si.materialID, (float)si.materialID, (float)si.materialID);
    else if (rendererType == 7) {                                               // LTC_BASELINE
        color = si.isLight ? si.emit : ltc_direct(si);
    } else if (rendererType == 8) {                                             // RATIO
        if (si.isLight) {
            color = si.emit;
            put4(b->ltc, o, color.x, color.y, color.z, 1.f);
            put4(b->sto_direct, o, color.x, color.y, color.z, 1.f);
            put4(b->sto_no_vis, o, color.x, color.y, color.z, 1.f);
        } else {
            const V3 ltc = ltc_direct(si);
            const int n = S.mode == 1 ? S.shadow_spp : 4;  // deviceCode.cu:128 / patch P3
            V3 c0 = mk(0.f, 0.f, 0.f), c1 = mk(0.f, 0.f, 0.f);
            for (int i = 0; i < n; i++) {
                V3 a, u;
                ratio_based_shadow(si, rng, st, a, u);
                c0 = c0 + a;
                c1 = c1 + u;
            }
            c0 = c0 / (float)n;
            c1 = c1 / (float)n;
            color = ltc;
            put4(b->ltc, o, ltc.x, ltc.y, ltc.z, 1.f);
            const float gd = (c0.x + c0.y + c0.z) / 3.f, gn = (c1.x + c1.y + c1.z) / 3.f;
            put4(b->sto_direct, o, gd, gd, gd, 1.f);
            put4(b->sto_no_vis, o, gn, gn, gn, 1.f);
        }
    } else if (rendererType == 9) {                                             // PATH
        color = si.isLight ? si.emit : estimate_path_tracing(si, rng, st, S.mode == 1 ? S.max_depth : 1);
    } else {
        color = mk(1.f, 0.f, 0.f);
    }
    if (accumId > 0 && b->accum) color = color + mk(b->accum + 4 * o);
    put4(b->accum, o, color.x, color.y, color.z, 1.f);
    put4(b->position, o, si.p.x, si.p.y, si.p.z, 1.f);
    put4(b->normal, o, si.n_geom.x, si.n_geom.y, si.n_geom.z, 1.f);
    put4(b->albedo, o, si.diffuse.x, si.diffuse.y, si.diffuse.z, 1.f);
    put4(b->alpha, o, si.alpha, 0.f, 0.f, 1.f);
    put4(b->uv, o, si.u, si.v, si.diffuse.z, 1.f);
    put4(b->materialID, o, (float)si.materialID, 0.f, 0.f, 1.f);
    color = (1.f / (accumId + 1.f)) * color;
    if (b->rgba8) b->rgba8[o] = make_rgba(color);
}

template <class F>
void parallel_rows(int y0, int y1, int nthreads, F&& body) {
    if (nthreads < 1) nthreads = 1;
    std::atomic<int> next(y0);
    auto worker = [&]() {
        for (;;) {
            const int y = next.fetch_add(1);
            if (y >= y1) break;
            body(y);
        }
    };
    if (nthreads == 1) {
        worker();
        return;
    }
    std::vector<std::thread> th;
    for (int i = 0; i < nthreads; i++) th.emplace_back(worker);
    for (auto& t : th) t.join();
}

}  // namespace

// ================================================================================================
// C interface (oracle_api.h)
// ================================================================================================
extern "C" {

const char* ora_kind(void) { return "restatement"; }

void ora_reset(void) { S = Scene(); }

int ora_add_mesh(const float* vtx, const float* nrm, const float* uv, int nVtx, const int32_t* idx, int nTri,
                 const float* diffuse3, float alpha, const float* emit3, int isLight, uint32_t materialID,
                 const uint8_t* tex, int texW, int texH) {
    S.meshes.emplace_back();
    Mesh& m = S.meshes.back();
    m.vtx.assign(vtx, vtx + 3 * (size_t)nVtx);
    m.nrm.assign(nrm, nrm + 3 * (size_t)nVtx);
    m.uv.assign(uv, uv + 2 * (size_t)nVtx);
    m.idx.assign(idx, idx + 3 * (size_t)nTri);
    if (tex) {
        m.tex.assign(tex, tex + 4 * (size_t)texW * texH);
        m.texW = texW;
        m.texH = texH;
    }
    m.diffuse = mk(diffuse3);
    m.emit = mk(emit3);
    m.alpha = alpha;
    m.isLight = isLight != 0;
    m.materialID = materialID;
    const int mi = (int)S.meshes.size() - 1;
    for (int p = 0; p < nTri; p++) {  // global prim = triangles of earlier meshes + local index
        S.primMesh.push_back(mi);
        S.primLocal.push_back(p);
        for (int k = 0; k < 3; k++)
            for (int a = 0; a < 3; a++) S.soup.v.push_back(vtx[3 * (size_t)idx[3 * p + k] + a]);
    }
    return mi;
}

void ora_set_lights(const float* t, int n) {
    S.lights.resize(n);
    for (int i = 0; i < n; i++) {
        const float* p = t + 19 * (size_t)i;
        Light& l = S.lights[i];
        l.v1 = mk(p); l.v2 = mk(p + 3); l.v3 = mk(p + 6); l.cg = mk(p + 9); l.normal = mk(p + 12); l.emit = mk(p + 15);
        l.area = p[18];
    }
}

void ora_set_ltc_lut(const float* m1, const float* m2, const float* m3, int w, int h, int quantise) {
    const float* src[3] = {m1, m2, m3};
    for (int i = 0; i < 3; i++) S.lut[i].assign(src[i], src[i] + 4 * (size_t)w * h);
    S.lutW = w;
    S.lutH = h;
    S.lutQuant = quantise;
}

void ora_commit(int use_bvh) {
    S.use_bvh = use_bvh != 0;
    if (S.use_bvh) S.soup.build();
}

void ora_set_camera(const float* pos, const float* d00, const float* du, const float* dv) {
    S.camPos = mk(pos);
    S.d00 = mk(d00);
    S.du = mk(du);
    S.dv = mk(dv);
}

int ora_set_options(int mode, int shadow_spp, int max_depth) {
    if (mode != 0 && mode != 1) return -1;
    if (mode == 0 && (shadow_spp != 4 || max_depth != 1)) return -1;  // literal fixes both (deviceCode.cu:128,151)
    S.mode = mode;
    S.shadow_spp = shadow_spp;
    S.max_depth = max_depth;
    return 0;
}

long long ora_render(int W, int H, int rendererType, int accumId, int y0, int y1, const OraBuffers* b, int nthreads) {
    std::atomic<long long> rays(0);
    parallel_rows(y0, y1, nthreads, [&](int y) {
        TraceStats row;
        for (int x = 0; x < W; x++) {
            TraceStats st;
            ray_gen(x, y, W, H, rendererType, accumId, b, st);
            const size_t o = (size_t)y * W + x;
            if (b->hit_prim) b->hit_prim[o] = st.first_prim;
            if (b->hit_uv) {
                b->hit_uv[2 * o] = st.first_u;
                b->hit_uv[2 * o + 1] = st.first_v;
            }
            row.rays += st.rays;
        }
        rays += row.rays;
    });
    return rays.load();
}

void ora_ltc_gbuffer(int n, const float* position, const float* normal, const float* albedo, const float* alpha,
                     const float* cam_pos, float* out, int nthreads) {
    const int chunk = 1024;
    parallel_rows(0, (n + chunk - 1) / chunk, nthreads, [&](int c) {
        const int e = std::min(n, (c + 1) * chunk);
        for (int i = c * chunk; i < e; i++) {
            const V3 nrm = mk(normal + 4 * (size_t)i);
            V3 col = mk(0.f, 0.f, 0.f);
            if (!(nrm.x == 0.f && nrm.y == 0.f && nrm.z == 0.f)) {
                Surf si;  // what rayGen sets up before the mode switch, deviceCode.cu:78-85
                si.hit = true;
                si.p = mk(position + 4 * (size_t)i);
                si.n_geom = nrm;
                si.diffuse = mk(albedo + 4 * (size_t)i);
                si.alpha = alpha[4 * (size_t)i];
                si.wo = normalize(mk(cam_pos) - si.p);
                onb(si.n_geom, si.to_local, si.to_world);
                si.wo_local = normalize(apply_mat(si.to_local, si.wo));
                col = ltc_direct(si);
            }
            put4(out, (size_t)i, col.x, col.y, col.z, 1.f);
        }
    });
}

// ltc_decision_margin() over a saved G-buffer of n pixels; out[n] floats, +inf for miss pixels (normal == 0).
void ora_ltc_margin(int n, const float* position, const float* normal, const float* alpha, const float* cam_pos,
                    float* out, int nthreads) {
    const int chunk = 1024;
    parallel_rows(0, (n + chunk - 1) / chunk, nthreads, [&](int c) {
        const int e = std::min(n, (c + 1) * chunk);
        for (int i = c * chunk; i < e; i++) {
            const V3 nrm = mk(normal + 4 * (size_t)i);
            out[i] = INFINITY;
            if (nrm.x == 0.f && nrm.y == 0.f && nrm.z == 0.f) continue;
            Surf si;
            si.hit = true;
            si.p = mk(position + 4 * (size_t)i);
            si.n_geom = nrm;
            si.alpha = alpha[4 * (size_t)i];
            si.wo = normalize(mk(cam_pos) - si.p);
            onb(si.n_geom, si.to_local, si.to_world);
            out[i] = ltc_decision_margin(si);
        }
    });
}

// ---- units ----
void ora_u_rng(int frame_id, uint32_t px, uint32_t py, uint32_t W, uint32_t H, int n, uint32_t* state0, uint32_t* raw,
               float* flt) {
    (void)H;
    Rng r = rng_make(frame_id, px, py, W);
    *state0 = r.s;
    for (int i = 0; i < n; i++) raw[i] = rng_u32(r);
    for (int i = 0; i < n; i++) flt[i] = rng_f(r);
}

int ora_u_clip(int vertex_count, float* v15) {
    V3 v[5];
    for (int i = 0; i < 5; i++) v[i] = mk(v15 + 3 * i);
    const int vc = clip_polygon(vertex_count, v);
    for (int i = 0; i < 5; i++) {
        v15[3 * i] = v[i].x;
        v15[3 * i + 1] = v[i].y;
        v15[3 * i + 2] = v[i].z;
    }
    return vc;
}

float ora_u_integrate_edge(const float* a, const float* b) { return integrate_edge(mk(a), mk(b)); }

static void store9(const M3& m, float* o) {
    for (int i = 0; i < 3; i++) {
        o[3 * i] = m.r[i].x;
        o[3 * i + 1] = m.r[i].y;
        o[3 * i + 2] = m.r[i].z;
    }
}

void ora_u_fetch_ltc(float alpha, float theta, float* m9_amp) {
    M3 m;
    float amp = 1.f;
    fetch_ltc(alpha, theta, m, amp);
    store9(m, m9_amp);
    m9_amp[9] = amp;
}

void ora_u_mat_inverse(const float* m9, float* out9) {
    M3 m;
    for (int i = 0; i < 3; i++) m.r[i] = mk(m9 + 3 * i);
    store9(inverse(m), out9);
}

void ora_u_onb(const float* n, float* mat9_inv9) {
    M3 m, inv;
    onb(mk(n), m, inv);
    store9(m, mat9_inv9);
    store9(inv, mat9_inv9 + 9);
}

void ora_u_sample_direction(const float* wo, float u1, float u2, const float* base, float metal, float alpha,
                            float* wi_pdf4) {
    (void)metal;
    float p = 0.f;
    const V3 wi = sample_direction(mk(wo), u1, u2, &p, mk(base), alpha);
    wi_pdf4[0] = wi.x;
    wi_pdf4[1] = wi.y;
    wi_pdf4[2] = wi.z;
    wi_pdf4[3] = p;
}

void ora_u_evaluate(const float* wi, const float* wo, const float* base, float alpha, float* out3) {
    const V3 r = evaluate(mk(wi), mk(wo), mk(base), alpha);
    out3[0] = r.x;
    out3[1] = r.y;
    out3[2] = r.z;
}

float ora_u_pdf(const float* wi, const float* wo, const float* base, float metal, float alpha) {
    (void)metal;
    return material_pdf(mk(wi), mk(wo), mk(base), alpha);
}

void ora_u_sample_tri(const float* v1, const float* v2, const float* v3, float u1, float u2, float* out3) {
    const V3 r = sample_point_on_triangle(mk(v1), mk(v2), mk(v3), u1, u2);
    out3[0] = r.x;
    out3[1] = r.y;
    out3[2] = r.z;
}

float ora_u_pdf_a2w(float pdf, float dist2, float cos_theta_) { return pdf_a2w(pdf, dist2, cos_theta_); }

uint32_t ora_u_make_rgba(const float* c3) { return make_rgba(mk(c3)); }

// ltc_ratio_estimator.py:8  colors = ltc * (stoDirect / stoNoVis)  (numpy float32, element-wise)
void ora_ratio_combine(long long n, const float* ltc, const float* sto_direct, const float* sto_no_vis, float* out) {
    for (long long i = 0; i < n; i++) out[i] = ltc[i] * (sto_direct[i] / sto_no_vis[i]);
}

}  // extern "C"

// ================================================================================================
// Denoise -- DEFINITION, not a restatement: the reference has no denoiser
// (cuda_include/denoise/bilateral.cuh is a 0-byte file; ltc_ratio_estimator.py:5-6 reads
// stodirectDenoise.exr / stoNoVisDenoise.exr that nothing in the repository writes).
// Edge-avoiding a-trous wavelet filter over the two grey sto_* images:
//   value  v(q) = (sto_direct.x, sto_no_vis.x)
//   guide  g(q) = (n.xyz, depth = |p - cam|); pixels with materialID == 0 or n == 0 (lights,
//                 misses) are passed through unfiltered and never used as a tap
//   weight w(q) = B3(dx) B3(dy) * max(0, n_c . n_q)^sigma_n (sigma_n <= 0: factor 1) * exp(-|d_c - d_q| / (sigma_p * step))
//                 (* exp(-|v_c - v_q| / sigma_c) per image when sigma_c > 0); NaN taps skipped
//   passes step = 1, 2, 4, ... (iterations of them), 5x5 taps, B3 = (1/16, 1/4, 3/8, 1/4, 1/16)
// ================================================================================================
extern "C" void ora_denoise(int W, int H, const float* normal, const float* position, const float* materialID,
                            const float* sto_direct, const float* sto_no_vis, const float* cam_pos, int iterations,
                            float sigma_n, float sigma_p, float sigma_c, float* out_direct, float* out_no_vis,
                            int nthreads) {
    const size_t n = (size_t)W * H;
    std::vector<float> guide(4 * n), a(2 * n), b(2 * n);
    for (size_t i = 0; i < n; i++) {
        const V3 nn = mk(normal + 4 * i);
        const bool skip = (nn.x == 0.f && nn.y == 0.f && nn.z == 0.f) || materialID[4 * i] == 0.f;
        guide[4 * i] = nn.x;
        guide[4 * i + 1] = nn.y;
        guide[4 * i + 2] = nn.z;
        guide[4 * i + 3] = skip ? -1.f : length(mk(position + 4 * i) - mk(cam_pos));
        a[2 * i] = sto_direct[4 * i];
        a[2 * i + 1] = sto_no_vis[4 * i];
    }
    static const float b3[5] = {1.f / 16.f, 1.f / 4.f, 3.f / 8.f, 1.f / 4.f, 1.f / 16.f};
    float* vin = a.data();
    float* vout = b.data();
    for (int it = 0; it < iterations; it++) {
        const int step = 1 << it;
        const float isp = 1.f / (sigma_p * (float)step);
        const float isc = sigma_c > 0.f ? 1.f / sigma_c : 0.f;
        parallel_rows(0, H, nthreads, [&](int y) {
            for (int x = 0; x < W; x++) {
                const size_t o = (size_t)y * W + x;
                const float* gc = &guide[4 * o];
                float r0 = vin[2 * o], r1 = vin[2 * o + 1];
                if (gc[3] >= 0.f) {
                    float sd = 0.f, sn = 0.f, wd = 0.f, wn = 0.f;
                    for (int j = -2; j <= 2; j++) {
                        const int yy = y + j * step;
                        if (yy < 0 || yy >= H) continue;
                        for (int i = -2; i <= 2; i++) {
                            const int xx = x + i * step;
                            if (xx < 0 || xx >= W) continue;
                            const size_t q = (size_t)yy * W + xx;
                            const float* g = &guide[4 * q];
                            if (g[3] < 0.f) continue;
                            const float v0 = vin[2 * q], v1 = vin[2 * q + 1];
                            const float nd = fmaxf(0.f, gc[0] * g[0] + gc[1] * g[1] + gc[2] * g[2]);
                            const float w = b3[i + 2] * b3[j + 2] * (sigma_n > 0.f ? powf(nd, sigma_n) : 1.f) * expf(-fabsf(gc[3] - g[3]) * isp);
                            float w0 = w, w1 = w;
                            if (isc > 0.f) {
                                w0 *= expf(-fabsf(r0 - v0) * isc);
                                w1 *= expf(-fabsf(r1 - v1) * isc);
                            }
                            if (v0 == v0 && w0 == w0) { sd += w0 * v0; wd += w0; }
                            if (v1 == v1 && w1 == w1) { sn += w1 * v1; wn += w1; }
                        }
                    }
                    if (wd > 0.f) r0 = sd / wd;
                    if (wn > 0.f) r1 = sn / wn;
                }
                vout[2 * o] = r0;
                vout[2 * o + 1] = r1;
            }
        });
        std::swap(vin, vout);
    }
    for (size_t i = 0; i < n; i++) {
        if (guide[4 * i + 3] >= 0.f) {
            const float d = vin[2 * i], s = vin[2 * i + 1];
            out_direct[4 * i] = out_direct[4 * i + 1] = out_direct[4 * i + 2] = d;
            out_no_vis[4 * i] = out_no_vis[4 * i + 1] = out_no_vis[4 * i + 2] = s;
            out_direct[4 * i + 3] = out_no_vis[4 * i + 3] = 1.f;
        } else {  // lights / misses: the unfiltered float4 (emit on light pixels, deviceCode.cu:121-123)
            for (int k = 0; k < 4; k++) {
                out_direct[4 * i + k] = sto_direct[4 * i + k];
                out_no_vis[4 * i + k] = sto_no_vis[4 * i + k];
            }
        }
    }
}

// ltc.cuh -- analytic polygonal area-light shading with linearly transformed cosines.
//
// Device-side restatement (own code, per-thread, everything in registers) of
//   cuda_src/deviceCode.cu:24-56            ltcDirectLighingBaseline   -> ltc_direct()
//   cuda_include/ltc/ltc_utils.cuh:10-23    fetchLtcMat                -> LtcLut::fetch()
//   cuda_include/ltc/ltc_utils.cuh:26-44    integrateEdge              -> edge_integral()
//   cuda_include/ltc/ltc_utils.cuh:47-127   integrateOverPolygon       -> integrate_tri_light()
//   cuda_include/ltc/polygon_utils.cuh:7-120 iz0 / clipPolygon         -> horizon_cut() / clip_tri()
//   cuda_include/utils.cuh:77-138,168-190   matrixInverse, orthonormalBasis
//
// Design notes (B200):
//  * The reference clips through a 25-entry switch over a 5-slot array that supports quads.
//    On this path the input is ALWAYS a triangle (first clip), or that triangle with its first
//    vertex repeated when the first clip returned 4 (second clip), so the kernel
//    uses a 3-bit mask and builds the polygon directly in five named registers -- no local
//    memory, no dynamic indexing.  Vertex ORDER per case is kept as in the reference so the
//    edge sum is added in the same order; zero-length edges (a,a) of the degenerate quad add
//    an exact 0 in the reference and are skipped here.
//  * The LUT lives in shared memory (LtcLut) and is filtered in software, with the texture
//    unit's 1.8 fixed-point weight quantisation selectable.
//  * Double promotions of the reference (double PI) are kept where they happen once per pixel
//    (LUT coordinate, basis); the per-edge `0.5*(1/sqrt(max(1-x*x,1e-7)))` runs in fp32 --
//    measured effect on the final colour < 1e-6 relative (tests/test_hostcheck.py::test_frames, tests/test_gpu_frames.py).
#pragma once
#include "vec.cuh"

namespace orb {

// ---- TriLight: bit-compatible with cuda_include/common.cuh:64-72 (76 bytes) ------------------
struct TriLight {
    float3 v1, v2, v3, cg, normal, emit;
    float area;
};
static_assert(sizeof(TriLight) == 76, "TriLight layout");

// ---- orthonormalBasis, utils.cuh:168-190: rows (c1, c2, n) = world -> local ------------------
ORB_HD Mat3 onb_rows(float3 n) {
    float3 c1, c2;
    if (n.z < -0.999999f) {
        c1 = v3(0.f, -1.f, 0.f);
        c2 = v3(-1.f, 0.f, 0.f);
    } else {
        const float a = (float)(1.0 / (1.0 + (double)n.z));  // `1. / (1. + n.z)` is double
        const float b = -n.x * n.y * a;
        c1 = normalize(v3((float)(1.0 - (double)(n.x * n.x * a)), b, -n.x));
        c2 = normalize(v3(b, (float)(1.0 - (double)(n.y * n.y * a)), -n.y));
    }
    Mat3 m;
    m.r0 = c1;
    m.r1 = c2;
    m.r2 = n;
    return m;
}

// ---- matrixInverse, utils.cuh:77-138: Gauss-Jordan with full pivoting on a 3x3 ---------------
// Written on nine scalars with compile-time indices only; the run-time pivot position is
// handled with selects so nothing is indexed dynamically.
struct M9 {
    float a[3][3];
};
ORB_HD float m9_get(const M9& m, int r, int c) {
    float v = 0.f;
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++)
            if (i == r && j == c) v = m.a[i][j];
    return v;
}
// The isotropic LTC tables (include/ltc/ltc_isotropic.h:4-8) only populate m00, m02, m11, m20, m22 -- and a bilinear
// fetch keeps exact zeros -- so the matrix splits into the (x, z) 2x2 block and the y scale: a cofactor inverse of a
// dozen instructions instead of the select-heavy full-pivot elimination below (~250).  Agrees with the elimination to
// a few ulp (golden `minv` units, tests/_parity.py check_scalar_units).  Any other matrix takes the general path.
ORB_HD bool mat3_is_xz_block(const Mat3& m) { return m.r0.y == 0.f && m.r1.x == 0.f && m.r1.z == 0.f && m.r2.y == 0.f; }
ORB_HD Mat3 mat3_inverse_xz_block(const Mat3& m) {
    const float inv_det = 1.f / (m.r0.x * m.r2.z - m.r0.z * m.r2.x);
    Mat3 r;
    r.r0 = v3(m.r2.z * inv_det, 0.f, -m.r0.z * inv_det);
    r.r1 = v3(0.f, 1.f / m.r1.y, 0.f);
    r.r2 = v3(-m.r2.x * inv_det, 0.f, m.r0.x * inv_det);
    return r;
}

ORB_HD Mat3 mat3_inverse_general(const Mat3& in);
ORB_HD Mat3 mat3_inverse(const Mat3& in) { return mat3_is_xz_block(in) ? mat3_inverse_xz_block(in) : mat3_inverse_general(in); }

ORB_HD Mat3 mat3_inverse_general(const Mat3& in) {
    M9 m;
    m.a[0][0] = in.r0.x; m.a[0][1] = in.r0.y; m.a[0][2] = in.r0.z;
    m.a[1][0] = in.r1.x; m.a[1][1] = in.r1.y; m.a[1][2] = in.r1.z;
    m.a[2][0] = in.r2.x; m.a[2][1] = in.r2.y; m.a[2][2] = in.r2.z;
    int ipiv0 = 0, ipiv1 = 0, ipiv2 = 0;
    int indxr[3], indxc[3];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        int irow = 0, icol = 0;
        float big = 0.f;
#pragma unroll
        for (int j = 0; j < 3; j++) {
            const int pj = j == 0 ? ipiv0 : (j == 1 ? ipiv1 : ipiv2);
            if (pj != 1) {
#pragma unroll
                for (int k = 0; k < 3; k++) {
                    const int pk = k == 0 ? ipiv0 : (k == 1 ? ipiv1 : ipiv2);
                    if (pk == 0) {
                        const float av = fabsf(m.a[j][k]);
                        if (av >= big) {
                            big = av;
                            irow = j;
                            icol = k;
                        }
                    }
                }
            }
        }
        ipiv0 += icol == 0;
        ipiv1 += icol == 1;
        ipiv2 += icol == 2;
        // swap rows irow <-> icol
#pragma unroll
        for (int r = 0; r < 3; r++)
#pragma unroll
            for (int c = r + 1; c < 3; c++)
                if ((irow == r && icol == c) || (irow == c && icol == r)) {
#pragma unroll
                    for (int k = 0; k < 3; k++) {
                        const float t = m.a[r][k];
                        m.a[r][k] = m.a[c][k];
                        m.a[c][k] = t;
                    }
                }
        indxr[i] = irow;
        indxc[i] = icol;
        const float pivinv = 1.f / m9_get(m, icol, icol);
        // row icol: set pivot to 1 then scale
#pragma unroll
        for (int r = 0; r < 3; r++)
            if (r == icol) {
#pragma unroll
                for (int k = 0; k < 3; k++) {
                    float v = (k == icol) ? 1.f : m.a[r][k];
                    m.a[r][k] = v * pivinv;
                }
            }
        float piv_row[3];
#pragma unroll
        for (int k = 0; k < 3; k++) piv_row[k] = m9_get(m, icol, k);
#pragma unroll
        for (int r = 0; r < 3; r++)
            if (r != icol) {
                const float save = m9_get(m, r, icol);
#pragma unroll
                for (int k = 0; k < 3; k++) {
                    const float v = (k == icol) ? 0.f : m.a[r][k];
                    // reference: minv[j][k] -= minv[icol][k] * save   (contracts to one fma)
                    m.a[r][k] = v - piv_row[k] * save;
                }
            }
    }
    // undo the column permutation, last pivot first
#pragma unroll
    for (int j = 2; j >= 0; j--) {
        const int a = indxr[j], b = indxc[j];
        if (a != b) {
#pragma unroll
            for (int c0 = 0; c0 < 3; c0++)
#pragma unroll
                for (int c1 = c0 + 1; c1 < 3; c1++)
                    if ((a == c0 && b == c1) || (a == c1 && b == c0)) {
#pragma unroll
                        for (int k = 0; k < 3; k++) {
                            const float t = m.a[k][c0];
                            m.a[k][c0] = m.a[k][c1];
                            m.a[k][c1] = t;
                        }
                    }
        }
    }
    Mat3 out;
    out.r0 = v3(m.a[0][0], m.a[0][1], m.a[0][2]);
    out.r1 = v3(m.a[1][0], m.a[1][1], m.a[1][2]);
    out.r2 = v3(m.a[2][0], m.a[2][1], m.a[2][2]);
    return out;
}

// ---- LUT -------------------------------------------------------------------------------------
// Three RGBA32F tables of w x h texels (reference: 8 x 8, include/ltc/ltc_isotropic.h), sampled
// like a CUDA texture with normalised coordinates, linear filter, clamp addressing
// (include/viewer.hpp:322-327): texel centre at +0.5.  `data` may point to shared or global
// memory; layout [table][row=alpha][col=theta] float4.
struct LtcLut {
    const float4* data;
    int w, h;
    int quantise;  // 1: filter weights in 1/256 units exactly as the B200 texture unit forms them

    ORB_HD float4 sample(int table, float x, float y) const {
        const float xb = x * (float)w - 0.5f, yb = y * (float)h - 0.5f;
        const float fx = floorf(xb), fy = floorf(yb);
        const float a = xb - fx, b = yb - fy;
        float w00 = (1.f - a) * (1.f - b), w10 = a * (1.f - b), w01 = (1.f - a) * b, w11 = a * b;
        if (quantise) {
            // the B200 texture unit, as measured (tools/probe_tex.py, profiles/r1n_texture_unit.json): both
            // fractions rounded to 1/256, w11 = round(a b) in 1/256 units, the other three by subtraction
            // so that the four weights sum to exactly 1 -- bit-exact on 98.5 % of 60 k probes, <= 1 ulp else
            const float qa = floorf(a * 256.f + 0.5f), qb = floorf(b * 256.f + 0.5f);
            const float q11 = floorf(qa * qb * (1.f / 256.f) + 0.5f);
            w11 = q11 * (1.f / 256.f);
            w10 = (qa - q11) * (1.f / 256.f);
            w01 = (qb - q11) * (1.f / 256.f);
            w00 = (256.f - qa - qb + q11) * (1.f / 256.f);
        }
        int i0 = (int)fx, j0 = (int)fy;
        int i1 = i0 + 1, j1 = j0 + 1;
        i0 = imin(imax(i0, 0), w - 1);
        i1 = imin(imax(i1, 0), w - 1);
        j0 = imin(imax(j0, 0), h - 1);
        j1 = imin(imax(j1, 0), h - 1);
        const float4* t = data + (size_t)table * w * h;
        const float4 t00 = t[j0 * w + i0], t10 = t[j0 * w + i1], t01 = t[j1 * w + i0], t11 = t[j1 * w + i1];
        float4 r;
        r.x = w00 * t00.x + w10 * t10.x + w01 * t01.x + w11 * t11.x;
        r.y = w00 * t00.y + w10 * t10.y + w01 * t01.y + w11 * t11.y;
        r.z = w00 * t00.z + w10 * t10.z + w01 * t01.z + w11 * t11.z;
        r.w = w00 * t00.w + w10 * t10.w + w01 * t01.w + w11 * t11.w;
        return r;
    }

    // fetchLtcMat: x = theta*0.99f/(0.5*PI) evaluated in double, y = alpha.
    ORB_HD void fetch(float alpha, float theta, Mat3& m, float& amplitude) const {
        const float x = (float)((double)(theta * 0.99f) / (0.5 * ORB_PI_D));
        const float4 r1 = sample(0, x, alpha), r2 = sample(1, x, alpha), r3 = sample(2, x, alpha);
        m.r0 = v3(r1);
        m.r1 = v3(r2);
        m.r2 = v3(r3);
        amplitude = r3.w;
    }
};

// ---- edge integral (z component of the vector irradiance of one great-circle edge) ------------
// Branch-free: both forms of theta/sin(theta) are evaluated and one is selected, so lanes whose
// edges have opposite signs of x stay converged.
ORB_HD float edge_integral(float3 p, float3 q) {
    const float x = dot(p, q);
    const float y = fabsf(x);
    const float num = 0.8543985f + (0.4965155f + 0.0145206f * y) * y;
    const float den = 3.4175940f + (4.1616724f + y) * y;
    const float v = orb_fdiv(num, den);  // den in [3.4, 8.6]
    const float w = 0.5f * orb_rsqrt(fmaxf(1.0f - x * x, 1e-7f)) - v;
    const float theta_sintheta = (x > 0.0f) ? v : w;
    return (p.x * q.y - p.y * q.x) * theta_sintheta;
}

// Intersection of segment (lhs, rhs) with the horizon z = 0, pushed back onto the unit sphere.
ORB_HD float3 horizon_cut(float3 lhs, float3 rhs) {
    const float s = orb_fdiv(lhs.z, lhs.z - rhs.z);
    const float cx = -s * lhs.x + lhs.x, cy = -s * lhs.y + lhs.y;
    return normalize(v3(s * rhs.x + cx, s * rhs.y + cy, 0.f));
}

ORB_HD float3 sel3(bool c, float3 a, float3 b) { return v3(c ? a.x : b.x, c ? a.y : b.y, c ? a.z : b.z); }

// Clip triangle (a,b,c) -- or, when `quad` is set, the degenerate quad (a,b,c,a) the reference
// feeds its second clip -- against z > 0 and return the absolute edge sum.  n_out receives the
// vertex count the reference's clipPolygon would have returned (0, 3, 4 or 5).
//
// The reference's 25-way switch (polygon_utils.cuh:33-120) is a large block of code per case;
// inlined twice per light it made k_ltc 92 KB of SASS and the kernel stalled on instruction
// fetch (ncu profiles/r1a: stall_no_instruction 11.8 per issue).  Here ALL cases share one
// instruction stream: the three possible horizon cuts are computed, the (up to four) polygon
// vertices are picked with selects in the reference's vertex order per case, and at most four
// edge integrals are summed in that order.  Unused cuts may be inf/NaN; they are never selected.
//   mask (a,b,c above)   polygon                 reference case
//   7                    a  b  c                 59
//   1                    a  ab ca                11
//   2                    ab b  bc                19
//   4                    ca bc c                 35
//   3                    a  b  bc ca             27
//   6                    ab b  c  ca             51
//   5                    a  ab bc c              43      (quad: ab bc c a, case 108)
ORB_HD float clipped_form_factor(float3 a, float3 b, float3 c, bool quad, int& n_out) {
    const int mask = (a.z > 0.f ? 1 : 0) | (b.z > 0.f ? 2 : 0) | (c.z > 0.f ? 4 : 0);
    if (mask == 0) {
        n_out = 0;
        return 0.f;
    }
    if (mask == 7) {  // nothing to clip (the common case: the light is above the horizon): edges ab, bc, ca
        float s = edge_integral(a, b);
        s += edge_integral(b, c);
        s += edge_integral(c, a);
        n_out = quad ? 4 : 3;
        return fabsf(s);
    }
    const float3 ab = horizon_cut(a, b), bc = horizon_cut(b, c), ca = horizon_cut(c, a);
    const bool m5q = mask == 5 && quad;
    const bool four = mask == 3 || mask == 6 || mask == 5;
    const float3 p0 = sel3(mask == 2 || mask == 6 || m5q, ab, sel3(mask == 4, ca, a));
    const float3 p1 = sel3(mask == 1 || (mask == 5 && !quad), ab, sel3(mask == 4 || m5q, bc, b));
    const float3 p2 = sel3(mask == 1, ca, sel3(mask == 2 || mask == 3 || (mask == 5 && !quad), bc, c));
    const float3 p3 = sel3(!four, p0, sel3(mask == 5, sel3(quad, a, c), ca));
    float s = edge_integral(p0, p1);
    s += edge_integral(p1, p2);
    s += edge_integral(p2, p3);
    if (four) s += edge_integral(p3, p0);
    const bool plus1 = quad && (mask == 7 || mask == 1 || mask == 3 || mask == 5);
    n_out = (four ? 4 : 3) + (plus1 ? 1 : 0);
    return fabsf(s);
}

// ---- per-pixel shading frame -----------------------------------------------------------------
struct LtcFrame {
    float3 p;          // shading point
    float3 diffuse;    // si.diffuse
    Mat3 to_local;     // rows c1, c2, n
    Mat3 iso;          // rotation about z that puts wo_local in the xz-plane
    Mat3 world_to_iso; // iso * to_local: both are rotations, applied as one (see integrate_tri_light)
    Mat3 minv;         // inverse of the fetched LTC matrix
    float amplitude;
    bool valid;        // false: wo_local.z < 0 -> colour 0 (deviceCode.cu:27-28)
    bool xz_block;     // minv has the isotropic sparsity: five multiply-adds per vertex instead of nine
};

ORB_HD float3 apply_minv(const LtcFrame& f, float3 v) {
    if (f.xz_block) return v3(f.minv.r0.x * v.x + f.minv.r0.z * v.z, f.minv.r1.y * v.y, f.minv.r2.x * v.x + f.minv.r2.z * v.z);
    return apply(f.minv, v);
}

// Everything ltcDirectLighingBaseline does before its light loop.  wo_local as rayGen computes
// it (deviceCode.cu:84); the function re-normalises it like the reference (:26).
ORB_HD void ltc_setup(LtcFrame& f, const LtcLut& lut, float3 p, float3 diffuse, float alpha, const Mat3& to_local,
                     float3 wo) {
    f.p = p;
    f.diffuse = diffuse;
    f.to_local = to_local;
    const float3 wl = normalize(apply(to_local, wo));
    f.valid = !(wl.z < 0.f);
    const float theta = acosf(wl.z);  // sphericalTheta, utils.cuh:202
    Mat3 m;
    lut.fetch(alpha, theta, m, f.amplitude);
    f.xz_block = mat3_is_xz_block(m);
    f.minv = mat3_inverse(m);
    const float3 x = normalize(v3(wl.x, wl.y, 0.f));
    const float3 z = v3(0.f, 0.f, 1.f);
    f.iso.r0 = x;
    f.iso.r2 = z;
    f.iso.r1 = normalize(cross(z, x));
    const Mat3 tl_t = transpose(to_local);  // columns of to_local
    f.world_to_iso.r0 = apply(tl_t, f.iso.r0);  // row i of iso * to_local = to_local^T * iso.row_i
    f.world_to_iso.r1 = apply(tl_t, f.iso.r1);
    f.world_to_iso.r2 = apply(tl_t, f.iso.r2);
}

// integrateOverPolygon for one triangle light.
ORB_HD float3 integrate_tri_light(const LtcFrame& f, float3 lv1, float3 lv2, float3 lv3, float3 lnormal,
                                 float3 lemit) {
    // The reference normalises after each of its two rotations (ltc_utils.cuh:64-76); a rotation of
    // a unit vector is a unit vector to within an ulp, so the two are applied as ONE matrix with no
    // re-normalisation (measured effect on the colour < 2e-6 relative, tests/test_hostcheck.py), and
    // the back-face test uses the un-normalised centroid direction (same sign).
    float3 a = normalize(lv1 - f.p), b = normalize(lv2 - f.p), c = normalize(lv3 - f.p);
    if (dot(a + b + c, lnormal) > 0.f) return v3(0.f);  // dot(-normalize(a+b+c), n) < 0
    a = apply(f.world_to_iso, a);
    b = apply(f.world_to_iso, b);
    c = apply(f.world_to_iso, c);
    int n_diff;
    const float diffuse_shading = clipped_form_factor(a, b, c, false, n_diff);
    float ggx_shading = 0.f;
    if (n_diff != 0) {
        const float3 ta = normalize(apply_minv(f, a)), tb = normalize(apply_minv(f, b)), tc = normalize(apply_minv(f, c));
        int n_ggx;
        ggx_shading = clipped_form_factor(ta, tb, tc, n_diff == 4, n_ggx);
    }
    return (f.diffuse * lemit * diffuse_shading) + (f.amplitude * lemit * ggx_shading);
}

// Distance of a pixel's LTC evaluation to its nearest SIGN DECISION: wo_local.z < 0 (deviceCode.cu:27), the
// back-face test (ltc_utils.cuh:58-60), the above-horizon bit of every vertex fed to the two clips
// (polygon_utils.cuh:27-31).  A pixel below ~1e-4 takes another case of clipPolygon's 25-case table when a direction
// differs in its last ulps between two implementations: north_star's "case flip" pixels, counted by a checked render
// (ORB_RENDER_CHECKED -> OrbCounters.ltc_near_decision).
ORB_HD float ltc_decision_margin(const LtcFrame& f, float wl_z, const TriLight* lights, int n_lights) {
    float m = fabsf(wl_z);
    if (!f.valid) return m;
    for (int l = 0; l < n_lights; l++) {
        const TriLight& L = lights[l];
        float3 a = normalize(L.v1 - f.p), b = normalize(L.v2 - f.p), c = normalize(L.v3 - f.p);
        const float bf = -dot(normalize(a + b + c), L.normal);
        m = fminf(m, fabsf(bf));
        if (bf < 0.f) continue;
        a = apply(f.world_to_iso, a);
        b = apply(f.world_to_iso, b);
        c = apply(f.world_to_iso, c);
        m = fminf(m, fminf(fabsf(a.z), fminf(fabsf(b.z), fabsf(c.z))));
        if (!(a.z > 0.f || b.z > 0.f || c.z > 0.f)) continue;
        const float3 ta = normalize(apply_minv(f, a)), tb = normalize(apply_minv(f, b)), tc = normalize(apply_minv(f, c));
        m = fminf(m, fminf(fabsf(ta.z), fminf(fabsf(tb.z), fabsf(tc.z))));
    }
    return m;
}

// (Measured and removed, round 2: evaluating two TriLights that are the halves of one quad -- the fan (a, b, c),
// (a, c, d) -- as ONE four-edge polygon in their common unclipped case.  The diagonal's two edge integrals cancel
// exactly, so four vertices and four edges per domain instead of six and six; but k_ltc went 0.533 -> 0.589 ms at 32
// lights (profiles/r2l_ab_ltc.jsonl): the fast path and its fall-back double the loop body, lanes of a warp split
// between them, and the kernel was instruction-fetch sensitive already in round 1.)
// The light loop.  `lights` normally points to shared memory (all lanes read the same light: a broadcast).
ORB_HD float3 ltc_direct(const LtcFrame& f, const TriLight* lights, int n_lights) {
    float3 colour = v3(0.f);
    if (!f.valid) return colour;
#pragma unroll 1
    for (int l = 0; l < n_lights; l++) {
        const TriLight& L = lights[l];
        colour += integrate_tri_light(f, L.v1, L.v2, L.v3, L.normal, L.emit);
    }
    return colour;
}

}  // namespace orb

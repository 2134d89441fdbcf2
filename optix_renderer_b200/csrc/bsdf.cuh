// bsdf.cuh -- the reference's "Frostbite" material: 0.5 Lambert + 0.5 GGX microfacet reflection.
//
// Device-side restatement of the REACHABLE parts of
//   cuda_include/frostbite.cuh:14-113,160-169,208-243
//   cuda_include/material.cuh:7-104
//   cuda_include/path/path.cuh:7-33     (sampleLight / sampleLightPdf / pdfA2W)
//   cuda_include/utils.cuh:194-226      (samplePointOnTriangle, balanceHeuristic, checkPositive)
// including their literal quirks: roughness squared twice, the
// height-correlated G2 that vanishes unless both tan^2 <= 1e-5, the forced diffuse lobe whose
// pdf still carries the specular term.  All directions are in the local shading frame (z = n).
#pragma once
#include "vec.cuh"

namespace orb {

ORB_HD float cos_theta(float3 w) { return w.z; }
ORB_HD float cos2_theta(float3 w) { return w.z * w.z; }
ORB_HD float sin2_theta(float3 w) { return fmaxf(0.0f, 1.0f - cos2_theta(w)); }
ORB_HD float tan2_theta(float3 w) { return sin2_theta(w) / cos2_theta(w); }
ORB_HD bool same_hemisphere(float3 a, float3 b) { return a.z * b.z > 0.0f; }

// clampDot(a, b, false): CUDA's max(float,float) = fmaxf, so a NaN dot gives EPS.
ORB_HD float clamp_dot_eps(float3 a, float3 b) { return fmaxf(dot(a, b), ORB_EPS); }

ORB_HD float3 fresnel_schlick(float cos_i, float3 f0) {
    const float a = fmaxf(0.0f, 1.0f - cos_i);
    const float a2 = a * a;
    const float a5 = a2 * a2 * a;
    return f0 + (v3(1.0f) - f0) * a5;
}

ORB_HD float ggx_d(float3 wh, float alpha) {
    const float alpha2 = alpha * alpha;
    const float a = 1.0f + cos2_theta(wh) * (alpha2 - 1.0f);
    return alpha2 / (ORB_PI_F * a * a);
}

ORB_HD float ggx_g1(float3 w, float alpha) {
    const float t2 = tan2_theta(w);
    if ((double)t2 > 1e5) return 0.0f;
    const float alpha2 = alpha * alpha;
    const float lambda = (-1.0f + sqrtf(alpha2 * t2 + 1.0f)) / 2.0f;
    return 1.0f / (1.0f + lambda);
}

// frostbite.cuh:62-72 -- note the literal `> 1e-5` (double) early-out.
ORB_HD float ggx_g2_height_correlated(float3 wi, float3 wo, float alpha) {
    const float t2o = tan2_theta(wo), t2i = tan2_theta(wi);
    if ((double)t2o > 1e-5 || (double)t2i > 1e-5) return 0.f;
    const float alpha2 = alpha * alpha;
    const float lo = (-1.0f + sqrtf(alpha2 * t2o + 1.0f)) / 2.0f;
    const float li = (-1.0f + sqrtf(alpha2 * t2i + 1.0f)) / 2.0f;
    return 1.0f / (1.0f + lo + li);
}

ORB_HD float3 lambert(float3 wi, float3 wo, float3 colour) {
    if (!same_hemisphere(wi, wo)) return v3(0.0f);
    return colour / ORB_PI_F;
}

// microfacetReflection_GGX with eta = 0 (the only value material.cuh passes): the eta < 1 branch.
ORB_HD float3 ggx_reflection(float3 wi, float3 wo, float3 f0, float alpha) {
    if (!same_hemisphere(wi, wo) || cos_theta(wi) == 0.0f || cos_theta(wo) == 0.0f) return v3(0.0f);
    float3 wh = wi + wo;
    if (wh.x == 0.0f && wh.y == 0.0f && wh.z == 0.0f) return v3(0.0f);
    wh = normalize(wh);
    const float ct = dot(wi, wh);
    const float3 F = (ct * ct > 0.0f) ? fresnel_schlick(fabsf(ct), f0) : v3(1.0f);
    const float G = ggx_g2_height_correlated(wi, wo, alpha);
    const float D = ggx_d(wh, alpha);
    return F * G * D / (4.0f * fabsf(cos_theta(wi)) * fabsf(cos_theta(wo)));
}

// evaluate(), material.cuh:7-22.
ORB_HD float3 bsdf_evaluate(float3 wi, float3 wo, float3 base, float alpha) {
    const float alpha2 = alpha * alpha;
    const float3 d = lambert(wi, wo, base);
    const float3 s = ggx_reflection(wi, wo, base, alpha2);
    return 0.5f * d + 0.5f * s;
}

// computeLobeProbabilities(), material.cuh:26-41 (the Fresnel term it computes is unused).
ORB_HD void lobe_probabilities(float3 base, float& p_diffuse, float& p_specular) {
    p_diffuse = fmaxf(base.x, fmaxf(base.y, base.z)) * 0.5f;
    p_specular = fmaxf(base.x, fmaxf(base.y, base.z));
    const float norm = 1.0f / (p_diffuse + p_specular);
    p_diffuse *= norm;
    p_specular *= norm;
}

ORB_HD float pdf_cosine(float3 wi, float3 wo) { return same_hemisphere(wi, wo) ? cos_theta(wi) / ORB_PI_F : 0.0f; }

ORB_HD float pdf_ggx_vndf_reflection(float3 wi, float3 wo, float alpha) {
    if (!same_hemisphere(wi, wo)) return 0.0f;
    const float3 wh = normalize(wi + wo);
    const float pdf_h = ggx_g1(wo, alpha) * ggx_d(wh, alpha) * fabsf(dot(wh, wo)) / fabsf(cos_theta(wo));
    const float dwh_dwi = 1.0f / (4.0f * dot(wi, wh));
    return pdf_h * dwh_dwi;
}

// pdf(), material.cuh:94-104.
ORB_HD float bsdf_pdf(float3 wi, float3 wo, float3 base, float alpha) {
    float pd, ps;
    lobe_probabilities(base, pd, ps);
    return pd * pdf_cosine(wi, wo) + ps * pdf_ggx_vndf_reflection(wi, wo, alpha);
}

ORB_HD float3 sample_cosine_hemisphere(float u1, float u2) {
    const float c = sqrtf(fmaxf(0.0f, 1.0f - u1));
    const float s = sqrtf(u1);
    const float phi = 2.0f * ORB_PI_F * u2;
    return v3(s * cosf(phi), s * sinf(phi), c);
}

// sampleGGX_VNDF, frostbite.cuh:208-232 (Heitz 2018).
ORB_HD float3 sample_ggx_vndf(float3 wo, float alpha, float u1, float u2) {
    const float3 wh_ = normalize(v3(alpha * wo.x, alpha * wo.y, wo.z));
    const float len2 = wh_.x * wh_.x + wh_.y * wh_.y;
    const float3 b1 = len2 > 0.0f ? v3(-wh_.y, wh_.x, 0.0f) * (1.0f / sqrtf(len2)) : v3(1.0f, 0.0f, 0.0f);
    const float3 b2 = cross(wh_, b1);
    const float r = sqrtf(u1);
    const float phi = 2.0f * ORB_PI_F * u2;
    const float t1 = r * cosf(phi);
    float t2 = r * sinf(phi);
    const float s = 0.5f * (1.0f + wh_.z);
    t2 = (1.0f - s) * sqrtf(1.0f - t1 * t1) + s * t2;
    const float3 h = t1 * b1 + t2 * b2 + sqrtf(fmaxf(0.0f, 1.0f - t1 * t1 - t2 * t2)) * wh_;
    return normalize(v3(alpha * h.x, alpha * h.y, fmaxf(0.0f, h.z)));
}

// helper_math clamp (fmaxf(lo, fminf(f, hi))) as used by remap(), material.cuh:44-47.
ORB_HD float remap_clamped(float value, float low1, float high1, float low2, float high2) {
    const float r = low2 + (value - low1) * (high2 - low2) / (high1 - low1);
    return fmaxf(low2, fminf(r, high2));
}

// sample_direction(), material.cuh:50-91, with metalness unused as in the reference.
ORB_HD float3 bsdf_sample(float3 wo, float u1, float u2, float& pdf_out, float3 base, float alpha) {
    float pd, ps;
    lobe_probabilities(base, pd, ps);
    pd = 1.f;  // material.cuh:57
    float3 wi;
    if (u1 < pd) {
        u1 = remap_clamped(u1, 0.0f, pd - ORB_EPS, 0.0f, 1 - ORB_EPS);
        wi = copysignf(1.f, cos_theta(wo)) * sample_cosine_hemisphere(u1, u2);
        wi = normalize(wi);
    } else {  // only reachable when u1 == 1.0f exactly
        u1 = remap_clamped(u1, pd, pd + ps - ORB_EPS, 0.0f, 1 - ORB_EPS);
        const float sg = copysignf(1.f, cos_theta(wo));
        const float3 wo_upper = sg * wo;
        const float3 wh = sg * sample_ggx_vndf(wo_upper, alpha, u1, u2);
        if (dot(wo, wh) < 0.0f) return v3(0.0f);  // pdf left untouched, as in the reference
        wi = 2.0f * wh * dot(wh, wo) - wh;
        if (!same_hemisphere(wi, wo)) return v3(0.0f);
    }
    pdf_out = pd * pdf_cosine(wi, wo) + ps * pdf_ggx_vndf_reflection(wi, wo, alpha);
    return wi;
}

// ---- light sampling helpers --------------------------------------------------------------------
ORB_HD float3 sample_point_on_triangle(float3 v1, float3 v2, float3 v3_, float u1, float u2) {
    const float su1 = sqrtf(u1);
    return (1 - su1) * v1 + su1 * ((1 - u2) * v2 + u2 * v3_);
}

ORB_HD float pdf_area_to_solid_angle(float pdf, float dist2, float cos_t) {
    const float ac = fabsf(cos_t);
    if ((double)ac < 1e-8) return 0.0f;
    return pdf * dist2 / ac;
}

ORB_HD float balance_heuristic(float f_pdf, float g_pdf) { return f_pdf / (f_pdf + g_pdf); }

ORB_HD float3 check_positive(float3 c) { return v3(owl_max(c.x, 0.f), owl_max(c.y, 0.f), owl_max(c.z, 0.f)); }

}  // namespace orb

// pixel.cuh -- what ONE pixel (or one ray of it) computes in each stage of the wavefront.
//
// The reference does all of this in one raygen thread per pixel (cuda_src/deviceCode.cu:59-175).
// Here every stage is a separate, stateless function of (pixel, frame, G-buffer record): the
// RNG stream position is re-derived instead of carried, so stages only exchange the frame
// buffers the reference already defines plus a 16-byte hit record.  The __global__ wrappers in
// orb200.cu add the memory traffic; tests/hostcheck compiles these same functions for the CPU.
//
//   primary_ray()         deviceCode.cu:65-73       jittered pinhole ray
//   ltc_pixel()           deviceCode.cu:80-84,24-56 frame set-up + LTC light loop
//   ratio_shadow_sample() ratio/ratio.cuh:11-126    one light sample, {shadowed, unshadowed}
//   path_bounce()         path/path.cuh:145-258     one iteration of the bounce loop
//   make_rgba()           owl::make_rgba
#pragma once
#include "bsdf.cuh"
#include "ltc.cuh"
#include "rng.cuh"
#include "scene.cuh"

namespace orb {

struct Camera {  // LaunchParams.camera, cuda_include/common.cuh:113-118
    float3 pos, d00, du, dv;
};

#define ORB_MODE_LITERAL 0
#define ORB_MODE_INTENDED 1

// Draws 0 and 1 of the pixel's stream; direction in exact arithmetic (see vec.cuh).
ORB_HD void primary_ray(const Camera& cam, int frame_id, int px, int py, int W, int H, float3& o, float3& d) {
    Rng rng = rng_seed(frame_id, (uint32_t)px, (uint32_t)py, (uint32_t)W);
    const float jx = rng_nextf(rng), jy = rng_nextf(rng);
    const float su = xdiv(xadd((float)px, jx), (float)W);
    const float sv = xdiv(xadd((float)py, jy), (float)H);
    o = cam.pos;
    d = xnormalize(xadd(xadd(cam.d00, xmul(su, cam.du)), xmul(sv, cam.dv)));
}

// si.wo / to_local / wo_local as rayGen sets them (deviceCode.cu:80-84), then the LTC sum.
ORB_HD float3 ltc_pixel(const LtcLut& lut, const TriLight* lights, int n_lights, float3 cam_pos, float3 p, float3 n,
                        float3 diffuse, float alpha) {
    const float3 wo = normalize(cam_pos - p);
    const Mat3 to_local = onb_rows(n);
    LtcFrame f;
    ltc_setup(f, lut, p, diffuse, alpha, to_local, wo);
    return ltc_direct(f, lights, n_lights);
}

// One call of ratio_based_shadow(si, rng, 1): five draws, one closest-hit "shadow" ray.  Split
// in two so the wavefront can put a ray-stream traversal kernel between them:
//   shadow_ray_setup()   draws + light sample -> the ray and what the resolve step needs
//   shadow_ray_resolve() closest hit -> colours[0] (shadowed) and colours[1] (unshadowed)
struct ShadowSample {
    float3 origin, L;   // the ray: p + 1e-3 n_geom, towards the sampled light point
    float3 emit;        // the SELECTED light's emit (unshadowed term)
    float dist2;        // squared distance to the sampled point
    float light_pdf;    // area pdf of the selection, 1 / (area * numTriLights)
    bool valid;         // false: no lights
};

ORB_HD ShadowSample shadow_ray_setup(const SceneView& s, float3 p, float3 n, Rng& rng) {
    ShadowSample q;
    const float r1x = rng_nextf(rng), r1y = rng_nextf(rng);
    rng_nextf(rng);  // rand2 is drawn and unused (ratio.cuh:30)
    rng_nextf(rng);
    const float sel = rng_nextf(rng);
    q.valid = s.n_lights > 0;
    q.origin = q.L = q.emit = v3(0.f);
    q.dist2 = q.light_pdf = 0.f;
    if (!q.valid) return q;
    // int(rand * (numTriLights - 1)): the last light is never selected
    const int idx = imin(imax((int)(sel * (float)(s.n_lights - 1)), 0), s.n_lights - 1);
    const TriLight lt = s.lights[idx];
    q.emit = lt.emit;
    q.light_pdf = 1.f / (lt.area * (float)s.n_lights);
    // samplePointOnTriangle, utils.cuh:194-199 -- defines the ray, so exact arithmetic
    const float su1 = xsqrt(r1x);
    const float3 inner = xadd(xmul(xsub(1.f, r1y), lt.v2), xmul(r1y, lt.v3));
    const float3 new_pos = xadd(xmul(xsub(1.f, su1), lt.v1), xmul(su1, inner));
    const float3 diff = xsub(new_pos, p);
    q.L = xnormalize(diff);
    const float dist = xlength(diff);
    q.dist2 = dist * dist;
    q.origin = xadd(p, xmul(n, 1e-3f));
    return q;
}

ORB_HD void shadow_ray_resolve(const SceneView& s, const ShadowSample& q, const Hit& h, float3 cam_pos, float3 p,
                               float3 n, float3 diffuse, float alpha, int mode, float3& direct, float3& novis) {
    direct = novis = v3(0.f);
    if (!q.valid) return;
    const Surface ls = shade_hit(s, h.prim, h.u, h.v);
    const float light_pdf_w = pdf_area_to_solid_angle(q.light_pdf, q.dist2, dot(-q.L, ls.n));
    float3 brdf = v3(0.f);  // committed source (ratio.cuh:58)
    if (mode == ORB_MODE_INTENDED) {  // patch P1: evaluate() as in the commented tail of that line
        const float3 wo = normalize(cam_pos - p);
        const Mat3 to_local = onb_rows(n);
        const float3 wo_local = normalize(apply(to_local, wo));
        const float3 wi_local = normalize(apply(to_local, q.L));
        brdf = bsdf_evaluate(wi_local, wo_local, diffuse, alpha);
    }
    const float mis_w = light_pdf_w / (light_pdf_w + 0.f);  // brdfPdf = 0
    const float cd = clamp_dot_eps(n, q.L);
    const float3 tp = v3(1.f);
    if (ls.is_light) direct = check_positive(check_positive(mis_w * ls.emit * tp * brdf * cd / light_pdf_w));
    novis = check_positive(check_positive(mis_w * q.emit * tp * brdf * cd / light_pdf_w));
}

// Both halves around a caller-supplied tracer (megakernel form: tests/hostcheck, k_unit).
// Returns false when no ray was traced (no lights).
template <class Tracer>
ORB_HD bool ratio_shadow_sample(const SceneView& s, Tracer&& trace, float3 cam_pos, float3 p, float3 n, float3 diffuse,
                                float alpha, Rng& rng, int mode, float3& direct, float3& novis) {
    const ShadowSample q = shadow_ray_setup(s, p, n, rng);
    if (!q.valid) {
        direct = novis = v3(0.f);
        return false;
    }
    const Hit h = trace(q.origin, q.L);
    shadow_ray_resolve(s, q, h, cam_pos, p, n, diffuse, alpha, mode, direct, novis);
    return true;
}

// ---- path tracer -----------------------------------------------------------------------------
struct PathState {  // the part of current_si / loop state that survives an iteration
    float3 p, n, diffuse, V, tp;
    float alpha;
};

#define ORB_PATH_CONTINUE 0
#define ORB_PATH_DONE 1

// One iteration of estimatePathTracing's loop for a path whose current vertex is a non-light
// hit, split like the shadow sample so a ray-stream traversal can sit in the middle:
//   path_ray_setup()  frame, five draws, sample_direction -> the bounce ray (+ what resolve needs)
//   path_resolve()    closest hit -> colour (ORB_PATH_DONE) or the next vertex (ORB_PATH_CONTINUE)
// mode literal = the committed debug stub (path.cuh:216-217): colour = sampled local direction,
// no ray.  On ORB_PATH_DONE `colour` holds the value the function would return (max(colour, EPS)
// applied).
struct PathSample {
    float3 origin, dir;       // p + EPS n_geom, to_world * wi_local (normalised)
    float3 wi_local, wo_local;
    float bsdf_pdf;
};

ORB_HD PathSample path_ray_setup(const PathState& st, Rng& rng) {
    PathSample q;
    // orthonormalBasis(n, to_world, to_local): the outputs are swapped w.r.t. rayGen
    const Mat3 to_world = onb_rows(st.n);
    const Mat3 to_local = transpose(to_world);
    q.wo_local = normalize(apply(to_local, st.V));
    rng_nextf(rng);  // rand1, unused
    rng_nextf(rng);
    const float r2x = rng_nextf(rng), r2y = rng_nextf(rng);
    rng_nextf(rng);  // light selection draw, unused
    q.bsdf_pdf = 0.f;
    q.wi_local = bsdf_sample(q.wo_local, r2x, r2y, q.bsdf_pdf, st.diffuse, st.alpha);
    q.origin = xadd(st.p, xmul(ORB_EPS, st.n));
    q.dir = normalize(apply(to_world, q.wi_local));
    return q;
}

ORB_HD float3 path_literal_colour(const PathSample& q) {
    return v3(owl_max(q.wi_local.x, ORB_EPS), owl_max(q.wi_local.y, ORB_EPS), owl_max(q.wi_local.z, ORB_EPS));
}

ORB_HD int path_resolve(const SceneView& s, PathState& st, const PathSample& q, const Hit& h, bool last, float3& colour) {
    colour = v3(ORB_EPS);
    if (h.prim < 0) return ORB_PATH_DONE;
    const Surface hs = shade_hit(s, h.prim, h.u, h.v);
    const float3 brdf = bsdf_evaluate(q.wi_local, q.wo_local, st.diffuse, st.alpha);
    if (hs.is_light) {
        const float light_pdf = 1.f / (hs.area * (float)s.n_lights);
        float dist = length(hs.p - st.p);
        dist *= dist;
        const float light_pdf_w = pdf_area_to_solid_angle(light_pdf, dist, cos_theta(q.wi_local));
        const float mis_w = q.bsdf_pdf / (q.bsdf_pdf + light_pdf_w);
        const float3 c = mis_w * hs.emit * st.tp / q.bsdf_pdf;
        colour = v3(owl_max(c.x, ORB_EPS), owl_max(c.y, ORB_EPS), owl_max(c.z, ORB_EPS));
        return ORB_PATH_DONE;
    }
    if (last) return ORB_PATH_DONE;
    st.tp *= cos_theta(q.wi_local) * brdf / q.bsdf_pdf;
    st.V = normalize(st.p - hs.p);
    st.p = hs.p;
    st.n = hs.n;
    st.diffuse = hs.diffuse;
    st.alpha = 0.f;  // path.cuh:251 (not re-clamped)
    return ORB_PATH_CONTINUE;
}

// Both halves around a caller-supplied tracer (megakernel form: tests/hostcheck, k_path_bounce).
template <class Tracer>
ORB_HD int path_bounce(const SceneView& s, Tracer&& trace, PathState& st, Rng& rng, int mode, bool last, float3& colour,
                       bool& traced) {
    traced = false;
    const PathSample q = path_ray_setup(st, rng);
    if (mode == ORB_MODE_LITERAL) {
        colour = path_literal_colour(q);
        return ORB_PATH_DONE;
    }
    const Hit h = trace(q.origin, q.dir);
    traced = true;
    return path_resolve(s, st, q, h, last, colour);
}

// owl::make_rgba: min(255, max(0, int(f * 256))) per channel, alpha 0xff.
ORB_HD uint32_t make_rgba(float3 c) {
    const float fx = c.x * 256.f, fy = c.y * 256.f, fz = c.z * 256.f;
    const int r = (fx != fx) ? 0 : (fx >= 255.f ? 255 : (fx <= 0.f ? 0 : (int)fx));
    const int g = (fy != fy) ? 0 : (fy >= 255.f ? 255 : (fy <= 0.f ? 0 : (int)fy));
    const int b = (fz != fz) ? 0 : (fz >= 255.f ? 255 : (fz <= 0.f ? 0 : (int)fz));
    return (uint32_t)r | ((uint32_t)g << 8) | ((uint32_t)b << 16) | 0xff000000u;
}

}  // namespace orb

// ref_driver.cpp -- TEST INFRASTRUCTURE ONLY (oracle/_ref build).
//
// Compiles the REFERENCE's own device translation unit (REF_DEVICE_TU, normally
// /root/reference/cuda_src/deviceCode.cu, which pulls in cuda_include/**) as host C++ behind
// owl_shim.h and drives it pixel by pixel, implementing oracle_api.h.  No reference source is
// copied: the TU is #included from where it lies (oracle/Makefile passes the path).
//
// Hooks supplied here = the un-vendored OWL/OptiX layer:
// traceRay -> ocpu::TriSoup closest hit; tex2D -> ocpu bilinear / nearest emulation.
#include "owl_shim.h"

#include <atomic>
#include <thread>
#include <vector>

#include "../oracle_api.h"
#include "../trace_cpu.h"

int shim_shadow_spp = 4;  // used by the P3-patched copies only
int shim_max_depth = 1;

// --- P0: tag every draw with its sequence number so vec2f(RandF,RandF) can put
// the earlier draw in .x as nvcc device code does.  lcg_random.cuh is included first under a
// rename; its #pragma once keeps the reference TU from including it again.
#define lcg_randomf lcg_randomf_raw
#include <lcg_random.cuh>
#undef lcg_randomf
static thread_local unsigned long long tl_seq = 0;
inline RandF lcg_randomf(LCGRand& rng) {
    float v = lcg_randomf_raw(rng);
    RandF r;
    r.v = v;
    r.seq = tl_seq++;
    return r;
}

#include REF_DEVICE_TU

using owl::common::vec2i;
using owl::common::vec3f;

namespace {

struct MeshStore {
    std::vector<float> vtx, nrm, uv;
    std::vector<int32_t> idx;
    std::vector<uint8_t> tex;
    int texW = 0, texH = 0;
};

struct Texture {
    int kind;  // 0 = float4 linear clamp, 1 = rgba8 nearest clamp
    const void* data;
    int w, h;
};

struct State {
    std::vector<MeshStore*> store;
    std::vector<TriangleMeshData> meshes;
    std::vector<int> primMesh, primLocal;  // global prim -> (mesh, local prim)
    ocpu::TriSoup soup;
    bool use_bvh = true;
    std::vector<Texture> textures;  // handle = index + 1
    std::vector<float> lut[3];
    int lutW = 0, lutH = 0, lutQuant = 0;
    std::vector<TriLight> lights;
    MissProgData miss;
} G;

struct Tls {
    const void* programData = nullptr;
    void* prd = nullptr;
    vec2i launch;
    float2 bary;
    unsigned int prim = 0;
    long long rays = 0;
    int tracesThisPixel = 0;
    int firstPrim = -1;
    float firstU = 0.f, firstV = 0.f;
};
thread_local Tls T;

}  // namespace

namespace owl {
void* shim_program_data() { return (void*)T.programData; }
void* shim_prd() { return T.prd; }
common::vec2i shim_launch_index() { return T.launch; }
void shim_trace(const common::vec3f& org, const common::vec3f& dir, float tmin, float tmax, void* prd) {
    const float o[3] = {org.x, org.y, org.z}, d[3] = {dir.x, dir.y, dir.z};
    ocpu::Hit h = G.use_bvh ? G.soup.closest(o, d, tmin, tmax) : G.soup.brute(o, d, tmin, tmax);
    T.rays++;
    if (T.tracesThisPixel++ == 0) {
        T.firstPrim = h.prim;
        T.firstU = h.u;
        T.firstV = h.v;
    }
    const void* savedPD = T.programData;
    void* savedPRD = T.prd;
    T.prd = prd;
    if (h.prim >= 0) {
        T.bary = make_float2(h.u, h.v);
        T.prim = (unsigned)G.primLocal[h.prim];
        T.programData = &G.meshes[G.primMesh[h.prim]];
        __closesthit__triangleMeshCH();
    } else {
        T.programData = &G.miss;
        __miss__miss();
    }
    T.programData = savedPD;
    T.prd = savedPRD;
}
}  // namespace owl

float2 optixGetTriangleBarycentrics() { return T.bary; }
unsigned int optixGetPrimitiveIndex() { return T.prim; }
float4 shim_tex2d(cudaTextureObject_t tex, float x, float y) {
    const Texture& t = G.textures[(size_t)tex - 1];
    float out[4];
    if (t.kind == 0)
        ocpu::tex2d_linear_clamp_f4((const float*)t.data, t.w, t.h, x, y, G.lutQuant, out);
    else
        ocpu::tex2d_nearest_clamp_rgba8((const uint8_t*)t.data, t.w, t.h, x, y, out);
    return make_float4(out[0], out[1], out[2], out[3]);
}

static vec3f v3(const float* p) { return vec3f(p[0], p[1], p[2]); }

extern "C" {

const char* ora_kind(void) {
#ifdef REF_INTENDED
    return "reference-intended";
#else
    return "reference-literal";
#endif
}

void ora_reset(void) {
    for (auto* s : G.store) delete s;
    G.store.clear();
    G.meshes.clear();
    G.primMesh.clear();
    G.primLocal.clear();
    G.soup.clear();
    G.textures.clear();
    G.lights.clear();
    for (auto& l : G.lut) l.clear();
    G.miss.const_color = vec3f(0.f);  // include/viewer.hpp:530
    memset((void*)&optixLaunchParams, 0, sizeof(optixLaunchParams));
    shim_shadow_spp = 4;
    shim_max_depth = 1;
}

int ora_add_mesh(const float* vtx, const float* nrm, const float* uv, int nVtx, const int32_t* idx, int nTri,
                 const float* diffuse3, float alpha, const float* emit3, int isLight, uint32_t materialID,
                 const uint8_t* tex, int texW, int texH) {
    MeshStore* s = new MeshStore();
    s->vtx.assign(vtx, vtx + 3 * (size_t)nVtx);
    s->nrm.assign(nrm, nrm + 3 * (size_t)nVtx);
    s->uv.assign(uv, uv + 2 * (size_t)nVtx);
    s->idx.assign(idx, idx + 3 * (size_t)nTri);
    if (tex) {
        s->tex.assign(tex, tex + 4 * (size_t)texW * texH);
        s->texW = texW;
        s->texH = texH;
    }
    G.store.push_back(s);
    TriangleMeshData m;
    memset((void*)&m, 0, sizeof(m));
    m.vertex = (vec3f*)s->vtx.data();
    m.normal = (vec3f*)s->nrm.data();
    m.index = (owl::common::vec3i*)s->idx.data();
    m.texCoord = (owl::common::vec2f*)s->uv.data();
    m.isLight = isLight != 0;
    m.emit = v3(emit3);
    m.materialID = materialID;
    m.diffuse = v3(diffuse3);
    m.alpha = alpha;
    m.hasDiffuseTexture = tex != nullptr;
    if (tex) {
        Texture t = {1, s->tex.data(), texW, texH};
        G.textures.push_back(t);
        m.diffuse_texture = (cudaTextureObject_t)G.textures.size();
    }
    int meshIdx = (int)G.meshes.size();
    G.meshes.push_back(m);
    for (int p = 0; p < nTri; p++) {
        G.primMesh.push_back(meshIdx);
        G.primLocal.push_back(p);
        for (int k = 0; k < 3; k++) {
            int vi = idx[3 * p + k];
            for (int a = 0; a < 3; a++) G.soup.v.push_back(vtx[3 * (size_t)vi + a]);
        }
    }
    return meshIdx;
}

void ora_set_lights(const float* triLights, int n) {
    static_assert(sizeof(TriLight) == 76, "TriLight layout");
    G.lights.resize(n);
    if (n) memcpy((void*)G.lights.data(), triLights, sizeof(TriLight) * (size_t)n);
    optixLaunchParams.triLights = G.lights.data();
    optixLaunchParams.numTriLights = n;
}

void ora_set_ltc_lut(const float* m1, const float* m2, const float* m3, int w, int h, int quantise) {
    const float* src[3] = {m1, m2, m3};
    cudaTextureObject_t* dst[3] = {&optixLaunchParams.ltc_1, &optixLaunchParams.ltc_2, &optixLaunchParams.ltc_3};
    G.lutW = w;
    G.lutH = h;
    G.lutQuant = quantise;
    for (int i = 0; i < 3; i++) {
        G.lut[i].assign(src[i], src[i] + 4 * (size_t)w * h);
        Texture t = {0, G.lut[i].data(), w, h};
        G.textures.push_back(t);
        *dst[i] = (cudaTextureObject_t)G.textures.size();
    }
}

void ora_commit(int use_bvh) {
    G.use_bvh = use_bvh != 0;
    if (G.use_bvh) G.soup.build();
    // textures were registered with pointers into vectors that no longer move
}

void ora_set_camera(const float* pos, const float* d00, const float* du, const float* dv) {
    optixLaunchParams.camera.pos = v3(pos);
    optixLaunchParams.camera.dir_00 = v3(d00);
    optixLaunchParams.camera.dir_du = v3(du);
    optixLaunchParams.camera.dir_dv = v3(dv);
}

int ora_set_options(int mode, int shadow_spp, int max_depth) {
#ifdef REF_INTENDED
    if (mode != 1) return -1;
    shim_shadow_spp = shadow_spp;
    shim_max_depth = max_depth;
    return 0;
#else
    if (mode != 0 || shadow_spp != 4 || max_depth != 1) return -1;
    return 0;
#endif
}

long long ora_render(int W, int H, int rendererType, int accumId, int y0, int y1, const OraBuffers* b,
                     int nthreads) {
    size_t npx = (size_t)W * H;
    std::vector<std::vector<float>> scratch;
    auto buf = [&](float* p) -> float4* {
        if (p) return (float4*)p;
        scratch.emplace_back(4 * npx, 0.f);
        return (float4*)scratch.back().data();
    };
    LaunchParams& lp = optixLaunchParams;
    lp.position_screen_buffer = buf(b->position);
    lp.normal_screen_buffer = buf(b->normal);
    lp.uv_screen_buffer = buf(b->uv);
    lp.albedo_screen_buffer = buf(b->albedo);
    lp.alpha_screen_buffer = buf(b->alpha);
    lp.materialID_screen_buffer = buf(b->materialID);
    lp.bounce0_screen_buffer = buf(b->bounce0);
    lp.bounce1_screen_buffer = buf(b->bounce1);
    lp.bounce2_screen_buffer = buf(b->bounce2);
    lp.ltc_screen_buffer = buf(b->ltc);
    lp.sto_direct_ratio_screen_buffer = buf(b->sto_direct);
    lp.sto_no_vis_ratio_screen_buffer = buf(b->sto_no_vis);
    lp.accum_screen_buffer = buf(b->accum);
    lp.accumId = accumId;
    lp.rendererType = rendererType;
    std::vector<uint32_t> fbScratch;
    uint32_t* fb = b->rgba8;
    if (!fb) {
        fbScratch.resize(npx);
        fb = fbScratch.data();
    }
    RayGenData rg;
    rg.frameBuffer = fb;
    rg.frameBufferSize = vec2i(W, H);

    if (nthreads < 1) nthreads = 1;
    std::atomic<long long> rays(0);
    std::atomic<int> nextRow(y0);
    auto worker = [&]() {
        T = Tls();
        T.programData = &rg;
        for (;;) {
            int y = nextRow.fetch_add(1);
            if (y >= y1) break;
            for (int x = 0; x < W; x++) {
                T.launch = vec2i(x, y);
                T.tracesThisPixel = 0;
                T.firstPrim = -1;
                T.firstU = T.firstV = 0.f;
                __raygen__rayGen();
                size_t o = (size_t)y * W + x;
                if (b->hit_prim) b->hit_prim[o] = T.firstPrim;
                if (b->hit_uv) {
                    b->hit_uv[2 * o] = T.firstU;
                    b->hit_uv[2 * o + 1] = T.firstV;
                }
            }
        }
        rays += T.rays;
    };
    if (nthreads == 1) {
        worker();
    } else {
        std::vector<std::thread> th;
        for (int i = 0; i < nthreads; i++) th.emplace_back(worker);
        for (auto& t : th) t.join();
    }
    return rays.load();
}

void ora_ltc_gbuffer(int n, const float* position, const float* normal, const float* albedo, const float* alpha,
                     const float* cam_pos, float* out, int nthreads) {
    if (nthreads < 1) nthreads = 1;
    std::atomic<int> next(0);
    const int chunk = 1024;
    auto worker = [&]() {
        for (;;) {
            int s = next.fetch_add(chunk);
            if (s >= n) break;
            int e = std::min(n, s + chunk);
            for (int i = s; i < e; i++) {
                vec3f nrm = v3(normal + 4 * (size_t)i);
                vec3f c(0.f);
                if (!(nrm.x == 0.f && nrm.y == 0.f && nrm.z == 0.f)) {
                    // what rayGen sets up before the mode switch, cuda_src/deviceCode.cu:78-85
                    SurfaceInteraction si;
                    si.hit = true;
                    si.p = v3(position + 4 * (size_t)i);
                    si.n_geom = nrm;
                    si.diffuse = v3(albedo + 4 * (size_t)i);
                    si.alpha = alpha[4 * (size_t)i];
                    si.wo = owl::normalize(v3(cam_pos) - si.p);
                    orthonormalBasis(si.n_geom, si.to_local, si.to_world);
                    si.wo_local = normalize(apply_mat(si.to_local, si.wo));
                    LCGRand rng;
                    rng.state = 0;
                    c = ltcDirectLighingBaseline(si, rng);
                }
                out[4 * (size_t)i + 0] = c.x;
                out[4 * (size_t)i + 1] = c.y;
                out[4 * (size_t)i + 2] = c.z;
                out[4 * (size_t)i + 3] = 1.f;
            }
        }
    };
    if (nthreads == 1) {
        worker();
    } else {
        std::vector<std::thread> th;
        for (int i = 0; i < nthreads; i++) th.emplace_back(worker);
        for (auto& t : th) t.join();
    }
}

// ---- unit entry points ------------------------------------------------------------------------
void ora_u_rng(int frame_id, uint32_t px, uint32_t py, uint32_t W, uint32_t H, int n, uint32_t* state0,
               uint32_t* raw, float* flt) {
    LCGRand rng = get_rng(frame_id, make_uint2(px, py), make_uint2(W, H));
    *state0 = rng.state;
    for (int i = 0; i < n; i++) raw[i] = lcg_random(rng);
    for (int i = 0; i < n; i++) flt[i] = lcg_randomf_raw(rng);
}

int ora_u_clip(int vertex_count, float* v15) {
    vec3f v[5];
    for (int i = 0; i < 5; i++) v[i] = v3(v15 + 3 * i);
    int vc = clipPolygon(vertex_count, v);
    for (int i = 0; i < 5; i++) {
        v15[3 * i] = v[i].x;
        v15[3 * i + 1] = v[i].y;
        v15[3 * i + 2] = v[i].z;
    }
    return vc;
}

float ora_u_integrate_edge(const float* a, const float* b) { return integrateEdge(v3(a), v3(b)); }

void ora_u_fetch_ltc(float alpha, float theta, float* m9_amp) {
    vec3f m[3];
    float amp = 1.f;
    fetchLtcMat(alpha, theta, m, amp);
    for (int i = 0; i < 3; i++) {
        m9_amp[3 * i] = m[i].x;
        m9_amp[3 * i + 1] = m[i].y;
        m9_amp[3 * i + 2] = m[i].z;
    }
    m9_amp[9] = amp;
}

void ora_u_mat_inverse(const float* m9, float* out9) {
    vec3f m[3] = {v3(m9), v3(m9 + 3), v3(m9 + 6)}, inv[3];
    matrixInverse(m, inv);
    for (int i = 0; i < 3; i++) {
        out9[3 * i] = inv[i].x;
        out9[3 * i + 1] = inv[i].y;
        out9[3 * i + 2] = inv[i].z;
    }
}

void ora_u_onb(const float* n, float* mat9_inv9) {
    vec3f m[3], inv[3];
    orthonormalBasis(v3(n), m, inv);
    for (int i = 0; i < 3; i++) {
        mat9_inv9[3 * i] = m[i].x;
        mat9_inv9[3 * i + 1] = m[i].y;
        mat9_inv9[3 * i + 2] = m[i].z;
        mat9_inv9[9 + 3 * i] = inv[i].x;
        mat9_inv9[9 + 3 * i + 1] = inv[i].y;
        mat9_inv9[9 + 3 * i + 2] = inv[i].z;
    }
}

void ora_u_sample_direction(const float* wo, float u1, float u2, const float* base, float metal, float alpha,
                            float* wi_pdf4) {
    vec3f w = v3(wo);
    float p = 0.f;
    vec3f wi = sample_direction(w, u1, u2, &p, v3(base), metal, alpha);
    wi_pdf4[0] = wi.x;
    wi_pdf4[1] = wi.y;
    wi_pdf4[2] = wi.z;
    wi_pdf4[3] = p;
}

void ora_u_evaluate(const float* wi, const float* wo, const float* base, float alpha, float* out3) {
    vec3f a = v3(wi), b = v3(wo), c = v3(base);
    vec3f r = evaluate(a, b, c, alpha, vec3f(0.f));
    out3[0] = r.x;
    out3[1] = r.y;
    out3[2] = r.z;
}

float ora_u_pdf(const float* wi, const float* wo, const float* base, float metal, float alpha) {
    vec3f a = v3(wi), b = v3(wo);
    return pdf(a, b, v3(base), metal, alpha);
}

void ora_u_sample_tri(const float* v1, const float* v2, const float* v3_, float u1, float u2, float* out3) {
    vec3f r = samplePointOnTriangle(v3(v1), v3(v2), v3(v3_), u1, u2);
    out3[0] = r.x;
    out3[1] = r.y;
    out3[2] = r.z;
}

float ora_u_pdf_a2w(float pdf_, float dist2, float cos_theta) { return pdfA2W(pdf_, dist2, cos_theta); }

uint32_t ora_u_make_rgba(const float* c3) { return owl::make_rgba(v3(c3)); }

void ora_ratio_combine(long long n, const float* ltc, const float* sto_direct, const float* sto_no_vis, float* out) {
    // ltc_ratio_estimator.py:8 (Python; restated, not compiled from the reference)
    for (long long i = 0; i < n; i++) out[i] = ltc[i] * (sto_direct[i] / sto_no_vis[i]);
}

}  // extern "C"

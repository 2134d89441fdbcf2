// hostcheck.cpp -- TEST INFRASTRUCTURE ONLY.  Never shipped, never loaded by the product.
//
// Compiles the product's per-pixel device functions (optix_renderer_b200/csrc/*.cuh, which are
// __host__ __device__) with g++ so that tests can compare them with the oracle on a machine
// without a GPU.  It re-creates only the GLUE of the CUDA kernels in orb200.cu (loops over
// pixels, buffer reads/writes) in plain C++; the arithmetic is the product's own headers.
// A small median-split builder emits the same 64-byte node layout the GPU LBVH builder emits,
// so trace_closest() runs unchanged.
//   g++ -O2 -mfma -ffp-contract=off -std=c++17 -shared -fPIC -I/usr/local/cuda/include
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../optix_renderer_b200/csrc/bvhq.cuh"
#include "../../optix_renderer_b200/csrc/pixel.cuh"

using namespace orb;

namespace {
struct State {
    std::vector<float> pos, nrm, uv;
    std::vector<int> prim_mesh;
    std::vector<MeshRec> meshes;
    std::vector<std::vector<uint8_t>> tex;
    std::vector<TriLight> lights;
    std::vector<float4> lut;
    int lut_w = 0, lut_h = 0, lut_q = 0;
    std::vector<float4> tri;
    std::vector<float2> tri_b;
    std::vector<BvhNode> nodes;
    int root_ref = 0;
    std::vector<QNode> qnodes;  // bvhq.cuh: the same hierarchy on the 16-bit grid
    QGrid qgrid;
    Camera cam;
    SceneView view;
} G;

struct Box {
    float lo[3], hi[3];
};

Box bounds(const std::vector<int>& idx, int first, int count) {
    Box b;
    for (int a = 0; a < 3; a++) { b.lo[a] = INFINITY; b.hi[a] = -INFINITY; }
    for (int i = first; i < first + count; i++)
        for (int k = 0; k < 3; k++)
            for (int a = 0; a < 3; a++) {
                const float v = G.pos[9 * (size_t)idx[i] + 3 * k + a];
                b.lo[a] = std::min(b.lo[a], v);
                b.hi[a] = std::max(b.hi[a], v);
            }
    return b;
}

// returns a child reference for [first, first+count)
int build_rec(std::vector<int>& idx, int first, int count, int max_leaf) {
    if (count <= max_leaf) return leaf_ref(first, count);
    Box b = bounds(idx, first, count);
    int axis = 0;
    for (int a = 1; a < 3; a++)
        if (b.hi[a] - b.lo[a] > b.hi[axis] - b.lo[axis]) axis = a;
    const int mid = first + count / 2;
    std::nth_element(idx.begin() + first, idx.begin() + mid, idx.begin() + first + count, [&](int x, int y) {
        const float cx = G.pos[9 * (size_t)x + axis] + G.pos[9 * (size_t)x + 3 + axis] + G.pos[9 * (size_t)x + 6 + axis];
        const float cy = G.pos[9 * (size_t)y + axis] + G.pos[9 * (size_t)y + 3 + axis] + G.pos[9 * (size_t)y + 6 + axis];
        return cx < cy || (cx == cy && x < y);
    });
    const int me = (int)G.nodes.size();
    G.nodes.push_back(BvhNode());
    const int l = build_rec(idx, first, mid - first, max_leaf);
    const int r = build_rec(idx, mid, first + count - mid, max_leaf);
    const Box bl = bounds(idx, first, mid - first), br = bounds(idx, mid, first + count - mid);
    const float pad = 1e-5f;
    BvhNode& n = G.nodes[me];
    n.a = make_float4(bl.lo[0] - pad, bl.lo[1] - pad, bl.lo[2] - pad, bl.hi[0] + pad);
    n.b = make_float4(bl.hi[1] + pad, bl.hi[2] + pad, br.lo[0] - pad, br.lo[1] - pad);
    n.c = make_float4(br.lo[2] - pad, br.hi[0] + pad, br.hi[1] + pad, br.hi[2] + pad);
    n.c0 = l;
    n.c1 = r;
    n.pad0 = n.pad1 = 0;
    return me;
}

void refresh() {
    SceneView& v = G.view;
    v.tri = G.tri.data();
    v.tri_b = G.tri_b.data();
    v.nodes = G.nodes.data();
    v.root_ref = G.root_ref;
    v.n_tris = (int)G.tri_b.size();
    v.pos = G.pos.data();
    v.nrm = G.nrm.data();
    v.uv = G.uv.data();
    v.prim_mesh = G.prim_mesh.data();
    v.meshes = G.meshes.data();
    v.lights = G.lights.data();
    v.n_lights = (int)G.lights.size();
    v.lut.data = G.lut.data();
    v.lut.w = G.lut_w;
    v.lut.h = G.lut_h;
    v.lut.quantise = G.lut_q;
}

struct HostTracer {
    mutable long long rays = 0;
    mutable unsigned long long nodes = 0, tris = 0;  // modes 0 and 5: node records fetched / triangles tested
    int brute = 0;
    Hit operator()(float3 o, float3 d) const {
        rays++;
        if (brute == 1) return trace_brute(G.view, o, d);
        unsigned nn = 0, nt = 0;
        if (brute == 6 || brute == 7) {  // the hybrid stack of the GPU kernels: first 3 entries in a "shared" column, rest local
            unsigned long long col[3];
            HybridStack<3> hs;
            hs.s = col;
            hs.stride = 1;
            const Hit hh = (brute == 7 && G.root_ref >= 0 && !G.qnodes.empty())
                               ? trace_closest_q(G.view, G.qnodes.data(), G.qgrid, o, d, hs, &nn, &nt)
                               : trace_closest(G.view, o, d, hs, &nn, &nt);
            nodes += nn;
            tris += nt;
            return hh;
        }
        int stk[128];
        float stk_t[128];
        Hit h;
        if (brute == 5 && G.root_ref >= 0 && !G.qnodes.empty())  // 32-byte quantised nodes (a leaf root has no nodes at all)
            h = trace_closest_q(G.view, G.qnodes.data(), G.qgrid, o, d, stk, stk_t, 1, &nn, &nt);
        else
            h = trace_closest(G.view, o, d, stk, stk_t, 1, &nn, &nt);
        nodes += nn;
        tris += nt;
        return h;
    }
};
}  // namespace

extern "C" {

void hc_reset() { G = State(); }

void hc_add_mesh(const float* vtx, const float* nrm, const float* uv, int nVtx, const int32_t* idx, int nTri,
                 const float* diffuse3, float alpha, const float* emit3, int isLight, uint32_t materialID,
                 const uint8_t* tex, int texW, int texH) {
    (void)nVtx;
    const int mesh = (int)G.meshes.size();
    for (int t = 0; t < nTri; t++) {
        for (int k = 0; k < 3; k++) {
            const int vi = idx[3 * t + k];
            for (int a = 0; a < 3; a++) {
                G.pos.push_back(vtx[3 * (size_t)vi + a]);
                G.nrm.push_back(nrm[3 * (size_t)vi + a]);
            }
        }
        for (int k = 0; k < 3; k++) {
            const int vi = idx[3 * t + k];
            G.uv.push_back(uv[2 * (size_t)vi]);
            G.uv.push_back(uv[2 * (size_t)vi + 1]);
        }
        G.prim_mesh.push_back(mesh);
    }
    MeshRec r;
    memset(&r, 0, sizeof(r));
    r.diffuse = v3(diffuse3[0], diffuse3[1], diffuse3[2]);
    r.alpha = alpha;
    r.emit = v3(emit3[0], emit3[1], emit3[2]);
    r.is_light = isLight != 0;
    r.material_id = materialID;
    G.tex.emplace_back();
    if (tex) {
        G.tex.back().assign(tex, tex + 4 * (size_t)texW * texH);
        r.tex_w = texW;
        r.tex_h = texH;
    }
    G.meshes.push_back(r);
}

void hc_set_lights(const float* l19, int n) {
    G.lights.resize(n);
    if (n) memcpy((void*)G.lights.data(), l19, sizeof(TriLight) * (size_t)n);
    refresh();
}

void hc_set_ltc_lut(const float* m1, const float* m2, const float* m3, int w, int h, int quantise) {
    const size_t n = (size_t)w * h;
    G.lut.resize(3 * n);
    memcpy((void*)G.lut.data(), m1, 16 * n);
    memcpy((void*)(G.lut.data() + n), m2, 16 * n);
    memcpy((void*)(G.lut.data() + 2 * n), m3, 16 * n);
    G.lut_w = w;
    G.lut_h = h;
    G.lut_q = quantise;
    refresh();
}

void hc_commit(int max_leaf) {
    for (size_t m = 0; m < G.meshes.size(); m++)
        G.meshes[m].tex = G.tex[m].empty() ? nullptr : (const uchar4*)G.tex[m].data();
    const int n = (int)G.prim_mesh.size();
    std::vector<int> idx(n);
    for (int i = 0; i < n; i++) idx[i] = i;
    G.nodes.clear();
    G.root_ref = n ? build_rec(idx, 0, n, max_leaf > 0 ? max_leaf : 4) : 0;
    G.tri.resize(2 * (size_t)n);
    G.tri_b.resize((size_t)n);
    for (int k = 0; k < n; k++) {
        float a[8], b[2];
        tri_record(a, b, &G.pos[9 * (size_t)idx[k]], idx[k]);  // the product's own record writer (scene.cuh)
        G.tri[2 * (size_t)k] = make_float4(a[0], a[1], a[2], a[3]);
        G.tri[2 * (size_t)k + 1] = make_float4(a[4], a[5], a[6], a[7]);
        G.tri_b[k] = make_float2(b[0], b[1]);
    }
    G.qnodes.clear();
    if (n && G.root_ref >= 0 && !G.nodes.empty()) {
        G.qgrid = qgrid_from_root(G.nodes[G.root_ref]);
        G.qnodes.resize(G.nodes.size());
        for (size_t i = 0; i < G.nodes.size(); i++) G.qnodes[i] = q_encode(G.nodes[i], G.qgrid);
    }
    refresh();
}

void hc_set_camera(const float* pos, const float* d00, const float* du, const float* dv) {
    G.cam.pos = v3(pos[0], pos[1], pos[2]);
    G.cam.d00 = v3(d00[0], d00[1], d00[2]);
    G.cam.du = v3(du[0], du[1], du[2]);
    G.cam.dv = v3(dv[0], dv[1], dv[2]);
}

void hc_trace(const float* org, const float* dir, int n, int brute, int32_t* prim, float* t, float* u, float* v) {
    HostTracer tr;
    tr.brute = brute;
    for (int i = 0; i < n; i++) {
        const Hit h = tr(v3(org[3 * i], org[3 * i + 1], org[3 * i + 2]), v3(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]));
        prim[i] = h.prim;
        t[i] = h.t;
        u[i] = h.u;
        v[i] = h.v;
    }
}

// hc_trace plus the traversal's fetch counts (modes 0 and 5)
void hc_trace_counted(const float* org, const float* dir, int n, int mode, int32_t* prim, float* t, float* u, float* v,
                      unsigned long long* nodes, unsigned long long* tris) {
    HostTracer tr;
    tr.brute = mode;
    for (int i = 0; i < n; i++) {
        const Hit h = tr(v3(org[3 * i], org[3 * i + 1], org[3 * i + 2]), v3(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]));
        prim[i] = h.prim;
        t[i] = h.t;
        u[i] = h.u;
        v[i] = h.v;
    }
    *nodes = tr.nodes;
    *tris = tr.tris;
}

// Same buffer set and order as OraBuffers (oracle/oracle_api.h).
struct HcBuffers {
    float *position, *normal, *uv, *albedo, *alpha, *materialID, *bounce0, *bounce1, *bounce2, *ltc, *sto_direct,
        *sto_no_vis, *accum;
    uint32_t* rgba8;
    int32_t* hit_prim;
    float* hit_uv;
};

static void put(float* b, size_t o, float x, float y, float z) {
    b[4 * o] = x;
    b[4 * o + 1] = y;
    b[4 * o + 2] = z;
    b[4 * o + 3] = 1.f;
}

// The wavefront of orb200.cu, serialised per pixel.  renderer = OrbRendererType values.
long long hc_render(int W, int H, int renderer, int accum_id, int mode, int spp, int depth, const HcBuffers* b) {
    HostTracer tr;
    const int frame_id = accum_id + 10007;
    if (mode == 0) { spp = 4; depth = 1; }
    for (int py = 0; py < H; py++)
        for (int px = 0; px < W; px++) {
            const size_t o = (size_t)py * W + px;
            float3 org, dir;
            primary_ray(G.cam, frame_id, px, py, W, H, org, dir);
            const Hit h = tr(org, dir);
            const Surface si = shade_hit(G.view, h.prim, h.u, h.v);
            b->hit_prim[o] = h.prim;
            b->hit_uv[2 * o] = h.u;
            b->hit_uv[2 * o + 1] = h.v;
            put(b->position, o, si.p.x, si.p.y, si.p.z);
            put(b->normal, o, si.n.x, si.n.y, si.n.z);
            put(b->albedo, o, si.diffuse.x, si.diffuse.y, si.diffuse.z);
            put(b->alpha, o, si.alpha, 0.f, 0.f);
            put(b->uv, o, si.u, si.v, si.diffuse.z);
            put(b->materialID, o, (float)si.material_id, 0.f, 0.f);
            float3 c = v3(0.f);
            if (!si.hit) c = si.n;
            else if (renderer == 5) c = v3(1.f);
            else if (renderer == 4) c = si.p;
            else if (renderer == 0) c = si.diffuse;
            else if (renderer == 1) c = v3(si.alpha);
            else if (renderer == 2) c = si.n;
            else if (renderer == 3) c = v3(0.f);
            else if (renderer == 6) c = v3((float)si.material_id);
            else if (si.is_light) {
                c = si.emit;
                if (renderer == 8) {
                    put(b->ltc, o, c.x, c.y, c.z);
                    put(b->sto_direct, o, c.x, c.y, c.z);
                    put(b->sto_no_vis, o, c.x, c.y, c.z);
                }
            } else if (renderer == 7 || renderer == 8) {
                c = ltc_pixel(G.view.lut, G.view.lights, G.view.n_lights, G.cam.pos, si.p, si.n, si.diffuse, si.alpha);
                if (renderer == 8) {
                    put(b->ltc, o, c.x, c.y, c.z);
                    Rng rng = rng_seed(frame_id, px, py, W);
                    rng_next(rng);
                    rng_next(rng);
                    float3 sd = v3(0.f), sn = v3(0.f);
                    for (int k = 0; k < spp; k++) {
                        float3 d, nv;
                        ratio_shadow_sample(G.view, tr, G.cam.pos, si.p, si.n, si.diffuse, si.alpha, rng, mode, d, nv);
                        sd += d;
                        sn += nv;
                    }
                    sd = sd / (float)spp;
                    sn = sn / (float)spp;
                    const float gd = (sd.x + sd.y + sd.z) / 3.f, gn = (sn.x + sn.y + sn.z) / 3.f;
                    put(b->sto_direct, o, gd, gd, gd);
                    put(b->sto_no_vis, o, gn, gn, gn);
                }
            } else if (renderer == 9) {
                PathState st;
                st.p = si.p;
                st.n = si.n;
                st.diffuse = si.diffuse;
                st.alpha = si.alpha;
                st.V = normalize(G.cam.pos - si.p);
                st.tp = v3(1.f);
                c = v3(ORB_EPS);
                for (int bo = 0; bo < depth; bo++) {
                    Rng rng = rng_seed(frame_id, px, py, W);
                    rng_skip(rng, 2 + 5 * bo);
                    bool traced;
                    if (path_bounce(G.view, tr, st, rng, mode, bo == depth - 1, c, traced) == ORB_PATH_DONE) break;
                }
            } else {
                c = v3(1.f, 0.f, 0.f);
            }
            if (accum_id > 0) c = c + v3(b->accum[4 * o], b->accum[4 * o + 1], b->accum[4 * o + 2]);
            put(b->accum, o, c.x, c.y, c.z);
            b->rgba8[o] = make_rgba((1.f / ((float)accum_id + 1.f)) * c);
        }
    return tr.rays;
}

void hc_ltc_gbuffer(int n, const float* position, const float* normal, const float* albedo, const float* alpha,
                    const float* cam_pos, float* out) {
    for (int i = 0; i < n; i++) {
        const float* nn = normal + 4 * (size_t)i;
        float3 c = v3(0.f);
        if (!(nn[0] == 0.f && nn[1] == 0.f && nn[2] == 0.f)) {
            const float* p = position + 4 * (size_t)i;
            const float* d = albedo + 4 * (size_t)i;
            c = ltc_pixel(G.view.lut, G.view.lights, G.view.n_lights, v3(cam_pos[0], cam_pos[1], cam_pos[2]),
                          v3(p[0], p[1], p[2]), v3(nn[0], nn[1], nn[2]), v3(d[0], d[1], d[2]), alpha[4 * (size_t)i]);
        }
        put(out, i, c.x, c.y, c.z);
    }
}

// ---- unit entry points (same meaning as the orb_unit_* functions of include/orb200.h) ----------
void hc_u_rng(int frame_id, uint32_t px, uint32_t py, uint32_t W, int n, uint32_t* state0, uint32_t* raw, float* flt) {
    Rng r = rng_seed(frame_id, px, py, W);
    *state0 = r.s;
    for (int i = 0; i < n; i++) raw[i] = rng_next(r);
    for (int i = 0; i < n; i++) flt[i] = rng_nextf(r);
}
uint32_t hc_u_rng_skip(int frame_id, uint32_t px, uint32_t py, uint32_t W, int n) {
    Rng r = rng_seed(frame_id, px, py, W);
    rng_skip(r, n);
    return r.s;
}
float hc_u_clip(const float* abc9, int quad, int* n_out) {
    return clipped_form_factor(v3(abc9[0], abc9[1], abc9[2]), v3(abc9[3], abc9[4], abc9[5]), v3(abc9[6], abc9[7], abc9[8]),
                               quad != 0, *n_out);
}
float hc_u_integrate_edge(const float* a, const float* b) {
    return edge_integral(v3(a[0], a[1], a[2]), v3(b[0], b[1], b[2]));
}
static void put9(float* o, const Mat3& m) {
    o[0] = m.r0.x; o[1] = m.r0.y; o[2] = m.r0.z;
    o[3] = m.r1.x; o[4] = m.r1.y; o[5] = m.r1.z;
    o[6] = m.r2.x; o[7] = m.r2.y; o[8] = m.r2.z;
}
void hc_u_fetch_ltc(float alpha, float theta, float* m9_amp) {
    Mat3 m;
    float amp;
    G.view.lut.fetch(alpha, theta, m, amp);
    put9(m9_amp, m);
    m9_amp[9] = amp;
}
void hc_u_mat_inverse(const float* m9, float* out9) {
    Mat3 m;
    m.r0 = v3(m9[0], m9[1], m9[2]);
    m.r1 = v3(m9[3], m9[4], m9[5]);
    m.r2 = v3(m9[6], m9[7], m9[8]);
    put9(out9, mat3_inverse(m));
}
void hc_u_onb(const float* n, float* rows9) { put9(rows9, onb_rows(v3(n[0], n[1], n[2]))); }
void hc_u_sample_direction(const float* wo, float u1, float u2, const float* base, float alpha, float* wi_pdf4) {
    float pdf = 0.f;
    const float3 wi = bsdf_sample(v3(wo[0], wo[1], wo[2]), u1, u2, pdf, v3(base[0], base[1], base[2]), alpha);
    wi_pdf4[0] = wi.x; wi_pdf4[1] = wi.y; wi_pdf4[2] = wi.z; wi_pdf4[3] = pdf;
}
void hc_u_evaluate(const float* wi, const float* wo, const float* base, float alpha, float* out3) {
    const float3 r = bsdf_evaluate(v3(wi[0], wi[1], wi[2]), v3(wo[0], wo[1], wo[2]), v3(base[0], base[1], base[2]), alpha);
    out3[0] = r.x; out3[1] = r.y; out3[2] = r.z;
}
float hc_u_pdf(const float* wi, const float* wo, const float* base, float alpha) {
    return bsdf_pdf(v3(wi[0], wi[1], wi[2]), v3(wo[0], wo[1], wo[2]), v3(base[0], base[1], base[2]), alpha);
}
uint32_t hc_u_make_rgba(const float* c) { return make_rgba(v3(c[0], c[1], c[2])); }
void hc_u_sample_tri(const float* v9, float u1, float u2, float* out3) {
    const float3 r = sample_point_on_triangle(v3(v9[0], v9[1], v9[2]), v3(v9[3], v9[4], v9[5]), v3(v9[6], v9[7], v9[8]), u1, u2);
    out3[0] = r.x; out3[1] = r.y; out3[2] = r.z;
}
float hc_u_pdf_a2w(float pdf, float d2, float c) { return pdf_area_to_solid_angle(pdf, d2, c); }

}  // extern "C"

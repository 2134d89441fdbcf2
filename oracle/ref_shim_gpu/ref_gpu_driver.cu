// ref_gpu_driver.cu -- TEST INFRASTRUCTURE ONLY (oracle/_ref/libref_gpu_{literal,intended}.so).
//
// Compiles the REFERENCE's own device translation unit (REF_DEVICE_TU, normally
// /root/reference/cuda_src/deviceCode.cu, which pulls in cuda_include/**) with nvcc for sm_100a behind
// owl_shim_gpu.h and launches __raygen__rayGen once per pixel: the reference's literal megakernel on the
// B200 (SURVEY.md 8(d) "GPU baseline to beat").  No reference source is copied: the TU is
// #included from where it lies (oracle/Makefile passes the path).
//
// Hooks supplied here = the un-vendored OWL / OptiX layer:
//   traceRay  -> single-ray closest hit over THIS repository's BVH (scene.cuh trace_closest over a binned-SAH
//                BVH2, as SURVEY.md 8(d) prescribes), then __closesthit__triangleMeshCH / __miss__miss;
//   tex2D     -> the GPU's texture unit: float4 linear / clamp / normalised coordinates for the three LTC
//                tables (viewer.hpp:321-327), RGBA8 point / clamp for diffuse textures (:441-447).
// Implements the scene / frame part of oracle_api.h (no unit entry points, no denoise) plus ora_last_ms().
#include <cstdio>
#include <cstring>
#include <vector>

// the product's traversal and builder FIRST: their float3 operators (namespace orb) must be bound before the
// reference's helper_math.cuh puts the same operators into the global namespace
#include "../../optix_renderer_b200/csrc/scene.cuh"
#include "../../optix_renderer_b200/csrc/bvh_build.cuh"
#include "../../optix_renderer_b200/csrc/bvh_sah.cuh"
#include "../oracle_api.h"
#include "owl_shim_gpu.h"

#include REF_DEVICE_TU

using owl::common::vec2i;
using owl::common::vec3f;

// ---- device-side scene ------------------------------------------------------------------------------
struct GpuScene {
    orb::SceneView view;  // tri, tri_b, nodes, root_ref, n_tris are used by the traversal; the rest stays null
    const TriangleMeshData* meshes;
    const int* primMesh;
    const int* primLocal;
    MissProgData miss;
    RayGenData raygen;
    int32_t* hit_prim;
    float* hit_uv;
    unsigned long long* rays;
};
__constant__ GpuScene g_scene;

namespace owl {
__device__ void shim_trace(const common::vec3f& org, const common::vec3f& dir, void* prd) {
    int stk[64];
    float stk_t[64];
    const orb::Hit h = orb::trace_closest(g_scene.view, make_float3(org.x, org.y, org.z), make_float3(dir.x, dir.y, dir.z),
                                          stk, stk_t, 1);
    ShimCtx& c = shim_ctx();
    if (c.traces++ == 0) {
        c.firstPrim = h.prim;
        c.firstU = h.u;
        c.firstV = h.v;
    }
    const void* savedPD = c.programData;
    void* savedPRD = c.prd;
    c.prd = prd;
    if (h.prim >= 0) {
        c.bary = make_float2(h.u, h.v);
        c.prim = (unsigned)g_scene.primLocal[h.prim];
        c.programData = &g_scene.meshes[g_scene.primMesh[h.prim]];
        __closesthit__triangleMeshCH();
    } else {
        c.programData = &g_scene.miss;
        __miss__miss();
    }
    c.programData = savedPD;
    c.prd = savedPRD;
}
}  // namespace owl

__global__ void __launch_bounds__(SHIM_BLOCK_X* SHIM_BLOCK_Y) k_ref_raygen(int W, int H) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    owl::ShimCtx& c = owl::shim_ctx();
    c.programData = &g_scene.raygen;
    c.prd = nullptr;
    c.traces = 0;
    c.firstPrim = -1;
    c.firstU = c.firstV = 0.f;
    if (x < W && y < H) {
        __raygen__rayGen();
        const size_t o = (size_t)y * W + x;
        g_scene.hit_prim[o] = c.firstPrim;
        g_scene.hit_uv[2 * o] = c.firstU;
        g_scene.hit_uv[2 * o + 1] = c.firstV;
    }
    unsigned n = (unsigned)c.traces;
    for (int o = 16; o; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
    if (((threadIdx.y * blockDim.x + threadIdx.x) & 31) == 0 && n) atomicAdd(g_scene.rays, (unsigned long long)n);
}

// ---- host-side state --------------------------------------------------------------------------------
namespace {
#define CKR(x)                                                                                   \
    do {                                                                                         \
        cudaError_t e_ = (x);                                                                    \
        if (e_ != cudaSuccess) {                                                                 \
            fprintf(stderr, "ref_gpu: %s failed: %s\n", #x, cudaGetErrorString(e_));             \
            abort();                                                                             \
        }                                                                                        \
    } while (0)

struct HostState {
    std::vector<TriangleMeshData> meshes;
    std::vector<int> primMesh, primLocal;
    std::vector<float> pos9;  // un-indexed corners, global primitive order
    std::vector<void*> owned;  // device allocations
    std::vector<cudaArray_t> arrays;
    std::vector<cudaTextureObject_t> texs;
    LaunchParams lp;
    GpuScene gs;
    orb::BvhBuildResult bvh;
    float4* buf[13] = {};
    uint32_t* fb = nullptr;
    int W = 0, H = 0;
    float last_ms = 0.f;
    bool inited = false;
} S;

template <class T> T* upload(const T* src, size_t n) {
    T* d = nullptr;
    CKR(cudaMalloc(&d, sizeof(T) * (n ? n : 1)));
    if (n) CKR(cudaMemcpy(d, src, sizeof(T) * n, cudaMemcpyHostToDevice));
    S.owned.push_back(d);
    return d;
}

cudaTextureObject_t make_texture(const void* data, int w, int h, bool float4_linear) {
    cudaChannelFormatDesc fmt = float4_linear ? cudaCreateChannelDesc<float4>() : cudaCreateChannelDesc<uchar4>();
    cudaArray_t arr;
    CKR(cudaMallocArray(&arr, &fmt, w, h));
    const size_t pitch = (size_t)w * (float4_linear ? 16 : 4);
    CKR(cudaMemcpy2DToArray(arr, 0, 0, data, pitch, pitch, h, cudaMemcpyHostToDevice));
    cudaResourceDesc rd;
    memset(&rd, 0, sizeof(rd));
    rd.resType = cudaResourceTypeArray;
    rd.res.array.array = arr;
    cudaTextureDesc td;
    memset(&td, 0, sizeof(td));
    td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;
    td.filterMode = float4_linear ? cudaFilterModeLinear : cudaFilterModePoint;
    td.readMode = float4_linear ? cudaReadModeElementType : cudaReadModeNormalizedFloat;
    td.normalizedCoords = 1;
    cudaTextureObject_t t = 0;
    CKR(cudaCreateTextureObject(&t, &rd, &td, nullptr));
    S.arrays.push_back(arr);
    S.texs.push_back(t);
    return t;
}

void free_all() {
    for (auto t : S.texs) cudaDestroyTextureObject(t);
    for (auto a : S.arrays) cudaFreeArray(a);
    for (auto p : S.owned) cudaFree(p);
    S.texs.clear();
    S.arrays.clear();
    S.owned.clear();
    cudaFree(S.bvh.tri);
    cudaFree(S.bvh.tri_b);
    cudaFree(S.bvh.nodes);
    S.bvh = orb::BvhBuildResult();
    for (auto& b : S.buf) { cudaFree(b); b = nullptr; }
    cudaFree(S.fb);
    S.fb = nullptr;
    cudaFree(S.gs.hit_prim);
    cudaFree(S.gs.hit_uv);
    S.gs.hit_prim = nullptr;
    S.gs.hit_uv = nullptr;
    S.W = S.H = 0;
}

vec3f v3(const float* p) { return vec3f(p[0], p[1], p[2]); }
}  // namespace

extern "C" {

const char* ora_kind(void) {
#ifdef REF_INTENDED
    return "reference-gpu-intended";
#else
    return "reference-gpu-literal";
#endif
}

void ora_reset(void) {
    free_all();
    S.meshes.clear();
    S.primMesh.clear();
    S.primLocal.clear();
    S.pos9.clear();
    memset((void*)&S.lp, 0, sizeof(S.lp));
    memset((void*)&S.gs, 0, sizeof(S.gs));
    S.gs.miss.const_color = vec3f(0.f);  // include/viewer.hpp:530
    const int spp = 4, depth = 1;
    CKR(cudaMemcpyToSymbol(shim_shadow_spp, &spp, 4));
    CKR(cudaMemcpyToSymbol(shim_max_depth, &depth, 4));
    CKR(cudaDeviceSetLimit(cudaLimitStackSize, 16384));
}

int ora_add_mesh(const float* vtx, const float* nrm, const float* uv, int nVtx, const int32_t* idx, int nTri,
                 const float* diffuse3, float alpha, const float* emit3, int isLight, uint32_t materialID,
                 const uint8_t* tex, int texW, int texH) {
    TriangleMeshData m;
    memset((void*)&m, 0, sizeof(m));
    m.vertex = (vec3f*)upload(vtx, 3 * (size_t)nVtx);
    m.normal = (vec3f*)upload(nrm, 3 * (size_t)nVtx);
    m.index = (owl::common::vec3i*)upload(idx, 3 * (size_t)nTri);
    m.texCoord = (owl::common::vec2f*)upload(uv, 2 * (size_t)nVtx);
    m.isLight = isLight != 0;
    m.emit = v3(emit3);
    m.materialID = materialID;
    m.diffuse = v3(diffuse3);
    m.alpha = alpha;
    m.hasDiffuseTexture = tex != nullptr;
    if (tex) m.diffuse_texture = make_texture(tex, texW, texH, false);
    const int meshIdx = (int)S.meshes.size();
    S.meshes.push_back(m);
    for (int p = 0; p < nTri; p++) {
        S.primMesh.push_back(meshIdx);
        S.primLocal.push_back(p);
        for (int k = 0; k < 3; k++) {
            const int vi = idx[3 * p + k];
            for (int a = 0; a < 3; a++) S.pos9.push_back(vtx[3 * (size_t)vi + a]);
        }
    }
    return meshIdx;
}

void ora_set_lights(const float* triLights, int n) {
    static_assert(sizeof(TriLight) == 76, "TriLight layout");
    S.lp.triLights = (TriLight*)upload((const char*)triLights, sizeof(TriLight) * (size_t)n);
    S.lp.numTriLights = n;
}

void ora_set_ltc_lut(const float* m1, const float* m2, const float* m3, int w, int h, int /*quantise*/) {
    S.lp.ltc_1 = make_texture(m1, w, h, true);
    S.lp.ltc_2 = make_texture(m2, w, h, true);
    S.lp.ltc_3 = make_texture(m3, w, h, true);
}

void ora_commit(int /*use_bvh*/) {
    const int n = (int)S.primMesh.size();
    float* d_pos = upload(S.pos9.data(), S.pos9.size());
    if (n) CKR(orb::build_bvh_sah(d_pos, n, 4, 0, S.bvh));
    memset((void*)&S.gs.view, 0, sizeof(S.gs.view));
    S.gs.view.tri = S.bvh.tri;
    S.gs.view.tri_b = S.bvh.tri_b;
    S.gs.view.nodes = S.bvh.nodes;
    S.gs.view.root_ref = S.bvh.root_ref;
    S.gs.view.n_tris = n;
    S.gs.meshes = upload(S.meshes.data(), S.meshes.size());
    S.gs.primMesh = upload(S.primMesh.data(), S.primMesh.size());
    S.gs.primLocal = upload(S.primLocal.data(), S.primLocal.size());
    unsigned long long zero = 0;
    S.gs.rays = upload(&zero, 1);
}

void ora_set_camera(const float* pos, const float* d00, const float* du, const float* dv) {
    S.lp.camera.pos = v3(pos);
    S.lp.camera.dir_00 = v3(d00);
    S.lp.camera.dir_du = v3(du);
    S.lp.camera.dir_dv = v3(dv);
}

int ora_set_options(int mode, int shadow_spp, int max_depth) {
#ifdef REF_INTENDED
    if (mode != 1) return -1;
    CKR(cudaMemcpyToSymbol(shim_shadow_spp, &shadow_spp, 4));
    CKR(cudaMemcpyToSymbol(shim_max_depth, &max_depth, 4));
    return 0;
#else
    if (mode != 0 || shadow_spp != 4 || max_depth != 1) return -1;
    return 0;
#endif
}

// Whole frames only (y0 / y1 / nthreads are ignored).  Buffers stay on the device between calls of the same
// size, so accum carries over like the reference's; b->accum, when given with accumId > 0, is uploaded first.
long long ora_render(int W, int H, int rendererType, int accumId, int, int, const OraBuffers* b, int) {
    const size_t npx = (size_t)W * H;
    if (W != S.W || H != S.H) {
        for (auto& p : S.buf) { cudaFree(p); p = nullptr; }
        cudaFree(S.fb);
        cudaFree(S.gs.hit_prim);
        cudaFree(S.gs.hit_uv);
        for (auto& p : S.buf) {
            CKR(cudaMalloc(&p, npx * 16));
            CKR(cudaMemset(p, 0, npx * 16));
        }
        CKR(cudaMalloc(&S.fb, npx * 4));
        CKR(cudaMalloc(&S.gs.hit_prim, npx * 4));
        CKR(cudaMalloc(&S.gs.hit_uv, npx * 8));
        S.W = W;
        S.H = H;
    }
    float* host[13] = {b->position, b->normal, b->uv, b->albedo, b->alpha, b->materialID, b->bounce0, b->bounce1,
                       b->bounce2, b->ltc, b->sto_direct, b->sto_no_vis, b->accum};
    if (accumId > 0 && b->accum) CKR(cudaMemcpy(S.buf[12], b->accum, npx * 16, cudaMemcpyHostToDevice));
    LaunchParams& lp = S.lp;
    lp.position_screen_buffer = S.buf[0];
    lp.normal_screen_buffer = S.buf[1];
    lp.uv_screen_buffer = S.buf[2];
    lp.albedo_screen_buffer = S.buf[3];
    lp.alpha_screen_buffer = S.buf[4];
    lp.materialID_screen_buffer = S.buf[5];
    lp.bounce0_screen_buffer = S.buf[6];
    lp.bounce1_screen_buffer = S.buf[7];
    lp.bounce2_screen_buffer = S.buf[8];
    lp.ltc_screen_buffer = S.buf[9];
    lp.sto_direct_ratio_screen_buffer = S.buf[10];
    lp.sto_no_vis_ratio_screen_buffer = S.buf[11];
    lp.accum_screen_buffer = S.buf[12];
    lp.accumId = accumId;
    lp.rendererType = rendererType;
    CKR(cudaMemcpyToSymbol(optixLaunchParams, &lp, sizeof(lp)));
    S.gs.raygen.frameBuffer = S.fb;
    S.gs.raygen.frameBufferSize = vec2i(W, H);
    CKR(cudaMemset(S.gs.rays, 0, 8));
    CKR(cudaMemcpyToSymbol(g_scene, &S.gs, sizeof(S.gs)));
    cudaEvent_t e0, e1;
    CKR(cudaEventCreate(&e0));
    CKR(cudaEventCreate(&e1));
    dim3 block(SHIM_BLOCK_X, SHIM_BLOCK_Y), grid((W + SHIM_BLOCK_X - 1) / SHIM_BLOCK_X, (H + SHIM_BLOCK_Y - 1) / SHIM_BLOCK_Y);
    CKR(cudaEventRecord(e0));
    k_ref_raygen<<<grid, block>>>(W, H);
    CKR(cudaEventRecord(e1));
    CKR(cudaGetLastError());
    CKR(cudaEventSynchronize(e1));
    CKR(cudaEventElapsedTime(&S.last_ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    for (int i = 0; i < 13; i++)
        if (host[i]) CKR(cudaMemcpy(host[i], S.buf[i], npx * 16, cudaMemcpyDeviceToHost));
    if (b->rgba8) CKR(cudaMemcpy(b->rgba8, S.fb, npx * 4, cudaMemcpyDeviceToHost));
    if (b->hit_prim) CKR(cudaMemcpy(b->hit_prim, S.gs.hit_prim, npx * 4, cudaMemcpyDeviceToHost));
    if (b->hit_uv) CKR(cudaMemcpy(b->hit_uv, S.gs.hit_uv, npx * 8, cudaMemcpyDeviceToHost));
    unsigned long long rays = 0;
    CKR(cudaMemcpy(&rays, S.gs.rays, 8, cudaMemcpyDeviceToHost));
    return (long long)rays;
}

// Texture-unit probe: a w x 1 float4 texture whose texel i holds the value i, sampled with the LTC tables'
// descriptor (linear, clamp, normalised) at the given x: out = i + filter weight, which shows how the unit
// quantises the weight ("texture filtering is outside the repo: validated on-device").
__global__ void k_probe_tex(cudaTextureObject_t t, const float* xs, int n, float* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = tex2D<float4>(t, xs[i], 0.5f).x;
}
void ora_probe_tex(const float* xs, int n, int w, float* out) {
    std::vector<float> texels(4 * (size_t)w);
    for (int i = 0; i < w; i++) texels[4 * i] = texels[4 * i + 1] = texels[4 * i + 2] = texels[4 * i + 3] = (float)i;
    cudaTextureObject_t t = make_texture(texels.data(), w, 1, true);
    float *dx = nullptr, *dout = nullptr;
    CKR(cudaMalloc(&dx, 4 * (size_t)n));
    CKR(cudaMalloc(&dout, 4 * (size_t)n));
    CKR(cudaMemcpy(dx, xs, 4 * (size_t)n, cudaMemcpyHostToDevice));
    k_probe_tex<<<(n + 255) / 256, 256>>>(t, dx, n, dout);
    CKR(cudaMemcpy(out, dout, 4 * (size_t)n, cudaMemcpyDeviceToHost));
    cudaFree(dx);
    cudaFree(dout);
}

// The same for a caller-supplied w x h float4 table at (x, y) pairs: what the unit returns for the real LTC tables.
__global__ void k_probe_tex2d(cudaTextureObject_t t, const float* xy, int n, float4* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = tex2D<float4>(t, xy[2 * i], xy[2 * i + 1]);
}
void ora_probe_tex2d(const float* texels, int w, int h, const float* xy, int n, float* out4) {
    cudaTextureObject_t t = make_texture(texels, w, h, true);
    float* dxy = nullptr;
    float4* dout = nullptr;
    CKR(cudaMalloc(&dxy, 8 * (size_t)n));
    CKR(cudaMalloc(&dout, 16 * (size_t)n));
    CKR(cudaMemcpy(dxy, xy, 8 * (size_t)n, cudaMemcpyHostToDevice));
    k_probe_tex2d<<<(n + 255) / 256, 256>>>(t, dxy, n, dout);
    CKR(cudaMemcpy(out4, dout, 16 * (size_t)n, cudaMemcpyDeviceToHost));
    cudaFree(dxy);
    cudaFree(dout);
}

// CUDA-event time of the last ora_render's kernel (the megakernel alone, no copies)
float ora_last_ms(void) { return S.last_ms; }

}  // extern "C"

/* oracle_api.h -- TEST INFRASTRUCTURE ONLY.
 *
 * One C interface implemented three times, so the tests can swap checkers:
 *   oracle/_ref/libref_literal.so   the reference's own device TU (cuda_src/deviceCode.cu and
 *                                   cuda_include/ **) compiled unmodified as host C++ behind
 *                                   oracle/ref_shim
 *   oracle/_ref/libref_intended.so  same, with patch set P1-P3 of SURVEY.md A.3 applied to a
 *                                   temporary copy at build time
 *   oracle/liboracle.so             this repo's CPU restatement (oracle/restate.cpp), literal and
 *                                   intended selected at run time
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load these libraries.
 */
#ifndef ORACLE_API_H
#define ORACLE_API_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* 13 float4 frame buffers of LaunchParams (cuda_include/common.cuh:84-100) + the RGBA8 frame
 * buffer (RayGenData, :125-128) + side outputs the checker adds.  NULL = not wanted
 * (the oracle then writes into scratch). */
typedef struct OraBuffers {
    float* position;
    float* normal;
    float* uv;
    float* albedo;
    float* alpha;
    float* materialID;
    float* bounce0;
    float* bounce1;
    float* bounce2;
    float* ltc;
    float* sto_direct;
    float* sto_no_vis;
    float* accum;      /* read when accumId > 0, then written */
    uint32_t* rgba8;   /* owl::make_rgba of accum/(accumId+1) */
    int32_t* hit_prim; /* side output: global primitive id of the primary hit, -1 = miss */
    float* hit_uv;     /* side output: float2 barycentrics of the primary hit */
} OraBuffers;

const char* ora_kind(void); /* "reference-literal" | "reference-intended" | "restatement" */

void ora_reset(void);
/* Mesh in reference layout (TriangleMeshData, cuda_include/common.cuh:130-151). tex = RGBA8
 * diffuse texture or NULL.  Returns the mesh index. */
int ora_add_mesh(const float* vtx, const float* nrm, const float* uv, int nVtx, const int32_t* idx, int nTri,
                 const float* diffuse3, float alpha, const float* emit3, int isLight, uint32_t materialID,
                 const uint8_t* tex, int texW, int texH);
/* TriLight[n], 19 floats each: v1 v2 v3 cg normal emit area (cuda_include/common.cuh:64-72). */
void ora_set_lights(const float* triLights, int n);
void ora_set_ltc_lut(const float* m1, const float* m2, const float* m3, int w, int h, int quantise);
void ora_commit(int use_bvh); /* 0 = brute-force closest hit */
void ora_set_camera(const float* pos, const float* d00, const float* du, const float* dv);
/* mode: 0 literal, 1 intended (the two _ref libraries accept only their own).  Literal fixes
 * shadow_spp = 4 and max_depth = 1 (cuda_src/deviceCode.cu:128,151).  Returns 0 on success. */
int ora_set_options(int mode, int shadow_spp, int max_depth);
/* One frame (rows y0..y1-1) of rayGen (cuda_src/deviceCode.cu:59-175). Returns rays traced. */
long long ora_render(int W, int H, int rendererType, int accumId, int y0, int y1, const OraBuffers* b,
                     int nthreads);
/* ltcDirectLighingBaseline (cuda_src/deviceCode.cu:24-56) over a saved G-buffer of n pixels
 * (float4 each); pixels whose normal is (0,0,0) (misses) give 0. out = float4 (rgb,1). */
void ora_ltc_gbuffer(int n, const float* position, const float* normal, const float* albedo,
                     const float* alpha, const float* cam_pos, float* out, int nthreads);

/* liboracle.so ONLY -- classification aid of the parity tests, not a reference function: per pixel of a saved
 * G-buffer, the distance of the LTC evaluation to its nearest sign decision (wo_local.z < 0, the back-face test,
 * the above-horizon bit of every clip vertex).  out[n]; +inf on miss pixels.  Pixels below ~1e-4 are where a
 * last-ulp difference moves the evaluation to another case of clipPolygon's table ("case flips"). */
void ora_ltc_margin(int n, const float* position, const float* normal, const float* alpha, const float* cam_pos,
                    float* out, int nthreads);

/* ---- unit entry points (known-answer tests) ---- */
void ora_u_rng(int frame_id, uint32_t px, uint32_t py, uint32_t W, uint32_t H, int n, uint32_t* state0,
               uint32_t* raw, float* flt); /* state0 = seed; raw[n] from lcg_random; flt[n] CONTINUES with lcg_randomf */
int ora_u_clip(int vertex_count, float* v15);
float ora_u_integrate_edge(const float* a, const float* b);
void ora_u_fetch_ltc(float alpha, float theta, float* m9_amp);
void ora_u_mat_inverse(const float* m9, float* out9);
void ora_u_onb(const float* n, float* mat9_inv9);
void ora_u_sample_direction(const float* wo, float u1, float u2, const float* base, float metal, float alpha,
                            float* wi_pdf4);
void ora_u_evaluate(const float* wi, const float* wo, const float* base, float alpha, float* out3);
float ora_u_pdf(const float* wi, const float* wo, const float* base, float metal, float alpha);
void ora_u_sample_tri(const float* v1, const float* v2, const float* v3, float u1, float u2, float* out3);
float ora_u_pdf_a2w(float pdf, float dist2, float cos_theta);
uint32_t ora_u_make_rgba(const float* c3);
/* ltc_ratio_estimator.py:8  colors = ltc * (stoDirect / stoNoVis), n float elements */
void ora_ratio_combine(long long n, const float* ltc, const float* sto_direct, const float* sto_no_vis, float* out);

/* Edge-avoiding a-trous filter over the two grey sto_* images (float4 buffers of W*H pixels).
 * liboracle.so ONLY: the reference has no denoiser (cuda_include/denoise/bilateral.cuh is empty),
 * so this is the definition the CUDA kernel is checked against, not a restatement. */
void ora_denoise(int W, int H, const float* normal, const float* position, const float* materialID,
                 const float* sto_direct, const float* sto_no_vis, const float* cam_pos, int iterations,
                 float sigma_n, float sigma_p, float sigma_c, float* out_direct, float* out_no_vis, int nthreads);

#ifdef __cplusplus
}
#endif
#endif

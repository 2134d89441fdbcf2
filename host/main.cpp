// orb_render -- headless host of the B200 hot path.
//
// Takes the place of src/hostCode.cpp:8-55 + the non-window parts of include/viewer.hpp: parses
// the reference's scene JSON, ingests its OBJ/MTL files, builds the light list and the camera
// vectors, drives liborb200.so through the C-ABI (include/orb200.h) where the reference drives
// OWL, and writes what the reference's mouse-click dump writes (viewer.hpp:715-800: raw float4
// `.btc` buffers under the same file names, consumed by save.py / ltc_ratio_estimator.py) plus
// a binary PPM of the RGBA8 frame.  No window, no GL, no ImGui: frames are rendered for a fixed
// count instead of until the window closes (Viewer::render, viewer.hpp:553-568, once per frame).
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>

#include <chrono>
#include <string>
#include <vector>

#include "scene_io.h"

using namespace orbhost;

static const char* kRendererNames[] = {"diffuse", "alpha", "normals", "shade_normals", "position", "mask",
                                       "material_id", "ltc_baseline", "ratio", "path"};

static void usage() {
    fprintf(stderr,
            "usage: orb_render SCENE.json [options]\n"
            "  --out DIR            output directory (default: saves)\n"
            "  --renderer N|NAME    0..9 or diffuse|alpha|normals|shade_normals|position|mask|material_id|\n"
            "                       ltc_baseline|ratio|path (default: the scene's renderers[0], else path like\n"
            "                       src/hostCode.cpp:32)\n"
            "  --frames N           frames to accumulate, accumId = 0..N-1 (default: the scene's spp)\n"
            "  --size WxH           frame size (default: the scene's width x height)\n"
            "  --camera I           index into the scene's cameras (default 0)\n"
            "  --mode literal|intended   committed source vs patches P1-P3 (default intended)\n"
            "  --shadow-spp N       RATIO shadow samples per pixel per frame (default 4, deviceCode.cu:128)\n"
            "  --depth N            PATH bounces (default 8; literal mode fixes 1)\n"
            "  --denoise            RATIO: a-trous denoise of stoDirect/stoNoVis + ratio combine\n"
            "  --lut FILE.npy       LTC table (default: optix_renderer_b200/data/ltc_isotropic_8x8.npy)\n"
            "  --exr                every .btc dump also as NAME.exr (what the reference's save.py produces)\n"
            "  --save-state FILE    write the accumulation checkpoint (accum buffer + frame count) after the last frame\n"
            "  --resume FILE        start from such a checkpoint: accumId continues where it stopped\n"
            "  --move-lights X,Y,Z  translate every area light by (X,Y,Z) AFTER the scene was committed: the light meshes go\n"
            "                       through orb_scene_update_mesh + orb_scene_refit (no rebuild), the TriLights are set again\n"
            "  --exact-lut          fetch the table with exact fp32 bilinear weights instead of the texture unit's\n"
            "                       1/256 weights (the reference reads it through tex2D, viewer.hpp:321-327)\n"
            "  --device D           CUDA device (default 0)\n"
            "  --dump-scene         print a JSON summary of the parsed scene and exit (no GPU needed)\n"
            "  --dump-texels        --dump-scene plus the texels of small textures (<= 4096) as hex RGBA8\n");
}

static bool write_file(const std::string& path, const void* data, size_t bytes) {
    FILE* fp = fopen(path.c_str(), "wb");
    if (!fp) return false;
    const bool ok = fwrite(data, 1, bytes, fp) == bytes;
    fclose(fp);
    return ok;
}

// Viewer::savebuffer, viewer.hpp:695-713: the raw float4 buffer, every float divided by avg_factor
// OpenEXR, scanline, uncompressed, three FLOAT channels: what the reference's save.py makes of a .btc dump
// (reshape to rows of float4, flipud, drop w, imageio.imwrite(name.exr): save.py:9-12) and what
// ltc_ratio_estimator.py:3-5 reads.  `rgba` is the float4 buffer, row 0 = bottom image row.
static bool write_exr(const std::string& path, int W, int H, const float* rgba) {
    std::vector<unsigned char> hd;
    auto put = [&](const void* p, size_t n) { hd.insert(hd.end(), (const unsigned char*)p, (const unsigned char*)p + n); };
    auto str = [&](const char* t) { put(t, strlen(t) + 1); };
    auto i32 = [&](int32_t v) { put(&v, 4); };
    auto f32 = [&](float v) { put(&v, 4); };
    auto attr = [&](const char* name, const char* type, int32_t size) { str(name); str(type); i32(size); };
    const unsigned char magic[4] = {0x76, 0x2f, 0x31, 0x01};
    put(magic, 4);
    i32(2);  // version 2, single-part scanline
    attr("channels", "chlist", 3 * (2 + 16) + 1);
    for (const char* c : {"B", "G", "R"}) {  // alphabetical, as the format requires
        str(c);
        i32(2);  // FLOAT
        const unsigned char lin[4] = {0, 0, 0, 0};
        put(lin, 4);
        i32(1);
        i32(1);
    }
    hd.push_back(0);
    attr("compression", "compression", 1);
    hd.push_back(0);
    attr("dataWindow", "box2i", 16);
    i32(0); i32(0); i32(W - 1); i32(H - 1);
    attr("displayWindow", "box2i", 16);
    i32(0); i32(0); i32(W - 1); i32(H - 1);
    attr("lineOrder", "lineOrder", 1);
    hd.push_back(0);
    attr("pixelAspectRatio", "float", 4);
    f32(1.f);
    attr("screenWindowCenter", "v2f", 8);
    f32(0.f); f32(0.f);
    attr("screenWindowWidth", "float", 4);
    f32(1.f);
    hd.push_back(0);
    const uint64_t line_bytes = 8 + (uint64_t)3 * W * 4;
    uint64_t ofs = hd.size() + (uint64_t)8 * H;
    for (int y = 0; y < H; y++, ofs += line_bytes) put(&ofs, 8);
    FILE* fp = fopen(path.c_str(), "wb");
    if (!fp) return false;
    bool ok = fwrite(hd.data(), 1, hd.size(), fp) == hd.size();
    std::vector<float> line((size_t)3 * W);
    for (int y = 0; y < H && ok; y++) {
        const float* src = rgba + 4 * (size_t)(H - 1 - y) * W;  // top image row first
        for (int x = 0; x < W; x++) {
            line[x] = src[4 * x + 2];
            line[(size_t)W + x] = src[4 * x + 1];
            line[(size_t)2 * W + x] = src[4 * x];
        }
        const int32_t head[2] = {y, (int32_t)(3 * W * 4)};
        ok = fwrite(head, 4, 2, fp) == 2 && fwrite(line.data(), 4, line.size(), fp) == line.size();
    }
    return fclose(fp) == 0 && ok;
}

static bool g_write_exr = false;  // --exr: every .btc dump also as NAME.exr (save.py's output)
static int g_W = 0, g_H = 0;      // frame size of the dumps

static bool save_btc(OrbContext* ctx, int buffer_id, const std::string& path, int avg_factor, std::vector<float>& tmp) {
    const size_t bytes = orb_buffer_bytes(ctx, buffer_id);
    tmp.resize(bytes / 4);
    if (orb_buffer_download(ctx, buffer_id, tmp.data(), bytes)) return false;
    for (auto& x : tmp) x = x / (float)avg_factor;
    if (!write_file(path, tmp.data(), bytes)) return false;
    if (g_write_exr) {
        return write_exr(path.substr(0, path.rfind('.')) + ".exr", g_W, g_H, tmp.data());
    }
    return true;
}

// Accumulation checkpoint: the progressive sum and the number of frames in it (Viewer::accumId); --save-state
// after the last frame, --resume before the first.
struct StateHeader {
    char magic[8];
    int32_t W, H, frames, renderer;
};
static bool save_state(OrbContext* ctx, const std::string& path, int W, int H, int frames, int renderer) {
    StateHeader h = {{'O', 'R', 'B', 'A', 'C', 'C', 'U', 'M'}, W, H, frames, renderer};
    std::vector<float> acc((size_t)4 * W * H);
    if (orb_buffer_download(ctx, ORB_BUF_ACCUM, acc.data(), acc.size() * 4)) return false;
    FILE* fp = fopen(path.c_str(), "wb");
    if (!fp) return false;
    const bool ok = fwrite(&h, sizeof h, 1, fp) == 1 && fwrite(acc.data(), 4, acc.size(), fp) == acc.size();
    return fclose(fp) == 0 && ok;
}
static int load_state(OrbContext* ctx, const std::string& path, int W, int H, int renderer) {  // frames done, or -1
    FILE* fp = fopen(path.c_str(), "rb");
    if (!fp) return -1;
    StateHeader h;
    std::vector<float> acc((size_t)4 * W * H);
    const bool ok = fread(&h, sizeof h, 1, fp) == 1 && memcmp(h.magic, "ORBACCUM", 8) == 0 && h.W == W && h.H == H &&
                    h.renderer == renderer && fread(acc.data(), 4, acc.size(), fp) == acc.size();
    fclose(fp);
    if (!ok || orb_buffer_upload(ctx, ORB_BUF_ACCUM, acc.data(), acc.size() * 4)) return -1;
    return h.frames;
}

static bool save_ppm(OrbContext* ctx, int W, int H, const std::string& path) {
    std::vector<uint32_t> px((size_t)W * H);
    if (orb_buffer_download(ctx, ORB_BUF_RGBA8, px.data(), px.size() * 4)) return false;
    std::vector<unsigned char> rgb((size_t)3 * W * H);
    for (int y = 0; y < H; y++)  // buffer row 0 is the bottom image row
        for (int x = 0; x < W; x++) {
            const uint32_t p = px[(size_t)(H - 1 - y) * W + x];
            unsigned char* d = &rgb[3 * ((size_t)y * W + x)];
            d[0] = p & 255; d[1] = (p >> 8) & 255; d[2] = (p >> 16) & 255;
        }
    char hd[64];
    const int n = snprintf(hd, sizeof hd, "P6\n%d %d\n255\n", W, H);
    FILE* fp = fopen(path.c_str(), "wb");
    if (!fp) return false;
    fwrite(hd, 1, n, fp);
    fwrite(rgb.data(), 1, rgb.size(), fp);
    fclose(fp);
    return true;
}

static unsigned long long fnv1a(const void* p, size_t n, unsigned long long h = 1469598103934665603ull) {
    const unsigned char* b = (const unsigned char*)p;
    for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 1099511628211ull; }
    return h;
}

static void dump_scene(const Scene& sc, int W, int H, int cam_i, bool texels = false) {
    const std::vector<OrbTriLight> lights = build_tri_lights(sc);
    float pos[3], d00[3], du[3], dv[3];
    camera_vectors(sc.cameras[cam_i], W, H, pos, d00, du, dv);
    printf("{\"spp\": %d, \"width\": %d, \"height\": %d, \"renderers\": [", sc.spp, sc.width, sc.height);
    for (size_t i = 0; i < sc.renderers.size(); i++) printf("%s%d", i ? ", " : "", sc.renderers[i]);
    printf("], \"n_cameras\": %zu, \"n_surface_meshes\": %zu, \"meshes\": [", sc.cameras.size(), sc.n_surface_meshes);
    for (size_t i = 0; i < sc.meshes.size(); i++) {
        const Mesh& m = sc.meshes[i];
        printf("%s{\"triangles\": %zu, \"material_id\": %u, \"is_light\": %s, \"diffuse\": [%.9g, %.9g, %.9g], \"alpha\": %.9g, "
               "\"emit\": [%.9g, %.9g, %.9g], \"tex\": [%d, %d], \"tex_hash\": \"%016llx\", \"vertex_hash\": \"%016llx\", "
               "\"normal_hash\": \"%016llx\", \"texcoord_hash\": \"%016llx\"",
               i ? ", " : "", m.index.size() / 3, m.material_id, m.is_light ? "true" : "false", m.diffuse[0], m.diffuse[1],
               m.diffuse[2], m.alpha, m.emit[0], m.emit[1], m.emit[2], m.tex_w, m.tex_h, fnv1a(m.tex.data(), m.tex.size()),
               fnv1a(m.vertex.data(), m.vertex.size() * 4), fnv1a(m.normal.data(), m.normal.size() * 4),
               fnv1a(m.texcoord.data(), m.texcoord.size() * 4));
        if (texels && !m.tex.empty() && m.tex.size() <= 4 * 4096) {  // --dump-texels: small textures as hex RGBA8, bottom row first
            printf(", \"texels\": \"");
            for (unsigned char b : m.tex) printf("%02x", b);
            printf("\"");
        }
        printf("}");
    }
    printf("], \"n_tri_lights\": %zu, \"tri_lights_hash\": \"%016llx\", \"camera\": {\"pos\": [%.9g, %.9g, %.9g], "
           "\"d00\": [%.9g, %.9g, %.9g], \"du\": [%.9g, %.9g, %.9g], \"dv\": [%.9g, %.9g, %.9g]}}\n",
           lights.size(), fnv1a(lights.data(), lights.size() * sizeof(OrbTriLight)), pos[0], pos[1], pos[2], d00[0], d00[1],
           d00[2], du[0], du[1], du[2], dv[0], dv[1], dv[2]);
}

#define ORB_CHECK(call)                                                        \
    do {                                                                       \
        if ((call) != ORB_OK) {                                                \
            fprintf(stderr, "orb_render: %s: %s\n", #call, orb_last_error()); \
            return 2;                                                          \
        }                                                                      \
    } while (0)

int main(int argc, char** argv) {
    std::string scene_file, out_dir = "saves", lut_path, renderer_arg, mode_arg = "intended";
    int frames = -1, W = 0, H = 0, cam_i = 0, shadow_spp = 4, depth = 8, device = 0;
    bool denoise = false, only_dump = false, dump_texels = false;
    std::string save_state_path, resume_path;
    float move_lights[3] = {0.f, 0.f, 0.f};
    bool have_move = false;
    int lut_quantise = 1;  // the texture unit's weight rule (measured on the B200, profiles/r1n_texture_unit.json)
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        auto next = [&](const char* what) -> const char* {
            if (i + 1 >= argc) {
                fprintf(stderr, "orb_render: %s needs a value\n", what);
                exit(1);
            }
            return argv[++i];
        };
        if (a == "--out") out_dir = next("--out");
        else if (a == "--renderer") renderer_arg = next("--renderer");
        else if (a == "--frames") frames = atoi(next("--frames"));
        else if (a == "--size") {
            if (sscanf(next("--size"), "%dx%d", &W, &H) != 2 || W <= 0 || H <= 0) { usage(); return 1; }
        }
        else if (a == "--camera") cam_i = atoi(next("--camera"));
        else if (a == "--mode") mode_arg = next("--mode");
        else if (a == "--shadow-spp") shadow_spp = atoi(next("--shadow-spp"));
        else if (a == "--depth") depth = atoi(next("--depth"));
        else if (a == "--denoise") denoise = true;
        else if (a == "--lut") lut_path = next("--lut");
        else if (a == "--exact-lut") lut_quantise = 0;
        else if (a == "--exr") g_write_exr = true;
        else if (a == "--save-state") save_state_path = next("--save-state");
        else if (a == "--resume") resume_path = next("--resume");
        else if (a == "--move-lights") {
            if (sscanf(next("--move-lights"), "%f,%f,%f", &move_lights[0], &move_lights[1], &move_lights[2]) != 3) { usage(); return 1; }
            have_move = true;
        }
        else if (a == "--selftest-exr") {  // writes a known image: tests/test_host.py reads it back
            const char* path = next("--selftest-exr");
            const int w = 7, h = 5;
            std::vector<float> img((size_t)4 * w * h);
            for (int y = 0; y < h; y++)
                for (int x = 0; x < w; x++)
                    for (int c = 0; c < 4; c++) img[4 * ((size_t)y * w + x) + c] = (float)(1000 * y + 10 * x + c) + 0.25f;
            return write_exr(path, w, h, img.data()) ? 0 : 3;
        }
        else if (a == "--device") device = atoi(next("--device"));
        else if (a == "--dump-scene") only_dump = true;
        else if (a == "--dump-texels") only_dump = dump_texels = true;
        else if (a == "-h" || a == "--help") { usage(); return 0; }
        else if (!a.empty() && a[0] != '-' && scene_file.empty()) scene_file = a;
        else { fprintf(stderr, "orb_render: unknown argument %s\n", a.c_str()); usage(); return 1; }
    }
    if (scene_file.empty()) { usage(); return 1; }
    if (mode_arg != "literal" && mode_arg != "intended") { usage(); return 1; }

    fprintf(stderr, "Loading scene %s\n", scene_file.c_str());
    Scene scene;
    std::string err;
    if (!parse_scene(scene_file, scene, &err)) {
        fprintf(stderr, "%s\nError loading scene\n", err.c_str());
        return -1;  // src/hostCode.cpp:23-27
    }
    if (W == 0) { W = scene.width; H = scene.height; }
    if (cam_i < 0 || cam_i >= (int)scene.cameras.size()) { fprintf(stderr, "orb_render: no camera %d\n", cam_i); return 1; }
    if (only_dump) { dump_scene(scene, W, H, cam_i, dump_texels); return 0; }

    int renderer = scene.renderers.empty() ? ORB_PATH : scene.renderers[0];
    if (!renderer_arg.empty()) {
        renderer = -1;
        for (int k = 0; k < 10; k++)
            if (renderer_arg == kRendererNames[k]) renderer = k;
        if (renderer < 0 && isdigit((unsigned char)renderer_arg[0])) renderer = atoi(renderer_arg.c_str());
    }
    if (renderer < 0 || renderer > ORB_PATH) { fprintf(stderr, "orb_render: bad renderer\n"); return 1; }
    if (frames <= 0) frames = scene.spp > 0 ? scene.spp : 1;

    // ---- LTC table ----
    if (lut_path.empty()) {
        std::string exe = argv[0];
        size_t k = exe.rfind('/');
        const std::string here = k == std::string::npos ? "./" : exe.substr(0, k + 1);
        lut_path = here + "../optix_renderer_b200/data/ltc_isotropic_8x8.npy";
    }
    std::vector<float> m1, m2, m3;
    int lw = 0, lh = 0;
    if (!load_ltc_lut_npy(lut_path, m1, m2, m3, lw, lh, &err)) { fprintf(stderr, "orb_render: %s\n", err.c_str()); return 1; }

    // ---- upload: what the Viewer constructor does through OWL (viewer.hpp:163-551) ----
    OrbContext* ctx = nullptr;
    ORB_CHECK(orb_create(device, &ctx));
    size_t n_tris = 0;
    for (const Mesh& m : scene.meshes) {
        OrbMeshDesc d;
        memset(&d, 0, sizeof d);
        d.vertex = m.vertex.data(); d.normal = m.normal.data(); d.texcoord = m.texcoord.data();
        d.n_vertices = (int32_t)(m.vertex.size() / 3);
        d.index = m.index.data();
        d.n_triangles = (int32_t)(m.index.size() / 3);
        for (int a = 0; a < 3; a++) { d.diffuse[a] = m.diffuse[a]; d.emit[a] = m.emit[a]; }
        d.alpha = m.alpha;
        d.is_light = m.is_light ? 1 : 0;
        d.material_id = m.material_id;
        if (!m.tex.empty()) { d.diffuse_tex_rgba8 = m.tex.data(); d.tex_w = m.tex_w; d.tex_h = m.tex_h; }
        ORB_CHECK(orb_scene_add_mesh(ctx, &d));
        n_tris += d.n_triangles;
    }
    const std::vector<OrbTriLight> lights = build_tri_lights(scene);
    ORB_CHECK(orb_scene_set_lights(ctx, lights.data(), (int32_t)lights.size()));
    ORB_CHECK(orb_scene_set_ltc_lut(ctx, m1.data(), m2.data(), m3.data(), lw, lh, lut_quantise));
    const auto t_build0 = std::chrono::steady_clock::now();
    ORB_CHECK(orb_scene_commit(ctx, nullptr));
    ORB_CHECK(orb_sync(ctx));
    const double build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_build0).count();
    if (have_move) {  // animated lights: the reference would rebuild both acceleration structures (viewer.hpp:494-517)
        const auto t_refit0 = std::chrono::steady_clock::now();
        for (size_t k = 0; k < scene.meshes.size(); k++) {
            Mesh& m = scene.meshes[k];
            if (!m.is_light) continue;
            for (size_t i = 0; i < m.vertex.size(); i++) m.vertex[i] += move_lights[i % 3];
            ORB_CHECK(orb_scene_update_mesh(ctx, (int32_t)k, m.vertex.data(), m.normal.data(), (int32_t)(m.vertex.size() / 3),
                                            m.index.data(), (int32_t)(m.index.size() / 3)));
        }
        ORB_CHECK(orb_scene_refit(ctx));
        const std::vector<OrbTriLight> moved = build_tri_lights(scene);
        ORB_CHECK(orb_scene_set_lights(ctx, moved.data(), (int32_t)moved.size()));
        fprintf(stderr, "lights moved by (%g, %g, %g): update + refit %.2f ms\n", move_lights[0], move_lights[1], move_lights[2],
                std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_refit0).count());
    }
    ORB_CHECK(orb_resize(ctx, W, H));
    g_W = W;
    g_H = H;
    float pos[3], d00[3], du[3], dv[3];
    camera_vectors(scene.cameras[cam_i], W, H, pos, d00, du, dv);
    ORB_CHECK(orb_set_camera(ctx, pos, d00, du, dv));
    fprintf(stderr, "framebuffer reso %d %d; %zu meshes, %zu triangles, %zu triangle lights; BVH build %.2f ms\n", W, H,
            scene.meshes.size(), n_tris, lights.size(), build_ms);

    // ---- frames: Viewer::render once per frame, accumId = 0, 1, ... ----
    OrbRenderConfig cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.renderer_type = renderer;
    cfg.mode = mode_arg == "intended" ? 1 : 0;
    cfg.shadow_spp = shadow_spp;
    cfg.max_depth = depth;
    int first_frame = 0;
    if (!resume_path.empty()) {
        first_frame = load_state(ctx, resume_path, W, H, renderer);
        if (first_frame < 0) { fprintf(stderr, "orb_render: cannot resume from %s (size / renderer mismatch?)\n", resume_path.c_str()); return 1; }
        fprintf(stderr, "resumed at accumId %d\n", first_frame);
    }
    const auto t0 = std::chrono::steady_clock::now();
    for (int f = first_frame; f < first_frame + frames; f++) {
        cfg.accum_id = f;
        ORB_CHECK(orb_render_async(ctx, &cfg));
    }
    if (renderer == ORB_RATIO && denoise) {
        ORB_CHECK(orb_denoise(ctx, nullptr, 0, 0));
        ORB_CHECK(orb_ratio_combine(ctx, 1, 0, 0));
    }
    ORB_CHECK(orb_sync(ctx));
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    OrbCounters cnt;
    ORB_CHECK(orb_get_counters(ctx, &cnt));
    const double rays = (double)cnt.rays_primary + (double)cnt.rays_shadow_total + (double)cnt.rays_bounce_total;
    fprintf(stderr, "%s: %d frame(s) in %.2f ms (%.2f ms/frame, %.1f Mrays/s), %llu NaN pixels in the last frame\n",
            kRendererNames[renderer], frames, ms, ms / frames, rays / (ms * 1e-3) / 1e6, (unsigned long long)cnt.nan_pixels);

    // ---- dumps: Viewer::mouseButtonLeft, viewer.hpp:715-800 (same names, same avg_factor) ----
    mkdir(out_dir.c_str(), 0777);
    const std::string o = out_dir + "/";
    std::vector<float> tmp;
    const int accum = first_frame + frames;  // Viewer::accumId after that many calls of render()
    if (!save_state_path.empty() && !save_state(ctx, save_state_path, W, H, accum, renderer)) {
        fprintf(stderr, "orb_render: cannot write %s\n", save_state_path.c_str());
        return 3;
    }
    bool ok = true;
    if (renderer == ORB_RATIO) {
        ok &= save_btc(ctx, ORB_BUF_STO_DIRECT, o + "stoDirect.btc", 1, tmp);
        ok &= save_btc(ctx, ORB_BUF_STO_NO_VIS, o + "stoNoVis.btc", 1, tmp);
        ok &= save_btc(ctx, ORB_BUF_LTC, o + "ltc.btc", 1, tmp);
        ok &= save_btc(ctx, ORB_BUF_NORMAL, o + "normal_screen_buffer.btc", 1, tmp);
        ok &= save_btc(ctx, ORB_BUF_MATERIAL_ID, o + "materialID_screen_buffer.btc", 1, tmp);
        if (denoise) {  // the inputs ltc_ratio_estimator.py:5-6 expects from an external denoiser, and its output
            ok &= save_btc(ctx, ORB_BUF_STO_DIRECT_DN, o + "stodirectDenoise.btc", 1, tmp);
            ok &= save_btc(ctx, ORB_BUF_STO_NO_VIS_DN, o + "stoNoVisDenoise.btc", 1, tmp);
            ok &= save_btc(ctx, ORB_BUF_RATIO_FINAL, o + "final_ratio_estimator.btc", 1, tmp);
        }
    } else if (renderer == ORB_PATH) {
        ok &= save_btc(ctx, ORB_BUF_POSITION, o + "position_screen_buffer.btc", 1, tmp);
        ok &= save_btc(ctx, ORB_BUF_NORMAL, o + "normal_screen_buffer.btc", 1, tmp);
        ok &= save_btc(ctx, ORB_BUF_UV, o + "uv_screen_buffer.btc", 1, tmp);
        ok &= save_btc(ctx, ORB_BUF_ALPHA, o + "alpha_screen_buffer.btc", 1, tmp);
        ok &= save_btc(ctx, ORB_BUF_ALBEDO, o + "albedo_screen_buffer.btc", 1, tmp);
        ok &= save_btc(ctx, ORB_BUF_MATERIAL_ID, o + "materialID_screen_buffer.btc", 1, tmp);
        ok &= save_btc(ctx, ORB_BUF_BOUNCE0, o + "bounce0_screen_buffer.btc", accum, tmp);
        ok &= save_btc(ctx, ORB_BUF_BOUNCE1, o + "bounce1_screen_buffer.btc", accum, tmp);
        ok &= save_btc(ctx, ORB_BUF_BOUNCE2, o + "bounce2_screen_buffer.btc", accum, tmp);
        ok &= save_btc(ctx, ORB_BUF_ACCUM, o + "path.btc", accum, tmp);
    } else if (renderer == ORB_LTC_BASELINE) {
        ok &= save_btc(ctx, ORB_BUF_ACCUM, o + "ltc_baseline.btc", 1, tmp);
    } else {  // debug AOVs: not dumped by the reference; the accumulated frame is the useful output
        ok &= save_btc(ctx, ORB_BUF_ACCUM, o + std::string(kRendererNames[renderer]) + ".btc", accum, tmp);
    }
    ok &= save_ppm(ctx, W, H, o + "rendered_output.ppm");
    orb_destroy(ctx);
    if (!ok) { fprintf(stderr, "orb_render: writing to %s failed\n", out_dir.c_str()); return 3; }
    fprintf(stderr, "wrote %s\n", out_dir.c_str());
    return 0;
}

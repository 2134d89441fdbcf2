// scene_io.h -- host-side scene ingest of the headless renderer (host/orb_render).
//
// Mirrors the data contract of the reference's host code (no code shared with it):
//   include/scene.h:8-34, src/scene.cpp:20-101   scene JSON: spp, width, height, renderers[],
//       cameras[{from,to,up,cos_fovy}], surface_geometry, area_lights           -> parse_scene()
//   src/Model.cpp:130-252                        OBJ + MTL -> one un-indexed mesh per (shape, material),
//       Kd -> diffuse, Ns -> alpha, Ke -> emit, map_Kd -> diffuse texture, materialID = mtl index + 1
//                                                                               -> load_obj()
//   src/scene.cpp:5-13                           light meshes appended, isLight = true, materialID = 0
//   include/viewer.hpp:236-265                   TriLight list                  -> build_tri_lights()
//   include/viewer.hpp:635-641                   camera vectors                 -> camera_vectors()
// Errors are reported by return value + message (parseScene returns bool, scene.cpp:20).
#pragma once
#include <stdint.h>

#include <string>
#include <vector>

#include "../include/orb200.h"

namespace orbhost {

struct Mesh {  // osc::TriangleMesh, include/Model.h:28-52 (fields the device path reads)
    std::vector<float> vertex, normal, texcoord;  // 3, 3, 2 floats per vertex; 3 vertices per face
    std::vector<int32_t> index;                   // (3k, 3k+1, 3k+2)
    float diffuse[3] = {0, 0, 0};
    float alpha = 0.f;
    float emit[3] = {0, 0, 0};
    uint32_t material_id = 0;
    bool is_light = false;
    std::vector<uint8_t> tex;  // RGBA8, already flipped like Model.cpp:108-116; empty = none
    int tex_w = 0, tex_h = 0;
    std::string name;
};

struct SceneCamera {  // include/scene.h:8-14
    float from[3], at[3], up[3];
    float cos_fovy;
};

struct Scene {  // include/scene.h:16-34
    std::string json_path;
    int spp = 1, width = 1024, height = 1024;
    std::vector<int> renderers;
    std::vector<SceneCamera> cameras;
    std::vector<Mesh> meshes;     // surface meshes, then light meshes (syncLights)
    size_t n_surface_meshes = 0;  // meshes[n_surface_meshes..] are the area lights
    float bounds_lo[3], bounds_hi[3];
};

// false + *err on failure (missing file, malformed JSON, no cameras, OBJ without vn/vt/material ...)
bool parse_scene(const std::string& json_file, Scene& scene, std::string* err);
bool load_obj(const std::string& obj_file, std::vector<Mesh>& out, std::string* err);
std::vector<OrbTriLight> build_tri_lights(const Scene& scene);
// pos, dir_00, dir_du, dir_dv for a W x H frame
void camera_vectors(const SceneCamera& cam, int width, int height, float pos[3], float d00[3], float du[3], float dv[3]);
// three w*h RGBA32F tables from the numpy file shipped in optix_renderer_b200/data (shape (3,h,w,4) f4)
bool load_ltc_lut_npy(const std::string& path, std::vector<float>& m1, std::vector<float>& m2, std::vector<float>& m3,
                      int& w, int& h, std::string* err);
// binary PPM (P6) reader for map_Kd textures; other formats are reported and ignored
bool load_ppm_rgba8(const std::string& path, std::vector<uint8_t>& rgba, int& w, int& h);
// map_Kd files as the reference's stb_image call reads them (src/Model.cpp:81-128): PNG, TGA, binary PPM -> RGBA8,
// bottom row first (image_io.cpp)
bool load_image_rgba8(const std::string& path, std::vector<uint8_t>& rgba, int& w, int& h);

}  // namespace orbhost

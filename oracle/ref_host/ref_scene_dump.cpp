// oracle/ref_host/ref_scene_dump.cpp -- TEST INFRASTRUCTURE.  The reference's own scene ingest -- src/scene.cpp
// (parseScene, Scene::syncLights) on top of src/Model.cpp (loadOBJ), with the json.h, tiny_obj_loader.h and stb_image.h
// it vendors, compiled where they lie -- printing what host/main.cpp --dump-scene prints for the same file.
// tests/golden/make_golden_obj.py records it; tests/test_host.py compares host/scene_io.cpp parse_scene with it.
#include <iostream>
#include <map>
#include <stdexcept>
#include <string>
#define OWL_TERMINAL_BLUE ""
#define OWL_TERMINAL_LIGHT_BLUE ""
#define OWL_TERMINAL_DEFAULT ""
#include "tiny_obj_loader.h"  // declarations only here; Model.cpp compiles the implementation part
namespace tinyobj {  // see ref_obj_dump.cpp: the ordering Model.cpp declares inside namespace std, where ADL looks
inline bool operator<(const index_t& a, const index_t& b) {
    if (a.vertex_index != b.vertex_index) return a.vertex_index < b.vertex_index;
    if (a.normal_index != b.normal_index) return a.normal_index < b.normal_index;
    return a.texcoord_index < b.texcoord_index;
}
}  // namespace tinyobj
#include REF_MODEL_CPP
#include REF_SCENE_CPP
#include <cstdio>
#include "owl_math.h"

// The two pieces of host arithmetic in include/viewer.hpp (which itself needs OpenGL / GLFW / OWL and cannot be
// compiled): the Makefile cuts the lines out of the file where it lies into temporary .inc files -- the light-list loop
// :236-265 and the camera vectors :635-641 -- and they are compiled here as the bodies of two functions.
using owl::common::vec2i;
using owl::common::vec3f;
using namespace owl::common;
struct TriLight { vec3f v1, v2, v3, cg, normal, emit; float area; };   // cuda_include/common.cuh:64-72
struct MeshLight { float flux; int triIdx; int triCount; };            // :74-78
struct ViewerPart {
    std::vector<TriLight> triLightList;
    std::vector<MeshLight> meshLightList;
    vec2i fbSize;
    void lights(Scene& scene) {
        Model* triLights = scene.triLights;
#include REF_LIGHTS_INC
    }
    void camera(const vec3f lookFrom, const vec3f lookAt, const vec3f lookUp, const float cosFovy, float out[12]) {
#include REF_CAMERA_INC
        const vec3f v[4] = {camera_pos, camera_d00, camera_ddu, camera_ddv};
        for (int i = 0; i < 4; i++) { out[3 * i] = v[i].x; out[3 * i + 1] = v[i].y; out[3 * i + 2] = v[i].z; }
    }
};

static unsigned long long fnv1a(const void* p, size_t n, unsigned long long h = 1469598103934665603ull) {
    const unsigned char* b = (const unsigned char*)p;
    for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 1099511628211ull; }
    return h;
}

int main(int argc, char** argv) {
    if (argc < 2) return 2;
    Scene sc;
    bool ok = false;
    try {
        ok = parseScene(argv[1], sc);
    } catch (const std::exception& e) {
        printf("\n@@DUMP@@{\"threw\": true}\n");
        fprintf(stderr, "parseScene threw: %s\n", e.what());
        return 0;
    }
    if (!ok) {
        printf("\n@@DUMP@@{\"ok\": false}\n");
        return 0;
    }
    printf("\n@@DUMP@@{\"ok\": true, \"spp\": %d, \"width\": %d, \"height\": %d, \"renderers\": [", sc.spp, sc.imgWidth, sc.imgHeight);
    for (size_t i = 0; i < sc.renderers.size(); i++) printf("%s%d", i ? ", " : "", sc.renderers[i]);
    printf("], \"cameras\": [");
    for (size_t i = 0; i < sc.cameras.size(); i++) {
        const SceneCamera& c = sc.cameras[i];
        printf("%s{\"from\": [%.9g, %.9g, %.9g], \"to\": [%.9g, %.9g, %.9g], \"up\": [%.9g, %.9g, %.9g], \"cos_fovy\": %.9g}", i ? ", " : "",
               c.from.x, c.from.y, c.from.z, c.at.x, c.at.y, c.at.z, c.up.x, c.up.y, c.up.z, c.cosFovy);
    }
    printf("], \"camera_vectors\": [");
    ViewerPart vp;
    vp.fbSize = vec2i(sc.imgWidth, sc.imgHeight);
    for (size_t i = 0; i < sc.cameras.size(); i++) {
        float o[12];
        vp.camera(sc.cameras[i].from, sc.cameras[i].at, sc.cameras[i].up, sc.cameras[i].cosFovy, o);
        printf("%s[", i ? ", " : "");
        for (int k = 0; k < 12; k++) printf("%s%.9g", k ? ", " : "", o[k]);
        printf("]");
    }
    vp.lights(sc);
    static_assert(sizeof(TriLight) == 76, "TriLight layout");
    printf("], \"n_tri_lights\": %zu, \"tri_lights_hash\": \"%016llx\"", vp.triLightList.size(),
           fnv1a(vp.triLightList.data(), vp.triLightList.size() * sizeof(TriLight)));
    printf(", \"meshes\": [");
    for (size_t i = 0; i < sc.model->meshes.size(); i++) {
        const osc::TriangleMesh& t = *sc.model->meshes[i];
        printf("%s{\"triangles\": %zu, \"material_id\": %u, \"is_light\": %s, \"emit\": [%.9g, %.9g, %.9g], \"vertex_hash\": \"%016llx\"}",
               i ? ", " : "", t.index.size(), t.materialID, t.isLight ? "true" : "false", t.emit.x, t.emit.y, t.emit.z,
               fnv1a(t.vertex.data(), t.vertex.size() * 12));
    }
    printf("]}\n");
    return 0;
}

// oracle/ref_host/ref_obj_dump.cpp -- TEST INFRASTRUCTURE.  The reference's own OBJ / MTL / texture ingest -- src/Model.cpp
// (loadOBJ, loadTexture) with the tiny_obj_loader.h and stb_image.h it vendors, compiled where they lie
// (-DREF_MODEL_CPP='"/root/reference/src/Model.cpp"', -I/root/reference/include) behind a stand-in for the OWL vector
// types -- printing, per mesh, what host/main.cpp --dump-scene prints for the meshes of its own loader
// (host/scene_io.cpp load_obj): tests/golden/make_golden_obj.py records it, tests/test_host.py compares.
#include <iostream>
#include <map>
#include <stdexcept>
#include <string>
#include "tiny_obj_loader.h"  // declarations only here; Model.cpp compiles the implementation part
// Model.cpp orders tinyobj::index_t with an operator< it declares inside namespace std, which g++'s two-phase lookup
// does not find from std::less (the reference is built with MSVC).  The same ordering where ADL looks:
namespace tinyobj {
inline bool operator<(const index_t& a, const index_t& b) {
    if (a.vertex_index != b.vertex_index) return a.vertex_index < b.vertex_index;
    if (a.normal_index != b.normal_index) return a.normal_index < b.normal_index;
    return a.texcoord_index < b.texcoord_index;
}
}  // namespace tinyobj
#include REF_MODEL_CPP
#include <cstdio>

static unsigned long long fnv1a(const void* p, size_t n, unsigned long long h = 1469598103934665603ull) {
    const unsigned char* b = (const unsigned char*)p;
    for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 1099511628211ull; }
    return h;
}

int main(int argc, char** argv) {
    if (argc < 2) return 2;
    osc::Model* m = nullptr;
    try {
        m = osc::loadOBJ(argv[1]);
    } catch (const std::exception& e) {
        fprintf(stderr, "loadOBJ threw: %s\n", e.what());
        return 1;
    }
    printf("\n@@DUMP@@{\"meshes\": [");
    for (size_t i = 0; i < m->meshes.size(); i++) {
        const osc::TriangleMesh& t = *m->meshes[i];
        int tw = 0, th = 0;
        unsigned long long thash = fnv1a(nullptr, 0);
        if (t.diffuseTextureID >= 0) {
            const osc::Texture* x = m->textures[t.diffuseTextureID];
            tw = x->resolution.x; th = x->resolution.y;
            thash = fnv1a(x->pixel, (size_t)4 * tw * th);
        }
        printf("%s{\"triangles\": %zu, \"material_id\": %u, \"diffuse\": [%.9g, %.9g, %.9g], \"alpha\": %.9g, "
               "\"emit\": [%.9g, %.9g, %.9g], \"tex\": [%d, %d], \"tex_hash\": \"%016llx\", \"vertex_hash\": \"%016llx\", "
               "\"normal_hash\": \"%016llx\", \"texcoord_hash\": \"%016llx\"}",
               i ? ", " : "", t.index.size(), t.materialID, t.diffuse.x, t.diffuse.y, t.diffuse.z, t.alpha, t.emit.x, t.emit.y, t.emit.z,
               tw, th, thash, fnv1a(t.vertex.data(), t.vertex.size() * 12), fnv1a(t.normal.data(), t.normal.size() * 12),
               fnv1a(t.texcoord.data(), t.texcoord.size() * 8));
    }
    printf("]}\n");
    return 0;
}

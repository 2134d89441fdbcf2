// oracle/ref_host/shim/owl/owl.h -- TEST INFRASTRUCTURE.  Stand-in for the un-vendored OWL header, just enough for the
// reference's HOST sources (src/Model.cpp, include/Model.h) to compile where they lie: the four vector PODs and the
// bounding box loadOBJ touches.  Layouts as OWL's (plain x, y, z members), so the reference's casts from float[3] hold.
#pragma once
#include <cstddef>
#include <cstdint>

namespace owl {
namespace common {
struct vec2i {
    int x, y;
    vec2i() : x(0), y(0) {}
    vec2i(int a) : x(a), y(a) {}
    vec2i(int a, int b) : x(a), y(b) {}
};
struct vec3i {
    int x, y, z;
    vec3i() : x(0), y(0), z(0) {}
    template <class A, class B, class C> vec3i(A a, B b, C c) : x((int)a), y((int)b), z((int)c) {}
};
struct vec2f {
    float x, y;
    vec2f() : x(0.f), y(0.f) {}
    vec2f(float a, float b) : x(a), y(b) {}
};
struct vec3f {
    float x, y, z;
    vec3f() : x(0.f), y(0.f), z(0.f) {}
    vec3f(float a, float b, float c) : x(a), y(b), z(c) {}
};
}  // namespace common
struct box3f {
    common::vec3f lower{1e30f, 1e30f, 1e30f}, upper{-1e30f, -1e30f, -1e30f};
    void extend(const common::vec3f& v) {
        lower.x = v.x < lower.x ? v.x : lower.x; lower.y = v.y < lower.y ? v.y : lower.y; lower.z = v.z < lower.z ? v.z : lower.z;
        upper.x = v.x > upper.x ? v.x : upper.x; upper.y = v.y > upper.y ? v.y : upper.y; upper.z = v.z > upper.z ? v.z : upper.z;
    }
};
}  // namespace owl

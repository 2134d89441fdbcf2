// oracle/ref_host/shim/owl_math.h -- TEST INFRASTRUCTURE.  The OWL vector operations the two host snippets of
// include/viewer.hpp use (light list :236-265, camera :635-641).  OWL is not vendored: these are
// its documented semantics -- component-wise operators, dot added left to right, normalize(v) = v * (1 / sqrt(dot(v, v))).
#pragma once
#include <math.h>
#include "owl/owl.h"

namespace owl {
namespace common {
inline vec3f operator+(const vec3f& a, const vec3f& b) { return vec3f(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3f operator-(const vec3f& a, const vec3f& b) { return vec3f(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3f operator*(const vec3f& a, float b) { return vec3f(a.x * b, a.y * b, a.z * b); }
inline vec3f operator*(float a, const vec3f& b) { return vec3f(a * b.x, a * b.y, a * b.z); }
inline vec3f operator/(const vec3f& a, float b) { return vec3f(a.x / b, a.y / b, a.z / b); }
inline vec3f& operator-=(vec3f& a, const vec3f& b) { a = a - b; return a; }
inline float dot(const vec3f& a, const vec3f& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline vec3f cross(const vec3f& a, const vec3f& b) { return vec3f(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline float length(const vec3f& a) { return sqrtf(dot(a, a)); }
inline vec3f normalize(const vec3f& a) { return a * (1.f / sqrtf(dot(a, a))); }
}  // namespace common
}  // namespace owl

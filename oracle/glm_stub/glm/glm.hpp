// TEST INFRASTRUCTURE ONLY (oracle build). Not part of the product path.
//
// Minimal stand-in for the parts of GLM that /root/reference/src/fluid_sim.cpp
// touches.  The reference depends on the system package libglm-dev
// (/root/reference/README.md:8, include/fluid_sim.h:5), which is un-vendored,
// unpinned and absent from this image.  Its hot-path call sites
// (src/fluid_sim.cpp:257,275-291) only use vec2 construction and componentwise
// + * / on float, which GLM implements as plain IEEE float operations, so this
// stand-in is arithmetically identical.  Off the hot path: length() (:464,:477)
// and the vec3 constructor (:497-515).
#pragma once
#include <cmath>

namespace glm {

struct vec2 {
    float x, y;
    vec2() : x(0.0f), y(0.0f) {}
    explicit vec2(float s) : x(s), y(s) {}
    vec2(float a, float b) : x(a), y(b) {}
};

struct vec3 {
    float x, y, z;
    vec3() : x(0.0f), y(0.0f), z(0.0f) {}
    explicit vec3(float s) : x(s), y(s), z(s) {}
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
};

inline vec2 operator+(const vec2 &a, const vec2 &b) { return vec2(a.x + b.x, a.y + b.y); }
inline vec2 operator-(const vec2 &a, const vec2 &b) { return vec2(a.x - b.x, a.y - b.y); }
inline vec2 operator*(const vec2 &a, float s) { return vec2(a.x * s, a.y * s); }
inline vec2 operator*(float s, const vec2 &a) { return vec2(s * a.x, s * a.y); }
inline vec2 operator/(const vec2 &a, float s) { return vec2(a.x / s, a.y / s); }

// GLM computes length(v) as sqrt(dot(v, v)) with dot = x*x + y*y in float.
inline float length(const vec2 &a) { return std::sqrt(a.x * a.x + a.y * a.y); }

}  // namespace glm

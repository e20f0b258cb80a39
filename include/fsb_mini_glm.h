// Minimal vec2 / vec3 value types with GLM's names, used by include/fluid_sim.h only when the real GLM
// headers are not installed (the reference returns glm::vec2 / glm::vec3 by value, include/fluid_sim.h:28-34).
#ifndef FSB_MINI_GLM_H
#define FSB_MINI_GLM_H
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
}  // namespace glm
#endif

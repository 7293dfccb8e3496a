// Test-infrastructure shim (NOT product code): the two glm names the reference
// uses (glm::vec3 at MTPC/pathTracing.cpp:301, glm::clamp at MTPC/MTPC.cpp:26-28).
#pragma once
namespace glm {
struct vec3 {
    float x, y, z;
    vec3() : x(0), y(0), z(0) {}
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
};
template <class T> inline T clamp(T v, T lo, T hi) { return v < lo ? lo : (v > hi ? hi : v); }
}

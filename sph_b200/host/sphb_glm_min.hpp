// Minimal stand-in for the one GLM type the SphUtils interface exposes (glm::vec3), used only when the
// real GLM is not on the include path (SPHB_SHIM_STANDALONE: headless builds on machines without the
// reference tree).  With GLM present the drop-in header uses the real thing.
#ifndef SPHB_GLM_MIN_HPP
#define SPHB_GLM_MIN_HPP
namespace glm {
struct vec3 {
    float x, y, z;
    vec3() : x(0.0f), y(0.0f), z(0.0f) {}
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
    explicit vec3(float s) : x(s), y(s), z(s) {}
};
}  // namespace glm
#endif

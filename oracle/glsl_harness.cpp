// glsl_harness.cpp -- TEST INFRASTRUCTURE, not product code.
//
// Compiles the REFERENCE's own compute shader (assets/shaders/ray_tracing.comp) as C++ on the host: the
// Makefile pipes the shader, where it lies under /root/reference, through a short sed (drop #version and
// the image binding, turn the uniform block into a struct, rename main, give float literals an `f`
// suffix so arithmetic stays in float as in GLSL) into _ref/ray_tracing_comp.inc, and this file supplies
// the few GLSL types and built-ins the shader uses.  Nothing of the shader is copied into the repo.
// -ffp-contract=off: every operation rounds once (GLSL without `precise` would allow either).
//
// Exports (C): ref_demo_compute() runs main() for every invocation of a width x height dispatch into
// an rgba8 buffer; ref_demo_hit() calls the shader's hit() on a caller-given ray and sphere.
#include <cmath>
#include <cstdint>
#include <cstddef>

namespace glsl {

struct vec3 {
    float x, y, z;
    vec3() : x(0), y(0), z(0) {}
    explicit vec3(float s) : x(s), y(s), z(s) {}
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
};
inline vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator*(float s, vec3 a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

struct vec4 {
    float x, y, z, w;
    vec4() : x(0), y(0), z(0), w(0) {}
    vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
};
struct ivec2 {
    int x, y;
    ivec2() : x(0), y(0) {}
    ivec2(unsigned a, unsigned b) : x(int(a)), y(int(b)) {}
    ivec2(int a, int b) : x(a), y(b) {}
};
struct uvec3 { unsigned x, y, z; };
struct mat4 { float m[16]; };

struct image2D { int width, height; uint8_t* texels; };

inline ivec2 imageSize(const image2D& im) { return ivec2(im.width, im.height); }

inline uint8_t unorm8(float v) {                          // rgba8 store: clamp to [0,1], scale, round to nearest even
    v = v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v);
    return uint8_t(std::nearbyintf(v * 255.0f));
}
inline void imageStore(image2D& im, ivec2 p, vec4 c) {
    if (p.x < 0 || p.y < 0 || p.x >= im.width || p.y >= im.height) return;      // out-of-bounds stores are discarded
    uint8_t* t = im.texels + (size_t(p.y) * im.width + p.x) * 4;
    t[0] = unorm8(c.x); t[1] = unorm8(c.y); t[2] = unorm8(c.z); t[3] = unorm8(c.w);
}

static thread_local uvec3 gl_GlobalInvocationID;
static thread_local image2D colorBuffer;

#include "_ref/ray_tracing_comp.inc"

}  // namespace glsl

extern "C" void ref_demo_compute(int width, int height, uint8_t* rgba8) {
    glsl::colorBuffer = glsl::image2D{width, height, rgba8};
    for (int y = 0; y < height; ++y)
        for (int x = 0; x < width; ++x) {
            glsl::gl_GlobalInvocationID = glsl::uvec3{unsigned(x), unsigned(y), 0u};
            glsl::shader_main();
        }
}

extern "C" int ref_demo_hit(const float origin[3], const float direction[3], const float center[3], float radius) {
    glsl::Ray r; r.origin = glsl::vec3(origin[0], origin[1], origin[2]); r.direction = glsl::vec3(direction[0], direction[1], direction[2]);
    glsl::Sphere s; s.center = glsl::vec3(center[0], center[1], center[2]); s.radius = radius; s.color = glsl::vec4(0, 0, 0, 0);
    return glsl::hit(r, s) ? 1 : 0;
}

// glsl_shim.h — the minimum of GLSL 4.30 needed to compile the reference's shaders/fragment.glsl as
// C++17 (oracle/build_ref.py).  TEST INFRASTRUCTURE ONLY; never included by the product.
//
// It defines the types and built-ins the shader uses with the semantics pinned in SURVEY.md Appendix
// B.2 (the same pins as oracle/rt_oracle.cpp and the CUDA kernels): fp32 without contraction (the
// build passes -ffp-contract=off), dot = (x*x'+y*y')+z*z', normalize = v / sqrt(dot(v,v)), NaN-ignoring
// min/max, GL_NEAREST texel fetch.  alignas(16) on vec3 reproduces the std430 strides (144 B / 48 B).
#pragma once
#include <cmath>
#include <cstdint>

typedef uint32_t uint;

struct vec2 {
    float x, y;
    vec2() : x(0), y(0) {}
    vec2(float a, float b) : x(a), y(b) {}
};

struct alignas(16) vec3 {
    float x, y, z;
    vec3() : x(0), y(0), z(0) {}
    vec3(float a) : x(a), y(a), z(a) {}
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
};

struct vec4 {
    float x, y, z, w;
    vec3 rgb;   // refreshed by texture(); the shader only reads .rgb of a texture() result
    vec4() : x(0), y(0), z(0), w(0) {}
    vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d), rgb(a, b, c) {}
    vec4(vec3 v, float d) : x(v.x), y(v.y), z(v.z), w(d), rgb(v) {}
};

inline vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator*(vec3 a, vec3 b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator*(vec3 a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, vec3 a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator*(int s, vec3 a) { return float(s) * a; }
inline vec3 operator/(vec3 a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline vec3 operator/(vec3 a, int s) { return a / float(s); }
inline vec3 operator/(float s, vec3 a) { return vec3(s / a.x, s / a.y, s / a.z); }
inline vec3 operator-(vec3 a) { return vec3(-a.x, -a.y, -a.z); }
inline vec3& operator+=(vec3& a, vec3 b) { a = a + b; return a; }
inline vec3& operator*=(vec3& a, vec3 b) { a = a * b; return a; }
inline vec3& operator*=(vec3& a, int s) { a = a * float(s); return a; }

inline float dot(vec3 a, vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline vec3 cross(vec3 a, vec3 b) {
    return vec3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
inline float length(vec3 v) { return std::sqrt(dot(v, v)); }
inline vec3 normalize(vec3 v) { return v / length(v); }
inline vec3 reflect(vec3 I, vec3 N) { return I - (2.0f * dot(N, I)) * N; }
inline vec3 mix(vec3 x, vec3 y, float a) { return x * (1.0f - a) + y * a; }
inline float abs(float v) { return std::fabs(v); }
inline float min(float a, float b) { return (a != a) ? b : ((b < a) ? b : a); }
inline float max(float a, float b) { return (a != a) ? b : ((a < b) ? b : a); }
inline vec3 min(vec3 a, vec3 b) { return vec3(min(a.x, b.x), min(a.y, b.y), min(a.z, b.z)); }
inline vec3 max(vec3 a, vec3 b) { return vec3(max(a.x, b.x), max(a.y, b.y), max(a.z, b.z)); }

// prevFrame: RGBA32F, GL_NEAREST (screen.py:42-44) => the texel under the fragment.
struct sampler2D {
    const float* texels;
    int width, height;
};
inline vec4 texture(const sampler2D& s, vec2 uv) {
    int px = int(uv.x * float(s.width));
    int py = int(uv.y * float(s.height));
    px = px < 0 ? 0 : (px >= s.width ? s.width - 1 : px);
    py = py < 0 ? 0 : (py >= s.height ? s.height - 1 : py);
    const float* t = s.texels + 4 * (size_t(py) * size_t(s.width) + size_t(px));
    return vec4(t[0], t[1], t[2], t[3]);
}

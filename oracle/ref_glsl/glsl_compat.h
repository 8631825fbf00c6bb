// glsl_compat.h — the minimum of GLSL needed to compile the REFERENCE's own
// src/shaders/brdf.glsl and the pcg2d/prng_seed part of src/shaders/rand.glsl as C++
// (test infrastructure; see oracle/Makefile target `ref`).  Semantics follow the GLSL 4.60 spec;
// where the spec leaves evaluation order or precision open (matrix*vector summation order,
// sin/cos accuracy) the arithmetic contract of DESIGN.md is used.
#pragma once
#include <cmath>
#include <cstdint>

namespace glsl {
typedef uint32_t uint;

struct vec2 {
    float x, y;
    vec2() : x(0), y(0) {}
    explicit vec2(float s) : x(s), y(s) {}
    vec2(float a, float b) : x(a), y(b) {}
};
struct uvec2 {
    uint x, y;
    uvec2() : x(0), y(0) {}
    explicit uvec2(uint s) : x(s), y(s) {}
    uvec2(uint a, uint b) : x(a), y(b) {}
    explicit uvec2(vec2 v) : x((uint)v.x), y((uint)v.y) {}
};
inline vec2 to_vec2(uvec2 u) { return vec2((float)u.x, (float)u.y); }
struct vec3 {
    float x, y, z;
    vec3() : x(0), y(0), z(0) {}
    explicit vec3(float s) : x(s), y(s), z(s) {}
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
    vec3(vec2 a, float c) : x(a.x), y(a.y), z(c) {}
};
inline vec2 xy(vec3 v) { return vec2(v.x, v.y); }

inline vec2 operator*(vec2 a, vec2 b) { return vec2(a.x * b.x, a.y * b.y); }
inline vec2 operator*(vec2 a, float s) { return vec2(a.x * s, a.y * s); }
inline vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator-(vec3 a) { return vec3(-a.x, -a.y, -a.z); }
inline vec3 operator*(vec3 a, vec3 b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator*(vec3 a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, vec3 a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator/(vec3 a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline vec3& operator+=(vec3& a, vec3 b) { a = a + b; return a; }

inline uvec2 operator*(uint s, uvec2 a) { return uvec2(s * a.x, s * a.y); }
inline uvec2 operator*(uvec2 a, uint s) { return uvec2(a.x * s, a.y * s); }
inline uvec2 operator+(uvec2 a, uint s) { return uvec2(a.x + s, a.y + s); }
inline uvec2 operator+(uvec2 a, uvec2 b) { return uvec2(a.x + b.x, a.y + b.y); }
inline uvec2 operator>>(uvec2 a, uint s) { return uvec2(a.x >> s, a.y >> s); }
inline uvec2& operator^=(uvec2& a, uvec2 b) { a.x ^= b.x; a.y ^= b.y; return a; }
inline vec2 operator*(vec2 a, double) = delete;

inline float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float sqrt(float x) { return std::sqrt(x); }
inline float sin(float x) { return (float)std::sin((double)x); }   // correctly rounded (contract)
inline float cos(float x) { return (float)std::cos((double)x); }
inline float min(float x, float y) { return (y < x) ? y : x; }     // GLSL 8.3
inline float max(float x, float y) { return (x < y) ? y : x; }
inline float mix(float a, float b, float t) { return a * (1.0f - t) + b * t; }
inline vec3 mix(vec3 a, vec3 b, float t) { return vec3(mix(a.x, b.x, t), mix(a.y, b.y, t), mix(a.z, b.z, t)); }
inline vec3 normalize(vec3 v) { return v * (1.0f / sqrt(dot(v, v))); }   // contract: one division
inline vec3 reflect(vec3 I, vec3 N) { return I - 2.0f * dot(N, I) * N; }

struct mat3 {
    vec3 c[3];  // columns
    mat3() {}
    mat3(vec3 a, vec3 b, vec3 d) { c[0] = a; c[1] = b; c[2] = d; }
};
inline mat3 transpose(mat3 m) {
    return mat3(vec3(m.c[0].x, m.c[1].x, m.c[2].x), vec3(m.c[0].y, m.c[1].y, m.c[2].y), vec3(m.c[0].z, m.c[1].z, m.c[2].z));
}
inline vec3 operator*(mat3 m, vec3 v) { return m.c[0] * v.x + m.c[1] * v.y + m.c[2] * v.z; }
}  // namespace glsl

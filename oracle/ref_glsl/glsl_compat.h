// glsl_compat.h — enough of GLSL 4.60 (+ GL_EXT_ray_tracing, GL_EXT_buffer_reference2, scalar block
// layout, explicit 16/64-bit integer types) to compile the REFERENCE's own shader sources
//   src/shaders/{rand,common,triangle,material,brdf}.glsl, ray_gen.rgen, closest_hit.rchit, miss.rmiss
// as C++ (test infrastructure; see oracle/Makefile target `ref` and ref_glsl/rewrite.py for the few
// textual rewrites C++ cannot express: float literal suffixes, `inout`, `layout(...)` declarations).
//
// Language semantics follow the GLSL spec (operators are component-wise, `v * M` is row-vector x
// matrix, matN(matMxK) takes the upper-left block, swizzles are l-values).  Where the spec leaves the
// precision or evaluation order open, the arithmetic contract of DESIGN.md §2 is used and marked
// [contract] below: matrix/vector products sum left to right, normalize(v) = v * (1 / sqrt(dot)),
// inverse(mat3) by cofactors with one division, sin/cos = include/b200pt_contract.h, texture() = the Vulkan
// bilinear formula in fp32.  What the Vulkan *driver* does behind traceRayEXT (acceleration-structure
// traversal, ray/triangle intersection) is not shader source: it is supplied by a callback.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

#include "../../include/b200pt_contract.h"

namespace glsl {
typedef uint32_t uint;

struct vec2; struct vec3; struct vec4; struct uvec2; struct uvec3;

// ---- swizzle proxies: plain storage aliases inside the vector unions (l-values, convertible)
template <class V, class T, int A, int B> struct swz2 {
    T v[(A > B ? A : B) + 1];
    operator V() const;
    swz2& operator=(const V& o);
};
template <class V, class T, int A, int B, int C> struct swz3 {
    T v[(A > B ? (A > C ? A : C) : (B > C ? B : C)) + 1];
    operator V() const;
    swz3& operator=(const V& o);
};

struct vec2 {
    union { struct { float x, y; }; struct { float r, g; }; swz2<vec2, float, 0, 1> xy; };
    vec2() : x(0), y(0) {}
    explicit vec2(float s) : x(s), y(s) {}
    vec2(float a, float b) : x(a), y(b) {}
};
struct uvec2 {
    union { struct { uint x, y; }; swz2<uvec2, uint, 0, 1> xy; };
    uvec2() : x(0), y(0) {}
    explicit uvec2(uint s) : x(s), y(s) {}
    uvec2(uint a, uint b) : x(a), y(b) {}
    explicit uvec2(const vec2& v) : x((uint)v.x), y((uint)v.y) {}
    operator vec2() const { return vec2((float)x, (float)y); }      // GLSL implicit conversion uvec2 -> vec2
};
struct ivec2 {
    int x, y;
    ivec2() : x(0), y(0) {}
    ivec2(int a, int b) : x(a), y(b) {}
    explicit ivec2(const uvec2& u) : x((int)u.x), y((int)u.y) {}
};
struct uvec3 {
    union { struct { uint x, y, z; }; swz2<uvec2, uint, 0, 1> xy; };
    uvec3() : x(0), y(0), z(0) {}
    uvec3(uint a, uint b, uint c) : x(a), y(b), z(c) {}
};
struct vec3 {
    union { struct { float x, y, z; }; struct { float r, g, b; }; swz2<vec2, float, 0, 1> xy; swz3<vec3, float, 0, 1, 2> xyz; swz3<vec3, float, 0, 1, 2> rgb; };
    vec3() : x(0), y(0), z(0) {}
    explicit vec3(float s) : x(s), y(s), z(s) {}
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
    vec3(vec2 a, float c) : x(a.x), y(a.y), z(c) {}
};
struct vec4 {
    union { struct { float x, y, z, w; }; struct { float r, g, b, a; }; swz2<vec2, float, 0, 1> xy; swz3<vec3, float, 0, 1, 2> xyz; swz3<vec3, float, 0, 1, 2> rgb; };
    vec4() : x(0), y(0), z(0), w(0) {}
    explicit vec4(float s) : x(s), y(s), z(s), w(s) {}
    vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
    vec4(vec3 v, float d) : x(v.x), y(v.y), z(v.z), w(d) {}
};
static_assert(sizeof(vec2) == 8 && sizeof(vec3) == 12 && sizeof(vec4) == 16, "scalar block layout strides");

template <class V, class T, int A, int B> swz2<V, T, A, B>::operator V() const { return V(v[A], v[B]); }
template <class V, class T, int A, int B> swz2<V, T, A, B>& swz2<V, T, A, B>::operator=(const V& o) { v[A] = o.x; v[B] = o.y; return *this; }
template <class V, class T, int A, int B, int C> swz3<V, T, A, B, C>::operator V() const { return V(v[A], v[B], v[C]); }
template <class V, class T, int A, int B, int C> swz3<V, T, A, B, C>& swz3<V, T, A, B, C>::operator=(const V& o) { v[A] = o.x; v[B] = o.y; v[C] = o.z; return *this; }

// ---- component-wise operators
inline vec2 operator+(vec2 a, vec2 b) { return vec2(a.x + b.x, a.y + b.y); }
inline vec2 operator-(vec2 a, float s) { return vec2(a.x - s, a.y - s); }
inline vec2 operator*(vec2 a, vec2 b) { return vec2(a.x * b.x, a.y * b.y); }
inline vec2 operator*(vec2 a, float s) { return vec2(a.x * s, a.y * s); }
inline vec2 operator/(vec2 a, vec2 b) { return vec2(a.x / b.x, a.y / b.y); }
inline vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator-(vec3 a) { return vec3(-a.x, -a.y, -a.z); }
inline vec3 operator*(vec3 a, vec3 b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator*(vec3 a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, vec3 a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator/(vec3 a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline vec3& operator+=(vec3& a, vec3 b) { a = a + b; return a; }
inline vec3& operator*=(vec3& a, vec3 b) { a = a * b; return a; }
inline vec3& operator/=(vec3& a, float s) { a = a / s; return a; }
inline vec4 operator+(vec4 a, vec4 b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline vec4 operator*(vec4 a, vec4 b) { return vec4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
inline vec4 operator*(vec4 a, float s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline vec4& operator*=(vec4& a, vec4 b) { a = a * b; return a; }

inline uvec2 operator*(uint s, uvec2 a) { return uvec2(s * a.x, s * a.y); }
inline uvec2 operator*(uvec2 a, uint s) { return uvec2(a.x * s, a.y * s); }
inline uvec2 operator+(uvec2 a, uint s) { return uvec2(a.x + s, a.y + s); }
inline uvec2 operator+(uvec2 a, uvec2 b) { return uvec2(a.x + b.x, a.y + b.y); }
inline uvec2 operator>>(uvec2 a, uint s) { return uvec2(a.x >> s, a.y >> s); }
inline uvec2& operator^=(uvec2& a, uvec2 b) { a.x ^= b.x; a.y ^= b.y; return a; }

// ---- built-in functions (GLSL 4.60 chapter 8)
inline float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }                 // [contract] left to right
inline vec3 cross(vec3 a, vec3 b) { return vec3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }  // 8.5
inline float sqrt(float x) { return std::sqrt(x); }
#ifdef GLSL_ALT_PRECISION
// A second legal reading of the same shaders (oracle/_ref/libref_glsl_alt.so): the textbook forms the GLSL spec prints,
// normalize(v) = v / length(v), and the C library's fp32 sin/cos (<= 1 ulp).  Used only to MEASURE how far two
// conforming implementations of the reference drift apart (tests/test_oracle_kat.py states the tolerance).
inline float sin(float x) { return std::sin(x); }
inline float cos(float x) { return std::cos(x); }
#else
inline float sin(float x) { float s, c; pt_contract_sincos(x, &s, &c); return s; }   // [contract] include/b200pt_contract.h
inline float cos(float x) { float s, c; pt_contract_sincos(x, &s, &c); return c; }
#endif
inline float min(float x, float y) { return (y < x) ? y : x; }     // 8.3
inline float max(float x, float y) { return (x < y) ? y : x; }
inline float mix(float a, float b, float t) { return a * (1.0f - t) + b * t; }                  // 8.3
inline vec3 mix(vec3 a, vec3 b, float t) { return vec3(mix(a.x, b.x, t), mix(a.y, b.y, t), mix(a.z, b.z, t)); }
#ifdef GLSL_ALT_PRECISION
inline vec3 normalize(vec3 v) { return v / sqrt(dot(v, v)); }
#else
// [contract] one division.  The Vulkan spec states normalize()'s precision as inherited from x * inversesqrt(dot(x, x))
// with inversesqrt() at 2 ulp; 1 / sqrt() with both operations correctly rounded is within that.
inline vec3 normalize(vec3 v) { return v * (1.0f / sqrt(dot(v, v))); }
#endif
inline vec3 reflect(vec3 I, vec3 N) { return I - 2.0f * dot(N, I) * N; }                        // 8.5
inline vec4 unpackUnorm4x8(uint p) {                                                            // 8.4: f / 255.0
    return vec4((float)(p & 0xffu) / 255.0f, (float)((p >> 8) & 0xffu) / 255.0f, (float)((p >> 16) & 0xffu) / 255.0f, (float)(p >> 24) / 255.0f);
}

// ---- matrices (column-major, as GLSL)
struct mat4x3 {       // 4 columns of vec3
    vec3 c[4];
};
struct mat3 {
    vec3 c[3];
    mat3() {}
    mat3(vec3 a, vec3 b, vec3 d) { c[0] = a; c[1] = b; c[2] = d; }
    explicit mat3(const mat4x3& m) { c[0] = m.c[0]; c[1] = m.c[1]; c[2] = m.c[2]; }            // 5.4.2: upper-left block
};
inline mat3 transpose(mat3 m) {
    return mat3(vec3(m.c[0].x, m.c[1].x, m.c[2].x), vec3(m.c[0].y, m.c[1].y, m.c[2].y), vec3(m.c[0].z, m.c[1].z, m.c[2].z));
}
inline vec3 operator*(mat3 m, vec3 v) { return m.c[0] * v.x + m.c[1] * v.y + m.c[2] * v.z; }   // [contract] left to right
inline vec3 operator*(vec3 v, mat3 m) { return vec3(dot(v, m.c[0]), dot(v, m.c[1]), dot(v, m.c[2])); }   // 5.10: row vector x matrix
// [contract] inverse(mat3): cofactors over one determinant, one division (element (r, c) below is row r, column c)
inline mat3 inverse(mat3 m) {
    const float a00 = m.c[0].x, a01 = m.c[1].x, a02 = m.c[2].x;
    const float a10 = m.c[0].y, a11 = m.c[1].y, a12 = m.c[2].y;
    const float a20 = m.c[0].z, a21 = m.c[1].z, a22 = m.c[2].z;
    const float c00 = a11 * a22 - a12 * a21;
    const float c01 = a12 * a20 - a10 * a22;
    const float c02 = a10 * a21 - a11 * a20;
    const float det = a00 * c00 + a01 * c01 + a02 * c02;
    const float id = 1.0f / det;
    mat3 r;
    r.c[0] = vec3(c00 * id, c01 * id, c02 * id);
    r.c[1] = vec3((a02 * a21 - a01 * a22) * id, (a00 * a22 - a02 * a20) * id, (a01 * a20 - a00 * a21) * id);
    r.c[2] = vec3((a01 * a12 - a02 * a11) * id, (a02 * a10 - a00 * a12) * id, (a00 * a11 - a01 * a10) * id);
    return r;
}

// ---- opaque types and the pipeline interface
// sampler2D: RGBA8 UNORM image, linear mag/min, REPEAT, LOD 0 (RayTracingPipeline.zig:54-70, Image.zig:24-39).
struct sampler2D { const uint8_t* rgba; uint width, height; };
// [contract] texture(): Vulkan spec texel filtering in fp32: (u*W - 0.5, v*H - 0.5), i0 = floor, weights = fract, wrap mod W.
inline vec4 texture(const sampler2D& s, vec2 uv) {
    const float W = (float)s.width, H = (float)s.height;
    const float u = uv.x * W - 0.5f, v = uv.y * H - 0.5f;
    const float fu = std::floor(u), fv = std::floor(v);
    const float a = u - fu, b = v - fv;
    auto wrap = [](float f, uint n) -> uint {
        if (!(f >= -2147483648.0f && f < 2147483648.0f)) return 0u;
        int64_t i = (int64_t)f; int64_t m = i % (int64_t)n; if (m < 0) m += n; return (uint)m;
    };
    const uint x0 = wrap(fu, s.width), x1 = (x0 + 1 == s.width) ? 0 : x0 + 1;
    const uint y0 = wrap(fv, s.height), y1 = (y0 + 1 == s.height) ? 0 : y0 + 1;
    auto texel = [&](uint x, uint y) -> vec4 {
        uint p; std::memcpy(&p, s.rgba + 4ull * ((uint64_t)y * s.width + x), 4);
        return unpackUnorm4x8(p);
    };
    const vec4 t00 = texel(x0, y0), t10 = texel(x1, y0), t01 = texel(x0, y1), t11 = texel(x1, y1);
    const float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
    return vec4(t00.x * w00 + t10.x * w10 + t01.x * w01 + t11.x * w11, t00.y * w00 + t10.y * w10 + t01.y * w01 + t11.y * w11,
                t00.z * w00 + t10.z * w10 + t01.z * w01 + t11.z * w11, t00.w * w00 + t10.w * w10 + t01.w * w01 + t11.w * w11);
}
struct image2D { float* rgba; uint width, height; };
inline void imageStore(const image2D& im, ivec2 p, vec4 v) {
    float* q = im.rgba + 4ull * ((uint64_t)p.y * im.width + (uint)p.x);
    q[0] = v.x; q[1] = v.y; q[2] = v.z; q[3] = v.w;
}
struct accelerationStructureEXT { void* opaque; };
const uint gl_RayFlagsOpaqueEXT = 1u, gl_RayFlagsCullBackFacingTrianglesEXT = 16u;

// buffer_reference blocks with one unsized array member (scalar layout): a typed view of a device address
#define GLSL_BUFFER_REFERENCE(NAME, TYPE, FIELD)      \
    struct NAME {                                     \
        const TYPE* FIELD;                            \
        NAME() : FIELD(nullptr) {}                    \
        explicit NAME(uint64_t a) : FIELD(reinterpret_cast<const TYPE*>(a)) {} \
    };
}  // namespace glsl

// pt_math.cuh — device vector algebra with GLSL semantics and a FIXED evaluation order.
//
// Arithmetic contract: IEEE binary32; this translation unit is compiled with
// --fmad=false so a*b+c is never contracted (FMAs are written explicitly with __fmaf_rn where they
// are wanted: the conservative box tests); expressions are evaluated left to right as written;
// sqrtf and / are the correctly rounded CUDA defaults (-prec-sqrt/-prec-div); sin/cos are the
// explicit double-precision evaluation of include/b200pt_contract.h, rounded to float; min/max follow GLSL 8.3 (min(x,y) = y<x ? y : x).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "../../include/b200pt_contract.h"

namespace pt {

struct V2 { float x, y; };
struct V3 { float x, y, z; };
struct V4 { float x, y, z, w; };
struct U2 { uint32_t x, y; };

#define PT_HD __host__ __device__ __forceinline__
#define PT_D __device__ __forceinline__

PT_HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
PT_HD V3 v3s(float s) { V3 r; r.x = s; r.y = s; r.z = s; return r; }
PT_HD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
PT_HD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
PT_HD V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
PT_HD V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
PT_HD V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
PT_HD V3 operator/(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
PT_HD V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
PT_HD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
PT_HD V3 cross(V3 a, V3 b) { return v3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }
// normalize(v) = v * (1 / sqrt(dot(v, v))): one correctly rounded division (arithmetic contract)
PT_HD V3 normalize(V3 a) { return a * (1.0f / sqrtf(dot(a, a))); }
PT_HD float gmin(float x, float y) { return (y < x) ? y : x; }
PT_HD float gmax(float x, float y) { return (x < y) ? y : x; }
PT_HD float mixf(float a, float b, float t) { return a * (1.0f - t) + b * t; }
PT_HD V3 mix(V3 a, V3 b, float t) { return v3(mixf(a.x, b.x, t), mixf(a.y, b.y, t), mixf(a.z, b.z, t)); }
PT_HD float comp(V3 a, int k) { return k == 0 ? a.x : (k == 1 ? a.y : a.z); }

// sin / cos of the arithmetic contract (include/b200pt_contract.h): explicit IEEE double arithmetic, the same bits here, in
// the CPU oracle and in the reference shaders compiled as C++ — no library call on any side
PT_HD void sincos_cr(float x, float& s, float& c) { pt_contract_sincos(x, &s, &c); }

// row-major 3x3
struct M3 { float m[3][3]; };

PT_HD V3 mul(const M3& a, V3 v) {
    return v3(a.m[0][0] * v.x + a.m[0][1] * v.y + a.m[0][2] * v.z,
              a.m[1][0] * v.x + a.m[1][1] * v.y + a.m[1][2] * v.z,
              a.m[2][0] * v.x + a.m[2][1] * v.y + a.m[2][2] * v.z);
}

// inverse(mat3): cofactors / determinant, fixed order (GLSL inverse(), triangle.glsl:79)
PT_HD M3 inverse(const M3& a) {
    const float a00 = a.m[0][0], a01 = a.m[0][1], a02 = a.m[0][2];
    const float a10 = a.m[1][0], a11 = a.m[1][1], a12 = a.m[1][2];
    const float a20 = a.m[2][0], a21 = a.m[2][1], a22 = a.m[2][2];
    const float c00 = a11 * a22 - a12 * a21;
    const float c01 = a12 * a20 - a10 * a22;
    const float c02 = a10 * a21 - a11 * a20;
    const float det = a00 * c00 + a01 * c01 + a02 * c02;
    const float id = 1.0f / det;
    M3 r;
    r.m[0][0] = c00 * id;
    r.m[0][1] = (a02 * a21 - a01 * a22) * id;
    r.m[0][2] = (a01 * a12 - a02 * a11) * id;
    r.m[1][0] = c01 * id;
    r.m[1][1] = (a00 * a22 - a02 * a20) * id;
    r.m[1][2] = (a02 * a10 - a00 * a12) * id;
    r.m[2][0] = c02 * id;
    r.m[2][1] = (a01 * a20 - a00 * a21) * id;
    r.m[2][2] = (a00 * a11 - a01 * a10) * id;
    return r;
}

}  // namespace pt

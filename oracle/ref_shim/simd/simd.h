/* <simd/simd.h> stand-in for compiling the reference's plain-C++ host code with g++ (oracle/_ref; TEST INFRASTRUCTURE ONLY).
 * Same sizes / alignments / member spellings as Apple's types: simd_float2 8/8, simd_float3 16/16, simd_float4 16/16,
 * column-major matrices with .columns[]. */
#pragma once
#include <cmath>
#include <cstddef>
#include <cstdint>
struct alignas(8) simd_float2 { float x, y; float& operator[](int i) { return (&x)[i]; } const float& operator[](int i) const { return (&x)[i]; } };
struct alignas(16) simd_float3 { float x, y, z, _w; float& operator[](int i) { return (&x)[i]; } const float& operator[](int i) const { return (&x)[i]; } };
struct alignas(16) simd_float4 { float x, y, z, w; float& operator[](int i) { return (&x)[i]; } const float& operator[](int i) const { return (&x)[i]; } };
struct simd_float3x3 { simd_float3 columns[3]; };
struct simd_float4x4 { simd_float4 columns[4]; };
typedef simd_float3x3 matrix_float3x3;
typedef simd_float4x4 matrix_float4x4;
typedef simd_float2 vector_float2;
typedef simd_float3 vector_float3;
typedef simd_float4 vector_float4;
static inline simd_float2 simd_make_float2(float x, float y) { return simd_float2{x, y}; }
static inline simd_float3 simd_make_float3(float x, float y, float z) { return simd_float3{x, y, z, 0.0f}; }
static inline simd_float4 simd_make_float4(float x, float y, float z, float w) { return simd_float4{x, y, z, w}; }
static inline simd_float4 simd_make_float4(simd_float3 v, float w) { return simd_float4{v.x, v.y, v.z, w}; }
static inline simd_float3 operator+(simd_float3 a, simd_float3 b) { return simd_float3{a.x + b.x, a.y + b.y, a.z + b.z, 0.0f}; }
static inline simd_float3 operator-(simd_float3 a, simd_float3 b) { return simd_float3{a.x - b.x, a.y - b.y, a.z - b.z, 0.0f}; }
static inline simd_float3 operator*(simd_float3 a, float s) { return simd_float3{a.x * s, a.y * s, a.z * s, 0.0f}; }
static inline simd_float3 operator*(float s, simd_float3 a) { return a * s; }
static inline simd_float3 operator/(simd_float3 a, float s) { return simd_float3{a.x / s, a.y / s, a.z / s, 0.0f}; }
static inline simd_float3& operator+=(simd_float3& a, simd_float3 b) { a = a + b; return a; }
/* matrix * vector: sum of columns scaled by the vector's components, evaluated left to right */
static inline simd_float3 simd_mul(simd_float3x3 m, simd_float3 v) {
    return simd_float3{m.columns[0].x * v.x + m.columns[1].x * v.y + m.columns[2].x * v.z,
                       m.columns[0].y * v.x + m.columns[1].y * v.y + m.columns[2].y * v.z,
                       m.columns[0].z * v.x + m.columns[1].z * v.y + m.columns[2].z * v.z, 0.0f};
}
static inline float simd_length(simd_float3 a) { return sqrtf(a.x * a.x + a.y * a.y + a.z * a.z); }
static inline float simd_dot(simd_float3 a, simd_float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline simd_float3 simd_normalize(simd_float3 a) { return a / simd_length(a); }

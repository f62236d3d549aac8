// hjo_vec.h — TEST INFRASTRUCTURE (CPU oracle). Not part of the product path.
//
// Minimal CUDA-style vector types and the ~30 operators the reference's device headers use.
// Stands in for the OptiX SDK's sutil/vec_math.h, which the reference includes everywhere
// (/root/reference/include/kernel/math.h:4) but which is NOT in the reference snapshot.
// Semantics follow the published OptiX 7.7 SDK header: normalize = v * (1/sqrtf(dot)), vec/scalar =
// multiply by reciprocal, reflect(i,n) = i - 2 n dot(n,i), lerp(a,b,t) = a + t (b-a),
// clamp(f,a,b) = fmaxf(a, fminf(f,b)). Everything is plain IEEE float; build with -ffp-contract=off.
#pragma once
#include <cmath>
#include <cfloat>
#include <cstdint>
#include <algorithm>

struct float2 { float x, y; };
struct float3 { float x, y, z; };
struct float4 { float x, y, z, w; };
struct uint3 { unsigned int x, y, z; };
struct uint4 { unsigned int x, y, z, w; };
struct uchar4 { unsigned char x, y, z, w; };

#ifndef M_PIf
#define M_PIf 3.14159265358979323846f
#endif

static inline float2 make_float2(float x, float y) { return float2{x, y}; }
static inline float3 make_float3(float x, float y, float z) { return float3{x, y, z}; }
static inline float3 make_float3(float s) { return float3{s, s, s}; }
static inline float3 make_float3(const float2& a, float z) { return float3{a.x, a.y, z}; }
static inline float3 make_float3(const float4& a) { return float3{a.x, a.y, a.z}; }
static inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }
static inline float4 make_float4(const float3& a, float w) { return float4{a.x, a.y, a.z, w}; }
static inline float4 make_float4(const float2& a, const float2& b) { return float4{a.x, a.y, b.x, b.y}; }
static inline float4 make_float4(float s) { return float4{s, s, s, s}; }
static inline uint4 make_uint4(unsigned x, unsigned y, unsigned z, unsigned w) { return uint4{x, y, z, w}; }
static inline uint3 make_uint3(unsigned x, unsigned y, unsigned z) { return uint3{x, y, z}; }
static inline uchar4 make_uchar4(unsigned char x, unsigned char y, unsigned char z, unsigned char w) { return uchar4{x, y, z, w}; }

static inline float3 operator-(const float3& a) { return float3{-a.x, -a.y, -a.z}; }
static inline float3 operator+(const float3& a, const float3& b) { return float3{a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline float3 operator-(const float3& a, const float3& b) { return float3{a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline float3 operator*(const float3& a, const float3& b) { return float3{a.x * b.x, a.y * b.y, a.z * b.z}; }
static inline float3 operator*(const float3& a, float s) { return float3{a.x * s, a.y * s, a.z * s}; }
static inline float3 operator*(float s, const float3& a) { return float3{s * a.x, s * a.y, s * a.z}; }
static inline float3 operator/(const float3& a, const float3& b) { return float3{a.x / b.x, a.y / b.y, a.z / b.z}; }
static inline float3 operator/(const float3& a, float s) { float inv = 1.0f / s; return a * inv; }
static inline float3 operator/(float s, const float3& a) { return float3{s / a.x, s / a.y, s / a.z}; }
static inline float3 operator+(const float3& a, float b) { return float3{a.x + b, a.y + b, a.z + b}; }
static inline float3 operator+(float a, const float3& b) { return float3{a + b.x, a + b.y, a + b.z}; }
static inline float3 operator-(const float3& a, float b) { return float3{a.x - b, a.y - b, a.z - b}; }
static inline float3 operator-(float a, const float3& b) { return float3{a - b.x, a - b.y, a - b.z}; }
static inline void operator+=(float3& a, const float3& b) { a.x += b.x; a.y += b.y; a.z += b.z; }
static inline void operator-=(float3& a, const float3& b) { a.x -= b.x; a.y -= b.y; a.z -= b.z; }
static inline void operator*=(float3& a, const float3& b) { a.x *= b.x; a.y *= b.y; a.z *= b.z; }
static inline void operator*=(float3& a, float s) { a.x *= s; a.y *= s; a.z *= s; }
static inline void operator/=(float3& a, float s) { float inv = 1.0f / s; a *= inv; }

static inline float2 operator+(const float2& a, const float2& b) { return float2{a.x + b.x, a.y + b.y}; }
static inline float2 operator-(const float2& a, const float2& b) { return float2{a.x - b.x, a.y - b.y}; }
static inline float2 operator*(const float2& a, float s) { return float2{a.x * s, a.y * s}; }
static inline float2 operator*(float s, const float2& a) { return float2{s * a.x, s * a.y}; }

static inline float4 operator+(const float4& a, const float4& b) { return float4{a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
static inline float4 operator-(const float4& a, const float4& b) { return float4{a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w}; }
static inline float4 operator*(const float4& a, float s) { return float4{a.x * s, a.y * s, a.z * s, a.w * s}; }
static inline float4 operator*(float s, const float4& a) { return float4{s * a.x, s * a.y, s * a.z, s * a.w}; }
static inline void operator+=(float4& a, const float4& b) { a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }

static inline float dot(const float3& a, const float3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline float dot(const float4& a, const float4& b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
static inline float3 cross(const float3& a, const float3& b) {
    return float3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
static inline float length(const float3& v) { return sqrtf(dot(v, v)); }
static inline float3 normalize(const float3& v) { float invLen = 1.0f / sqrtf(dot(v, v)); return v * invLen; }
static inline float3 reflect(const float3& i, const float3& n) { return i - 2.0f * n * dot(n, i); }
static inline float clamp(float f, float a, float b) { return fmaxf(a, fminf(f, b)); }
static inline float3 clamp(const float3& v, float a, float b) { return float3{clamp(v.x, a, b), clamp(v.y, a, b), clamp(v.z, a, b)}; }
static inline float3 lerp(const float3& a, const float3& b, float t) { return a + t * (b - a); }
static inline float4 lerp(const float4& a, const float4& b, float t) { return a + t * (b - a); }

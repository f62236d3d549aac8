// hj_vec.cuh — device vector operators for the sm_100a kernels.
//
// The reference includes the OptiX SDK's sutil/vec_math.h everywhere (/root/reference/include/kernel/math.h:4),
// which is not in the snapshot; these are the operators its device headers use, with the SDK semantics:
// normalize = v * (1/sqrtf(dot)), vec/scalar = multiply by the reciprocal, reflect(i,n) = i - 2 n dot(n,i),
// lerp(a,b,t) = a + t (b-a), clamp = fmaxf(a, fminf(f,b)). Compiled with -fmad=false: every expression
// rounds exactly as written, so results match a host build with -ffp-contract=off operation for operation.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define HJ_DI __device__ __forceinline__
#define HJ_NI __device__ __noinline__     // big shading routines: one copy each keeps k_shade inside the instruction cache

#ifndef M_PIf
#define M_PIf 3.14159265358979323846f
#endif

HJ_DI float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
HJ_DI float3 f3(float s) { return make_float3(s, s, s); }
HJ_DI float3 f3(const float4& a) { return make_float3(a.x, a.y, a.z); }
HJ_DI float4 f4(const float3& a, float w) { return make_float4(a.x, a.y, a.z, w); }

HJ_DI float3 operator-(const float3& a) { return f3(-a.x, -a.y, -a.z); }
HJ_DI float3 operator+(const float3& a, const float3& b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
HJ_DI float3 operator-(const float3& a, const float3& b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
HJ_DI float3 operator*(const float3& a, const float3& b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
HJ_DI float3 operator*(const float3& a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
HJ_DI float3 operator*(float s, const float3& a) { return f3(s * a.x, s * a.y, s * a.z); }
HJ_DI float3 operator/(const float3& a, const float3& b) { return f3(a.x / b.x, a.y / b.y, a.z / b.z); }
HJ_DI float3 operator/(const float3& a, float s) { float inv = 1.0f / s; return a * inv; }
HJ_DI float3 operator+(const float3& a, float b) { return f3(a.x + b, a.y + b, a.z + b); }
HJ_DI float3 operator+(float a, const float3& b) { return f3(a + b.x, a + b.y, a + b.z); }
HJ_DI float3 operator-(float a, const float3& b) { return f3(a - b.x, a - b.y, a - b.z); }
HJ_DI void operator+=(float3& a, const float3& b) { a.x += b.x; a.y += b.y; a.z += b.z; }
HJ_DI void operator*=(float3& a, const float3& b) { a.x *= b.x; a.y *= b.y; a.z *= b.z; }
HJ_DI void operator*=(float3& a, float s) { a.x *= s; a.y *= s; a.z *= s; }
HJ_DI void operator/=(float3& a, float s) { float inv = 1.0f / s; a *= inv; }

HJ_DI float2 operator+(const float2& a, const float2& b) { return make_float2(a.x + b.x, a.y + b.y); }
HJ_DI float2 operator*(float s, const float2& a) { return make_float2(s * a.x, s * a.y); }

HJ_DI float4 operator+(const float4& a, const float4& b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
HJ_DI float4 operator-(const float4& a, const float4& b) { return make_float4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
HJ_DI float4 operator*(float s, const float4& a) { return make_float4(s * a.x, s * a.y, s * a.z, s * a.w); }

HJ_DI float dot(const float3& a, const float3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
HJ_DI float dot(const float4& a, const float4& b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
HJ_DI float3 cross(const float3& a, const float3& b) {
    return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
HJ_DI float length(const float3& v) { return sqrtf(dot(v, v)); }
HJ_DI float3 normalize(const float3& v) { float invLen = 1.0f / sqrtf(dot(v, v)); return v * invLen; }
HJ_DI float3 reflect(const float3& i, const float3& n) { return i - 2.0f * n * dot(n, i); }
HJ_DI float clampf(float f, float a, float b) { return fmaxf(a, fminf(f, b)); }
HJ_DI float3 clamp3(const float3& v, float a, float b) { return f3(clampf(v.x, a, b), clampf(v.y, a, b), clampf(v.z, a, b)); }
HJ_DI float3 lerp3(const float3& a, const float3& b, float t) { return a + t * (b - a); }
HJ_DI float4 lerp4(const float4& a, const float4& b, float t) { return a + t * (b - a); }
HJ_DI float comp(const float3& v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : v.z); }
HJ_DI bool is_zero3(const float3& v) { return v.x == 0.0f && v.y == 0.0f && v.z == 0.0f; }

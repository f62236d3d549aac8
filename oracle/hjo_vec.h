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

// Elementary functions of the shading code. Default build (libhjo.so): the portable implementations of the product's
// henjou-renderer_b200/csrc/hj_libm.h — one source for kernel and oracle, so the GPU comparison is bit-exact.
// -DHJO_GLIBC_LIBM (libhjo_glibc.so): glibc's, as the reference-headers build (_ref/libhjref.so) uses them; this build is the
// bridge: it matches _ref to float rounding, and differs from the default build ONLY in these seven functions.
#ifdef HJO_GLIBC_LIBM
static inline float hjo_sinf(float x) { return sinf(x); }
static inline float hjo_cosf(float x) { return cosf(x); }
static inline float hjo_acosf(float x) { return acosf(x); }
static inline float hjo_atan2f(float y, float x) { return atan2f(y, x); }
static inline float hjo_logf(float x) { return logf(x); }
static inline float hjo_expf(float x) { return expf(x); }
static inline float hjo_powf(float x, float y) { return powf(x, y); }
#else
#include "../henjou-renderer_b200/csrc/hj_libm.h"
static inline float hjo_sinf(float x) { return hjm_sinf(x); }
static inline float hjo_cosf(float x) { return hjm_cosf(x); }
static inline float hjo_acosf(float x) { return hjm_acosf(x); }
static inline float hjo_atan2f(float y, float x) { return hjm_atan2f(y, x); }
static inline float hjo_logf(float x) { return hjm_logf(x); }
static inline float hjo_expf(float x) { return hjm_expf(x); }
static inline float hjo_powf(float x, float y) { return hjm_powf(x, y); }
#endif

// ---- decision trace (hjo_path_events, tools/parity_margins.py): every discrete decision of ONE path together with the
// two values it compared, so that two builds of this oracle that differ only in libm can be aligned event by event and the
// first diverging decision of a mismatching path shown to sit within a few ulp of its threshold. Null outside that tool.
struct hjo_event { uint32_t kind, outcome; float a, b; };
struct HjoEventLog { hjo_event* ev; uint32_t cap, n; };
static thread_local HjoEventLog* g_hjo_events = nullptr;
enum { HJO_EV_ROULETTE = 1, HJO_EV_LOBE_DIFFUSE = 2, HJO_EV_LOBE_SPECULAR = 3, HJO_EV_BELOW_HORIZON = 4, HJO_EV_GLASS_INSIDE = 5,
       HJO_EV_GLASS_FRESNEL = 6, HJO_EV_GLASS_TIR = 7, HJO_EV_GGX_UP = 8, HJO_EV_GGX_DOWN = 9, HJO_EV_GGX_GRAZING = 10,
       HJO_EV_GGX_ESCAPE = 11, HJO_EV_GGX_BELOW = 12, HJO_EV_LIGHT_PICK = 13, HJO_EV_HIT = 20, HJO_EV_SHADOW = 21 };
static inline void hjo_note(uint32_t kind, uint32_t outcome, float a, float b) {
    HjoEventLog* L = g_hjo_events;
    if (L && L->n < L->cap) L->ev[L->n] = hjo_event{kind, outcome, a, b};
    if (L) L->n++;
}
static inline bool hjo_lt(uint32_t kind, float a, float b) { const bool r = a < b; if (g_hjo_events) hjo_note(kind, r, a, b); return r; }
static inline bool hjo_gt(uint32_t kind, float a, float b) { const bool r = a > b; if (g_hjo_events) hjo_note(kind, r, a, b); return r; }

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

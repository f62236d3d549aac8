// hj_sampling.cuh — device restatement of the reference's RNG, sampling/frame helpers and tonemaps:
// /root/reference/include/kernel/cmj.h, math.h, color.h and the host sRGB quantiser
// include/renderer/renderer.h:73-101. Same draw order and expression trees as the reference (quirks kept,
// SURVEY §8a): each function cites the lines it follows.
#pragma once
#include "hj_vec.cuh"
#include "hj_libm.h"
#include <float.h>

#define HJ_PI 3.14159265358979323846f       // common/constant.h:3
#define HJ_TWO_PI 6.28318530717958647692f   // common/constant.h:4
#define HJ_INV_PI 0.31830988618379067154f   // common/constant.h:6

// The elementary functions are the shared ones of hj_libm.h (same bits as the CPU oracle build, which includes the same
// header). One out-of-line copy of each: they are 60-150 instructions apiece, the shading kernels call them from ~40 sites
// and must stay inside the instruction cache.
static HJ_NI float hj_powf(float a, float b) { return hjm_powf(a, b); }
static HJ_NI float hj_sinf(float a) { return hjm_sinf(a); }
static HJ_NI float hj_cosf(float a) { return hjm_cosf(a); }
static HJ_NI void hj_sincosf(float a, float& s, float& c) { hjm_sincosf(a, s, c); }
static HJ_NI float hj_acosf(float a) { return hjm_acosf(a); }
static HJ_NI float hj_logf(float a) { return hjm_logf(a); }
static HJ_NI float hj_expf(float a) { return hjm_expf(a); }
static HJ_NI float hj_atan2f(float y, float x) { return hjm_atan2f(y, x); }

// cmj.h:38-51 xxhash32(uint4)
HJ_DI uint32_t xxhash32_u4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
    const uint32_t P2 = 2246822519u, P3 = 3266489917u, P4 = 668265263u, P5 = 374761393u;
    uint32_t h = w + P5 + x * P3;
    h = P4 * ((h << 17) | (h >> 15));
    h += y * P3;
    h = P4 * ((h << 17) | (h >> 15));
    h += z * P3;
    h = P4 * ((h << 17) | (h >> 15));
    h = P2 * (h ^ (h >> 15));
    h = P3 * (h ^ (h >> 13));
    return h ^ (h >> 16);
}

// cmj.h:53-58 CMJState. n_spp = global sample index, scramble = frame, image_idx = y*W+x.
struct Sampler {
    uint32_t n_spp;      // the reference's u64 is truncated to 32 bits inside the hash (cmj.h:123)
    uint32_t scramble;
    uint32_t depth;
    uint32_t image_idx;
};

// cmj.h:60-91 (Kensler's permute)
HJ_DI uint32_t cmj_permute(uint32_t i, uint32_t l, uint32_t p) {
    uint32_t w = l - 1;
    w |= w >> 1; w |= w >> 2; w |= w >> 4; w |= w >> 8; w |= w >> 16;
    do {
        i ^= p;             i *= 0xe170893d;
        i ^= p >> 16;
        i ^= (i & w) >> 4;
        i ^= p >> 8;        i *= 0x0929eb3f;
        i ^= p >> 23;
        i ^= (i & w) >> 1;  i *= 1 | p >> 27;
        i *= 0x6935fa69;
        i ^= (i & w) >> 11; i *= 0x74dcb303;
        i ^= (i & w) >> 2;  i *= 0x9e501cc3;
        i ^= (i & w) >> 2;  i *= 0xc860a3df;
        i &= w;
        i ^= i >> 5;
    } while (i >= l);
    return (i + p) % l;
}
// cmj.h:93-106 (4294967808 divisor, sic)
HJ_DI float cmj_randfloat(uint32_t i, uint32_t p) {
    i ^= p;
    i ^= i >> 17; i ^= i >> 10; i *= 0xb36534e5;
    i ^= i >> 12; i ^= i >> 21; i *= 0x93fc4795;
    i ^= 0xdf6e307f;
    i ^= i >> 17; i *= 1 | p >> 18;
    return i * (1.0f / 4294967808.0f);
}
// cmj.h:108-117, M = N = 4
HJ_DI float2 cmj_point(uint32_t index, uint32_t scramble) {
    index = cmj_permute(index, 16, scramble * 0x51633e2d);
    uint32_t sx = cmj_permute(index % 4, 4, scramble * 0xa511e9b3);
    uint32_t sy = cmj_permute(index / 4, 4, scramble * 0x63d83595);
    float jx = cmj_randfloat(index, scramble * 0xa399d265);
    float jy = cmj_randfloat(index, scramble * 0x711ad6a5);
    return make_float2((index % 4 + (sy + jx) / 4) / 4, (index / 4 + (sx + jy) / 4) / 4);
}
// cmj.h:119-128: one call = one dimension
static HJ_NI float2 draw2(Sampler& s) {
    const uint32_t index = s.n_spp % 16u;
    const uint32_t scr = xxhash32_u4(s.n_spp / 16u, s.image_idx, s.depth, s.scramble);
    const float2 r = cmj_point(index, scr);
    s.depth++;
    return r;
}
// cmj.h:130-133: cmj_1d burns a full 2-D draw and keeps x. Only the terms x depends on are evaluated (the sx permute
// and the jy jitter feed y alone): same bits as draw2(s).x.
static HJ_NI float draw1(Sampler& s) {
    const uint32_t scr = xxhash32_u4(s.n_spp / 16u, s.image_idx, s.depth, s.scramble);
    const uint32_t index = cmj_permute(s.n_spp % 16u, 16, scr * 0x51633e2d);
    const uint32_t sy = cmj_permute(index / 4, 4, scr * 0x63d83595);
    const float jx = cmj_randfloat(index, scr * 0xa399d265);
    s.depth++;
    return (index % 4 + (sy + jx) / 4) / 4;
}
// a draw whose value is never used (rt.h:266 / 426): only the dimension counter moves
HJ_DI void draw_skip(Sampler& s) { s.depth++; }

// ------------------------------------------------------------------ math.h
// math.h:26-29
HJ_DI float3 schlick3(const float3& F0, const float3& w, const float3& n) {
    float term1 = 1.0f - dot(w, n);
    return (1.0f - F0) * hj_powf(term1, 5.0f) + F0;
}
// math.h:31-37
HJ_DI float schlick_ior(float no, float ni, const float3& w, const float3& n) {
    float F0 = (no - ni) / (no + ni);
    F0 = F0 * F0;
    float term1 = 1.0f - dot(w, n);
    return F0 + (1.0f - F0) * hj_powf(term1, 5.0f);
}
// math.h:43-51
HJ_DI void onb(const float3& n, float3& t, float3& b) {
    float sign = copysignf(1.0f, n.z);
    const float a = -1.0f / (sign + n.z);
    const float bb = n.x * n.y * a;
    t = f3(1.0f + sign * n.x * n.x * a, sign * bb, -sign * n.x);
    b = f3(bb, sign + n.y * n.y * a, -n.y);
}
// math.h:53-59 / 61-71 : y-up local frame, (t, n, b)
HJ_DI float3 to_local(const float3& v, const float3& t, const float3& n, const float3& b) {
    return f3(dot(v, t), dot(v, n), dot(v, b));
}
HJ_DI float3 to_world(const float3& v, const float3& t, const float3& n, const float3& b) {
    return f3(v.x * t.x + v.y * n.x + v.z * b.x,
              v.x * t.y + v.y * n.y + v.z * b.y,
              v.x * t.z + v.y * n.z + v.z * b.z);
}
// math.h:88-90
HJ_DI float norm2(const float3& v) { return v.x * v.x + v.y * v.y + v.z * v.z; }
// math.h:92-103
HJ_DI bool refract_dir(const float3& v, const float3& n, float ior1, float ior2, float3& r) {
    const float3 t_h = -ior1 / ior2 * (v - dot(v, n) * n);
    if (norm2(t_h) > 1.0f) return false;
    const float3 t_p = -sqrtf(fmaxf(1.0f - norm2(t_h), 0.0f)) * n;
    r = t_h + t_p;
    return true;
}
// math.h:105-107
HJ_DI float absdot(const float3& a, const float3& b) { return fabsf(dot(a, b)); }
// math.h:109-111
HJ_DI float lerp1(float a, float b, float t) { return (1 - t) * a + t * b; }
// math.h:113-120
HJ_DI float smoothstep1(float a, float b, float t) {
    float x = clampf((t - a) / (b - a), 0.0f, 1.0f);
    return x * x * (3.0f - 2.0f * x);
}
HJ_DI float step1(float a, float x) { return float(a < x); }

// ------------------------------------------------------------------ color.h
// color.h:10-29 (double-precision literals kept: 1.0 - ...)
HJ_DI float uchimura1(float x, float P, float a, float m, float l, float c, float b) {
    float l0 = ((P - m) * l) / a;
    float S0 = m + l0;
    float S1 = m + a * l0;
    float C2 = (a * P) / (P - S1);
    float CP = -C2 / P;
    float w0 = 1.0 - smoothstep1(0.0, m, x);
    float w2 = step1(m + l0, x);
    float w1 = 1.0 - w0 - w2;
    float T = m * hj_powf(x / m, c) + b;
    float S = P - (P - S1) * hj_expf(CP * (x - S0));
    float L = m + a * (x - m);
    return T * w0 + L * w1 + S * w2;
}
// color.h:40-53
HJ_DI float3 tonemap_uchimura(const float3& x) {
    const float P = 1.0, a = 1.0, m = 0.22, l = 0.4, c = 1.33, b = 0.0;
    return f3(uchimura1(x.x, P, a, m, l, c, b), uchimura1(x.y, P, a, m, l, c, b), uchimura1(x.z, P, a, m, l, c, b));
}
// color.h:55-63
HJ_DI float3 tonemap_aces(float3 x) {
    float a = 2.51f, b = 0.03f, c = 2.43f, d = 0.59f, e = 0.14f;
    return clamp3((x * (a * x + b)) / (x * (c * x + d) + e), 0.0f, 1.0f);
}
// renderer.h:73-85 toSRGB + quantizeUnsignedChar (host code in the reference; done on the device here)
HJ_DI float3 to_srgb(const float3& col) {
    float ig = 1.0f / 2.4f;
    float3 p = f3(hj_powf(col.x, ig), hj_powf(col.y, ig), hj_powf(col.z, ig));
    return f3(col.x < 0.0031308f ? 12.92f * col.x : 1.055f * p.x - 0.055f,
              col.y < 0.0031308f ? 12.92f * col.y : 1.055f * p.y - 0.055f,
              col.z < 0.0031308f ? 12.92f * col.z : 1.055f * p.z - 0.055f);
}
HJ_DI unsigned char quantize8(float x) {
    if (!(x > 0.0f)) return 0;
    float s = x * 256.0f;
    if (s >= 255.0f) return 255;
    return (unsigned char)(unsigned int)s;
}

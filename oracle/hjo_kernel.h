// hjo_kernel.h — TEST INFRASTRUCTURE (CPU oracle). Not part of the product path.
//
// CPU restatement of the reference's device library /root/reference/include/kernel/*.h:
// same draw order, same expression trees, same quirks (SURVEY §8a "bug-for-bug notes").
// Each function cites the reference lines it follows. Build: g++ -O2 -ffp-contract=off.
// Pinned against (a) the KAT block of SURVEY §8c and (b) oracle/_ref, which compiles the
// reference's own headers unmodified (tests/test_oracle_*.py).
#pragma once
#include "hjo_common.h"

namespace hjo {

constexpr float kPi = 3.14159265358979323846f;      // common/constant.h:3
constexpr float kTwoPi = 6.28318530717958647692f;   // common/constant.h:4
constexpr float kInvPi = 0.31830988618379067154f;   // common/constant.h:6

// ------------------------------------------------------------------ RNG: cmj.h
// cmj.h:14-23
static inline uint32_t xxhash32_u(uint32_t p) {
    const uint32_t P2 = 2246822519u, P3 = 3266489917u, P4 = 668265263u, P5 = 374761393u;
    uint32_t h = p + P5;
    h = P4 * ((h << 17) | (h >> 15));
    h = P2 * (h ^ (h >> 15));
    h = P3 * (h ^ (h >> 13));
    return h ^ (h >> 16);
}
// cmj.h:38-51
static inline uint32_t xxhash32_u4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
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

// cmj.h:53-58
struct Sampler {
    unsigned long long n_spp = 0;
    uint32_t scramble = 0;
    uint32_t depth = 0;
    uint32_t image_idx = 0;
};

// cmj.h:60-91 (Kensler's permute)
static inline uint32_t permute(uint32_t i, uint32_t l, uint32_t p) {
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
// cmj.h:93-106 (note the 4294967808 divisor, sic)
static inline float randfloat(uint32_t i, uint32_t p) {
    i ^= p;
    i ^= i >> 17; i ^= i >> 10; i *= 0xb36534e5;
    i ^= i >> 12; i ^= i >> 21; i *= 0x93fc4795;
    i ^= 0xdf6e307f;
    i ^= i >> 17; i *= 1 | p >> 18;
    return i * (1.0f / 4294967808.0f);
}
// cmj.h:108-117, M = N = 4
static inline float2 cmj_point(uint32_t index, uint32_t scramble) {
    index = permute(index, 16, scramble * 0x51633e2d);
    uint32_t sx = permute(index % 4, 4, scramble * 0xa511e9b3);
    uint32_t sy = permute(index / 4, 4, scramble * 0x63d83595);
    float jx = randfloat(index, scramble * 0xa399d265);
    float jy = randfloat(index, scramble * 0x711ad6a5);
    return make_float2((index % 4 + (sy + jx) / 4) / 4, (index / 4 + (sx + jy) / 4) / 4);
}
// cmj.h:119-128: one call = one dimension; n_spp/16 is truncated to 32 bits by make_uint4
static inline float2 draw2(Sampler& s) {
    const uint32_t index = (uint32_t)(s.n_spp % 16);
    const uint32_t scr = xxhash32_u4((uint32_t)(s.n_spp / 16), s.image_idx, s.depth, s.scramble);
    const float2 r = cmj_point(index, scr);
    s.depth++;
    return r;
}
// cmj.h:130-133: burns a full 2-D draw
static inline float draw1(Sampler& s) { return draw2(s).x; }

// ------------------------------------------------------------------ math.h
// math.h:7-15
static inline float3 cosine_sample(float u, float v, float& pdf) {
    float phi = 2.0f * M_PIf * v;
    float theta = 0.5f * hjo_acosf(1.0f - 2.0f * u);
    float ct = hjo_cosf(theta), st = hjo_sinf(theta);
    pdf = ct / M_PIf;
    return make_float3(hjo_cosf(phi) * st, ct, hjo_sinf(phi) * st);
}
// math.h:26-29
static inline float3 schlick3(const float3& F0, const float3& w, const float3& n) {
    float term1 = 1.0f - dot(w, n);
    return (1.0f - F0) * hjo_powf(term1, 5.0f) + F0;
}
// math.h:31-37
static inline float schlick_ior(float no, float ni, const float3& w, const float3& n) {
    float F0 = (no - ni) / (no + ni);
    F0 = F0 * F0;
    float term1 = 1.0f - dot(w, n);
    return F0 + (1.0f - F0) * hjo_powf(term1, 5.0f);
}
// math.h:43-51
static inline void onb(const float3& n, float3& t, float3& b) {
    float sign = copysignf(1.0f, n.z);
    const float a = -1.0f / (sign + n.z);
    const float bb = n.x * n.y * a;
    t = make_float3(1.0f + sign * n.x * n.x * a, sign * bb, -sign * n.x);
    b = make_float3(bb, sign + n.y * n.y * a, -n.y);
}
// math.h:53-59 / 61-71 : y-up local frame, (t, n, b)
static inline float3 to_local(const float3& v, const float3& t, const float3& n, const float3& b) {
    return make_float3(dot(v, t), dot(v, n), dot(v, b));
}
static inline float3 to_world(const float3& v, const float3& t, const float3& n, const float3& b) {
    return make_float3(v.x * t.x + v.y * n.x + v.z * b.x,
                       v.x * t.y + v.y * n.y + v.z * b.y,
                       v.x * t.z + v.y * n.z + v.z * b.z);
}
// math.h:88-90
static inline float norm2(const float3& v) { return v.x * v.x + v.y * v.y + v.z * v.z; }
// math.h:92-103
static inline bool refract_dir(const float3& v, const float3& n, float ior1, float ior2, float3& r) {
    const float3 t_h = -ior1 / ior2 * (v - dot(v, n) * n);
    if (hjo_gt(HJO_EV_GLASS_TIR, norm2(t_h), 1.0f)) return false;
    const float3 t_p = -sqrtf(fmaxf(1.0f - norm2(t_h), 0.0f)) * n;
    r = t_h + t_p;
    return true;
}
// math.h:105-107
static inline float absdot(const float3& a, const float3& b) { return fabsf(dot(a, b)); }
// math.h:109-111 (scalar lerp of the reference's own header)
static inline float lerp1(float a, float b, float t) { return (1 - t) * a + t * b; }
// math.h:113-116, :118-120
static inline float smoothstep1(float a, float b, float t) {
    float x = clamp((t - a) / (b - a), 0.0f, 1.0f);
    return x * x * (3.0f - 2.0f * x);
}
static inline float step1(float a, float x) { return float(a < x); }

// ------------------------------------------------------------------ color.h
// color.h:10-29: note the double-precision literals (1.0 - ...) in the reference
static inline float uchimura1(float x, float P, float a, float m, float l, float c, float b) {
    float l0 = ((P - m) * l) / a;
    float L0 = m - m / a; (void)L0;
    float L1 = m + (1.0 - m) / a; (void)L1;
    float S0 = m + l0;
    float S1 = m + a * l0;
    float C2 = (a * P) / (P - S1);
    float CP = -C2 / P;
    float w0 = 1.0 - smoothstep1(0.0, m, x);
    float w2 = step1(m + l0, x);
    float w1 = 1.0 - w0 - w2;
    float T = m * hjo_powf(x / m, c) + b;
    float S = P - (P - S1) * hjo_expf(CP * (x - S0));
    float L = m + a * (x - m);
    return T * w0 + L * w1 + S * w2;
}
// color.h:40-53
static inline float3 tonemap_uchimura(const float3& x) {
    const float P = 1.0, a = 1.0, m = 0.22, l = 0.4, c = 1.33, b = 0.0;
    return make_float3(uchimura1(x.x, P, a, m, l, c, b), uchimura1(x.y, P, a, m, l, c, b), uchimura1(x.z, P, a, m, l, c, b));
}
// color.h:55-63
static inline float3 tonemap_aces(float3 x) {
    float a = 2.51f, b = 0.03f, c = 2.43f, d = 0.59f, e = 0.14f;
    return clamp((x * (a * x + b)) / (x * (c * x + d) + e), 0.0f, 1.0f);
}
// renderer.h:73-85 host toSRGB + quantizeUnsignedChar
static inline float3 to_srgb(const float3& col) {
    float ig = 1.0f / 2.4f;
    float3 p = make_float3(hjo_powf(col.x, ig), hjo_powf(col.y, ig), hjo_powf(col.z, ig));
    return make_float3(col.x < 0.0031308f ? 12.92f * col.x : 1.055f * p.x - 0.055f,
                       col.y < 0.0031308f ? 12.92f * col.y : 1.055f * p.y - 0.055f,
                       col.z < 0.0031308f ? 12.92f * col.z : 1.055f * p.z - 0.055f);
}
static inline unsigned char quantize8(float x) {
    // (unsigned)(negative or NaN float) is UB on the host; the device cvt saturates to 0. Define: 0.
    if (!(x > 0.0f)) return 0;
    float s = x * 256.0f;
    if (s >= 255.0f) return 255;
    return (unsigned char)(unsigned int)s;
}

}  // namespace hjo

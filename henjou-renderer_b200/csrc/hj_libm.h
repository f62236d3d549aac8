// hj_libm.h — the single-precision elementary functions of the shading path (sinf, cosf, acosf, atan2f, logf, expf,
// powf), written once in portable IEEE arithmetic so that the sm_100a kernels and a host build produce THE SAME BITS.
//
// Why: the reference's device code (/root/reference/include/kernel/disneyBRDF.h:30-38,64-80,93-100,131-139,
// BSDFs.h:551-586,616-632, math.h:26-37, color.h:10-29, renderer.h:73-85) calls libm. CUDA's device libm and glibc
// differ by <= 2 ulp, and a path whose discrete decision (lobe select, Russian roulette, Fresnel branch) sits within
// that of its threshold takes another branch — round 1 had to tolerate 0.3-1 % of such paths in its parity tests.
// With one implementation on both sides the per-path comparison against the CPU restatement is exact.
//
// How: every function converts to double, reduces the range with explicit fma()/rint() (correctly rounded on both
// sides) and evaluates a near-minimax polynomial (tools/libm_coeffs.py: interpolation at Chebyshev nodes, mpmath at 50
// digits) by Horner steps of explicit fma(); only +, -, *, /, sqrt, fma, rint and integer bit operations are used, all
// of which are IEEE-exact under `nvcc -fmad=false -prec-div=true -prec-sqrt=true` and `g++ -ffp-contract=off`. The
// double result is within ~1e-13 relative of the true value, so the final conversion to float is the correctly
// rounded result except in ~1 of 10^5 arguments (tests/test_libm.py measures <= 0.5000x ulp against mpmath and <= 1 ulp
// against glibc). Special cases follow C99 Annex F (what glibc and CUDA both implement). No float->int conversion of
// an unbounded value (undefined on the host, saturating on the device) appears anywhere.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define HJM_FN __host__ __device__ __forceinline__
#else
#define HJM_FN static inline
#endif

HJM_FN uint64_t hjm_bits(double d) {
#if defined(__CUDA_ARCH__)
    return (uint64_t)__double_as_longlong(d);
#else
    uint64_t u; memcpy(&u, &d, 8); return u;
#endif
}
HJM_FN double hjm_from_bits(uint64_t u) {
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)u);
#else
    double d; memcpy(&d, &u, 8); return d;
#endif
}
HJM_FN uint32_t hjm_bitsf(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
HJM_FN float hjm_from_bitsf(uint32_t u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}
HJM_FN float hjm_nanf() { return hjm_from_bitsf(0x7fc00000u); }
HJM_FN bool hjm_isnanf(float x) { return (hjm_bitsf(x) & 0x7fffffffu) > 0x7f800000u; }
HJM_FN bool hjm_isinff(float x) { return (hjm_bitsf(x) & 0x7fffffffu) == 0x7f800000u; }
HJM_FN bool hjm_signf(float x) { return (hjm_bitsf(x) >> 31) != 0u; }

// ---------------------------------------------------------------------------------------------- sin / cos
// r = x - k pi/2 (two-constant Cody-Waite in double: exact enough for every float below ~1e9), quadrant from k mod 4
// computed in double. sin(r) = r + r^3 S(r^2), cos(r) = 1 - r^2/2 + r^4 C(r^2) on |r| <= pi/4.
HJM_FN void hjm_sincos_reduced(double xd, double& s, double& c, int& q) {
    const double k = rint(xd * 0.6366197723675814);
    double r = fma(-k, 1.5707963267948966, xd);
    r = fma(-k, 6.123233995736766e-17, r);
    const double kq = k - 4.0 * rint(k * 0.25);          // in {-2,...,2}: safe to convert
    q = (int)kq & 3;
    const double z = r * r;
    double ps = 1.5917771759341917e-10;
    ps = fma(ps, z, -2.5051126876085338e-08); ps = fma(ps, z, 2.755731607868067e-06); ps = fma(ps, z, -0.00019841269836712495);
    ps = fma(ps, z, 0.008333333333330918); ps = fma(ps, z, -0.16666666666666666);
    s = fma(r * z, ps, r);
    double pc = -1.1382408754153254e-11;
    pc = fma(pc, z, 2.087614315990944e-09); pc = fma(pc, z, -2.755731725679604e-07); pc = fma(pc, z, 2.4801587298736863e-05);
    pc = fma(pc, z, -0.0013888888888887378); pc = fma(pc, z, 0.041666666666666664);
    c = fma(z * z, pc, fma(-0.5, z, 1.0));
}
HJM_FN void hjm_sincosf(float x, float& sn, float& cs) {
    if (hjm_isnanf(x) || hjm_isinff(x)) { sn = cs = hjm_nanf(); return; }
    double s, c; int q;
    hjm_sincos_reduced((double)x, s, c, q);
    const double ss = (q & 1) ? c : s, cc = (q & 1) ? s : c;
    sn = (float)((q & 2) ? -ss : ss);
    cs = (float)(((q + 1) & 2) ? -cc : cc);
    if (x == 0.0f) sn = x;                               // sin(-0) = -0
}
HJM_FN float hjm_sinf(float x) { float s, c; hjm_sincosf(x, s, c); return s; }
HJM_FN float hjm_cosf(float x) { float s, c; hjm_sincosf(x, s, c); return c; }

// ---------------------------------------------------------------------------------------------- acos
// asin(t) = t + t z R(z), z = t^2 <= 0.25. |x| <= 0.5: acos = pi/2 - asin(x); else with z = (1-|x|)/2, t = sqrt(z):
// acos(|x|) = 2 asin(t), acos(-|x|) = pi - 2 asin(t).
HJM_FN double hjm_asin_kernel(double t, double z) {
    double p = 0.028187667142318098;
    p = fma(p, z, -0.010773195325585556); p = fma(p, z, 0.01604912458494932); p = fma(p, z, 0.0077986264075622334);
    p = fma(p, z, 0.011876344296279416); p = fma(p, z, 0.013929545950485803); p = fma(p, z, 0.01735526856395352);
    p = fma(p, z, 0.022372047201808885); p = fma(p, z, 0.030381947379559158); p = fma(p, z, 0.04464285710323956);
    p = fma(p, z, 0.07500000000020869); p = fma(p, z, 0.1666666666666665);
    return fma(t * z, p, t);
}
HJM_FN float hjm_acosf(float x) {
    if (hjm_isnanf(x)) return hjm_nanf();
    const double xd = (double)x, ax = fabs(xd);
    if (ax > 1.0) return hjm_nanf();
    if (ax <= 0.5) return (float)(1.5707963267948966 - (hjm_asin_kernel(xd, xd * xd) - 6.123233995736766e-17));
    const double z = (1.0 - ax) * 0.5, t = sqrt(z);
    const double a = 2.0 * hjm_asin_kernel(t, z);
    return (float)(xd > 0.0 ? a : (3.141592653589793 - (a - 1.2246467991473532e-16)));
}

// ---------------------------------------------------------------------------------------------- atan2
// atan(u) = u + u z T(z) on |u| <= sqrt2 - 1; (sqrt2-1, 1]: atan(t) = pi/4 + atan((t-1)/(t+1)).
HJM_FN double hjm_atan01(double t) {                     // t in [0, 1]
    double u = t, base = 0.0;
    if (t > 0.41421356237309503) { u = (t - 1.0) / (t + 1.0); base = 0.7853981633974483; }
    const double z = u * u;
    double p = -0.019168248850065444;
    p = fma(p, z, 0.03922460271658807); p = fma(p, z, -0.0508520544113668); p = fma(p, z, 0.058581022151876085);
    p = fma(p, z, -0.066645060702545); p = fma(p, z, 0.07692182808979683); p = fma(p, z, -0.09090904561682452);
    p = fma(p, z, 0.11111111014851602); p = fma(p, z, -0.1428571428466151); p = fma(p, z, 0.19999999999995496);
    p = fma(p, z, -0.3333333333333333);
    return base + fma(u * z, p, u);
}
HJM_FN float hjm_atan2f(float y, float x) {
    if (hjm_isnanf(x) || hjm_isnanf(y)) return hjm_nanf();
    const double ax = fabs((double)x), ay = fabs((double)y);
    double a;
    if (ax == 0.0 && ay == 0.0) a = 0.0;
    else if (hjm_isinff(x) && hjm_isinff(y)) a = 0.7853981633974483;
    else if (ay <= ax) a = hjm_atan01(ay / ax);
    else a = 1.5707963267948966 - hjm_atan01(ax / ay);
    if (hjm_signf(x)) a = 3.141592653589793 - a;
    float r = (float)a;
    return hjm_signf(y) ? -r : r;
}

// ---------------------------------------------------------------------------------------------- log / exp / pow
// log of a positive finite double that came from a float (always a normal double): x = 2^e m, m in [sqrt(1/2), sqrt 2),
// log m = 2 atanh(s) = 2 s (1 + z L(z)), s = (m-1)/(m+1), z = s^2 <= 0.0295.
HJM_FN double hjm_log_pos(double x) {
    const uint64_t u = hjm_bits(x);
    int e = (int)((u >> 52) & 0x7ffu) - 1023;
    double m = hjm_from_bits((u & 0x000fffffffffffffull) | 0x3ff0000000000000ull);
    if (m > 1.4142135623730951) { m *= 0.5; e += 1; }
    const double s = (m - 1.0) / (m + 1.0), z = s * s;
    double p = 0.07309671726756307;
    p = fma(p, z, 0.07665743617932867); p = fma(p, z, 0.09091448091741543); p = fma(p, z, 0.11111105518908294);
    p = fma(p, z, 0.14285714313285103); p = fma(p, z, 0.19999999999949095); p = fma(p, z, 0.3333333333333335);
    const double lm = 2.0 * fma(s * z, p, s);
    const double ed = (double)e;
    return fma(ed, 0.6931471803691238, fma(ed, 1.9082149292705877e-10, lm));     // ln2 split: hi has 21 trailing zero bits
}
// exp of a finite double, |t| clamped to 200 (far outside the float range either way, well inside the double range):
// t = k ln2 + r, exp(r) = 1 + r + r^2 E(r), scaled by 2^k in two exact steps.
HJM_FN double hjm_exp_d(double t) {
    t = t > 200.0 ? 200.0 : (t < -200.0 ? -200.0 : t);
    const double k = rint(t * 1.4426950408889634);
    double r = fma(-k, 0.6931471803691238, t);
    r = fma(-k, 1.9082149292705877e-10, r);
    double p = 2.0914701283103285e-09;
    p = fma(p, r, 2.51052370507603e-08); p = fma(p, r, 2.7557273608709007e-07); p = fma(p, r, 2.7557255352032673e-06);
    p = fma(p, r, 2.4801587325574868e-05); p = fma(p, r, 0.00019841269874858612); p = fma(p, r, 0.0013888888888883742);
    p = fma(p, r, 0.008333333333326124); p = fma(p, r, 0.04166666666666667); p = fma(p, r, 0.1666666666666667);
    p = fma(p, r, 0.5);
    const double v = fma(r * r, p, r) + 1.0;
    const int ki = (int)k;                               // |k| <= 289
    const int k1 = ki / 2, k2 = ki - k1;
    const double s1 = hjm_from_bits((uint64_t)(k1 + 1023) << 52), s2 = hjm_from_bits((uint64_t)(k2 + 1023) << 52);
    return (v * s1) * s2;
}
HJM_FN float hjm_logf(float x) {
    if (hjm_isnanf(x)) return hjm_nanf();
    if (x == 0.0f) return hjm_from_bitsf(0xff800000u);   // -inf
    if (hjm_signf(x)) return hjm_nanf();
    if (hjm_isinff(x)) return x;
    return (float)hjm_log_pos((double)x);
}
HJM_FN float hjm_expf(float x) {
    if (hjm_isnanf(x)) return hjm_nanf();
    return (float)hjm_exp_d((double)x);
}
// C99 F.9.4.4 ladder, then exp(y log|x|) in double.
HJM_FN float hjm_powf(float x, float y) {
    const float inf = hjm_from_bitsf(0x7f800000u);
    if (y == 0.0f || x == 1.0f) return 1.0f;
    if (hjm_isnanf(x) || hjm_isnanf(y)) return hjm_nanf();
    const float ax = fabsf(x), ay = fabsf(y);
    if (hjm_isinff(y)) {
        if (ax == 1.0f) return 1.0f;
        return ((ax > 1.0f) == (y > 0.0f)) ? inf : 0.0f;
    }
    // integer / odd test without an out-of-range conversion: floats >= 2^24 are even integers
    bool y_int, y_odd;
    if (ay >= 16777216.0f) { y_int = true; y_odd = false; }
    else { const float fl = floorf(ay); y_int = fl == ay; y_odd = y_int && (((uint32_t)fl) & 1u); }
    const bool neg = hjm_signf(x);
    if (hjm_isinff(x) || x == 0.0f) {
        const bool big = hjm_isinff(x) == (y > 0.0f);    // |result| is inf (else 0)
        const float mag = big ? inf : 0.0f;
        return (neg && y_odd) ? -mag : mag;
    }
    if (neg && !y_int) return hjm_nanf();
    if (y_int && ay <= 8.0f) {
        // small integer exponents (the Schlick terms pow(1 - cos, 5) of math.h:26-37 are most of the calls): repeated multiplication
        // in double — three or four roundings at 1e-16, then one rounding to float; no log / exp, the sign comes out by itself
        const double xd = (double)x, x2 = xd * xd, x4 = x2 * x2;
        const uint32_t k = (uint32_t)ay;
        double p = (k & 1u) ? xd : 1.0;
        if (k & 2u) p *= x2;
        if (k & 4u) p *= x4;
        if (k & 8u) p *= x4 * x4;
        return (float)(y > 0.0f ? p : 1.0 / p);
    }
    const float r = (float)hjm_exp_d((double)y * hjm_log_pos((double)ax));
    return (neg && y_odd) ? -r : r;
}

// hj_bsdf.cuh — device restatement of /root/reference/include/kernel/disneyBRDF.h and BSDFs.h: the three
// lobes the dispatcher (BSDFs.h:979-1038) can reach — DisneyBRDF (+ thin-film LUT), MetaMaterialGlass
// (negative-IOR BTDF) and EnagyConservationGGX (multiple-scattering random walk). IdealGlass / GGX /
// Lambert / FastMultipleGGX are dead code behind that dispatcher and are not built here.
// Local shading frame is y-up; "wo" is the direction back along the incoming ray.
#pragma once
#include "hj_scene.cuh"

// ------------------------------------------------------------------ DisneyBRDF (disneyBRDF.h:16-327)
struct Disney {
    float3 basecolor; float alpha, metallic, sheen, clearcoat, subsurface, cc_alpha; bool thinfilm;

    // disneyBRDF.h:165-177
    HJ_DI void init(const Surface& s) {
        basecolor = s.basecolor;
        alpha = clampf(s.roughness * s.roughness, 0.01f, 1.0f);
        subsurface = 0.0f;
        metallic = s.metallic; sheen = s.sheen; clearcoat = s.clearcoat;
        cc_alpha = lerp1(0.1f, 0.001f, 1.0f);   // clearcoatGloss hard-wired to 1 (:174-175)
        thinfilm = s.is_thinfilm;
    }
    HJ_DI float pdf_diffuse(const float3& wi) const { return fabsf(wi.y) * HJ_INV_PI; }            // :40-42
    HJ_DI float D(const float3& wm) const {                                                         // :44-48
        float t1 = wm.x * wm.x / (alpha * alpha) + wm.z * wm.z / (alpha * alpha) + wm.y * wm.y;
        float t2 = HJ_PI * alpha * alpha * t1 * t1;
        return 1.0f / t2;
    }
    HJ_DI float Lambda(const float3& w) const {                                                     // :58-61
        float delta = 1.0f + (alpha * alpha * w.x * w.x + alpha * alpha * w.z * w.z) / (w.y * w.y);
        return (-1.0f + sqrtf(delta)) * 0.5f;
    }
    HJ_DI float G1(const float3& w) const { return 1.0f / (1.0f + Lambda(w)); }                   // :50-52
    HJ_DI float G2(const float3& wi, const float3& wo) const { return 1.0f / (1.0f + Lambda(wi) + Lambda(wo)); }  // :54-56
    // :64-80 spherical-cap VNDF (arXiv 2306.05044); the fma is explicit in the reference
    HJ_DI float3 sample_vndf(const float2& uv, const float3& wo) const {
        float3 swo = normalize(f3(wo.x * alpha, wo.y, wo.z * alpha));
        float phi = 2.0f * HJ_PI * uv.x;
        float z = fmaf((1.0f - uv.y), (1.0f + swo.y), -swo.y);
        float st = sqrtf(clampf(1.0f - z * z, 0.0f, 1.0f));
        float sp, cp; hj_sincosf(phi, sp, cp);        // one range reduction for both: same bits as cosf(phi), sinf(phi)
        float x = st * cp, y = st * sp;
        float3 h = f3(x, z, y) + swo;
        return normalize(f3(h.x * alpha, h.y, h.z * alpha));
    }
    HJ_DI float pdf_specular(const float3& wm, const float3& wo) const {                            // :88-90
        return 0.25f * D(wm) * G1(wo) * absdot(wo, wm) / (absdot(wm, wo) * fabsf(wo.y));
    }
    HJ_DI static float cc_D(const float3& wm, float a) {                                            // :131-139
        float a2 = a * a;
        float t = 1.0f + (a2 - 1.0f) * wm.y * wm.y;
        return (a2 - 1.0f) / (HJ_PI * hj_logf(a2) * t);
    }
    HJ_DI static float cc_Lambda(const float3& w, float a) {                                        // :126-129
        float t1 = 1.0f + (a * a * w.x * w.x + a * a * w.z * w.z) / (w.y * w.y);
        return 0.5f * (-1.0f + sqrtf(t1));
    }
    HJ_DI float pdf_clearcoat(const float3& wm, const float3& wo) const {                           // :102-104
        return cc_D(wm, cc_alpha) * fabsf(wm.y) / (4.0f * fabsf(dot(wm, wo)));
    }
    HJ_DI float3 sample_clearcoat(const float2& uv, const float3& wo, float& pdf) const {           // :93-100
        float ct = sqrtf(fmaxf((1.0f - hj_powf(cc_alpha * cc_alpha, 1.0f - uv.x)) / (1.0f - cc_alpha * cc_alpha), 0.0f));
        float st = sqrtf(fmaxf(1.0f - ct * ct, 0.0f));
        float phi = HJ_TWO_PI * uv.y;
        float sp, cp; hj_sincosf(phi, sp, cp);
        float3 wm = f3(cp * st, ct, sp * st);
        pdf = pdf_clearcoat(wm, wo);
        return wm;
    }
    HJ_DI static float ft_schlick(float wn, float F90) {                                            // :106-109
        float d = fmaxf(1.0f - wn, 0.0f);
        return 1.0f + (F90 - 1.0f) * d * d * d * d * d;
    }
    HJ_DI float3 specular(const float3& wo, const float3& wi, const float3& F0) const {             // :112-120
        float3 wm = normalize(wo + wi);
        float d = D(wm), g = G2(wi, wo);
        float3 f = schlick3(F0, wo, wm);
        return 0.25f * f * d * g / (fabsf(wo.y) * fabsf(wi.y));
    }
    HJ_DI float3 clearcoat_lobe(const float3& wo, const float3& wi, float a) const {                // :142-150
        const float3 wm = normalize(wo + wi);
        float d = cc_D(wm, a);
        float g = 1.0f / (1.0f + cc_Lambda(wi, 0.25f) + cc_Lambda(wo, 0.25f));
        float3 f = schlick3(f3(0.04f), wo, wm);
        return 0.25f * d * g * f / (fabsf(wo.y) * fabsf(wi.y));
    }
    // :179-235
    HJ_NI float3 eval(const DevScene& S, const float3& wo, const float3& wi) const {
        float3 wm = normalize(wo + wi);
        float dot_wi_n = fabsf(wi.y);
        float dot_wo_n = fabsf(wi.y);               // sic (:189)
        float cd = absdot(wi, wm);
        float FD90 = 0.5f + 2.0f * alpha * cd * cd;
        float fi = ft_schlick(dot_wi_n, FD90), fo = ft_schlick(dot_wo_n, FD90);
        float3 f_diffuse = basecolor * fi * fo * HJ_INV_PI;
        float deltacos = 1.0f / (dot_wi_n + dot_wo_n) - 0.5f;
        float3 f_ss = basecolor * HJ_INV_PI * 1.25f * (fi * fo * deltacos + 0.5f);
        float3 F0 = lerp3(f3(0.08f), basecolor, metallic);
        if (thinfilm) {                             // :213-217
            float thickness = basecolor.x;
            float cosine = absdot(wi, wm);
            F0 = lut_fetch(S, thickness, cosine);
        }
        float3 f_spec = specular(wo, wi, F0);
        float d5 = fmaxf(1.0f - absdot(wi, wm), 0.0f);
        float3 f_sheen = sheen * f3(1.0f) * d5 * d5 * d5 * d5 * d5;
        float3 f_cc = 0.25f * clearcoat_lobe(wo, wi, cc_alpha);
        return (lerp3(f_diffuse, f_ss, subsurface) + f_sheen) * (1.0f - metallic) + f_spec + f_cc * clearcoat;
    }
    // :237-307 — two RNG calls: lobe select, then the 2-D sample
    HJ_NI float3 sample(const DevScene& S, const float3& wo, float3& wi, float& pdf, Sampler& st) const {
        float dwt = 1.0f * (1.0f - metallic), swt = 0.5f, cwt = 0.0f;
        float sum = dwt + swt + cwt;
        float dw = dwt / sum, sw = swt / sum, cw = cwt / sum;
        float sel = draw1(st);
        float pd = 1.0f, ps = 1.0f, pc = 1.0f;
        float2 xi = draw2(st);
        if (sel < dw) {
            float theta = 0.5f * hj_acosf(1.0f - 2.0f * xi.x);    // sampleDiffuse :30-38
            float phi = 2.0f * M_PIf * xi.y;
            float ct, sn, cp, sp; hj_sincosf(theta, sn, ct); hj_sincosf(phi, sp, cp);
            wi = f3(cp * sn, ct, sp * sn);
            pd = pdf_diffuse(wi);
            float3 wm = normalize(wi + wo);
            ps = pdf_specular(wm, wo);
            pc = pdf_clearcoat(wm, wo);
        } else if (sel < dw + sw) {
            float3 wm = sample_vndf(xi, wo);
            ps = pdf_specular(wm, wo);
            wi = reflect(-wo, wm);
            pd = pdf_diffuse(wi);
            pc = pdf_clearcoat(wm, wo);
        } else {
            float3 wm = sample_clearcoat(xi, wo, pc);
            wi = reflect(-wo, wm);
            pd = pdf_diffuse(wi);
            ps = pdf_specular(wm, wo);
        }
        pdf = dw * pd + sw * ps + cw * pc;
        if (wi.y < 0.0) { pdf = 1.0f; return f3(0.0f); }
        return eval(S, wo, wi);
    }
    // :309-326
    HJ_DI float pdf(const float3& wo, const float3& wi) const {
        float dwt = 1.0f * (1.0f - metallic), swt = 0.5f, cwt = 0.0f;
        float sum = dwt + swt + cwt;
        float dw = dwt / sum, sw = swt / sum;
        float3 wm = normalize(wo + wi);
        return dw * pdf_diffuse(wi) + sw * pdf_specular(wm, wo);
    }
};

// ------------------------------------------------------------------ MetaMaterialGlass::sampleBSDF (BSDFs.h:419-469)
// Schlick-Fresnel choice between reflection and NEGATIVE refraction: the refracted direction is mirrored
// about the normal, wi = reflect(-t,(0,-1,0)) (:454). Inside test is wo.y < 0 (:433). One RNG call.
HJ_DI float3 metaglass_sample(const float3& rho, float ior, const float3& wo, float3& wi, float& pdf, Sampler& st) {
    float ior_o = 1.0, ior_i = ior, sign = 1.0;
    float3 lwo = wo, lwi = f3(0.0f);
    float3 n = f3(0, 1, 0);
    if (wo.y < 0.0) { ior_o = ior; ior_i = 1.0; lwo.y = -lwo.y; sign = -1.0; }
    const float fr = schlick_ior(ior_o, ior_i, lwo, n);
    float3 f;
    float p = draw1(st);
    if (p < fr) {
        lwi = reflect(-lwo, n); pdf = 1; f = rho / fabsf(lwi.y);
    } else {
        float3 t;
        if (refract_dir(lwo, n, ior_o, ior_i, t)) {
            lwi = reflect(-t, f3(0, -1, 0));
            pdf = 1; f = rho / fabsf(lwi.y);
        } else {
            lwi = reflect(-lwo, n); pdf = 1; f = rho / fabsf(lwi.y);
        }
    }
    wi = lwi; wi.y = sign * wi.y;
    return f;
}

// ------------------------------------------------------------------ EnagyConservationGGX (BSDFs.h:483-852)
struct MultiScatterGGX {
    float3 F0; float alpha;
    HJ_DI void init(const float3& f0, float rough) { F0 = f0; alpha = clampf(rough * rough, 0.0001f, 1.0f); }   // :827-830
    HJ_DI static float C1(float h) { return fminf(1.0f, fmaxf(0.0f, 0.5f * (h + 1.0f))); }                       // :494-500
    HJ_DI static float invC1(float U) { return fmaxf(-1.0f, fminf(1.0f, 2.0f * U - 1.0f)); }                     // :502-505
    HJ_DI static float sgn(float x) { return (x > 0.0f) ? 1.0f : -1.0f; }                                        // :521-523
    HJ_DI float Lambda(const float3& v) const {                                                                   // :525-532
        if (v.y > 0.9999f) return 0.0f;
        if (v.y < -0.9999f) return -1.0f;
        float delta = 1.0f + (alpha * alpha * v.x * v.x + alpha * alpha * v.z * v.z) / (v.y * v.y);
        return (-1.0 + sgn(v.y) * sqrtf(delta)) / 2.0f;     // double arithmetic in the reference
    }
    HJ_DI float G1_height(const float3& w, float h0) const {                                                      // :551-563
        if (w.y > 0.9999f) return 1.0f;
        if (w.y <= 0.0f) return 0.0f;
        return hj_powf(C1(h0), Lambda(w));
    }
    HJ_DI float sample_height(const float3& wr, float hr, float U) const {                                        // :566-586
        if (wr.y > 0.9999f) return FLT_MAX;
        if (wr.y < -0.9999f) return invC1(U * C1(hr));
        if (fabsf(wr.y) < 0.0001f) return hr;
        const float g = G1_height(wr, hr);
        if (U > 1.0f - g) return FLT_MAX;
        return invC1(C1(hr) / hj_powf((1.0f - U), 1.0f / Lambda(wr)));
    }
    HJ_DI float3 sample_vndf(const float2& uv, const float3& wo) const {                                          // :616-632
        float3 swo = normalize(f3(wo.x * alpha, wo.y, wo.z * alpha));
        float phi = 2.0f * HJ_PI * uv.x;
        float z = fmaf((1.0f - uv.y), (1.0f + swo.y), -swo.y);
        float st = sqrtf(clampf(1.0f - z * z, 0.0f, 1.0f));
        float sp, cp; hj_sincosf(phi, sp, cp);        // one range reduction for both: same bits as cosf(phi), sinf(phi)
        float x = st * cp, y = st * sp;
        float3 h = f3(x, z, y) + swo;
        return normalize(f3(h.x * alpha, h.y, h.z * alpha));
    }
    HJ_DI float3 sample_phase(const float3& wi, Sampler& st, float3& weight) const {                              // :737-746
        const float2 uv = draw2(st);
        float3 wm = sample_vndf(uv, wi);
        const float3 wo = -wi + 2.0f * wm * dot(wi, wm);
        weight = schlick3(F0, wi, wm);
        return wo;
    }
    // :784-819. `wo` is left untouched on the NaN exit (the reference leaves it unset): callers pre-set (0,1,0).
    HJ_NI float3 walk(const float3& wi, float3& wo, int& order, Sampler& st) const {
        float3 wr = -wi;
        float hr = 1.0f + invC1(0.999f);
        order = 0;
        float3 weight = f3(1.0f);
        while (true) {
            float U = draw1(st);
            hr = sample_height(wr, hr, U);
            if (hr == FLT_MAX) break;
            else order++;
            if (order > 5) { wo = f3(0, 0, 1); return f3(0, 0, 0); }
            float3 w1;
            wr = sample_phase(-wr, st, w1);
            weight *= w1;
            if ((hr != hr) || (wr.z != wr.z)) return f3(0, 0, 1);
        }
        wo = wr;
        return weight;
    }
    // :843-851 — pdf is NOT written on the early-out
    HJ_DI float3 sample(const float3& wo, float3& wi, Sampler& st, float& pdf) const {
        int order;
        float3 f = walk(wo, wi, order, st);
        if (wi.y < 0.0 || order > 5) return f3(0.0f);
        pdf = fabsf(wi.y);
        return f;
    }
};

// ------------------------------------------------------------------ dispatcher (BSDFs.h:979-1038)
struct SurfaceBsdf {
    bool is_specular, is_ggx; float ior;
    Disney disney; MultiScatterGGX eggx;
    HJ_DI void init(const Surface& s) {                                                                           // :995-1004
        is_specular = s.is_specular; is_ggx = s.metallic > 0.5; ior = s.ior;
        disney.init(s); eggx.init(s.basecolor, s.roughness);
    }
    HJ_DI float3 eval(const DevScene& S, const float3& wo, const float3& wi) const {                              // :1006-1013
        if (is_specular) return f3(0.0f);
        return disney.eval(S, wo, wi);
    }
    HJ_DI float3 sample(const DevScene& S, const float3& wo, float3& wi, float& pdf, Sampler& st) const {         // :1015-1030
        if (is_specular) return metaglass_sample(f3(1.0f), ior, wo, wi, pdf, st);
        if (!is_ggx) return disney.sample(S, wo, wi, pdf, st);
        return eggx.sample(wo, wi, st, pdf);
    }
    HJ_DI float pdf(const float3& wo, const float3& wi) const {                                                   // :1032-1037
        if (is_specular) return 0;
        return disney.pdf(wo, wi);
    }
};

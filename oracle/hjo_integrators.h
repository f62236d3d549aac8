// hjo_integrators.h — TEST INFRASTRUCTURE (CPU oracle). Not part of the product path.
//
// CPU restatement of /root/reference/include/kernel/light_sample.h and rt.h (Pathtrace :85-159,
// NEE :162-281, MIS :284-440). MaxDepth (literal 10 in the reference, rt.h:89,166,288) is a
// run-time parameter here because BASELINE.json's configs ask for 8 and 32.
// Defined behaviour for the reference's undefined spots (SURVEY §7 "hard parts"):
//   * light_sample with no lights leaves `emission` unset (light_sample.h:12-16) -> 0 here;
//   * EC-GGX early-out leaves pt_pdf / local_wi unset in MIS (rt.h:384-387) -> 1.0 / (0,1,0) here.
#pragma once
#include "hjo_bsdf.h"
#include "hjo_trace.h"

namespace hjo {

// light_sample.h:9-75
static inline float3 sample_light(Sampler& st, float& pdf, float3& normal, float3& emission) {
    const Scene& S = *g_trace.scene;
    const int count = (int)S.light_prim_ids.size();
    if (count < 1) { pdf = -1.0; normal = make_float3(0.0f); emission = make_float3(0.0f); return make_float3(0.0f); }
    float p = draw1(st);
    int index = (int)(p * count);
    if (index == count) index--;
    if (g_hjo_events) hjo_note(HJO_EV_LIGHT_PICK, (uint32_t)index, p * count, (float)index);
    unsigned int prim = S.light_prim_ids[index];
    // binary search of the owning instance (:26-38), kept verbatim in behaviour incl. unsigned wrap
    unsigned int left = 0U, right = (unsigned int)S.prim_offsets.size() - 1, middle = (left + right) / 2U;
    while (left <= right) {
        if (S.prim_offsets[middle] <= prim) left = middle + 1;
        else { if (middle == 0) { break; } right = middle - 1; }
        middle = (left + right) / 2;
    }
    float select_pdf = 1.0f / count;
    unsigned int inst = middle;
    const Mat3x4& xf = S.xf[inst]; const Mat3x4& ixf = S.inv_xf[inst];
    const unsigned i0 = S.indices[prim * 3 + 0], i1 = S.indices[prim * 3 + 1], i2 = S.indices[prim * 3 + 2];
    const float3 v0 = xform_point(xf, S.vertices[i0]), v1 = xform_point(xf, S.vertices[i1]), v2 = xform_point(xf, S.vertices[i2]);
    const float3 n0 = xform_normal(ixf, S.normals[i0]), n1 = xform_normal(ixf, S.normals[i1]), n2 = xform_normal(ixf, S.normals[i2]);
    const float area = length(cross(v1 - v0, v2 - v0)) * 0.5f;
    float2 xi = draw2(st);
    float f1 = 1.0f - sqrtf(xi.x);
    float f2 = sqrtf(xi.x) * (1.0f - xi.y);
    float f3 = sqrtf(xi.x) * xi.y;
    const float3 pos = v0 * f1 + v1 * f2 + v2 * f3;
    const float3 nrm = normalize(n0 * f1 + n1 * f2 + n2 * f3);
    pdf = 1.0 / area;           // double division, rounded to float (light_sample.h:69)
    pdf *= select_pdf;
    normal = nrm;
    emission = S.light_prim_emission[index];
    return pos;
}

// light_sample.h:77-92
static inline float light_pdf_area(unsigned int prim, unsigned int inst) {
    const Scene& S = *g_trace.scene;
    const unsigned i0 = S.indices[prim * 3 + 0], i1 = S.indices[prim * 3 + 1], i2 = S.indices[prim * 3 + 2];
    const Mat3x4& xf = S.xf[inst];
    const float3 v0 = xform_point(xf, S.vertices[i0]), v1 = xform_point(xf, S.vertices[i1]), v2 = xform_point(xf, S.vertices[i2]);
    float area = length(cross(v1 - v0, v2 - v0)) * 0.5f;
    return 1.0f / (area * (int)S.light_prim_ids.size());
}

static inline bool is_zero(const float3& v) { return v.x == 0.0f && v.y == 0.0f && v.z == 0.0f; }

// rt.h:85-159
static inline float3 integrate_pt(float3 o, float3 d, Sampler st, float3& aov_albedo, float3& aov_normal, int max_depth) {
    float3 L = {0, 0, 0}, thr = {1, 1, 1};
    for (int depth = 0; depth < max_depth; depth++) {
        float rp = fmaxf(thr.x, fmaxf(thr.y, thr.z));
        if (hjo_lt(HJO_EV_ROULETTE, rp, draw1(st))) break;
        thr /= rp;
        g_trace.traced_extend++;
        Surface prd = trace_extend(o, d, 0.001f, 1e16f);
        if (depth == 0) { aov_albedo = prd.basecolor; aov_normal = prd.normal; }
        if (!prd.is_hit) { L += thr * prd.emission; break; }
        if (prd.is_light) { L += thr * prd.emission; break; }
        SurfaceBsdf bsdf(prd);
        float3 t, b, n = prd.normal;
        onb(n, t, b);
        float pdf = 1.0;
        float3 lwo = to_local(-d, t, n, b);
        float3 lwi = {0.0, 1.0, 0.0};
        float3 f = bsdf.sample(lwo, lwi, pdf, st);
        float3 wi = to_world(lwi, t, n, b);
        thr *= f * fabsf(dot(wi, n)) / pdf;
        o = prd.position; d = wi;
    }
    return L;
}

// rt.h:162-281
static inline float3 integrate_nee(float3 o, float3 d, Sampler st, float3& aov_albedo, float3& aov_normal, int max_depth) {
    float3 L = {0, 0, 0}, thr = {1, 1, 1};
    for (int depth = 0; depth < max_depth; depth++) {
        float rp = fmaxf(thr.x, fmaxf(thr.y, thr.z));
        if (hjo_lt(HJO_EV_ROULETTE, rp, draw1(st))) break;
        thr /= rp;
        g_trace.traced_extend++;
        Surface prd = trace_extend(o, d, 0.001f, 1e16f);
        if (depth == 0) { aov_albedo = prd.basecolor; aov_normal = prd.normal; }
        if (!prd.is_hit) { if (depth == 0) L += thr * prd.emission; break; }
        if (prd.is_light) { if (depth == 0) L += thr * prd.emission; break; }
        SurfaceBsdf bsdf(prd);
        float3 t, b, n = prd.normal;
        onb(n, t, b);
        float3 lwo = to_local(-d, t, n, b);
        {
            float lpdf; float3 lcol, lnrm;
            float3 lpos = sample_light(st, lpdf, lnrm, lcol);
            float3 so = prd.position;
            float3 sd = normalize(lpos - so);
            float dist = length(lpos - so);
            // wavefront build: contribution computed first, shadow ray only if it is non-zero
            float c1 = absdot(n, sd), c2 = absdot(lnrm, -sd);
            float3 lwi = to_local(sd, t, n, b);
            float3 f = bsdf.eval(lwo, lwi);
            float G = c2 / (dist * dist);
            float3 contrib = thr * (f * G * c1 / lpdf) * lcol;
            bool useful = !g_trace.scene->light_prim_ids.empty() && !is_zero(contrib);
            if (useful) {
                g_trace.traced_shadow++;
                if (!trace_shadow(so, sd, 0.001f, dist - 0.001f)) L += contrib;
            } else g_trace.rays_shadow++;   // the reference traces it; its result cannot change L
        }
        float pdf = 1.0;
        float3 lwi = {0.0, 1.0, 0.0};
        (void)draw2(st);            // dead draw, rt.h:266
        float3 f = bsdf.sample(lwo, lwi, pdf, st);
        float3 wi = to_world(lwi, t, n, b);
        thr *= f * fabsf(dot(wi, n)) / pdf;
        o = prd.position; d = wi;
    }
    return L;
}

// rt.h:284-440
static inline float3 integrate_mis(const float3& o0, const float3& d0, Sampler& st, float3& aov_albedo, float3& aov_normal, int max_depth) {
    float3 L = {0, 0, 0}, thr = {1, 1, 1};
    float3 o = o0, d = d0;
    const bool have_lights = !g_trace.scene->light_prim_ids.empty();
    for (int depth = 0; depth < max_depth; depth++) {
        float rp = fmaxf(thr.x, fmaxf(thr.y, thr.z));
        if (hjo_lt(HJO_EV_ROULETTE, rp, draw1(st))) break;
        thr /= rp;
        g_trace.traced_extend++;
        Surface prd = trace_extend(o, d, 0.001f, 1e16f);
        if (depth == 0) { aov_albedo = prd.basecolor; aov_normal = prd.normal; }
        if (!prd.is_hit) { if (depth == 0) L += thr * prd.emission; break; }
        if (prd.is_light) { if (depth == 0) L += thr * prd.emission; break; }
        SurfaceBsdf bsdf(prd);
        float3 t, b, n = prd.normal;
        onb(n, t, b);
        float3 lwo = to_local(-d, t, n, b);
        {   // light sampling, rt.h:341-380
            float lpdf; float3 lnrm, lem;
            float3 lpos = sample_light(st, lpdf, lnrm, lem);
            float3 ld = lpos - prd.position;
            float dist = length(ld);
            ld = normalize(ld);
            float c1 = absdot(n, ld), c2 = absdot(lnrm, -ld);
            float3 lwi = to_local(ld, t, n, b);
            float3 f = bsdf.eval(lwo, lwi);
            float G = c2 / (dist * dist);
            float ppdf = bsdf.pdf(lwo, lwi) * G;
            float w = lpdf / (lpdf + ppdf);
            float3 contrib = thr * (f * G * c1 / lpdf) * w * lem;
            bool useful = have_lights && !is_zero(contrib);
            if (useful) {
                g_trace.traced_shadow++;
                if (!trace_shadow(prd.position, ld, 0.001f, dist - 0.001f)) L += contrib;
            } else g_trace.rays_shadow++;   // the reference traces it; its result cannot change L
        }
        if (g_trace.sky_sampling) {   // EXTENSION (hj_frame_params::sky_sampling): a direction drawn from the sky's luminance table, MIS against the BSDF pdf
            const float2 xi = draw2(st);
            float spdf;
            const float3 sd = sky_sample(*g_trace.scene, xi, spdf);
            bool useful = false; float3 contrib = make_float3(0.0f);
            if (spdf > 0.0f && !prd.is_specular) {
                const float c1 = absdot(n, sd);
                const float3 lwi = to_local(sd, t, n, b);
                const float3 f = bsdf.eval(lwo, lwi);
                const float w = spdf / (spdf + bsdf.pdf(lwo, lwi));
                contrib = thr * (f * c1 / spdf) * w * sky_radiance(*g_trace.scene, sd, g_trace.ibl_intensity);
                useful = !is_zero(contrib);
            }
            if (useful) {
                g_trace.traced_shadow++;
                if (!trace_shadow(prd.position, sd, 0.001f, 1e16f)) L += contrib;
            }
        }
        {   // BSDF sampling towards emitters, rt.h:383-420
            float ppdf = 1.0f;                  // defined (reference: uninitialised)
            float3 lwi = {0.0f, 1.0f, 0.0f};    // defined (reference: uninitialised)
            float3 f = bsdf.sample(lwo, lwi, ppdf, st);
            float3 wi = to_world(lwi, t, n, b);
            float c1 = absdot(wi, n);
            bool useful = !is_zero(f * c1);
            if (!useful) g_trace.rays_extend++;   // the reference traces it; a zero BSDF weight cannot change L
            else {
            g_trace.traced_extend++;
            Surface lh = trace_extend(prd.position, wi, 0.001f, 1e16f);
            if (lh.is_hit) {
                if (lh.is_light) {
                    float c2 = absdot(-wi, lh.normal);
                    float dist = length(lh.position - prd.position);
                    float invG = dist * dist / c2;
                    float lpdf = prd.is_specular ? 0.0f : light_pdf_area(lh.primitive_id, lh.instance_id) * invG;
                    float w = ppdf / (ppdf + lpdf);
                    L += thr * w * c1 * lh.emission * f / ppdf;
                }
            } else if (g_trace.sky_sampling && !prd.is_specular) {      // extension: balance heuristic against the sky's light-sampling pdf
                const float w = ppdf / (ppdf + sky_pdf(*g_trace.scene, wi));
                L += thr * w * f * c1 * lh.emission / ppdf;
            } else {
                L += thr * f * c1 * lh.emission / ppdf;
            }
            }
        }
        float pdf = 1.0;
        float3 lwi = {0.0, 1.0, 0.0};
        (void)draw2(st);            // dead draw, rt.h:426
        float3 f = bsdf.sample(lwo, lwi, pdf, st);
        float3 wi = to_world(lwi, t, n, b);
        thr *= f * fabsf(dot(wi, n)) / pdf;
        o = prd.position; d = wi;
    }
    return L;
}

}  // namespace hjo

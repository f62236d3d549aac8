// hjo_trace.h — TEST INFRASTRUCTURE (CPU oracle). Not part of the product path.
//
// Thread-local trace context shared by both oracle builds: the stand-ins for optixTrace
// (/root/reference/include/kernel/rt.h:15-69) and for the tex2D fetch of the thin-film LUT
// (/root/reference/include/kernel/disneyBRDF.h:11-14).
#pragma once
#include "hjo_common.h"

namespace hjo {

// the scene the texture fetches read (set by the render entry points); lut_const overrides the LUT (KAT)
struct ShadeCtx { const Scene* scene = nullptr; };
static thread_local ShadeCtx g_shade;
static bool g_lut_const_on = false;
static float3 g_lut_const = {0.5f, 0.5f, 0.5f};

static inline float3 lut_lookup(float u, float v) {
    if (g_lut_const_on) return g_lut_const;
    const Scene* S = g_shade.scene;
    if (!S || !S->has_lut) return make_float3(0.0f);
    return make_float3(bilinear8(S->lut, u, v));
}

struct TraceCtx {
    const Scene* scene = nullptr;
    float ibl_intensity = 1.0f;
    bool sky_sampling = false;      // hj_frame_params::sky_sampling with the MIS integrator (extension)
    PathDebug dbg;
    uint64_t rays_extend = 0, rays_shadow = 0;       // every RayTrace / TraceOcculution of the reference
    uint64_t traced_extend = 0, traced_shadow = 0;   // rays the wavefront build actually traces (zero-weight rays skipped)
};
static thread_local TraceCtx g_trace;

// rt.h:43-69 RayTrace -> closest hit + reconstructed __closesthit__ch / __miss__ms
static inline Surface trace_extend(const float3& o, const float3& d, float tmin, float tmax) {
    Hit h = closest_hit(*g_trace.scene, o, d, tmin, tmax);
    debug_note_extend(g_trace.dbg, h);
    if (g_hjo_events) hjo_note(HJO_EV_HIT, h.hit ? h.prim : 0xffffffffu, h.hit ? fminf(fminf(h.b1, h.b2), 1.0f - h.b1 - h.b2) : 0.0f, h.hit ? h.t : 0.0f);
    g_trace.rays_extend++;
    return make_surface(*g_trace.scene, h, d, g_trace.ibl_intensity);
}
// rt.h:15-41 TraceOcculution -> __closesthit__shadow sets is_hit
static inline bool trace_shadow(const float3& o, const float3& d, float tmin, float tmax) {
    g_trace.dbg.rays_shadow++;
    g_trace.rays_shadow++;
    const bool occluded = any_hit(*g_trace.scene, o, d, tmin, tmax);
    if (g_hjo_events) hjo_note(HJO_EV_SHADOW, occluded ? 1u : 0u, tmin, tmax);
    return occluded;
}

}  // namespace hjo

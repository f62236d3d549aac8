// hjo_render.h — TEST INFRASTRUCTURE (CPU oracle). Not part of the product path.
//
// The reconstructed __raygen__rg (decisions D1, D2, D7) written once and shared by both
// oracle builds through a small policy type:
//     struct Policy { using State = ...;
//                     static State make(uint64 n_spp, uint32 scramble, uint32 image_idx);
//                     static float2 draw2(State&);
//                     static float3 integrate(int mode, float3 o, float3 d, State&, float3& alb, float3& nrm, int max_depth); };
// Camera model: stale PTX computeCameraRay, /root/reference/lib/ptx/HenjouRenderer_generated_henjouRendererCU.cu.ptx:49-104:
//     u = (2x - W)/H, v = (2y - H)/H, d = normalize(camera_f*dir + u*right + v*up), row 0 = bottom of the picture.
// Write-out: mean radiance -> optional tonemap (kernel/color.h) -> float4 AOV; RGBA8 = host float4ConvertColor
// (include/renderer/renderer.h:73-101).
#pragma once
#include "hjo_kernel.h"
#include "hjo_trace.h"
#include "hjo_api.h"
#include <chrono>
#include <omp.h>

namespace hjo {

struct Context {
    Scene scene;
    bool have_scene = false, have_accel = false;
};

static inline float3 apply_tonemap(int mode, const float3& c) {
    if (mode == HJ_TONEMAP_ACES) return tonemap_aces(c);
    if (mode == HJ_TONEMAP_UCHIMURA) return tonemap_uchimura(c);
    return c;
}

template <class Policy>
static int render_rows(Context& C, const hj_frame_params& fp, int n_threads, uint32_t y0, uint32_t y1, hjo_render_out& out) {
    if (!C.have_scene || !C.have_accel) return HJ_ERR_NO_ACCEL;
    const uint32_t W = fp.width, H = fp.height;
    if (W == 0 || H == 0 || fp.sample_end < fp.sample_begin || fp.spp == 0) return HJ_ERR_INVALID;
    if (y1 == 0 || y1 > H) y1 = H;
    const uint32_t ns = fp.sample_end - fp.sample_begin;
    const float3 cpos = make_float3(fp.camera_pos[0], fp.camera_pos[1], fp.camera_pos[2]);
    const float3 cdir = make_float3(fp.camera_dir[0], fp.camera_dir[1], fp.camera_dir[2]);
    const float3 cup = make_float3(fp.camera_up[0], fp.camera_up[1], fp.camera_up[2]);
    const float3 cright = make_float3(fp.camera_right[0], fp.camera_right[1], fp.camera_right[2]);
    uint64_t re = 0, rs = 0, te = 0, ts = 0;
    if (n_threads < 1) n_threads = omp_get_max_threads();
    auto t_begin = std::chrono::steady_clock::now();
#pragma omp parallel num_threads(n_threads) reduction(+ : re, rs, te, ts)
    {
        g_trace = TraceCtx();
        g_trace.scene = &C.scene;
        g_trace.ibl_intensity = fp.ibl_intensity;
        g_trace.sky_sampling = fp.sky_sampling != 0 && fp.integrator == HJ_INTEGRATOR_MIS;
        g_shade.scene = &C.scene;
#pragma omp for schedule(dynamic, 1)
        for (uint32_t y = y0; y < y1; y++) {
            for (uint32_t x = 0; x < W; x++) {
                const uint32_t pix = y * W + x;
                float3 sum = make_float3(0.0f);
                float3 alb0 = make_float3(0.0f), nrm0 = make_float3(0.0f);
                for (uint32_t s = fp.sample_begin; s < fp.sample_end; s++) {
                    typename Policy::State st = Policy::make((uint64_t)s, fp.frame, pix);   // D2
                    float jx = 0.0f, jy = 0.0f;
                    if (fp.pixel_jitter) { float2 xi = Policy::draw2(st); jx = xi.x; jy = xi.y; }   // D1
                    const float u = (2.0f * ((float)x + jx) - (float)W) / (float)H;
                    const float v = (2.0f * ((float)y + jy) - (float)H) / (float)H;
                    const float3 d = normalize(fp.camera_f * cdir + u * cright + v * cup);
                    g_trace.dbg = PathDebug();
                    float3 alb = make_float3(0.0f), nrm = make_float3(0.0f);
                    const float3 L = Policy::integrate(fp.integrator, cpos, d, st, alb, nrm, fp.max_depth);
                    sum += L;
                    if (s == 0) { alb0 = alb; nrm0 = nrm; }
                    const size_t pi = (size_t)(s - fp.sample_begin) * W * H + pix;
                    if (out.path_radiance) {
                        out.path_radiance[4 * pi + 0] = L.x; out.path_radiance[4 * pi + 1] = L.y; out.path_radiance[4 * pi + 2] = L.z;
                        out.path_radiance[4 * pi + 3] = (float)(g_trace.dbg.rays_extend + g_trace.dbg.rays_shadow);
                    }
                    if (out.path_ids) { out.path_ids[2 * pi] = g_trace.dbg.first_prim; out.path_ids[2 * pi + 1] = g_trace.dbg.hash; }
                }
                if (out.accum) { out.accum[4 * pix] = sum.x; out.accum[4 * pix + 1] = sum.y; out.accum[4 * pix + 2] = sum.z; out.accum[4 * pix + 3] = (float)ns; }
                const float inv = 1.0f / (float)fp.spp;
                const float3 col = apply_tonemap(fp.tonemap, sum * inv);
                if (out.color) { out.color[4 * pix] = col.x; out.color[4 * pix + 1] = col.y; out.color[4 * pix + 2] = col.z; out.color[4 * pix + 3] = 1.0f; }
                if (fp.sample_begin == 0) {
                    if (out.albedo) { out.albedo[4 * pix] = alb0.x; out.albedo[4 * pix + 1] = alb0.y; out.albedo[4 * pix + 2] = alb0.z; out.albedo[4 * pix + 3] = 1.0f; }
                    if (out.normal) { out.normal[4 * pix] = nrm0.x; out.normal[4 * pix + 1] = nrm0.y; out.normal[4 * pix + 2] = nrm0.z; out.normal[4 * pix + 3] = 1.0f; }
                }
                if (out.rgba8) {
                    const float3 sc = to_srgb(col);
                    out.rgba8[4 * pix] = quantize8(sc.x); out.rgba8[4 * pix + 1] = quantize8(sc.y); out.rgba8[4 * pix + 2] = quantize8(sc.z); out.rgba8[4 * pix + 3] = 255;
                }
            }
        }
        re += g_trace.rays_extend; rs += g_trace.rays_shadow; te += g_trace.traced_extend; ts += g_trace.traced_shadow;
    }
    out.seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
    out.rays_extend = re; out.rays_shadow = rs; out.traced_extend = te; out.traced_shadow = ts;
    return HJ_OK;
}

// One path (pixel, global sample index) with the decision trace switched on: the events of hjo_vec.h in order, and the
// path's radiance. Returns the number of events the path produced (may exceed cap; only the first cap are stored).
template <class Policy>
static int path_events(Context& C, const hj_frame_params& fp, uint32_t pix, uint32_t s, hjo_event* ev, uint32_t cap, float* radiance3) {
    if (!C.have_scene || !C.have_accel) return HJ_ERR_NO_ACCEL;
    const uint32_t W = fp.width, H = fp.height;
    if (W == 0 || H == 0 || pix >= W * H) return HJ_ERR_INVALID;
    const float3 cpos = make_float3(fp.camera_pos[0], fp.camera_pos[1], fp.camera_pos[2]);
    const float3 cdir = make_float3(fp.camera_dir[0], fp.camera_dir[1], fp.camera_dir[2]);
    const float3 cup = make_float3(fp.camera_up[0], fp.camera_up[1], fp.camera_up[2]);
    const float3 cright = make_float3(fp.camera_right[0], fp.camera_right[1], fp.camera_right[2]);
    g_trace = TraceCtx(); g_trace.scene = &C.scene; g_trace.ibl_intensity = fp.ibl_intensity; g_shade.scene = &C.scene;
    g_trace.sky_sampling = fp.sky_sampling != 0 && fp.integrator == HJ_INTEGRATOR_MIS;
    const uint32_t x = pix % W, y = pix / W;
    typename Policy::State st = Policy::make((uint64_t)s, fp.frame, pix);
    float jx = 0.0f, jy = 0.0f;
    if (fp.pixel_jitter) { float2 xi = Policy::draw2(st); jx = xi.x; jy = xi.y; }
    const float u = (2.0f * ((float)x + jx) - (float)W) / (float)H;
    const float v = (2.0f * ((float)y + jy) - (float)H) / (float)H;
    const float3 d = normalize(fp.camera_f * cdir + u * cright + v * cup);
    HjoEventLog log{ev, cap, 0u};
    g_hjo_events = &log;
    float3 alb = make_float3(0.0f), nrm = make_float3(0.0f);
    const float3 L = Policy::integrate(fp.integrator, cpos, d, st, alb, nrm, fp.max_depth);
    g_hjo_events = nullptr;
    if (radiance3) { radiance3[0] = L.x; radiance3[1] = L.y; radiance3[2] = L.z; }
    return (int)log.n;
}

static inline int trace_batch(Context& C, const float* rays, uint64_t n, int anyhit, uint32_t* prim, float* tuv) {
    if (!C.have_accel) return HJ_ERR_NO_ACCEL;
#pragma omp parallel for schedule(dynamic, 256)
    for (long long i = 0; i < (long long)n; i++) {
        const float* r = rays + 8 * i;
        const float3 o = make_float3(r[0], r[1], r[2]), d = make_float3(r[4], r[5], r[6]);
        if (anyhit) {
            bool h = any_hit(C.scene, o, d, r[3], r[7]);
            if (prim) prim[i] = h ? 1u : 0xffffffffu;
        } else {
            Hit h = closest_hit(C.scene, o, d, r[3], r[7]);
            if (prim) prim[i] = h.hit ? h.prim : 0xffffffffu;
            if (tuv) { tuv[3 * i] = h.hit ? h.t : 0.0f; tuv[3 * i + 1] = h.hit ? h.b1 : 0.0f; tuv[3 * i + 2] = h.hit ? h.b2 : 0.0f; }
        }
    }
    return HJ_OK;
}

}  // namespace hjo

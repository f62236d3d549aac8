// hjo_api.cpp — TEST INFRASTRUCTURE (CPU oracle). Not part of the product path.
//
// Build "restatement": the CPU oracle proper. Every arithmetic routine is the hjo_* restatement of the
// reference's device headers; nothing from /root/reference is read at build or run time, so this library
// travels to the GPU box. Pinned against the SURVEY §8c KAT block and against oracle/_ref (tests/test_oracle_*.py).
#include "hjo_integrators.h"
#include "hjo_render.h"
#include "hjo_api.h"

using namespace hjo;

struct hjo_ctx { Context c; };

namespace {
struct RestatedPolicy {
    using State = Sampler;
    static State make(uint64_t n_spp, uint32_t scramble, uint32_t image_idx) {
        Sampler s; s.n_spp = n_spp; s.scramble = scramble; s.depth = 0; s.image_idx = image_idx; return s;
    }
    static float2 draw2(State& s) { return hjo::draw2(s); }
    static float3 integrate(int mode, const float3& o, const float3& d, State& s, float3& alb, float3& nrm, int max_depth) {
        if (mode == HJ_INTEGRATOR_PT) return integrate_pt(o, d, s, alb, nrm, max_depth);
        if (mode == HJ_INTEGRATOR_NEE) return integrate_nee(o, d, s, alb, nrm, max_depth);
        return integrate_mis(o, d, s, alb, nrm, max_depth);
    }
};
Surface kat_surface(const hjo_kat_surface* k) {
    Surface s; s.is_hit = true;
    s.basecolor = make_float3(k->basecolor[0], k->basecolor[1], k->basecolor[2]);
    s.metallic = k->metallic; s.roughness = k->roughness; s.sheen = k->sheen; s.clearcoat = k->clearcoat; s.ior = k->ior;
    s.is_specular = k->is_specular != 0; s.is_thinfilm = k->is_thinfilm != 0;
    return s;
}
}  // namespace

extern "C" {

int hjo_build_kind(void) { return 0; }
hjo_ctx* hjo_create(void) { return new hjo_ctx(); }
void hjo_destroy(hjo_ctx* c) { delete c; }

int hjo_scene_upload(hjo_ctx* c, const hj_scene_desc* d) {
    if (!c || !load_scene(c->c.scene, d)) return HJ_ERR_INVALID;
    c->c.have_scene = true; c->c.have_accel = false;
    return HJ_OK;
}
int hjo_accel_build(hjo_ctx* c, const float* xf, const float* ixf, uint32_t n, int use_bvh) {
    if (!c || !c->c.have_scene) return HJ_ERR_NO_SCENE;
    if (n != c->c.scene.prim_offsets.size()) return HJ_ERR_INVALID;
    c->c.scene.use_bvh = use_bvh != 0;
    set_transforms(c->c.scene, xf, ixf);
    c->c.have_accel = true;
    return HJ_OK;
}
int hjo_render(hjo_ctx* c, const hj_frame_params* fp, int n_threads, uint32_t y0, uint32_t y1, hjo_render_out* out) {
    if (!c || !fp || !out) return HJ_ERR_INVALID;
    return render_rows<RestatedPolicy>(c->c, *fp, n_threads, y0, y1, *out);
}
int hjo_trace(hjo_ctx* c, const float* rays, uint64_t n, int anyhit, uint32_t* prim, float* tuv) {
    if (!c) return HJ_ERR_INVALID;
    return trace_batch(c->c, rays, n, anyhit, prim, tuv);
}

int hjo_path_events(hjo_ctx* c, const hj_frame_params* fp, uint32_t pixel, uint32_t sample, void* events, uint32_t cap, float* radiance3) {
    if (!c || !fp) return HJ_ERR_INVALID;
    return path_events<RestatedPolicy>(c->c, *fp, pixel, sample, (hjo_event*)events, cap, radiance3);
}
int hjo_kat_libm(int fn, const float* a, const float* b, float* out, uint64_t n) {
    for (uint64_t i = 0; i < n; i++) {
        switch (fn) {
            case 0: out[i] = hjo_sinf(a[i]); break;
            case 1: out[i] = hjo_cosf(a[i]); break;
            case 2: out[i] = hjo_acosf(a[i]); break;
            case 3: out[i] = hjo_logf(a[i]); break;
            case 4: out[i] = hjo_expf(a[i]); break;
            case 5: out[i] = hjo_powf(a[i], b[i]); break;
            case 6: out[i] = hjo_atan2f(a[i], b[i]); break;
            default: return HJ_ERR_INVALID;
        }
    }
    return HJ_OK;
}
#ifdef HJO_GLIBC_LIBM
int hjo_libm_kind(void) { return 1; }
#else
int hjo_libm_kind(void) { return 0; }
#endif

uint32_t hjo_kat_xxhash32(uint32_t p) { return xxhash32_u(p); }
uint32_t hjo_kat_xxhash32_u4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { return xxhash32_u4(x, y, z, w); }
uint32_t hjo_kat_permute(uint32_t i, uint32_t l, uint32_t p) { return permute(i, l, p); }
float hjo_kat_randfloat(uint32_t i, uint32_t p) { return randfloat(i, p); }
uint32_t hjo_kat_cmj2d(uint64_t n_spp, uint32_t image_idx, uint32_t scramble, uint32_t depth0, int n, float* out) {
    Sampler s = RestatedPolicy::make(n_spp, scramble, image_idx); s.depth = depth0;
    for (int i = 0; i < n; i++) { float2 r = draw2(s); out[2 * i] = r.x; out[2 * i + 1] = r.y; }
    return s.depth;
}
void hjo_kat_onb(const float* n, float* t, float* b) {
    float3 tt, bb; onb(make_float3(n[0], n[1], n[2]), tt, bb);
    t[0] = tt.x; t[1] = tt.y; t[2] = tt.z; b[0] = bb.x; b[1] = bb.y; b[2] = bb.z;
}
int hjo_kat_refract(const float* v, const float* n, float ior1, float ior2, float* r) {
    float3 rr = make_float3(0.0f);
    bool ok = refract_dir(make_float3(v[0], v[1], v[2]), make_float3(n[0], n[1], n[2]), ior1, ior2, rr);
    r[0] = rr.x; r[1] = rr.y; r[2] = rr.z; return ok ? 1 : 0;
}
void hjo_kat_cosine(float u, float v, float* dir, float* pdf) {
    float3 d = cosine_sample(u, v, *pdf); dir[0] = d.x; dir[1] = d.y; dir[2] = d.z;
}
void hjo_kat_tonemap(int mode, const float* in3, float* out3) {
    float3 c = apply_tonemap(mode, make_float3(in3[0], in3[1], in3[2])); out3[0] = c.x; out3[1] = c.y; out3[2] = c.z;
}
void hjo_kat_to_srgb8(const float* in3, uint8_t* out3) {
    float3 c = to_srgb(make_float3(in3[0], in3[1], in3[2])); out3[0] = quantize8(c.x); out3[1] = quantize8(c.y); out3[2] = quantize8(c.z);
}
void hjo_kat_set_lut_const(int enabled, float r, float g, float b) { g_lut_const_on = enabled != 0; g_lut_const = make_float3(r, g, b); }

void hjo_kat_bsdf_sample(const hjo_kat_surface* k, const float* wo, uint64_t n_spp, uint32_t image_idx, uint32_t scramble, float* out8) {
    Surface s = kat_surface(k);
    SurfaceBsdf bsdf(s);
    Sampler st = RestatedPolicy::make(n_spp, scramble, image_idx);
    float3 wi = make_float3(0.0f, 1.0f, 0.0f); float pdf = 1.0f;
    float3 f = bsdf.sample(make_float3(wo[0], wo[1], wo[2]), wi, pdf, st);
    out8[0] = wi.x; out8[1] = wi.y; out8[2] = wi.z; out8[3] = f.x; out8[4] = f.y; out8[5] = f.z; out8[6] = pdf; out8[7] = (float)st.depth;
}
void hjo_kat_bsdf_eval(const hjo_kat_surface* k, const float* wo, const float* wi, float* out4) {
    Surface s = kat_surface(k);
    SurfaceBsdf bsdf(s);
    float3 o = make_float3(wo[0], wo[1], wo[2]), i = make_float3(wi[0], wi[1], wi[2]);
    float3 f = bsdf.eval(o, i);
    out4[0] = f.x; out4[1] = f.y; out4[2] = f.z; out4[3] = bsdf.pdf(o, i);
}
void hjo_kat_ideal_glass(float ior, const float* wo, uint64_t n_spp, uint32_t image_idx, uint32_t scramble, float* out7) {
    Sampler st = RestatedPolicy::make(n_spp, scramble, image_idx);
    float3 wi; float pdf;
    float3 f = glass_sample(make_float3(1.0f), ior, false, make_float3(wo[0], wo[1], wo[2]), wi, pdf, st);
    out7[0] = wi.x; out7[1] = wi.y; out7[2] = wi.z; out7[3] = f.x; out7[4] = f.y; out7[5] = f.z; out7[6] = pdf;
}
int hjo_kat_light_sample(hjo_ctx* c, uint64_t n_spp, uint32_t image_idx, uint32_t scramble, float* out11) {
    if (!c || !c->c.have_accel) return HJ_ERR_NO_ACCEL;
    g_trace.scene = &c->c.scene;
    Sampler st = RestatedPolicy::make(n_spp, scramble, image_idx);
    float pdf; float3 nrm, em;
    float3 p = sample_light(st, pdf, nrm, em);
    out11[0] = p.x; out11[1] = p.y; out11[2] = p.z; out11[3] = nrm.x; out11[4] = nrm.y; out11[5] = nrm.z;
    out11[6] = em.x; out11[7] = em.y; out11[8] = em.z; out11[9] = pdf; out11[10] = (float)st.depth;
    return HJ_OK;
}
float hjo_kat_light_pdf(hjo_ctx* c, uint32_t prim, uint32_t inst) {
    if (!c || !c->c.have_accel) return -1.0f;
    g_trace.scene = &c->c.scene;
    return light_pdf_area(prim, inst);
}

}  // extern "C"

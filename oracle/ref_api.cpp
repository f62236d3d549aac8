// ref_api.cpp — TEST INFRASTRUCTURE (CPU oracle, "reference" build -> oracle/_ref/libhjref.so).
// Not part of the product path.
//
// Compiles the reference's OWN device headers (/root/reference/include/kernel/*.h, unmodified, where they
// lie) on the host behind oracle/ref_shim and exposes them through the same C entry points as the
// restatement (hjo_api.h). Purpose: pin the restatement — tests/test_oracle_ref.py demands identical
// integers and float agreement between the two builds on KATs and on whole rendered frames.
// What is NOT reference code here (the snapshot has no source for it): the raygen loop
// (hjo_render.h), the closest-hit/miss payload fill and the traversal (hjo_common.h). MaxDepth is the
// literal 10 inside the reference's integrators (rt.h:89,166,288); max_depth != 10 is rejected.
#include "ref_shim/ref_shim_base.h"
#include <kernel/rt.h>
#include <kernel/color.h>
#include "hjo_render.h"
#include "hjo_api.h"

struct hjo_ctx { hjo::Context c; };

// ---- the OptiX device intrinsics the headers call --------------------------------------------------
static thread_local unsigned int tl_p0, tl_p1;
unsigned int optixGetPayload_0() { return tl_p0; }
unsigned int optixGetPayload_1() { return tl_p1; }

void optixTrace(OptixTraversableHandle, float3 o, float3 d, float tmin, float tmax, float, OptixVisibilityMask, unsigned int,
                unsigned int sbt_offset, unsigned int, unsigned int, unsigned int& p0, unsigned int& p1) {
    Payload* prd = reinterpret_cast<Payload*>(unpack_ptr(p0, p1));
    if (sbt_offset == 1) {   // __closesthit__shadow: prd.is_hit = true (stale PTX :1300-1332)
        prd->is_hit = hjo::trace_shadow(o, d, tmin, tmax);
        return;
    }
    hjo::Surface s = hjo::trace_extend(o, d, tmin, tmax);
    prd->is_hit = s.is_hit;
    prd->ray_origin = o; prd->ray_direction = d;
    prd->position = s.position; prd->normal = s.normal; prd->texcoord = s.texcoord; prd->material_id = s.material_id;
    prd->basecolor = s.basecolor; prd->metallic = s.metallic; prd->roughness = s.roughness; prd->sheen = s.sheen;
    prd->clearcoat = s.clearcoat; prd->ior = s.ior; prd->transmission = s.transmission; prd->is_specular = s.is_specular;
    prd->emission = s.emission; prd->is_light = s.is_light; prd->is_thinfilm = s.is_thinfilm;
    prd->primitive_id = s.primitive_id; prd->instance_id = s.instance_id;
}

namespace {
void bind_params(hjo::Scene& S) {
    static_assert(sizeof(Matrix4x3) == sizeof(hjo::Mat3x4), "matrix layout");
    params.traversal_handle = 1;
    params.vertices = S.vertices.data(); params.indices = S.indices.data(); params.normals = S.normals.data();
    params.texcoords = S.texcoords.data(); params.colors = nullptr; params.prim_offsets = S.prim_offsets.data();
    params.transforms = reinterpret_cast<Matrix4x3*>(S.xf.data()); params.inv_transforms = reinterpret_cast<Matrix4x3*>(S.inv_xf.data());
    params.instance_count = (unsigned int)S.prim_offsets.size();
    params.light_prim_ids = S.light_prim_ids.data(); params.light_prim_count = (unsigned int)S.light_prim_ids.size();
    params.light_prim_emission = S.light_prim_emission.data();
    params.RAYTYPE = 2; params.lut_texture = 1;
}
struct RefPolicy {
    using State = CMJState;
    static State make(uint64_t n_spp, uint32_t scramble, uint32_t image_idx) {
        CMJState s; s.n_spp = n_spp; s.scramble = scramble; s.depth = 0; s.image_idx = image_idx; return s;
    }
    static float2 draw2(State& s) { return cmj_2d(s); }
    static float3 integrate(int mode, const float3& o, const float3& d, State& s, float3& alb, float3& nrm, int) {
        if (mode == HJ_INTEGRATOR_PT) return Pathtrace(o, d, s, alb, nrm);
        if (mode == HJ_INTEGRATOR_NEE) return NEE(o, d, s, alb, nrm);
        return MIS(o, d, s, alb, nrm);
    }
};
Payload kat_payload(const hjo_kat_surface* k) {
    Payload p; p.is_hit = true;
    p.basecolor = make_float3(k->basecolor[0], k->basecolor[1], k->basecolor[2]);
    p.metallic = k->metallic; p.roughness = k->roughness; p.sheen = k->sheen; p.clearcoat = k->clearcoat; p.ior = k->ior;
    p.is_specular = k->is_specular != 0; p.is_thinfilm = k->is_thinfilm != 0;
    return p;
}
}  // namespace

extern "C" {

int hjo_build_kind(void) { return 1; }
hjo_ctx* hjo_create(void) { return new hjo_ctx(); }
void hjo_destroy(hjo_ctx* c) { delete c; }

int hjo_scene_upload(hjo_ctx* c, const hj_scene_desc* d) {
    if (!c || !hjo::load_scene(c->c.scene, d)) return HJ_ERR_INVALID;
    c->c.have_scene = true; c->c.have_accel = false;
    return HJ_OK;
}
int hjo_accel_build(hjo_ctx* c, const float* xf, const float* ixf, uint32_t n, int use_bvh) {
    if (!c || !c->c.have_scene) return HJ_ERR_NO_SCENE;
    if (n != c->c.scene.prim_offsets.size()) return HJ_ERR_INVALID;
    c->c.scene.use_bvh = use_bvh != 0;
    hjo::set_transforms(c->c.scene, xf, ixf);
    c->c.have_accel = true;
    return HJ_OK;
}
int hjo_render(hjo_ctx* c, const hj_frame_params* fp, int n_threads, uint32_t y0, uint32_t y1, hjo_render_out* out) {
    if (!c || !fp || !out) return HJ_ERR_INVALID;
    if (fp->max_depth != 10 || fp->sky_sampling) return HJ_ERR_INVALID;    // the reference's integrators: depth 10, no sky sampling
    bind_params(c->c.scene);
    params.image_width = fp->width; params.image_height = fp->height; params.spp = fp->spp; params.frame = fp->frame;
    params.ibl_intensity = fp->ibl_intensity;
    return hjo::render_rows<RefPolicy>(c->c, *fp, n_threads, y0, y1, *out);
}
int hjo_trace(hjo_ctx* c, const float* rays, uint64_t n, int anyhit, uint32_t* prim, float* tuv) {
    if (!c) return HJ_ERR_INVALID;
    return hjo::trace_batch(c->c, rays, n, anyhit, prim, tuv);
}

int hjo_path_events(hjo_ctx* c, const hj_frame_params* fp, uint32_t pixel, uint32_t sample, void* events, uint32_t cap, float* radiance3) {
    if (!c || !fp || fp->max_depth != 10) return HJ_ERR_INVALID;
    bind_params(c->c.scene);
    params.image_width = fp->width; params.image_height = fp->height; params.spp = fp->spp; params.frame = fp->frame;
    params.ibl_intensity = fp->ibl_intensity;
    return hjo::path_events<RefPolicy>(c->c, *fp, pixel, sample, (hjo_event*)events, cap, radiance3);
}
int hjo_kat_libm(int fn, const float* a, const float* b, float* out, uint64_t n) {      // glibc, as the reference's headers call it
    for (uint64_t i = 0; i < n; i++) {
        switch (fn) {
            case 0: out[i] = sinf(a[i]); break;
            case 1: out[i] = cosf(a[i]); break;
            case 2: out[i] = acosf(a[i]); break;
            case 3: out[i] = logf(a[i]); break;
            case 4: out[i] = expf(a[i]); break;
            case 5: out[i] = powf(a[i], b[i]); break;
            case 6: out[i] = atan2f(a[i], b[i]); break;
            default: return HJ_ERR_INVALID;
        }
    }
    return HJ_OK;
}
int hjo_libm_kind(void) { return 1; }

uint32_t hjo_kat_xxhash32(uint32_t p) { return xxhash32(p); }
uint32_t hjo_kat_xxhash32_u4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { return xxhash32(make_uint4(x, y, z, w)); }
uint32_t hjo_kat_permute(uint32_t i, uint32_t l, uint32_t p) { return cmj_permute(i, l, p); }
float hjo_kat_randfloat(uint32_t i, uint32_t p) { return cmj_randfloat(i, p); }
uint32_t hjo_kat_cmj2d(uint64_t n_spp, uint32_t image_idx, uint32_t scramble, uint32_t depth0, int n, float* out) {
    CMJState s = RefPolicy::make(n_spp, scramble, image_idx); s.depth = depth0;
    for (int i = 0; i < n; i++) { float2 r = cmj_2d(s); out[2 * i] = r.x; out[2 * i + 1] = r.y; }
    return s.depth;
}
void hjo_kat_onb(const float* n, float* t, float* b) {
    float3 tt, bb; orthonormal_basis(make_float3(n[0], n[1], n[2]), tt, bb);
    t[0] = tt.x; t[1] = tt.y; t[2] = tt.z; b[0] = bb.x; b[1] = bb.y; b[2] = bb.z;
}
int hjo_kat_refract(const float* v, const float* n, float ior1, float ior2, float* r) {
    float3 rr = make_float3(0.0f);
    bool ok = refract(make_float3(v[0], v[1], v[2]), make_float3(n[0], n[1], n[2]), ior1, ior2, rr);
    r[0] = rr.x; r[1] = rr.y; r[2] = rr.z; return ok ? 1 : 0;
}
void hjo_kat_cosine(float u, float v, float* dir, float* pdf) {
    float3 d = cosineSampling(u, v, *pdf); dir[0] = d.x; dir[1] = d.y; dir[2] = d.z;
}
void hjo_kat_tonemap(int mode, const float* in3, float* out3) {
    float3 c = make_float3(in3[0], in3[1], in3[2]);
    if (mode == HJ_TONEMAP_ACES) c = ACESFilm(c);
    else if (mode == HJ_TONEMAP_UCHIMURA) c = Tonemap_Uchimura(c);
    out3[0] = c.x; out3[1] = c.y; out3[2] = c.z;
}
void hjo_kat_to_srgb8(const float* in3, uint8_t* out3) {
    // host code of renderer.h:73-85 needs sutil/glm/OptiX to compile; the restated copy is used in both builds
    float3 c = hjo::to_srgb(make_float3(in3[0], in3[1], in3[2]));
    out3[0] = hjo::quantize8(c.x); out3[1] = hjo::quantize8(c.y); out3[2] = hjo::quantize8(c.z);
}
void hjo_kat_set_lut_const(int enabled, float r, float g, float b) { hjo::g_lut_const_on = enabled != 0; hjo::g_lut_const = make_float3(r, g, b); }

void hjo_kat_bsdf_sample(const hjo_kat_surface* k, const float* wo, uint64_t n_spp, uint32_t image_idx, uint32_t scramble, float* out8) {
    Payload p = kat_payload(k);
    BSDF bsdf(p);
    CMJState st = RefPolicy::make(n_spp, scramble, image_idx);
    float3 wi = make_float3(0.0f, 1.0f, 0.0f); float pdf = 1.0f;
    float3 f = bsdf.sampleBSDF(make_float3(wo[0], wo[1], wo[2]), wi, pdf, st);
    out8[0] = wi.x; out8[1] = wi.y; out8[2] = wi.z; out8[3] = f.x; out8[4] = f.y; out8[5] = f.z; out8[6] = pdf; out8[7] = (float)st.depth;
}
void hjo_kat_bsdf_eval(const hjo_kat_surface* k, const float* wo, const float* wi, float* out4) {
    Payload p = kat_payload(k);
    BSDF bsdf(p);
    float3 o = make_float3(wo[0], wo[1], wo[2]), i = make_float3(wi[0], wi[1], wi[2]);
    float3 f = bsdf.evaluateBSDF(o, i);
    out4[0] = f.x; out4[1] = f.y; out4[2] = f.z; out4[3] = bsdf.getPDF(o, i);
}
void hjo_kat_ideal_glass(float ior, const float* wo, uint64_t n_spp, uint32_t image_idx, uint32_t scramble, float* out7) {
    CMJState st = RefPolicy::make(n_spp, scramble, image_idx);
    IdealGlass g(make_float3(1.0f), ior);
    float3 wi; float pdf;
    float3 f = g.sampleBSDF(make_float3(wo[0], wo[1], wo[2]), wi, pdf, st);
    out7[0] = wi.x; out7[1] = wi.y; out7[2] = wi.z; out7[3] = f.x; out7[4] = f.y; out7[5] = f.z; out7[6] = pdf;
}
int hjo_kat_light_sample(hjo_ctx* c, uint64_t n_spp, uint32_t image_idx, uint32_t scramble, float* out11) {
    if (!c || !c->c.have_accel) return HJ_ERR_NO_ACCEL;
    bind_params(c->c.scene);
    CMJState st = RefPolicy::make(n_spp, scramble, image_idx);
    float pdf; float3 nrm, em = make_float3(0.0f);
    float3 p = light_sample(st, pdf, nrm, em);
    out11[0] = p.x; out11[1] = p.y; out11[2] = p.z; out11[3] = nrm.x; out11[4] = nrm.y; out11[5] = nrm.z;
    out11[6] = em.x; out11[7] = em.y; out11[8] = em.z; out11[9] = pdf; out11[10] = (float)st.depth;
    return HJ_OK;
}
float hjo_kat_light_pdf(hjo_ctx* c, uint32_t prim, uint32_t inst) {
    if (!c || !c->c.have_accel) return -1.0f;
    bind_params(c->c.scene);
    return getLightPDF(prim, inst);
}

}  // extern "C"

/* hjo_api.h — TEST INFRASTRUCTURE (CPU oracle). Not part of the product path.
 *
 * C entry points shared by the two oracle builds:
 *   oracle/libhjo.so        — the CPU restatement of the reference's device headers (hjo_*.h);
 *   oracle/_ref/libhjref.so — the reference's OWN headers (/root/reference/include/kernel/ *.h),
 *                             compiled unmodified behind oracle/ref_shim (only where /root/reference exists).
 * Both consume the product ABI's plain structs (include/henjou_b200.h) so tests feed the oracle and the
 * CUDA path the same bytes. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load these libraries.
 */
#ifndef HJO_API_H
#define HJO_API_H
#include "../include/henjou_b200.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct hjo_ctx hjo_ctx;

/* 0 = restatement, 1 = reference headers */
int hjo_build_kind(void);

hjo_ctx* hjo_create(void);
void hjo_destroy(hjo_ctx*);
int hjo_scene_upload(hjo_ctx*, const hj_scene_desc*);
/* use_bvh 0: brute force over all triangles (closest-hit definition), 1: CPU BVH with the same triangle test */
int hjo_accel_build(hjo_ctx*, const float* xforms3x4, const float* inv_xforms3x4, uint32_t n_instances, int use_bvh);

typedef struct hjo_render_out {
    float* color;          /* float4[W*H] or NULL */
    float* albedo;         /* float4[W*H] or NULL */
    float* normal;         /* float4[W*H] or NULL */
    uint8_t* rgba8;        /* uchar4[W*H] or NULL */
    float* accum;          /* float4[W*H] raw sum over the sample slice, or NULL */
    float* path_radiance;  /* float4[(sample_end-sample_begin)*W*H] or NULL: rgb, w = rays of the path as the wavefront build traces them */
    uint32_t* path_ids;    /* uint32[2*(sample_end-sample_begin)*W*H] or NULL: first-hit prim | ~0, hash of closest-hit ids */
    uint64_t rays_extend, rays_shadow;       /* every RayTrace / TraceOcculution call of the reference's integrators */
    uint64_t traced_extend, traced_shadow;   /* rays the wavefront build traces (zero-weight queries skipped) */
    double seconds;                          /* wall time of the pixel loop only */
} hjo_render_out;

/* Renders pixel rows [y0,y1) (y1=0: all rows); untouched rows of the outputs are left as they were. */
int hjo_render(hjo_ctx*, const hj_frame_params*, int n_threads, uint32_t y0, uint32_t y1, hjo_render_out* out);

/* rays: 8 floats each (ox,oy,oz,tmin, dx,dy,dz,tmax). prim: ~0 on miss. tuv: 3 floats per ray (t,b1,b2). */
int hjo_trace(hjo_ctx*, const float* rays, uint64_t n, int any_hit, uint32_t* prim, float* tuv);

/* Decision trace of one path (pixel index, global sample index): events {kind, outcome, a, b} (4 x 32 bit each; kinds in
 * hjo_vec.h) in path order, at most cap stored; returns the number produced or a negative hj_status. The reference-headers
 * build records closest-hit / shadow events only (its BSDF code is the reference's, unmodified). */
int hjo_path_events(hjo_ctx*, const hj_frame_params*, uint32_t pixel, uint32_t sample, void* events, uint32_t cap, float* radiance3);
/* elementary functions as this build evaluates them: fn 0 sinf, 1 cosf, 2 acosf, 3 logf, 4 expf, 5 powf(a,b), 6 atan2f(a,b) */
int hjo_kat_libm(int fn, const float* a, const float* b, float* out, uint64_t n);
/* 0: the shared portable implementations of hj_libm.h (bit-identical to the CUDA kernels), 1: glibc */
int hjo_libm_kind(void);

/* ---- known-answer entry points (SURVEY §8c KAT block) ---- */
uint32_t hjo_kat_xxhash32(uint32_t p);
uint32_t hjo_kat_xxhash32_u4(uint32_t x, uint32_t y, uint32_t z, uint32_t w);
uint32_t hjo_kat_permute(uint32_t i, uint32_t l, uint32_t p);
float hjo_kat_randfloat(uint32_t i, uint32_t p);
/* n successive cmj_2d calls on CMJState{n_spp,scramble,depth0,image_idx}; out = 2n floats; returns final depth */
uint32_t hjo_kat_cmj2d(uint64_t n_spp, uint32_t image_idx, uint32_t scramble, uint32_t depth0, int n, float* out);
void hjo_kat_onb(const float* n, float* t, float* b);
int hjo_kat_refract(const float* v, const float* n, float ior1, float ior2, float* r);
void hjo_kat_cosine(float u, float v, float* dir, float* pdf);
void hjo_kat_tonemap(int mode, const float* in3, float* out3);
void hjo_kat_to_srgb8(const float* in3, uint8_t* out3);

typedef struct hjo_kat_surface {
    float basecolor[3]; float metallic, roughness, sheen, clearcoat, ior;
    int32_t is_specular, is_thinfilm;
} hjo_kat_surface;
/* constant LUT override for the thin-film KAT (enabled=0 restores the scene LUT) */
void hjo_kat_set_lut_const(int enabled, float r, float g, float b);
/* BSDF(Payload).sampleBSDF: out = wi[3], f[3], pdf, rng calls consumed (8 floats) */
void hjo_kat_bsdf_sample(const hjo_kat_surface*, const float* wo, uint64_t n_spp, uint32_t image_idx, uint32_t scramble, float* out8);
/* BSDF(Payload).evaluateBSDF / getPDF: out = f[3], pdf */
void hjo_kat_bsdf_eval(const hjo_kat_surface*, const float* wo, const float* wi, float* out4);
/* IdealGlass(rho=1, ior).sampleBSDF — dead code in the dispatcher, KAT only: out = wi[3], f[3], pdf */
void hjo_kat_ideal_glass(float ior, const float* wo, uint64_t n_spp, uint32_t image_idx, uint32_t scramble, float* out7);
/* light_sample on the uploaded scene: out = pos[3], normal[3], emission[3], pdf, rng calls (11 floats) */
int hjo_kat_light_sample(hjo_ctx*, uint64_t n_spp, uint32_t image_idx, uint32_t scramble, float* out11);
float hjo_kat_light_pdf(hjo_ctx*, uint32_t prim, uint32_t inst);

#ifdef __cplusplus
}
#endif
#endif

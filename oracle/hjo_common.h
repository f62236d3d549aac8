// hjo_common.h — TEST INFRASTRUCTURE (CPU oracle). Not part of the product path.
//
// The parts of the hot path for which the reference snapshot has NO source and
// which are therefore *defined* by this build (decisions D1-D7), stated once here on the CPU and used
// by BOTH oracle builds: the restatement (hjo_api.cpp) and the build that compiles the reference's own
// device headers (ref_api.cpp -> oracle/_ref/). Contents:
//   * scene container: world-space flattening of the SceneData arrays (closest-hit vertex fetch of the
//     stale PTX, /root/reference/lib/ptx/HenjouRenderer_generated_henjouRendererCU.cu.ptx:880-1287);
//   * texture fetches: manual float bilinear, wrap addressing, optional sRGB decode — the float
//     definition of the CUDA texture descriptors at include/renderer/renderer.h:774-789,835-846,882-893;
//   * ray/triangle test (watertight, Woop et al. 2013, no FMA) and the closest-hit tie rule
//     (equal t -> lower global primitive id) — stands in for the closed-source OptiX traversal
//     (call sites include/kernel/rt.h:25,53); brute force and a median-split CPU BVH;
//   * closest-hit / miss surface reconstruction (D4-D6).
// "Parity unpinned" for exactly these parts: the reference holds no test vector at this boundary.
#pragma once
#include "hjo_vec.h"
#include "../include/henjou_b200.h"
#include <vector>
#include <cstring>
#include <cstdio>

namespace hjo {

struct Mat3x4 { float4 r0, r1, r2; };

// math.h:73-76 transform_position and :78-87 transform_normal (transposed inverse), restated.
static inline float3 xform_point(const Mat3x4& m, const float3& p) {
    float4 q = make_float4(p, 1.0f);
    return make_float3(dot(m.r0, q), dot(m.r1, q), dot(m.r2, q));
}
static inline float3 xform_normal(const Mat3x4& inv, const float3& n) {
    float4 q = make_float4(n, 0.0f);
    float4 c0 = make_float4(inv.r0.x, inv.r1.x, inv.r2.x, 0.0f);
    float4 c1 = make_float4(inv.r0.y, inv.r1.y, inv.r2.y, 0.0f);
    float4 c2 = make_float4(inv.r0.z, inv.r1.z, inv.r2.z, 0.0f);
    return make_float3(dot(c0, q), dot(c1, q), dot(c2, q));
}

struct Image8 { std::vector<uint8_t> px; int w = 0, h = 0; int srgb = 0; };
struct ImageF { std::vector<float4> px; int w = 0, h = 0; };

// sRGB EOTF as the texture unit applies it per texel before filtering (renderer.h:785).
static inline float srgb_to_linear(float c) {
    return (c <= 0.04045f) ? c / 12.92f : hjo_powf((c + 0.055f) / 1.055f, 2.4f);
}
static inline int wrapi(int i, int n) { int r = i % n; return r < 0 ? r + n : r; }

static inline float4 texel8(const Image8& im, int x, int y) {
    const uint8_t* p = &im.px[4 * ((size_t)wrapi(y, im.h) * im.w + wrapi(x, im.w))];
    float4 c = make_float4(p[0] / 255.0f, p[1] / 255.0f, p[2] / 255.0f, p[3] / 255.0f);
    if (im.srgb) { c.x = srgb_to_linear(c.x); c.y = srgb_to_linear(c.y); c.z = srgb_to_linear(c.z); }
    return c;
}
// CUDA linear filtering with normalized coords: xB = u*W - 0.5, i = floor(xB), a = frac(xB).
static inline float4 bilinear8(const Image8& im, float u, float v) {
    float xb = u * (float)im.w - 0.5f, yb = v * (float)im.h - 0.5f;
    float fx = floorf(xb), fy = floorf(yb);
    float a = xb - fx, b = yb - fy;
    int i = (int)fx, j = (int)fy;
    float4 t00 = texel8(im, i, j), t10 = texel8(im, i + 1, j), t01 = texel8(im, i, j + 1), t11 = texel8(im, i + 1, j + 1);
    float4 top = lerp(t00, t10, a), bot = lerp(t01, t11, a);
    return lerp(top, bot, b);
}
static inline float4 texelF(const ImageF& im, int x, int y) { return im.px[(size_t)wrapi(y, im.h) * im.w + wrapi(x, im.w)]; }
static inline float4 bilinearF(const ImageF& im, float u, float v) {
    if (im.w == 1 && im.h == 1) return im.px[0];
    float xb = u * (float)im.w - 0.5f, yb = v * (float)im.h - 0.5f;
    float fx = floorf(xb), fy = floorf(yb);
    float a = xb - fx, b = yb - fy;
    int i = (int)fx, j = (int)fy;
    float4 t00 = texelF(im, i, j), t10 = texelF(im, i + 1, j), t01 = texelF(im, i, j + 1), t11 = texelF(im, i + 1, j + 1);
    float4 top = lerp(t00, t10, a), bot = lerp(t01, t11, a);
    return lerp(top, bot, b);
}

struct Hit { bool hit = false; float t = 0, b1 = 0, b2 = 0; uint32_t prim = 0xffffffffu; };

// What __closesthit__ch / __miss__ms leave in the Payload (Payload.h:12-42), reconstructed (D4-D6).
struct Surface {
    bool is_hit = false;
    float3 position{0, 0, 0}, normal{0, 0, 0};
    float2 texcoord{0, 0};
    int material_id = 0;
    float3 basecolor{0, 0, 0};
    float metallic = 0, roughness = 0, sheen = 0, clearcoat = 0, ior = 1.0f, transmission = 1.0f;
    bool is_specular = false, is_light = false, is_thinfilm = false;
    float3 emission{0, 0, 0};
    int primitive_id = 0, instance_id = 0;
};

struct BvhNode { float lo[3], hi[3]; uint32_t left, count; };  // count>0: leaf, left = first index into order[]

struct Scene {
    // copies of the SceneData arrays
    std::vector<float3> vertices, normals;
    std::vector<float2> texcoords;
    std::vector<uint32_t> indices, material_ids, prim_offsets, light_prim_ids;
    std::vector<float3> light_prim_emission;
    std::vector<hj_material> materials;
    std::vector<Image8> textures;
    ImageF ibl;
    std::vector<float> sky_row_cdf, sky_marg;   // sky_sampling extension: running sums of luminance x sin(theta_row) per row / of the row totals
    Image8 lut;
    bool has_lut = false;
    // per-frame state (≙ updateIASMatrix, renderer.h:257-291)
    std::vector<Mat3x4> xf, inv_xf;
    std::vector<uint32_t> tri_inst;           // owning instance per global triangle
    std::vector<float3> wpos;                 // 3 world positions per triangle
    std::vector<float3> wnrm;                 // 3 normalised world normals per triangle (D4)
    std::vector<BvhNode> bvh;
    std::vector<uint32_t> order;
    bool use_bvh = true;

    uint32_t n_tris() const { return (uint32_t)material_ids.size(); }
};

static inline bool load_scene(Scene& S, const hj_scene_desc* d) {
    if (!d || !d->vertices || !d->indices || !d->material_ids || !d->prim_offsets || d->n_instances == 0) return false;
    const float3* v = (const float3*)d->vertices;
    S.vertices.assign(v, v + d->n_vertices);
    if (d->normals) { const float3* n = (const float3*)d->normals; S.normals.assign(n, n + d->n_vertices); }
    else S.normals.assign(d->n_vertices, float3{0, 1, 0});
    if (d->texcoords) { const float2* t = (const float2*)d->texcoords; S.texcoords.assign(t, t + d->n_vertices); }
    else S.texcoords.assign(d->n_vertices, float2{0, 0});
    S.indices.assign(d->indices, d->indices + 3 * (size_t)d->n_triangles);
    S.material_ids.assign(d->material_ids, d->material_ids + d->n_triangles);
    S.prim_offsets.assign(d->prim_offsets, d->prim_offsets + d->n_instances);
    S.materials.assign(d->materials, d->materials + d->n_materials);
    S.textures.clear();
    for (uint32_t i = 0; i < d->n_textures; i++) {
        Image8 im; im.w = d->textures[i].width; im.h = d->textures[i].height; im.srgb = d->textures[i].srgb;
        im.px.assign(d->textures[i].rgba8, d->textures[i].rgba8 + 4 * (size_t)im.w * im.h);
        S.textures.push_back(std::move(im));
    }
    S.light_prim_ids.assign(d->light_prim_ids, d->light_prim_ids + d->n_light_prims);
    const float3* le = (const float3*)d->light_prim_emission;
    S.light_prim_emission.assign(le, le + d->n_light_prims);
    if (d->ibl_rgba32f && d->ibl_width > 0) {
        S.ibl.w = d->ibl_width; S.ibl.h = d->ibl_height;
        const float4* p = (const float4*)d->ibl_rgba32f; S.ibl.px.assign(p, p + (size_t)S.ibl.w * S.ibl.h);
    } else { S.ibl.w = S.ibl.h = 1; S.ibl.px.assign(1, float4{0, 0, 0, 0}); }
    {   // the table of the sky_sampling extension: the loops of k_sky_rows / k_sky_marginal (csrc/hj_api.cu), same float addition order
        const int W = S.ibl.w, H = S.ibl.h;
        S.sky_row_cdf.assign((size_t)W * H, 0.0f); S.sky_marg.assign((size_t)H, 0.0f);
        for (int y = 0; y < H; y++) {
            const float st = hjo_sinf(M_PIf * ((float)y + 0.5f) / (float)H);
            float run = 0.0f;
            for (int x = 0; x < W; x++) {
                const float4 c = S.ibl.px[(size_t)y * W + x];
                const float lum = 0.2126f * c.x + 0.7152f * c.y + 0.0722f * c.z;
                run += fmaxf(lum, 0.0f) * st;
                S.sky_row_cdf[(size_t)y * W + x] = run;
            }
        }
        float run = 0.0f;
        for (int y = 0; y < H; y++) { run += S.sky_row_cdf[(size_t)y * W + W - 1]; S.sky_marg[y] = run; }
    }
    S.has_lut = d->lut_rgba8 && d->lut_width > 0;
    if (S.has_lut) {
        S.lut.w = d->lut_width; S.lut.h = d->lut_height; S.lut.srgb = 0;
        S.lut.px.assign(d->lut_rgba8, d->lut_rgba8 + 4 * (size_t)S.lut.w * S.lut.h);
    }
    S.tri_inst.resize(d->n_triangles);
    for (uint32_t i = 0; i < d->n_instances; i++) {
        uint32_t b = S.prim_offsets[i], e = (i + 1 < d->n_instances) ? S.prim_offsets[i + 1] : d->n_triangles;
        for (uint32_t t = b; t < e; t++) S.tri_inst[t] = i;
    }
    return true;
}

// General 3x4 affine inverse (cofactors), float. The reference uses glm::inverse on the 4x4
// (renderer.h:272-284); for T*R*S matrices both give the same matrix up to rounding.
static inline Mat3x4 invert_affine(const Mat3x4& m) {
    float a = m.r0.x, b = m.r0.y, c = m.r0.z, d = m.r1.x, e = m.r1.y, f = m.r1.z, g = m.r2.x, h = m.r2.y, i = m.r2.z;
    float A = e * i - f * h, B = -(d * i - f * g), C = d * h - e * g;
    float det = a * A + b * B + c * C;
    float id = 1.0f / det;
    Mat3x4 r;
    r.r0 = make_float4(A * id, -(b * i - c * h) * id, (b * f - c * e) * id, 0);
    r.r1 = make_float4(B * id, (a * i - c * g) * id, -(a * f - c * d) * id, 0);
    r.r2 = make_float4(C * id, -(a * h - b * g) * id, (a * e - b * d) * id, 0);
    float3 t = make_float3(m.r0.w, m.r1.w, m.r2.w);
    r.r0.w = -(r.r0.x * t.x + r.r0.y * t.y + r.r0.z * t.z);
    r.r1.w = -(r.r1.x * t.x + r.r1.y * t.y + r.r1.z * t.z);
    r.r2.w = -(r.r2.x * t.x + r.r2.y * t.y + r.r2.z * t.z);
    return r;
}

// ---------------------------------------------------------------------------------------------
// Watertight ray/triangle test (Woop, Benthin, Wald 2013), single precision, no FMA, with the
// double-precision fallback on a zero edge function. Pure function of (ray, triangle): the CUDA
// kernel states the same operation sequence so closest-hit ids are bit-exact.
// ---------------------------------------------------------------------------------------------
struct RayPrep { int kx, ky, kz; float Sx, Sy, Sz; float3 o; };

static inline float comp(const float3& v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : v.z); }

static inline RayPrep prep_ray(const float3& o, const float3& d) {
    RayPrep r; r.o = o;
    float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    r.kz = (ax > ay) ? ((ax > az) ? 0 : 2) : ((ay > az) ? 1 : 2);
    r.kx = r.kz + 1; if (r.kx == 3) r.kx = 0;
    r.ky = r.kx + 1; if (r.ky == 3) r.ky = 0;
    if (comp(d, r.kz) < 0.0f) { int t = r.kx; r.kx = r.ky; r.ky = t; }
    float dz = comp(d, r.kz);
    r.Sx = comp(d, r.kx) / dz; r.Sy = comp(d, r.ky) / dz; r.Sz = 1.0f / dz;
    return r;
}

static inline bool tri_test(const RayPrep& r, const float3& p0, const float3& p1, const float3& p2,
                            float tmin, float tmax, float& t, float& b1, float& b2) {
    const float3 A = p0 - r.o, B = p1 - r.o, C = p2 - r.o;
    const float Az = comp(A, r.kz), Bz = comp(B, r.kz), Cz = comp(C, r.kz);
    const float Ax = comp(A, r.kx) - r.Sx * Az, Ay = comp(A, r.ky) - r.Sy * Az;
    const float Bx = comp(B, r.kx) - r.Sx * Bz, By = comp(B, r.ky) - r.Sy * Bz;
    const float Cx = comp(C, r.kx) - r.Sx * Cz, Cy = comp(C, r.ky) - r.Sy * Cz;
    float U = Cx * By - Cy * Bx, V = Ax * Cy - Ay * Cx, W = Bx * Ay - By * Ax;
    if (U == 0.0f || V == 0.0f || W == 0.0f) {
        double CxBy = (double)Cx * (double)By, CyBx = (double)Cy * (double)Bx; U = (float)(CxBy - CyBx);
        double AxCy = (double)Ax * (double)Cy, AyCx = (double)Ay * (double)Cx; V = (float)(AxCy - AyCx);
        double BxAy = (double)Bx * (double)Ay, ByAx = (double)By * (double)Ax; W = (float)(BxAy - ByAx);
    }
    if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return false;
    const float det = U + V + W;
    if (det == 0.0f) return false;
    const float T = U * (r.Sz * Az) + V * (r.Sz * Bz) + W * (r.Sz * Cz);
    const float rdet = 1.0f / det;
    const float tt = T * rdet;
    if (!(tt > tmin && tt < tmax)) return false;
    t = tt; b1 = V * rdet; b2 = W * rdet;
    return true;
}

// ---------------------------------------------------------------------------------------------
// per-frame flattening + CPU BVH (median split on the longest centroid axis, leaves <= 4)
// ---------------------------------------------------------------------------------------------
static inline void bvh_build_rec(Scene& S, std::vector<float3>& cen, uint32_t b, uint32_t e) {
    BvhNode n; for (int k = 0; k < 3; k++) { n.lo[k] = FLT_MAX; n.hi[k] = -FLT_MAX; }
    float clo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, chi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (uint32_t i = b; i < e; i++) {
        uint32_t t = S.order[i];
        for (int k = 0; k < 3; k++) {
            const float3& p = S.wpos[3 * (size_t)t + k];
            float c[3] = {p.x, p.y, p.z};
            for (int a = 0; a < 3; a++) { n.lo[a] = std::min(n.lo[a], c[a]); n.hi[a] = std::max(n.hi[a], c[a]); }
        }
        float c[3] = {cen[t].x, cen[t].y, cen[t].z};
        for (int a = 0; a < 3; a++) { clo[a] = std::min(clo[a], c[a]); chi[a] = std::max(chi[a], c[a]); }
    }
    // conservative padding so the float slab test never culls a hit the triangle test accepts
    for (int a = 0; a < 3; a++) {
        float m = std::max(fabsf(n.lo[a]), fabsf(n.hi[a]));
        float pad = m * 1e-5f + 1e-7f;
        n.lo[a] -= pad; n.hi[a] += pad;
    }
    size_t me = S.bvh.size();
    S.bvh.push_back(n);
    if (e - b <= 4) { S.bvh[me].left = b; S.bvh[me].count = e - b; return; }
    int ax = 0; float ext = chi[0] - clo[0];
    for (int a = 1; a < 3; a++) if (chi[a] - clo[a] > ext) { ext = chi[a] - clo[a]; ax = a; }
    uint32_t mid = (b + e) / 2;
    std::nth_element(S.order.begin() + b, S.order.begin() + mid, S.order.begin() + e, [&](uint32_t x, uint32_t y) {
        float cx = ax == 0 ? cen[x].x : (ax == 1 ? cen[x].y : cen[x].z);
        float cy = ax == 0 ? cen[y].x : (ax == 1 ? cen[y].y : cen[y].z);
        return cx < cy || (cx == cy && x < y);
    });
    S.bvh[me].count = 0;
    bvh_build_rec(S, cen, b, mid);
    S.bvh[me].left = (uint32_t)S.bvh.size();   // right child index; left child is me+1
    bvh_build_rec(S, cen, mid, e);
}

// xforms: n_instances row-major 3x4 (NULL = identity); inv likewise (NULL = computed).
static inline void set_transforms(Scene& S, const float* xforms, const float* inv_xforms) {
    size_t ni = S.prim_offsets.size();
    S.xf.resize(ni); S.inv_xf.resize(ni);
    for (size_t i = 0; i < ni; i++) {
        Mat3x4 m;
        if (xforms) { const float* p = xforms + 12 * i; m.r0 = {p[0], p[1], p[2], p[3]}; m.r1 = {p[4], p[5], p[6], p[7]}; m.r2 = {p[8], p[9], p[10], p[11]}; }
        else { m.r0 = {1, 0, 0, 0}; m.r1 = {0, 1, 0, 0}; m.r2 = {0, 0, 1, 0}; }
        S.xf[i] = m;
        if (inv_xforms) { const float* p = inv_xforms + 12 * i; S.inv_xf[i].r0 = {p[0], p[1], p[2], p[3]}; S.inv_xf[i].r1 = {p[4], p[5], p[6], p[7]}; S.inv_xf[i].r2 = {p[8], p[9], p[10], p[11]}; }
        else if (xforms) S.inv_xf[i] = invert_affine(m);
        else S.inv_xf[i] = m;
    }
    uint32_t nt = S.n_tris();
    S.wpos.resize(3 * (size_t)nt); S.wnrm.resize(3 * (size_t)nt);
    std::vector<float3> cen(nt);
    for (uint32_t t = 0; t < nt; t++) {
        const Mat3x4& m = S.xf[S.tri_inst[t]]; const Mat3x4& im = S.inv_xf[S.tri_inst[t]];
        for (int k = 0; k < 3; k++) {
            uint32_t vi = S.indices[3 * (size_t)t + k];
            S.wpos[3 * (size_t)t + k] = xform_point(m, S.vertices[vi]);
            S.wnrm[3 * (size_t)t + k] = normalize(xform_normal(im, S.normals[vi]));
        }
        cen[t] = (S.wpos[3 * (size_t)t] + S.wpos[3 * (size_t)t + 1] + S.wpos[3 * (size_t)t + 2]) * (1.0f / 3.0f);
    }
    S.order.resize(nt); for (uint32_t t = 0; t < nt; t++) S.order[t] = t;
    S.bvh.clear();
    if (nt > 0) bvh_build_rec(S, cen, 0, nt);
}

static inline bool slab(const BvhNode& n, const float3& o, const float3& inv, float tmax_cur) {
    float t0 = (n.lo[0] - o.x) * inv.x, t1 = (n.hi[0] - o.x) * inv.x;
    float tn = std::min(t0, t1), tf = std::max(t0, t1);
    t0 = (n.lo[1] - o.y) * inv.y; t1 = (n.hi[1] - o.y) * inv.y;
    tn = std::max(tn, std::min(t0, t1)); tf = std::min(tf, std::max(t0, t1));
    t0 = (n.lo[2] - o.z) * inv.z; t1 = (n.hi[2] - o.z) * inv.z;
    tn = std::max(tn, std::min(t0, t1)); tf = std::min(tf, std::max(t0, t1));
    // NaN (0*inf) comparisons fall through as "hit": conservative
    return !(tn * 0.99999f > tf * 1.00001f) && !(tf < 0.0f) && !(tn * 0.99999f > tmax_cur);
}

static inline void test_prim(const Scene& S, const RayPrep& rp, uint32_t p, float tmin, float tmax, Hit& best) {
    float t, b1, b2;
    if (!tri_test(rp, S.wpos[3 * (size_t)p], S.wpos[3 * (size_t)p + 1], S.wpos[3 * (size_t)p + 2], tmin, tmax, t, b1, b2)) return;
    if (!best.hit || t < best.t || (t == best.t && p < best.prim)) { best.hit = true; best.t = t; best.b1 = b1; best.b2 = b2; best.prim = p; }
}

// closest hit in (tmin,tmax); ties on t -> lower global primitive id
static inline Hit closest_hit(const Scene& S, const float3& o, const float3& d, float tmin, float tmax) {
    Hit best; RayPrep rp = prep_ray(o, d);
    if (!S.use_bvh) { for (uint32_t p = 0; p < S.n_tris(); p++) test_prim(S, rp, p, tmin, tmax, best); return best; }
    if (S.bvh.empty()) return best;
    float3 inv = make_float3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    uint32_t stack[128]; int sp = 0; stack[sp++] = 0;
    while (sp) {
        uint32_t ni = stack[--sp]; const BvhNode& n = S.bvh[ni];
        if (!slab(n, o, inv, best.hit ? best.t : tmax)) continue;
        if (n.count) { for (uint32_t i = 0; i < n.count; i++) test_prim(S, rp, S.order[n.left + i], tmin, tmax, best); }
        else { stack[sp++] = n.left; stack[sp++] = ni + 1; }
    }
    return best;
}
static inline bool any_hit(const Scene& S, const float3& o, const float3& d, float tmin, float tmax) {
    Hit best; RayPrep rp = prep_ray(o, d);
    if (!S.use_bvh) { for (uint32_t p = 0; p < S.n_tris() && !best.hit; p++) test_prim(S, rp, p, tmin, tmax, best); return best.hit; }
    if (S.bvh.empty()) return false;
    float3 inv = make_float3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    uint32_t stack[128]; int sp = 0; stack[sp++] = 0;
    while (sp) {
        uint32_t ni = stack[--sp]; const BvhNode& n = S.bvh[ni];
        if (!slab(n, o, inv, tmax)) continue;
        if (n.count) { for (uint32_t i = 0; i < n.count; i++) { test_prim(S, rp, S.order[n.left + i], tmin, tmax, best); if (best.hit) return true; } }
        else { stack[sp++] = n.left; stack[sp++] = ni + 1; }
    }
    return false;
}

// D6: equirect sky lookup, y-up. u = phi/2pi with phi = atan2(z,x) in [0,2pi), v = acos(y)/pi.
static inline float3 sky_radiance(const Scene& S, const float3& dir, float ibl_intensity) {
    float phi = hjo_atan2f(dir.z, dir.x);
    if (phi < 0.0f) phi += 2.0f * M_PIf;
    float theta = hjo_acosf(clamp(dir.y, -1.0f, 1.0f));
    float4 c = bilinearF(S.ibl, phi * (1.0f / (2.0f * M_PIf)), theta * (1.0f / M_PIf));
    return ibl_intensity * make_float3(c);
}

// ---- sky importance sampling (extension behind hj_frame_params::sky_sampling; the reference BSDF-samples the sky only, rt.h:417-419).
// Restated operation for operation from csrc/hj_scene.cuh (sky_pdf_texel / sky_search / sky_sample / sky_pdf).
static inline float sky_texel_weight(const Scene& S, int x, int y) {
    const float* row = S.sky_row_cdf.data() + (size_t)y * S.ibl.w;
    return row[x] - (x ? row[x - 1] : 0.0f);
}
static inline float sky_pdf_texel(const Scene& S, int x, int y, float sin_theta) {
    const float total = S.sky_marg[S.ibl.h - 1];
    if (!(total > 0.0f) || !(sin_theta > 0.0f)) return 0.0f;
    const float pdf_uv = sky_texel_weight(S, x, y) * (float)S.ibl.w * (float)S.ibl.h / total;
    return pdf_uv / (2.0f * M_PIf * M_PIf * sin_theta);
}
static inline int sky_search(const float* cdf, int n, float t) {
    int lo = 0, hi = n - 1;
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (cdf[mid] > t) hi = mid; else lo = mid + 1; }
    return lo;
}
static inline float3 sky_sample(const Scene& S, const float2& xi, float& pdf) {
    pdf = 0.0f;
    const int W = S.ibl.w, H = S.ibl.h;
    const float total = S.sky_marg[H - 1];
    if (!(total > 0.0f)) return make_float3(0.0f, 1.0f, 0.0f);
    const float ty = xi.x * total;
    const int y = sky_search(S.sky_marg.data(), H, ty);
    const float* row = S.sky_row_cdf.data() + (size_t)y * W;
    const float tx = xi.y * row[W - 1];
    const int x = sky_search(row, W, tx);
    const float px = x ? row[x - 1] : 0.0f, fx = row[x] - px;
    const float py = y ? S.sky_marg[y - 1] : 0.0f, fy = S.sky_marg[y] - py;
    const float du = fx > 0.0f ? clamp((tx - px) / fx, 0.0f, 0.99999994f) : 0.5f;
    const float dv = fy > 0.0f ? clamp((ty - py) / fy, 0.0f, 0.99999994f) : 0.5f;
    const float u = ((float)x + du) / (float)W, v = ((float)y + dv) / (float)H;
    const float pa = 2.0f * M_PIf * u, ta = M_PIf * v;
    const float sp = hjo_sinf(pa), cp = hjo_cosf(pa), st = hjo_sinf(ta), ct = hjo_cosf(ta);
    pdf = sky_pdf_texel(S, x, y, st);
    return make_float3(st * cp, ct, st * sp);
}
static inline float sky_pdf(const Scene& S, const float3& dir) {
    float phi = hjo_atan2f(dir.z, dir.x);
    if (phi < 0.0f) phi += 2.0f * M_PIf;
    const float theta = hjo_acosf(clamp(dir.y, -1.0f, 1.0f));
    const float u = phi * (1.0f / (2.0f * M_PIf)), v = theta * (1.0f / M_PIf);
    int x = (int)(u * (float)S.ibl.w), y = (int)(v * (float)S.ibl.h);
    x = x < 0 ? 0 : (x > S.ibl.w - 1 ? S.ibl.w - 1 : x); y = y < 0 ? 0 : (y > S.ibl.h - 1 ? S.ibl.h - 1 : y);
    return sky_pdf_texel(S, x, y, hjo_sinf(theta));
}

// D4/D5: what __closesthit__ch leaves in the payload; miss -> D6.
static inline Surface make_surface(const Scene& S, const Hit& h, const float3& dir, float ibl_intensity) {
    Surface s;
    if (!h.hit) {
        s.is_hit = false; s.is_light = true;
        s.emission = sky_radiance(S, dir, ibl_intensity);
        return s;
    }
    size_t p = h.prim;
    float b0 = 1.0f - h.b1 - h.b2;
    s.is_hit = true;
    s.position = b0 * S.wpos[3 * p] + h.b1 * S.wpos[3 * p + 1] + h.b2 * S.wpos[3 * p + 2];
    s.normal = normalize(b0 * S.wnrm[3 * p] + h.b1 * S.wnrm[3 * p + 1] + h.b2 * S.wnrm[3 * p + 2]);
    const float2 t0 = S.texcoords[S.indices[3 * p]], t1 = S.texcoords[S.indices[3 * p + 1]], t2 = S.texcoords[S.indices[3 * p + 2]];
    s.texcoord = b0 * t0 + h.b1 * t1 + h.b2 * t2;
    s.material_id = (int)S.material_ids[p];
    const hj_material& m = S.materials[s.material_id];
    s.basecolor = make_float3(m.basecolor[0], m.basecolor[1], m.basecolor[2]);
    s.metallic = m.metallic; s.roughness = m.roughness;
    if (m.basecolor_tex >= 0 && m.basecolor_tex < (int)S.textures.size())
        s.basecolor = make_float3(bilinear8(S.textures[m.basecolor_tex], s.texcoord.x, s.texcoord.y));
    if (m.roughness_tex >= 0 && m.roughness_tex < (int)S.textures.size()) {
        float4 mr = bilinear8(S.textures[m.roughness_tex], s.texcoord.x, s.texcoord.y);
        s.roughness = mr.y * m.roughness; s.metallic = mr.z * m.metallic;
    }
    s.sheen = m.sheen; s.clearcoat = m.clearcoat; s.ior = m.ior; s.transmission = m.transmission;
    s.is_specular = m.ideal_specular != 0; s.is_thinfilm = m.is_thinfilm != 0;
    s.is_light = m.is_light != 0;
    s.emission = make_float3(m.emission[0], m.emission[1], m.emission[2]);
    s.primitive_id = (int)h.prim; s.instance_id = (int)S.tri_inst[p];
    return s;
}

// per-path debug record (thread-local; reset by the render loop before each camera path)
struct PathDebug { uint32_t rays_extend = 0, rays_shadow = 0, first_prim = 0xffffffffu, hash = 0x811c9dc5u; bool first = true; };

static inline void debug_note_extend(PathDebug& d, const Hit& h) {
    d.rays_extend++;
    uint32_t id = h.hit ? h.prim : 0xffffffffu;
    if (d.first) { d.first_prim = id; d.first = false; }
    d.hash = (d.hash ^ (id + 1u)) * 0x01000193u;
}

}  // namespace hjo

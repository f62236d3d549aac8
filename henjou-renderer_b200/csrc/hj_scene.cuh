// hj_scene.cuh — device-side scene view and the reconstructed closest-hit / miss programs.
//
// The reference's __closesthit__ch / __miss__ms have no source in the snapshot; what they leave in the
// Payload (/root/reference/include/kernel/Payload.h:12-42) is reconstructed from the stale PTX
// (lib/ptx/HenjouRenderer_generated_henjouRendererCU.cu.ptx:880-1287, :813-878) and SURVEY decisions D4-D6.
// Textures are plain linear-memory arrays with manual float bilinear filtering and explicit sRGB decode
// (the float definition of the texture descriptors at include/renderer/renderer.h:774-789,835-846,882-893).
#pragma once
#include "hj_sampling.cuh"

struct DevTexture { const uchar4* px; int w, h, srgb; };

// HitGroupData (renderer.h:659-687) reduced to the fields the device code reads.
struct DevMaterial {
    float basecolor[3]; int basecolor_tex;
    float emission[3]; int is_light;
    float metallic, roughness; int roughness_tex; float sheen;
    float clearcoat, ior, transmission; int flags;   // bit0 ideal_specular, bit1 is_thinfilm
};

struct DevScene {
    // object-space SceneData arrays (include/renderer/scene.h:19-37)
    const float* vertices; const float* normals; const float* texcoords;
    const uint32_t* indices; const uint32_t* material_ids; const uint32_t* prim_offsets; const uint32_t* tri_inst;
    uint32_t n_vertices, n_tris, n_instances, n_materials, n_textures, n_lights;
    const float4* xf; const float4* inv_xf;              // 3 rows per instance (Matrix4x3, cu/matrix_4x3.h:12)
    const DevMaterial* materials; const DevTexture* textures; const float* srgb_table;
    const uint32_t* light_prim_ids; const float* light_prim_emission;
    const float4* ibl; int ibl_w, ibl_h;
    // sky importance sampling (option, frame parameter sky_sampling): unnormalised running sums of luminance x sin(theta) per row,
    // and of the row totals (built by k_sky_rows / k_sky_marginal at upload)
    const float* sky_row_cdf; const float* sky_marg;
    const uchar4* lut; int lut_w, lut_h, has_lut;
    // world-space flattening, rebuilt by hj_accel_build (≙ updateIASMatrix, renderer.h:257-291)
    float4* tri_pos;     // 3 per triangle, by global primitive id
    float4* tri_attr;    // 4 per triangle: (n0,u0) (n1,v0) (n2,u1) (v1,u2,v2,material id)
    // acceleration structure
    const uint4* nodes;  // 7 x 16 B (112 B) per 8-wide node, layout in hj_trace.cuh
    const float4* btris; // 3 per triangle in leaf order, v0.w = global primitive id
};

// math.h:73-76 transform_position, :78-87 transform_normal (transposed inverse)
HJ_DI float3 xform_point(const float4* m, const float3& p) {
    float4 q = f4(p, 1.0f);
    return f3(dot(m[0], q), dot(m[1], q), dot(m[2], q));
}
HJ_DI float3 xform_normal(const float4* inv, const float3& n) {
    float4 q = f4(n, 0.0f);
    float4 c0 = make_float4(inv[0].x, inv[1].x, inv[2].x, 0.0f);
    float4 c1 = make_float4(inv[0].y, inv[1].y, inv[2].y, 0.0f);
    float4 c2 = make_float4(inv[0].z, inv[1].z, inv[2].z, 0.0f);
    return f3(dot(c0, q), dot(c1, q), dot(c2, q));
}
HJ_DI float3 ld3(const float* p, uint32_t i) { return f3(p[3 * (size_t)i], p[3 * (size_t)i + 1], p[3 * (size_t)i + 2]); }

HJ_DI int wrapi(int i, int n) { int r = i % n; return r < 0 ? r + n : r; }

// texel fetch: 8-bit unorm -> float, sRGB decode through a 256-entry table (table[i] = EOTF(i/255), filled
// at context creation) applied per texel before filtering, as the texture unit does (renderer.h:785).
HJ_DI float4 texel8(const DevTexture& im, const float* srgb_table, int x, int y) {
    const uchar4 p = im.px[(size_t)wrapi(y, im.h) * im.w + wrapi(x, im.w)];
    if (im.srgb) return make_float4(srgb_table[p.x], srgb_table[p.y], srgb_table[p.z], p.w / 255.0f);
    return make_float4(p.x / 255.0f, p.y / 255.0f, p.z / 255.0f, p.w / 255.0f);
}
// CUDA linear filtering with normalised coords: xB = u*W - 0.5, i = floor(xB), a = frac(xB)
static HJ_NI float4 bilinear8(const DevTexture& im, const float* srgb_table, float u, float v) {
    float xb = u * (float)im.w - 0.5f, yb = v * (float)im.h - 0.5f;
    float fx = floorf(xb), fy = floorf(yb);
    float a = xb - fx, b = yb - fy;
    int i = (int)fx, j = (int)fy;
    float4 t00 = texel8(im, srgb_table, i, j), t10 = texel8(im, srgb_table, i + 1, j);
    float4 t01 = texel8(im, srgb_table, i, j + 1), t11 = texel8(im, srgb_table, i + 1, j + 1);
    float4 top = lerp4(t00, t10, a), bot = lerp4(t01, t11, a);
    return lerp4(top, bot, b);
}
HJ_DI float4 texelF(const DevScene& S, int x, int y) { return S.ibl[(size_t)wrapi(y, S.ibl_h) * S.ibl_w + wrapi(x, S.ibl_w)]; }

// D6: equirect sky lookup, y-up: u = phi/2pi with phi = atan2(z,x) in [0,2pi), v = acos(y)/pi.
static HJ_NI float3 sky_radiance(const DevScene& S, const float3& dir, float ibl_intensity) {
    float4 c;
    if (S.ibl_w == 1 && S.ibl_h == 1) c = S.ibl[0];
    else {
        float phi = hj_atan2f(dir.z, dir.x);
        if (phi < 0.0f) phi += 2.0f * M_PIf;
        float theta = hj_acosf(clampf(dir.y, -1.0f, 1.0f));
        float u = phi * (1.0f / (2.0f * M_PIf)), v = theta * (1.0f / M_PIf);
        float xb = u * (float)S.ibl_w - 0.5f, yb = v * (float)S.ibl_h - 0.5f;
        float fx = floorf(xb), fy = floorf(yb);
        float a = xb - fx, b = yb - fy;
        int i = (int)fx, j = (int)fy;
        float4 top = lerp4(texelF(S, i, j), texelF(S, i + 1, j), a), bot = lerp4(texelF(S, i, j + 1), texelF(S, i + 1, j + 1), a);
        c = lerp4(top, bot, b);
    }
    return ibl_intensity * f3(c);
}

// ---- sky importance sampling (the reference BSDF-samples the sky only, rt.h:417-419 — this is an extension behind
// hj_frame_params::sky_sampling; the CPU checker of the tests restates it operation for operation). Density on the unit (u, v) square:
// piecewise constant per texel, proportional to luminance x sin(theta_row); solid-angle density = that / (2 pi^2 sin theta).
HJ_DI float sky_texel_weight(const DevScene& S, int x, int y) {
    const float* row = S.sky_row_cdf + (size_t)y * S.ibl_w;
    return row[x] - (x ? row[x - 1] : 0.0f);
}
HJ_DI float sky_pdf_texel(const DevScene& S, int x, int y, float sin_theta) {
    const float total = S.sky_marg[S.ibl_h - 1];
    if (!(total > 0.0f) || !(sin_theta > 0.0f)) return 0.0f;
    const float pdf_uv = sky_texel_weight(S, x, y) * (float)S.ibl_w * (float)S.ibl_h / total;
    return pdf_uv / (2.0f * M_PIf * M_PIf * sin_theta);
}
// first index in [0, n) whose running sum exceeds t (the last one if none does)
HJ_DI int sky_search(const float* cdf, int n, float t) {
    int lo = 0, hi = n - 1;
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (cdf[mid] > t) hi = mid; else lo = mid + 1; }
    return lo;
}
static HJ_NI float3 sky_sample(const DevScene& S, const float2& xi, float& pdf) {
    pdf = 0.0f;
    const float total = S.sky_marg[S.ibl_h - 1];
    if (!(total > 0.0f)) return f3(0.0f, 1.0f, 0.0f);
    const float ty = xi.x * total;
    const int y = sky_search(S.sky_marg, S.ibl_h, ty);
    const float* row = S.sky_row_cdf + (size_t)y * S.ibl_w;
    const float tx = xi.y * row[S.ibl_w - 1];
    const int x = sky_search(row, S.ibl_w, tx);
    const float px = x ? row[x - 1] : 0.0f, fx = row[x] - px;
    const float py = y ? S.sky_marg[y - 1] : 0.0f, fy = S.sky_marg[y] - py;
    const float du = fx > 0.0f ? clampf((tx - px) / fx, 0.0f, 0.99999994f) : 0.5f;
    const float dv = fy > 0.0f ? clampf((ty - py) / fy, 0.0f, 0.99999994f) : 0.5f;
    const float u = ((float)x + du) / (float)S.ibl_w, v = ((float)y + dv) / (float)S.ibl_h;
    float sp, cp, st, ct;
    hj_sincosf(2.0f * M_PIf * u, sp, cp); hj_sincosf(M_PIf * v, st, ct);
    pdf = sky_pdf_texel(S, x, y, st);
    return f3(st * cp, ct, st * sp);
}
static HJ_NI float sky_pdf(const DevScene& S, const float3& dir) {
    float phi = hj_atan2f(dir.z, dir.x);
    if (phi < 0.0f) phi += 2.0f * M_PIf;
    const float theta = hj_acosf(clampf(dir.y, -1.0f, 1.0f));
    const float u = phi * (1.0f / (2.0f * M_PIf)), v = theta * (1.0f / M_PIf);
    int x = (int)(u * (float)S.ibl_w), y = (int)(v * (float)S.ibl_h);
    x = x < 0 ? 0 : (x > S.ibl_w - 1 ? S.ibl_w - 1 : x); y = y < 0 ? 0 : (y > S.ibl_h - 1 ? S.ibl_h - 1 : y);
    return sky_pdf_texel(S, x, y, hj_sinf(theta));
}

// disneyBRDF.h:11-14 — thin-film LUT: RGBA8, wrap, bilinear, non-sRGB
HJ_DI float3 lut_fetch(const DevScene& S, float u, float v) {
    if (!S.has_lut) return f3(0.0f);
    DevTexture t; t.px = S.lut; t.w = S.lut_w; t.h = S.lut_h; t.srgb = 0;
    return f3(bilinear8(t, S.srgb_table, u, v));
}

// What the closest-hit program leaves in the payload (Payload.h:12-42)
struct Surface {
    float3 position, normal, basecolor, emission;
    float metallic, roughness, sheen, clearcoat, ior;
    bool is_specular, is_light, is_thinfilm;
};

// D4/D5: barycentric interpolation of world-space position / normalised world normals / uv, material fetch.
HJ_DI Surface make_surface(const DevScene& S, uint32_t prim, float b1, float b2) {
    Surface s;
    const float4* P = S.tri_pos + 3 * (size_t)prim;
    const float4* A = S.tri_attr + 4 * (size_t)prim;
    const float4 a0 = A[0], a1 = A[1], a2 = A[2], a3 = A[3];
    const float b0 = 1.0f - b1 - b2;
    s.position = b0 * f3(P[0]) + b1 * f3(P[1]) + b2 * f3(P[2]);
    s.normal = normalize(b0 * f3(a0) + b1 * f3(a1) + b2 * f3(a2));
    const float2 uv = b0 * make_float2(a0.w, a1.w) + b1 * make_float2(a2.w, a3.x) + b2 * make_float2(a3.y, a3.z);
    const DevMaterial m = S.materials[__float_as_uint(a3.w)];
    s.basecolor = f3(m.basecolor[0], m.basecolor[1], m.basecolor[2]);
    s.metallic = m.metallic; s.roughness = m.roughness;
    if (m.basecolor_tex >= 0 && m.basecolor_tex < (int)S.n_textures)
        s.basecolor = f3(bilinear8(S.textures[m.basecolor_tex], S.srgb_table, uv.x, uv.y));
    if (m.roughness_tex >= 0 && m.roughness_tex < (int)S.n_textures) {
        float4 mr = bilinear8(S.textures[m.roughness_tex], S.srgb_table, uv.x, uv.y);
        s.roughness = mr.y * m.roughness; s.metallic = mr.z * m.metallic;
    }
    s.sheen = m.sheen; s.clearcoat = m.clearcoat; s.ior = m.ior;
    s.is_specular = (m.flags & 1) != 0; s.is_thinfilm = (m.flags & 2) != 0;
    s.is_light = m.is_light != 0;
    s.emission = f3(m.emission[0], m.emission[1], m.emission[2]);
    return s;
}

// hj_host_scene.h — host-side scene of the headless loader: the reference's SceneData
// (/root/reference/include/renderer/scene.h:19-37), its per-node TRS animation (include/renderer/animation.h:34-131,
// include/common/matrix.h) and render options (include/renderer/render_option.h), filled from render_option.json +
// glTF by hj_gltf.h with the loader semantics of include/loader/gltfloader.h:1068-1601. All arithmetic is float, in the
// reference's order (double literals where the reference has them).
#pragma once
#include <cmath>
#include <cstdint>
#include <string>
#include <vector>
#include "../../include/henjou_b200.h"

namespace hj {

struct Affine { float v[16]; };   // row-major 4x4, common/matrix.h:6-19

inline Affine translate_affine(const float t[3]) {                                   // matrix.h:21-24
    return Affine{{1, 0, 0, t[0], 0, 1, 0, t[1], 0, 0, 1, t[2], 0, 0, 0, 1}};
}
inline Affine scale_affine(const float s[3]) {                                       // matrix.h:26-29
    return Affine{{s[0], 0, 0, 0, 0, s[1], 0, 0, 0, 0, s[2], 0, 0, 0, 0, 1}};
}
inline Affine rotate_affine(const float q[4]) {                                      // matrix.h:32-58: 2.0 * a * b in double, rounded
    const float x = q[0], y = q[1], z = q[2], w = q[3];
    const float q2xy = 2.0 * x * y, q2xz = 2.0 * x * z, q2xw = 2.0 * x * w, q2yz = 2.0 * y * z, q2yw = 2.0 * y * w,
                q2zw = 2.0 * z * w, q2ww = 2.0 * w * w;
    return Affine{{q2ww + 2.0f * x * x - 1.0f, q2xy - q2zw, q2xz + q2yw, 0,
                   q2xy + q2zw, q2ww + 2.0f * y * y - 1.0f, q2yz - q2xw, 0,
                   q2xz - q2yw, q2yz + q2xw, q2ww + 2.0f * z * z - 1.0f, 0,
                   0, 0, 0, 1}};
}
inline Affine operator*(const Affine& a, const Affine& b) {                         // matrix.h:69-78
    Affine r;
    for (int j = 0; j < 4; j++)
        for (int i = 0; i < 4; i++)
            r.v[i + j * 4] = a.v[0 + j * 4] * b.v[i + 0 * 4] + a.v[1 + j * 4] * b.v[i + 1 * 4] + a.v[2 + j * 4] * b.v[i + 2 * 4] + a.v[3 + j * 4] * b.v[i + 3 * 4];
    return r;
}
inline void affine_mul4(const Affine& m, const float p[4], float out[4]) {           // matrix.h:60-67
    for (int r = 0; r < 4; r++) out[r] = p[0] * m.v[4 * r] + p[1] * m.v[4 * r + 1] + p[2] * m.v[4 * r + 2] + p[3] * m.v[4 * r + 3];
}
// Rows 0..2 of the inverse, by cofactors in float — what glm::inverse(mat4) computes for updateIASMatrix
// (include/renderer/renderer.h:272-284; glm's exact operation order is not reproduced: parity unpinned at 1 ulp).
inline void affine_inverse_rows(const Affine& a, float out12[12]) {
    const float* m = a.v;
    float inv[16];
    inv[0] = m[5] * m[10] * m[15] - m[5] * m[11] * m[14] - m[9] * m[6] * m[15] + m[9] * m[7] * m[14] + m[13] * m[6] * m[11] - m[13] * m[7] * m[10];
    inv[4] = -m[4] * m[10] * m[15] + m[4] * m[11] * m[14] + m[8] * m[6] * m[15] - m[8] * m[7] * m[14] - m[12] * m[6] * m[11] + m[12] * m[7] * m[10];
    inv[8] = m[4] * m[9] * m[15] - m[4] * m[11] * m[13] - m[8] * m[5] * m[15] + m[8] * m[7] * m[13] + m[12] * m[5] * m[11] - m[12] * m[7] * m[9];
    inv[12] = -m[4] * m[9] * m[14] + m[4] * m[10] * m[13] + m[8] * m[5] * m[14] - m[8] * m[6] * m[13] - m[12] * m[5] * m[10] + m[12] * m[6] * m[9];
    inv[1] = -m[1] * m[10] * m[15] + m[1] * m[11] * m[14] + m[9] * m[2] * m[15] - m[9] * m[3] * m[14] - m[13] * m[2] * m[11] + m[13] * m[3] * m[10];
    inv[5] = m[0] * m[10] * m[15] - m[0] * m[11] * m[14] - m[8] * m[2] * m[15] + m[8] * m[3] * m[14] + m[12] * m[2] * m[11] - m[12] * m[3] * m[10];
    inv[9] = -m[0] * m[9] * m[15] + m[0] * m[11] * m[13] + m[8] * m[1] * m[15] - m[8] * m[3] * m[13] - m[12] * m[1] * m[11] + m[12] * m[3] * m[9];
    inv[2] = m[1] * m[6] * m[15] - m[1] * m[7] * m[14] - m[5] * m[2] * m[15] + m[5] * m[3] * m[14] + m[13] * m[2] * m[7] - m[13] * m[3] * m[6];
    inv[6] = -m[0] * m[6] * m[15] + m[0] * m[7] * m[14] + m[4] * m[2] * m[15] - m[4] * m[3] * m[14] - m[12] * m[2] * m[7] + m[12] * m[3] * m[6];
    inv[10] = m[0] * m[5] * m[15] - m[0] * m[7] * m[13] - m[4] * m[1] * m[15] + m[4] * m[3] * m[13] + m[12] * m[1] * m[7] - m[12] * m[3] * m[5];
    inv[3] = -m[1] * m[6] * m[11] + m[1] * m[7] * m[10] + m[5] * m[2] * m[11] - m[5] * m[3] * m[10] - m[9] * m[2] * m[7] + m[9] * m[3] * m[6];
    inv[7] = m[0] * m[6] * m[11] - m[0] * m[7] * m[10] - m[4] * m[2] * m[11] + m[4] * m[3] * m[10] + m[8] * m[2] * m[7] - m[8] * m[3] * m[6];
    inv[11] = -m[0] * m[5] * m[11] + m[0] * m[7] * m[9] + m[4] * m[1] * m[11] - m[4] * m[3] * m[9] - m[8] * m[1] * m[7] + m[8] * m[3] * m[5];
    const float det = m[0] * inv[0] + m[1] * inv[4] + m[2] * inv[8] + m[3] * inv[12];
    const float r = 1.0f / det;
    for (int i = 0; i < 12; i++) out12[i] = inv[i] * r;
}

// AnimationData<T> + Animation (animation.h:20-94). Key 0 holds the node's static TRS (gltfloader.h:1313-1343).
struct NodeAnimation {
    std::vector<float> t_key, r_key, s_key;
    std::vector<float> t_val, r_val, s_val;   // 3 / 4 / 3 floats per key

    template <int N>
    static void interp(const std::vector<float>& val, const std::vector<float>& key, float time, float out[N]) {   // animation.h:39-79
        auto at = [&](size_t k, float o[N]) { for (int c = 0; c < N; c++) o[c] = val[N * k + c]; };
        if (key.size() == 1 || time < 0) { at(0, out); return; }
        int first = 0, len = (int)key.size();
        while (len > 0) {
            int half = len >> 1, middle = first + half;
            if (key[middle] <= time) { first = middle + 1; len -= half + 1; }
            else len = half;
        }
        int offset = first - 1;
        if ((int)key.size() - 1 <= offset) { at(key.size() - 1, out); return; }
        float time_offset = time - key[offset];
        float time_delta = key[offset + 1] - key[offset];
        float delta = time_offset / time_delta;
        float a[N], b[N];
        at((size_t)offset, a); at((size_t)offset + 1, b);
        for (int c = 0; c < N; c++) out[c] = a[c] * (1.0f - delta) + b[c] * (delta);
    }
    Affine affine(float time) const {                                                // getAnimationAffine, animation.h:81-94
        float t[3] = {0, 0, 0}, r[4] = {0, 0, 0, 0}, s[3] = {0, 0, 0};
        if (!t_key.empty()) interp<3>(t_val, t_key, time, t);
        if (!r_key.empty()) interp<4>(r_val, r_key, time, r);
        if (!s_key.empty()) interp<3>(s_val, s_key, time, s);
        return translate_affine(t) * rotate_affine(r) * scale_affine(s);
    }
    Affine rotate_only(float time) const {                                           // getRotateAnimationAffine, animation.h:96-103
        float r[4] = {0, 0, 0, 0};
        if (!r_key.empty()) interp<4>(r_val, r_key, time, r);
        return rotate_affine(r);
    }
};

struct HostTexture { int width = 0, height = 0, srgb = 0; std::vector<uint8_t> rgba; std::string name; };

struct RenderOption {                                                                // include/renderer/render_option.h
    int image_width = 0, image_height = 0; std::string image_name, image_directory; unsigned max_spp = 1;
    std::string render_mode = "Default";
    std::string gltf_path, gltf_name;
    bool allow_camera_animation = true;
    float camera_position[3] = {0, 0, 0}, camera_direction[3] = {0, 0, -1}; float camera_fov = 0.0f;
    int camera_animation_id = -1;
    std::string ptxfile_path;
    unsigned fps = 24; int start_frame = 0, end_frame = 1; float time_limit = 0.0f;
    std::string IBL_path; float IBL_intensity = 1.0f; bool use_IBL = false; float scene_sky_default[3] = {0, 0, 0};
    bool use_date = false, save_renderOption = false;
    std::string LUT_path;
    // optional section "Integrator" (absent in the reference's files): SURVEY decisions D3 / D7 as defaults
    int integrator = HJ_INTEGRATOR_MIS, max_depth = 10, tonemap = HJ_TONEMAP_NONE;
};

struct HostScene {
    std::vector<float> vertices, normals, texcoords;     // float3 / float3 / float2 per de-indexed vertex
    std::vector<uint32_t> indices, material_ids, prim_offsets;
    std::vector<hj_material> materials;
    std::vector<HostTexture> textures;
    std::vector<uint32_t> light_prim_ids; std::vector<float> light_prim_emission;
    std::vector<NodeAnimation> animations;               // one per glTF node
    std::vector<uint32_t> instance_node;                 // InstanceData::animation_id: node index of instance i
    std::vector<float> ibl; int ibl_w = 1, ibl_h = 1;    // float4 image
    HostTexture lut; bool has_lut = false;
    uint32_t n_triangles() const { return (uint32_t)material_ids.size(); }
    uint32_t n_instances() const { return (uint32_t)prim_offsets.size(); }
};

}  // namespace hj

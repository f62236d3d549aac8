// TEST INFRASTRUCTURE (oracle reference build). RECONSTRUCTED stand-in for the absent
// HenjouRenderer/henjouRenderer.h: the launch-parameter struct whose fields the host fills at
// /root/reference/include/renderer/renderer.h:1175-1227 (order of the first 17 confirmed by the stale PTX,
// SURVEY §3.3). Only the fields the kernel headers read matter for the oracle.
#pragma once
#include "../ref_shim_base.h"
#include <cu/matrix_4x3.h>

template <class T> T tex2D(cudaTextureObject_t tex, float u, float v);
template <> inline float4 tex2D<float4>(cudaTextureObject_t, float u, float v) {
    // the only tex2D in the kernel headers is the thin-film LUT, kernel/disneyBRDF.h:11-14
    return make_float4(hjo::lut_lookup(u, v), 1.0f);
}

struct Params {
    uchar4* image;
    unsigned int image_width, image_height;
    OptixTraversableHandle traversal_handle;
    unsigned int spp, frame;
    float3 camera_pos, camera_dir, camera_up, camera_right;
    float camera_f;
    float3* vertices;
    unsigned int* indices;
    float3* normals;
    float2* texcoords;
    float3* colors;
    unsigned int* prim_offsets;
    Matrix4x3* transforms;
    Matrix4x3* inv_transforms;
    unsigned int instance_count;
    cudaTextureObject_t* textures;
    cudaTextureObject_t ibl_texture;
    float ibl_intensity;
    unsigned int* light_prim_ids;
    unsigned int light_prim_count;
    float3* light_prim_emission;
    unsigned int RAYTYPE;
    float4* aov_albedo;
    float4* aov_color;
    float4* aov_normal;
    cudaTextureObject_t lut_texture;
};

/*
 * henjou_b200.h — C ABI of libhenjou_b200.so, the B200-native (sm_100a) replacement for the
 * device half of kinakomoti-321/Henjou-Renderer.
 *
 * What this boundary replaces in the reference (all paths relative to /root/reference):
 *   - the device module loaded from `PTX_File.ptxfile_path` (HenjouRenderer/render_option.json:28-30,
 *     include/renderer/renderer.h:508-519) exporting __raygen__rg / __miss__ms / __closesthit__ch /
 *     __closesthit__shadow (renderer.h:528-573);
 *   - the OptiX acceleration-structure calls buildGAS / buildIAS / updateIASMatrix
 *     (renderer.h:319-490, 257-291);
 *   - the launch itself: cudaMemcpy(Params) + optixLaunch(W,H,1) + sync (renderer.h:1229-1242);
 *   - the AOV read-back + sRGB quantisation (renderer.h:73-101, 1279-1283).
 *
 * Plain C: pointers + sizes only, no C++/torch types. All input pointers are HOST pointers that are
 * borrowed for the duration of the call (copied to the device inside). No exception crosses the ABI:
 * every entry point returns HJ_OK (0) or a negative hj_status; hj_last_error() gives the message.
 * A context is not thread-safe; it owns one CUDA stream and all of its device memory.
 */
#ifndef HENJOU_B200_H
#define HENJOU_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HJ_ABI_VERSION 3

typedef struct hj_ctx hj_ctx;

typedef enum hj_status {
    HJ_OK = 0,
    HJ_ERR_INVALID = -1,   /* bad argument / inconsistent scene description      */
    HJ_ERR_CUDA = -2,      /* a CUDA runtime call failed (message has the detail) */
    HJ_ERR_NO_SCENE = -3,  /* hj_accel_build / hj_render before hj_scene_upload   */
    HJ_ERR_NO_ACCEL = -4,  /* hj_render before hj_accel_build                     */
    HJ_ERR_NO_FRAME = -5,  /* hj_readback before hj_render                        */
    HJ_ERR_COMM = -6,      /* NCCL failure                                        */
    HJ_ERR_OOM = -7        /* device allocation failed                            */
} hj_status;

/* rt.h:85 (Pathtrace), rt.h:162 (NEE), rt.h:284 (MIS). SURVEY decision D3: default MIS. */
typedef enum hj_integrator { HJ_INTEGRATOR_PT = 0, HJ_INTEGRATOR_NEE = 1, HJ_INTEGRATOR_MIS = 2 } hj_integrator;

/* kernel/color.h:10-63. SURVEY decision D7: default none (linear mean radiance in the colour AOV). */
typedef enum hj_tonemap { HJ_TONEMAP_NONE = 0, HJ_TONEMAP_ACES = 1, HJ_TONEMAP_UCHIMURA = 2 } hj_tonemap;

typedef enum hj_aov {
    HJ_AOV_COLOR = 0,        /* float4[W*H]  Params::aov_color   (renderer.h:1222)                         */
    HJ_AOV_ALBEDO = 1,       /* float4[W*H]  Params::aov_albedo  (renderer.h:1221)                         */
    HJ_AOV_NORMAL = 2,       /* float4[W*H]  Params::aov_normal  (renderer.h:1223)                         */
    HJ_AOV_RGBA8 = 3,        /* uchar4[W*H]  Params::image; float4ConvertColor semantics (renderer.h:87)   */
    HJ_AOV_ACCUM = 4,        /* float4[W*H]  raw radiance SUM of this context's sample slice (multi-GPU)   */
    HJ_AOV_PATH_RADIANCE = 5,/* float4[n_samples*W*H] debug: per-sample radiance rgb, w = #rays of path     */
    HJ_AOV_OUTPUT = 7,       /* float4[OW*OH] result of hj_post: AOV_Output (renderer.h:1106), OW x OH = W x H or 2W x 2H */
    HJ_AOV_OUTPUT_RGBA8 = 8, /* uchar4[OW*OH] float4ConvertColor of HJ_AOV_OUTPUT: the pixels of the written PNG     */
    HJ_AOV_PATH_IDS = 6      /* uint32[2*n_samples*W*H] debug: {first-hit global prim id | 0xffffffff,
                                                               hash of every closest-hit id on the path}   */
} hj_aov;

typedef enum hj_accel_mode { HJ_ACCEL_BUILD = 0, HJ_ACCEL_REFIT = 1 } hj_accel_mode;

/* Mirrors HitGroupData as filled by Renderer::optixSBTBuild (renderer.h:659-687) from
 * Material (include/renderer/material.h:10-63). Texture ids index hj_scene_desc::textures, -1 = none. */
typedef struct hj_material {
    float basecolor[3];
    int32_t basecolor_tex;
    float emission[3];          /* already multiplied by KHR_materials_emissive_strength (gltfloader.h:1237-1245) */
    int32_t emission_tex;       /* always -1 from the loader (gltfloader.h:1159); kept for layout parity */
    int32_t is_light;
    float metallic;
    int32_t metallic_tex;       /* loader sets = roughness_tex (gltfloader.h:1156) */
    float roughness;
    int32_t roughness_tex;
    float specular[3];
    float sheen;
    float clearcoat;
    int32_t clearcoat_tex;
    int32_t bump_tex;
    int32_t normal_tex;
    float transmission;
    float ior;
    int32_t ideal_specular;     /* roughness==0 && transmission>0 (gltfloader.h:1260-1263) */
    int32_t is_thinfilm;
} hj_material;

/* Decoded 8-bit RGBA image as produced by Texture (include/renderer/texture.h:16-39). */
typedef struct hj_texture {
    const uint8_t* rgba8;       /* width*height*4 bytes, row 0 first */
    int32_t width, height;
    int32_t srgb;               /* 1: decode sRGB->linear on fetch (renderer.h:785-789), 0: NonColor */
} hj_texture;

/* The nine SceneData arrays (include/renderer/scene.h:19-37) + materials, textures, sky and LUT. */
typedef struct hj_scene_desc {
    const float* vertices;            /* float3[n_vertices], object space                          */
    const float* normals;             /* float3[n_vertices]; NULL: (0,1,0) everywhere — the reference LOADER substitutes geometric normals
                                         (gltfloader.h:1460-1482), a host that relies on that must compute them before the call   */
    const float* texcoords;           /* float2[n_vertices]; NULL: (0,0) everywhere (gltfloader.h:1476-1482)                        */
    const uint32_t* indices;          /* uint32[3*n_triangles] (loader emits the identity)         */
    const uint32_t* material_ids;     /* uint32[n_triangles]                                       */
    const uint32_t* prim_offsets;     /* uint32[n_instances], first global triangle of instance i  */
    uint32_t n_vertices, n_triangles, n_instances;
    const hj_material* materials;
    uint32_t n_materials;
    const hj_texture* textures;
    uint32_t n_textures;
    const uint32_t* light_prim_ids;   /* ascending global triangle ids                             */
    const float* light_prim_emission; /* float3[n_light_prims]                                     */
    uint32_t n_light_prims;
    const float* ibl_rgba32f;         /* float4[ibl_width*ibl_height]; 1x1 = constant sky (texture.h:58-65) */
    int32_t ibl_width, ibl_height;
    const uint8_t* lut_rgba8;         /* thin-film LUT, RGBA8, non-sRGB (renderer.h:854-898); may be NULL */
    int32_t lut_width, lut_height;
} hj_scene_desc;

/* Params (renderer.h:1175-1227) minus the device pointers. */
typedef struct hj_frame_params {
    uint32_t width, height;
    uint32_t spp;                 /* Params::spp — total samples per pixel of the frame (the normaliser)  */
    uint32_t sample_begin;        /* this context renders global sample indices [sample_begin,sample_end) */
    uint32_t sample_end;          /* single GPU: 0, spp                                                    */
    uint32_t frame;               /* Params::frame — CMJ scramble                              */
    float camera_pos[3];
    float camera_dir[3];
    float camera_up[3];
    float camera_right[3];
    float camera_f;               /* 2/tan(fov) (renderer.h:1147) */
    float ibl_intensity;
    int32_t integrator;           /* hj_integrator */
    int32_t max_depth;            /* literal 10 in the reference (rt.h:89,166,288) */
    int32_t tonemap;              /* hj_tonemap */
    int32_t pixel_jitter;         /* 1 = first CMJ draw of each sample jitters the pixel */
    int32_t debug_paths;          /* 1 = record HJ_AOV_PATH_* (needs n_samples*W*H*24 bytes) */
    int32_t sky_sampling;         /* 0 (default): the reference's behaviour — the sky is reached by BSDF sampling only (rt.h:417-419).
                                     1 (MIS integrator only): EXTENSION, SURVEY §8 f4 — at every non-specular vertex one direction is also drawn
                                     from the sky image's luminance x sin(theta) table (built on the GPU at hj_scene_upload), one more shadow ray,
                                     balance heuristic between the two; one extra RNG dimension per vertex. Restated in oracle/libhjo.so; the
                                     reference-headers build refuses it. */
} hj_frame_params;

typedef struct hj_stats {
    uint64_t rays_extend;         /* closest-hit queries traced in the last hj_render (RayTrace, rt.h:43)      */
    uint64_t rays_shadow;         /* occlusion queries traced (TraceOcculution, rt.h:15)                       */
    uint64_t paths;               /* camera paths started                                                       */
    uint64_t nodes_fetched;       /* BVH nodes fetched (only with hj_set_option("count_fetches",1))             */
    uint64_t tris_fetched;        /* triangles fetched (same)                                                   */
    uint64_t kernel_launches;     /* kernels launched by the last hj_render                                     */
    float ms_render;              /* device time of the last hj_render (CUDA events on the ctx stream)          */
    float ms_trace;               /* ... of which traversal kernels (only with option "time_stages")            */
    float ms_shade;
    float ms_accel;               /* device time of the last hj_accel_build                                     */
    uint32_t bvh_nodes;           /* nodes in the current acceleration structure                                */
    uint32_t bvh_node_bytes;      /* bytes per node                                                             */
    uint32_t bvh_depth;           /* levels of the 8-wide tree                                                  */
    uint32_t max_nodes_per_ray;   /* worst single ray of the last hj_render (only with "count_fetches")         */
    uint32_t max_tris_per_ray;
    float ms_post;                /* device time of the last hj_post                                            */
    uint32_t post_launches;       /* kernels launched by the last hj_post                                       */
    uint32_t passes;              /* wavefront passes of the last hj_render (each: max_depth x (1 extend + 1 shadow + 1 MIS) traversal launches) */
} hj_stats;

int hj_abi_version(void);

/* Creates a context on CUDA device `device`. Fails (HJ_ERR_CUDA) when no sm_100 device is present:
 * there is no CPU fallback. */
int hj_create(hj_ctx** out, int device);
void hj_destroy(hj_ctx* ctx);
const char* hj_last_error(const hj_ctx* ctx);   /* ctx may be NULL: last hj_create error */

/* Tuning / diagnostic options (none changes a result):
 *   "count_fetches" 0|1, "time_stages" 0|1       counters pass / serial event-pair timing of the stages (hj_stats)
 *   "paths_per_pass" N (0 = auto), "pipelines" 0 (auto: 2, or 8 beyond depth 16) .. 8, "bands" 0 (auto)..64 (pixel bands of a low-spp frame)
 *   "morton_bits" 4..21 per axis (default 10: 30-bit codes, 4 radix passes), "builder" 0 LBVH | 1 PLOC (default), "ploc_tail" 0|1 (last PLOC rounds in one block), "leaf_max" 1..3 (triangles per BVH leaf), "leaf_split" 0|1 (default 1: free
 *   slots of a wide node split multi-triangle leaves), "mis_emitter_first" 0|1, "trace_blocks_per_sm" 0 (auto)..32
 *   "nan_repair" 0|1 — the ONE option that can change a result, off by default: about one path in 1e8 ends in NaN inside the reference's
 *   BSDF code (reproduced on purpose); with 1 such samples are left out of their pixel's mean instead of poisoning it (a frame without
 *   non-finite samples is bit-identical either way)
 *   "verify_accel" 0|1 (also env HJ_VERIFY_ACCEL=1): every hj_accel_build checks the BVH it built (each leaf triangle
 *   inside its slot's box, every primitive in exactly one leaf) and fails with HJ_ERR_INVALID otherwise */
int hj_set_option(hj_ctx* ctx, const char* key, int64_t value);

/* ≙ Renderer::cpySceneDataToDevice + textureBind + setSky + setLUT (renderer.h:197-255, 740-898). */
int hj_scene_upload(hj_ctx* ctx, const hj_scene_desc* scene);

/* ≙ buildGAS+buildIAS (mode BUILD) / updateIASMatrix (renderer.h:257-291).
 * xforms3x4: n_instances row-major 3x4 instance-to-world matrices (Matrix4x3, include/cu/matrix_4x3.h:12);
 * inv_xforms3x4: rows of the inverse (renderer.h:272-284), NULL = computed here. NULL xforms = identity. */
int hj_accel_build(hj_ctx* ctx, const float* xforms3x4, const float* inv_xforms3x4, uint32_t n_instances, int mode);

/* The same build, enqueued on the context's low-priority build stream without a single host synchronisation (PLOC rounds and
 * collapse levels are driven from the device). It fills the second set of acceleration buffers while a frame enqueued by
 * hj_render_async reads the first, i.e. frame N+1's update runs under frame N's shading — what the reference's frame loop does
 * with updateIASMatrix (renderer.h:1126-1133, 257-291) minus its stall. The next hj_render / hj_render_async / hj_trace (or
 * hj_accel_wait) waits for it, validates it and switches to the new buffers; a failed build leaves the previous structure in place
 * and is reported there. The matrices are copied before the call returns. At most one build is in flight: a second call first
 * completes the pending one. */
int hj_accel_build_async(hj_ctx* ctx, const float* xforms3x4, const float* inv_xforms3x4, uint32_t n_instances, int mode);
int hj_accel_wait(hj_ctx* ctx);

/* ≙ cudaMemcpy(Params) + optixLaunch + CUDA_SYNC_CHECK (renderer.h:1229-1242). Synchronous. */
int hj_render(hj_ctx* ctx, const hj_frame_params* frame);

/* Same as hj_render but returns right after enqueueing on the context stream (bench / overlap). */
int hj_render_async(hj_ctx* ctx, const hj_frame_params* frame);
int hj_sync(hj_ctx* ctx);

/* Copies an AOV of the last rendered frame into caller memory; `bytes` must be the exact size. */
int hj_readback(hj_ctx* ctx, int aov, void* dst, size_t bytes);

/* The stage between the colour AOV and the written image — OptixDenoiserManager::denoise() (denoiser.h:143-177) as
 * driven per frame by renderer.h:1256-1283, plus float4ConvertColor (renderer.h:87-101). Reads the COLOR / ALBEDO /
 * NORMAL AOVs of the last frame, writes HJ_AOV_OUTPUT and HJ_AOV_OUTPUT_RGBA8.
 *   HJ_POST_PASSTHROUGH        Render_mode "Default" (blendFactor 1, renderer.h:1113-1115): output == colour AOV, exact
 *   HJ_POST_DENOISE            Render_mode "Denoise" / "Debug" (HDR model): edge-avoiding a-trous filter guided by albedo +
 *                              normal — the stand-in for the OptiX AI network, which has no source (parity unpinned)
 *   HJ_POST_DENOISE_UPSCALE2X  Render_mode "DenoiseUpScale2X" (renderer.h:1096-1099: the frame is rendered at W/2 x H/2):
 *                              the same filter, then a 2x enlargement; output is 2W x 2H of the rendered frame
 * Options (hj_set_option): "denoise_iterations" (0..8, default 5), "denoise_sigma_{color,normal,albedo}_milli" (1/1000; defaults 2000, 100, 50). */
typedef enum hj_post_mode { HJ_POST_PASSTHROUGH = 0, HJ_POST_DENOISE = 1, HJ_POST_DENOISE_UPSCALE2X = 2 } hj_post_mode;
int hj_post(hj_ctx* ctx, int mode);

/* Page-locks / releases caller memory (cudaHostRegister) so that hj_scene_upload / hj_readback DMA straight from / into it
 * instead of going through the driver's staging copies; for hosts without the CUDA headers. The reference keeps its
 * geometry in std::vector and copies it with cudaMemcpy (renderer.h:560-640, pageable): pinning is optional, the result
 * is the same bytes. Pin once, reuse across frames, unpin before freeing the memory. */
int hj_host_pin(hj_ctx* ctx, void* ptr, size_t bytes);
int hj_host_unpin(hj_ctx* ctx, void* ptr);

/* Device address of an AOV buffer (for torch.distributed / zero-copy consumers); NULL if absent. */
void* hj_device_ptr(hj_ctx* ctx, int aov);

/* Re-runs normalise+tonemap+quantise from HJ_AOV_ACCUM (after an external reduction wrote into it). */
int hj_finalize(hj_ctx* ctx);

int hj_get_stats(hj_ctx* ctx, hj_stats* out);

/* Batch ray query against the current acceleration structure — the two optixTrace call sites of the reference
 * (RayTrace rt.h:43-69: closest hit; TraceOcculution rt.h:15-41: any hit) exposed on their own for parity checks.
 * rays: 8 floats each (ox,oy,oz,tmin, dx,dy,dz,tmax), HOST memory. prim[i] = global primitive id of the closest
 * hit (ties on t -> lower id) or 0xffffffff; any_hit: 1 / 0xffffffff. tuv: 3 floats per ray (t,b1,b2), may be NULL. */
int hj_trace(hj_ctx* ctx, const float* rays, uint64_t n, int any_hit, uint32_t* prim, float* tuv);

/* Diagnostic: evaluates one elementary function of csrc/hj_libm.h ON THE DEVICE over n HOST floats (fn 0 sinf, 1 cosf,
 * 2 acosf, 3 logf, 4 expf, 5 powf(a,b), 6 atan2f(a,b); b may be NULL for the unary ones). The shading kernels call these
 * instead of CUDA's libm (the reference calls libm at disneyBRDF.h:30-38,64-80,93-100,131-139, BSDFs.h:551-586, math.h:26-37,
 * color.h:10-29, renderer.h:73-85); the same header compiled for the host gives the same bits (tests/test_gpu_libm.py). */
int hj_libm_eval(hj_ctx* ctx, int fn, const float* a, const float* b, float* out, uint64_t n);

/* hj_trace without any device allocation. NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it). Same arguments, same results
 * (bit for bit) as hj_trace, but the device staging of rays / results and the kernel's work-fetch cursor live in one fixed
 * `static __device__` array inside the library (about 3 MB; every GPU that loads the library holds it until the process exits) instead
 * of buffers the context grows with cudaMalloc. It calls no cudaMalloc, no cudaFree and no allocating library routine.
 *   - at most HJ_TRACE_NOALLOC_MAX_RAYS rays per call; more is refused with HJ_ERR_INVALID and a message (split the batch);
 *   - an acceleration structure must be in place and no hj_accel_build_async may be pending (HJ_ERR_NO_ACCEL otherwise: completing a
 *     pending build may allocate);
 *   - all calls on one GPU share the array: calls that overlap on the GPU (two contexts, streams or threads) race. The call is
 *     synchronous, so calls made one after the other never overlap. The array keeps the previous call's contents; what must start at
 *     zero is cleared by every call. */
#define HJ_TRACE_NOALLOC_MAX_RAYS 65536
int hj_trace_noalloc(hj_ctx* ctx, const float* rays, uint64_t n, int any_hit, uint32_t* prim, float* tuv);

/* Multi-GPU: sample-slice data parallelism, replicated scene, one sum-reduction of the
 * float4 accumulation buffer per frame over NCCL/NVLink. unique_id is the 128-byte ncclUniqueId. */
int hj_comm_unique_id(void* unique_id_128);
int hj_comm_init(hj_ctx* ctx, const void* unique_id_128, int rank, int nranks);
/* hj_scene_upload for the replicated scene of a multi-GPU job, called by every rank of the communicator with the same counts: the five
 * big SceneData arrays (vertices, normals, texcoords, indices, material_ids) cross PCIe ONCE, on `root`, and reach the other ranks by
 * ncclBroadcast over NVLink / NVSwitch; on the other ranks those five pointers are not read (they may be NULL). The small arrays
 * (prim_offsets, materials, textures, lights, sky, LUT) are read from every rank's own description. */
int hj_comm_scene_upload(hj_ctx* ctx, const hj_scene_desc* scene, int root);
/* all-reduces HJ_AOV_ACCUM over the communicator and finalises the colour/RGBA8 AOVs on every rank. */
int hj_comm_reduce(hj_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* HENJOU_B200_H */

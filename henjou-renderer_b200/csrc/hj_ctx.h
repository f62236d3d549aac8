// hj_ctx.h — the context behind the C ABI of include/henjou_b200.h: owns one CUDA stream and every device
// allocation (the role of Renderer's cuh::CUDevicePointer members, /root/reference/include/cu/cuda_buffer.h:18-98
// and include/renderer/renderer.h:138-196).
#pragma once
#include "../../include/henjou_b200.h"
#include "hj_scene.cuh"
#include <string>
#include <vector>
#include <cstdio>

struct DevBuf {
    void* p = nullptr; size_t cap = 0;
    // grow-only; contents are NOT preserved
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T* as() const { return (T*)p; }
};

// Everything a frame reads of the acceleration structure. Two sets: hj_accel_build_async fills the one the renderer is not
// reading (on its own stream, under the previous frame's shading) and the next hj_render flips to it (the
// reference updates its IAS inside the frame loop, renderer.h:257-291,1126-1133).
struct AccelSet {
    DevBuf xf, inv_xf;            // 3 float4 rows per instance and the rows of the inverse
    DevBuf tri_pos, tri_attr;     // world-space flattening (k_flatten)
    DevBuf nodes, btris;          // the traversal structure
    uint32_t n_nodes = 0, depth = 0;
    cudaEvent_t ev_last_read = nullptr;   // end of the last frame that read this set: a build into it waits for that
};
struct BuildStatus;

struct BvhState {
    // binary LBVH (Karras 2012), N leaves = triangles in Morton order
    DevBuf keys, keys_alt, vals, vals_alt, block_hist, digit_total;
    DevBuf blo, bhi;              // float4[2N-1] boxes: internal i -> i, leaf j -> N-1+j
    DevBuf left, right, parent;   // uint32
    DevBuf cnt, flags;            // uint32[N-1]: leaves below an internal node; fit counters
    DevBuf p_cid[2], p_lo[2], p_hi[2], p_nn, p_keep, p_bcount;   // PLOC cluster arrays (ping-pong)
    uint32_t root = 0; int ploc_rounds = 0;
    DevBuf bounds;                // 6 ordered-int atomics
    DevBuf queue_a, queue_b, counters, verr;   // collapse work queues: uint2 {binary node, wide node}
    DevBuf status;                // BuildStatus written by the build kernels; h_status = its pinned host mirror
    BuildStatus* h_status = nullptr;
    float* h_xf = nullptr; size_t h_xf_cap = 0;   // pinned staging of the instance matrices (2 x 12 floats per instance)
    bool have_topology = false;
    // the build in flight (enqueued on build_stream, not yet validated / flipped to)
    int pending_set = -1; int pending_mode = 0; bool pending_topo = false;
    cudaStream_t build_stream = nullptr; cudaEvent_t ev_b0 = nullptr, ev_b1 = nullptr;
    int coop_blocks_ploc = 0, coop_blocks_collapse = 0;
};

#define HJ_MAX_PIPES 8
// One wavefront pipeline: the path state of one pass in flight. Several pipelines run passes concurrently on their
// own streams so that the tail of one persistent traversal kernel overlaps the next pipeline's work.
struct WavefrontState {
    uint32_t capacity = 0;        // path slots
    cudaStream_t stream = nullptr; cudaEvent_t ev_acc = nullptr;
    DevBuf ray_o, ray_d, hit, thr, L, rng_depth, dbg;   // per-slot state (SoA)
    DevBuf q_extend[2], q_shadow, q_mis;                 // queues
    DevBuf q_cls[3], q_ns, sf[4];                        // shade stages: per-class queues, non-specular queue, surface records
    DevBuf sh_o, sh_d, sh_c;      // shadow ray payload per slot: origin+tmax, dir, contribution
    DevBuf q_sky, sk_d, sk_c;     // sky-sample shadow rays (frame parameter sky_sampling): queue, direction, contribution
    DevBuf mi_o, mi_d, mi_a, mi_t;   // MIS ray payload per slot
    DevBuf counters;              // device-side queue counters + ray statistics
    DevBuf* all[31] = {&q_sky, &sk_d, &sk_c, &ray_o, &ray_d, &hit, &thr, &L, &rng_depth, &dbg, &q_extend[0], &q_extend[1], &q_shadow, &q_mis,
                       &sh_o, &sh_d, &sh_c, &mi_o, &mi_d, &mi_a, &mi_t, &counters, &q_cls[0], &q_cls[1], &q_cls[2], &q_ns,
                       &sf[0], &sf[1], &sf[2], &sf[3], nullptr};
};

struct NcclApi;

struct hj_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_scene = nullptr;   // ev_scene: end of the last hj_scene_upload on `stream`
    std::string err;
    // options
    int opt_count_fetches = 0, opt_time_stages = 0, opt_leaf_max = 3, opt_trace_blocks_per_sm = 0, opt_pipelines = 0, opt_builder = 1, opt_mis_emitter_first = 1, opt_leaf_split = 1, opt_verify_accel = 0, opt_bands = 0, opt_ploc_tail = 1;   // builder: 0 Karras radix tree (LBVH), 1 PLOC
    int64_t opt_paths_per_pass = 0;
    // Morton bits per axis. 10 = the classic 30-bit code (Karras 2012): 4 sort passes instead of 8. r02 sweep 8..21 on C1/C2/C3/C5: the tree
    // quality moves by +-2 % with the tie structure and no setting wins everywhere (C2: 10 bits 370.6 ms, 21 bits 379.5; C5: 564.4 vs 560.0;
    // C3: 525.2 vs 526.2); PLOC searches a neighbourhood, ties keep the mesh order (stable sort), so the resolution is not what limits it.
    int opt_morton_bits = 10;
    int opt_nan_repair = 0;       // 0 (default, parity): a non-finite sample poisons its pixel as in the reference; 1: it is left out of the mean
    DevBuf trace_scratch[4];      // hj_trace: rays, primitive ids, (t,b1,b2), queue cursor
    // scene
    bool have_scene = false, have_accel = false, have_frame = false;
    DevScene S{};                 // device pointers (host copy of the struct, passed to kernels by value)
    DevBuf d_vertices, d_normals, d_texcoords, d_indices, d_material_ids, d_prim_offsets, d_tri_inst;
    DevBuf d_materials, d_textures, d_texels, d_srgb, d_light_ids, d_light_em, d_ibl, d_lut, d_sky_rows, d_sky_marg;
    AccelSet accel[2]; int accel_cur = 0;      // S points into accel[accel_cur]
    BvhState bvh;
    WavefrontState wf[HJ_MAX_PIPES];
    int n_pipes_used = 0;
    cudaEvent_t ev_frame = nullptr;
    // frame
    hj_frame_params fp{};
    DevBuf aov_color, aov_albedo, aov_normal, aov_rgba8, aov_accum, aov_path_rad, aov_path_ids;
    uint32_t fb_w = 0, fb_h = 0;
    // post stage (hj_post.cu): output plane of the last hj_post, a-trous ping-pong planes
    DevBuf aov_output, aov_output_rgba8, post_tmp[2];
    uint32_t out_w = 0, out_h = 0; bool have_output = false;
    cudaEvent_t ev_post0 = nullptr, ev_post1 = nullptr;
    int opt_denoise_iterations = 5, opt_denoise_sigma_color_milli = 2000, opt_denoise_sigma_normal_milli = 100, opt_denoise_sigma_albedo_milli = 50;   // r01-v sweep, tools/post_bench.py
    hj_stats stats{};
    // option "time_stages": event pairs around every traversal (kind 0) and shading (kind 1) launch of the last frame
    std::vector<cudaEvent_t> ev_pool; std::vector<int> ev_kind;
    // multi-GPU
    NcclApi* nccl = nullptr; void* comm = nullptr; int rank = 0, nranks = 1;
    bool frame_reduced = false;   // hj_comm_reduce already summed this frame's accumulation buffer (a second call only re-finalizes)
};

#define HJ_CUDA(ctx, call)                                                                         \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            char b_[512];                                                                          \
            snprintf(b_, sizeof b_, "%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            (ctx)->err = b_;                                                                       \
            (void)cudaGetLastError(); /* a recoverable error must not fail the next call's launch check */ \
            return (e_ == cudaErrorMemoryAllocation) ? HJ_ERR_OOM : HJ_ERR_CUDA;                   \
        }                                                                                          \
    } while (0)

// hj_bvh.cu: enqueue a build of accel[set] on the build stream (no host synchronisation); finish = wait, validate, flip
int hj_bvh_enqueue(hj_ctx* c, const float* xf3x4, const float* inv3x4, int mode);
int hj_bvh_finish(hj_ctx* c);
void hj_bvh_release(hj_ctx* c);
// hj_wavefront.cu
int hj_wavefront_render(hj_ctx* c, const hj_frame_params* fp);
int hj_wavefront_finalize(hj_ctx* c);
// hj_post.cu
int hj_post_run(hj_ctx* c, int mode);
int hj_wavefront_stats(hj_ctx* c);
int hj_trace_batch(hj_ctx* c, const float* rays_host, uint64_t n, int any_hit, uint32_t* prim_host, float* tuv_host);

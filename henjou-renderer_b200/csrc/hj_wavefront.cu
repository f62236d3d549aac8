// hj_wavefront.cu — wavefront path integrator: the B200 replacement for the reference's megakernel
// (__raygen__rg + __closesthit__ch + __miss__ms + __closesthit__shadow launched by optixLaunch(W,H,1),
// /root/reference/include/renderer/renderer.h:1241) and its integrators Pathtrace / NEE / MIS
// (include/kernel/rt.h:85-159, 162-281, 284-440).
//
// One path per slot; slot = s_local * W*H + pixel, so the RNG stream of a path is CMJState{n_spp = global sample
// index, scramble = frame, image_idx = pixel} regardless of how samples are sliced over passes or GPUs.
// Per bounce:   k_trace<0> (closest hit)  ->  k_surface (closest-hit / miss programs, sort into BSDF-class queues)  ->
// k_sample<class> (light sample + BSDF evaluation -> shadow queue; MIS sample -> MIS queue; continuation sample, Russian
// roulette -> next-bounce queue; warp-aggregated pushes)  ->  k_trace<1> (any hit)  ->  k_trace<3> (MIS emitter ray). Queue
// lengths live in device memory and every kernel is a grid-stride loop over its queue, so a frame is enqueued without a single
// host synchronisation.
// The draw order of every RNG dimension and the order of the float additions into the radiance follow the
// reference loop exactly; rays whose contribution is exactly zero are not traced (they cannot change the result).
#include "hj_ctx.h"
#include <cstdlib>
#include <mutex>
#include "hj_bsdf.cuh"
#include "hj_trace.cuh"

struct WfCounters {
    uint32_t n_extend[2]; uint32_t n_shadow, n_mis;
    uint32_t n_cls[3], n_ns;      // shade-stage queues: Disney / EC-GGX / glass surfaces, and all non-specular ones
    uint32_t n_sky;               // sky-sample shadow rays (sky_sampling)
    uint32_t next[4];             // work-fetch cursors of the persistent traversal kernels (extend, shadow, mis, sky)
    unsigned long long rays_extend, rays_shadow, nodes, tris;
    uint32_t max_nodes, max_tris;
    unsigned long long diag[4][8];   // count_fetches: lane-state census of the traversal loop per ray type
};

struct FrameK {
    uint32_t W, H, WH, spp, frame, sample_begin;
    uint32_t pix0, npix;      // this pass covers pixels [pix0, pix0 + npix) of its samples: slot = s_local * npix + (pixel - pix0)
    float3 cpos, cdir, cup, cright; float cf, ibl;
    int integrator, max_depth, tonemap, jitter, debug;
    int nan_repair;           // option nan_repair: non-finite samples are left out of the pixel's sum and of its sample count
    int sky;                  // hj_frame_params::sky_sampling (MIS only): light-sampled sky with MIS, an extension over the reference
};

struct WfPtrs {
    float4 *ray_o, *ray_d, *hit, *thr, *L; uint32_t* rng; uint2* dbg;
    uint32_t *q_ext[2], *q_shadow, *q_mis, *q_cls[3], *q_ns;
    float4 *sf0, *sf1, *sf2, *sf3;   // surface record per slot: (pos, roughness) (normal, metallic) (basecolor, sheen) (clearcoat, ior, flags, -)
    float4 *sh_o, *sh_d, *sh_c, *mi_o, *mi_d, *mi_a, *mi_t;
    uint32_t* q_sky; float4 *sk_d, *sk_c;
    WfCounters* cnt;
    float4 *aov_albedo, *aov_normal;
};

#define WF_THREADS 128

// Path state is touched once per stage and never reused from cache: with HJ_STREAM_STATE the traversal kernels read and write it with
// the evict-first hints (ld.global.cs / st.global.cs) so that it does not push BVH nodes and triangles out of L1 / L2.
#ifndef HJ_STREAM_STATE
#define HJ_STREAM_STATE 1      // r02: C2 380.4 -> 378.0 ms (three runs each), C3 / C5 unchanged
#endif
#if HJ_STREAM_STATE
#define LD_STATE(p) __ldcs(p)
#define ST_STATE(p, v) __stcs((p), (v))
#else
#define LD_STATE(p) (*(p))
#define ST_STATE(p, v) (*(p) = (v))
#endif

// warp-aggregated queue push: one atomic per warp; must be reached by all 32 lanes
HJ_DI void queue_push(uint32_t* counter, uint32_t* queue, bool pred, uint32_t value) {
    const uint32_t m = __ballot_sync(0xffffffffu, pred);
    if (!m) return;
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t base = 0;
    if (lane == (uint32_t)(__ffs(m) - 1)) base = atomicAdd(counter, (uint32_t)__popc(m));
    base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
    if (pred) queue[base + __popc(m & ((1u << lane) - 1u))] = value;
}
HJ_DI void warp_add64(unsigned long long* p, uint32_t v) {
    v = __reduce_add_sync(0xffffffffu, v);
    if ((threadIdx.x & 31u) == 0 && v) atomicAdd(p, (unsigned long long)v);
}

// the Russian-roulette step that opens every loop iteration (rt.h:96-102 / 173-179 / 295-301)
HJ_DI bool roulette(float3& thr, Sampler& st) {
    const float rp = fmaxf(thr.x, fmaxf(thr.y, thr.z));
    if (rp < draw1(st)) return false;
    thr /= rp;
    return true;
}

// ---------------------------------------------------------------------------------------------- raygen
// Reconstructed __raygen__rg: camera model of the stale PTX computeCameraRay
// (lib/ptx/HenjouRenderer_generated_henjouRendererCU.cu.ptx:49-104): u = (2x-W)/H, v = (2y-H)/H,
// d = normalize(camera_f*dir + u*right + v*up); D1: the first CMJ draw of a sample jitters the pixel.
__global__ void __launch_bounds__(WF_THREADS) k_raygen(FrameK F, WfPtrs P, uint32_t s0, uint32_t n_slots) {
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x; slot < n_slots; slot += stride) {
        const uint32_t pix = F.pix0 + slot % F.npix, s = s0 + slot / F.npix;
        const uint32_t x = pix % F.W, y = pix / F.W;
        Sampler st; st.n_spp = s; st.scramble = F.frame; st.depth = 0; st.image_idx = pix;
        float jx = 0.0f, jy = 0.0f;
        if (F.jitter) { const float2 xi = draw2(st); jx = xi.x; jy = xi.y; }
        const float u = (2.0f * ((float)x + jx) - (float)F.W) / (float)F.H;
        const float v = (2.0f * ((float)y + jy) - (float)F.H) / (float)F.H;
        const float3 d = normalize(F.cf * F.cdir + u * F.cright + v * F.cup);
        float3 thr = f3(1.0f);
        const bool alive = F.max_depth > 0 && roulette(thr, st);
        P.ray_o[slot] = f4(F.cpos, 0.0f); P.ray_d[slot] = f4(d, 0.0f);
        P.thr[slot] = f4(thr, 0.0f); P.L[slot] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        P.rng[slot] = st.depth;
        if (F.debug) P.dbg[slot] = make_uint2(0xffffffffu, 0x811c9dc5u);
        if (s == F.sample_begin) { P.aov_albedo[pix] = make_float4(0, 0, 0, 1); P.aov_normal[pix] = make_float4(0, 0, 0, 1); }   // guides: first sample of this context's slice (sample 0 on one GPU)
        // every first-iteration roulette passes (throughput 1 > any draw), so the queue is the identity
        P.q_ext[0][slot] = alive ? slot : 0xffffffffu;
    }
}

// ---------------------------------------------------------------------------------------------- light sampling
// light_sample.h:9-75: uniform emissive triangle, binary search of the owning instance, uniform barycentrics,
// world-space area pdf. Two RNG calls; none when the scene has no emissive triangle (:12-16).
HJ_DI float3 sample_light(const DevScene& S, Sampler& st, float& pdf, float3& normal, float3& emission) {
    const int count = (int)S.n_lights;
    float p = draw1(st);
    int index = (int)(p * count);
    if (index == count) index--;
    const unsigned int prim = S.light_prim_ids[index];
    unsigned int left = 0U, right = S.n_instances - 1, middle = (left + right) / 2U;
    while (left <= right) {
        if (S.prim_offsets[middle] <= prim) left = middle + 1;
        else { if (middle == 0) { break; } right = middle - 1; }
        middle = (left + right) / 2;
    }
    const float select_pdf = 1.0f / count;
    const unsigned int inst = middle;
    const float4* xf = S.xf + 3 * (size_t)inst; const float4* ixf = S.inv_xf + 3 * (size_t)inst;
    const unsigned i0 = S.indices[prim * 3 + 0], i1 = S.indices[prim * 3 + 1], i2 = S.indices[prim * 3 + 2];
    const float3 v0 = xform_point(xf, ld3(S.vertices, i0)), v1 = xform_point(xf, ld3(S.vertices, i1)), v2 = xform_point(xf, ld3(S.vertices, i2));
    const float3 n0 = xform_normal(ixf, ld3(S.normals, i0)), n1 = xform_normal(ixf, ld3(S.normals, i1)), n2 = xform_normal(ixf, ld3(S.normals, i2));
    const float area = length(cross(v1 - v0, v2 - v0)) * 0.5f;
    const float2 xi = draw2(st);
    const float f1 = 1.0f - sqrtf(xi.x);
    const float f2 = sqrtf(xi.x) * (1.0f - xi.y);
    const float f3_ = sqrtf(xi.x) * xi.y;
    const float3 pos = v0 * f1 + v1 * f2 + v2 * f3_;
    const float3 nrm = normalize(n0 * f1 + n1 * f2 + n2 * f3_);
    pdf = 1.0 / area;           // double division, rounded to float (light_sample.h:69)
    pdf *= select_pdf;
    normal = nrm;
    emission = ld3(S.light_prim_emission, (uint32_t)index);
    return pos;
}
// light_sample.h:77-92 (the world-space vertices are the flattened ones: same transform_position results)
HJ_DI float light_pdf_area(const DevScene& S, uint32_t prim) {
    const float4* P = S.tri_pos + 3 * (size_t)prim;
    const float3 v0 = f3(P[0]), v1 = f3(P[1]), v2 = f3(P[2]);
    const float area = length(cross(v1 - v0, v2 - v0)) * 0.5f;
    return 1.0f / (area * (int)S.n_lights);
}

// ---------------------------------------------------------------------------------------------- shade (staged)
// One loop iteration of Pathtrace / NEE / MIS after the closest-hit query (rt.h:104-156 / 181-278 / 303-438), followed
// by the Russian roulette that opens the NEXT iteration, as three small kernels instead of one 82 KB one: a monolithic
// k_shade ran at 0.32 IPC with instruction-cache misses as its top stall and 15 of 32 lanes (r01-g) — straight-line code
// executed once per path has no reuse in the 6 KB L0 / 32 KB L1.5 instruction caches. Stages:
//   k_surface          closest-hit / miss programs: sky, emitter hits, AOVs, surface record, sort into class queues
//   k_sample<INTEG,C>  light sample + BSDF evaluation -> shadow ray (non-specular classes; a kernel of its own, k_nee, until r02);
//                      MIS BSDF sample -> MIS ray; dead draw; continuation sample; roulette   (one class C per launch)
// Every stage derives its RNG dimension from the counter stored at the start of the iteration, so the draw order of
// SURVEY §8 "draw ledger" is unchanged: [light pick, light barycentrics] [MIS sample] [dead] [continuation] [roulette].
#ifndef HJ_SHADE_MINB
#define HJ_SHADE_MINB 8       // 64 registers. r02 (after the shared double-precision libm): C2 379.3 -> 376.1 ms against 6 blocks / 80 registers
#endif
#ifndef HJ_SURF_MINB
#define HJ_SURF_MINB HJ_SHADE_MINB
#endif
#ifndef HJ_NEE_MINB
#define HJ_NEE_MINB HJ_SHADE_MINB
#endif
#define HJ_CLS_DISNEY 0
#define HJ_CLS_GGX 1
#define HJ_CLS_GLASS 2

HJ_DI Surface load_surface(const WfPtrs& P, uint32_t slot) {
    const float4 r0 = P.sf0[slot], r1 = P.sf1[slot], r2 = P.sf2[slot], r3 = P.sf3[slot];
    Surface s;
    s.position = f3(r0); s.roughness = r0.w; s.normal = f3(r1); s.metallic = r1.w; s.basecolor = f3(r2); s.sheen = r2.w;
    s.clearcoat = r3.x; s.ior = r3.y;
    const uint32_t fl = __float_as_uint(r3.z);
    s.is_specular = (fl & 1u) != 0; s.is_thinfilm = (fl & 2u) != 0; s.is_light = false; s.emission = f3(0.0f);
    return s;
}

template <int INTEG>
__global__ void __launch_bounds__(WF_THREADS, HJ_SURF_MINB) k_surface(DevScene S, FrameK F, WfPtrs P, int cur, int depth, uint32_t s0) {
    const uint32_t n = P.cnt->n_extend[cur];
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; (i & ~31u) < n; i += stride) {
        const uint32_t slot = (i < n) ? P.q_ext[cur][i] : 0xffffffffu;
        // (r02: starting the NEXT iteration's gather chain with prefetch.global.L2 here was measured: shading 127 -> 132 ms)
        int cls = -1;
        if (slot != 0xffffffffu) {
            const float4 hq = P.hit[slot];
            const uint32_t prim = __float_as_uint(hq.w);
            const uint32_t pix = F.pix0 + slot % F.npix;
            const bool first_sample = (depth == 0 && s0 + slot / F.npix == F.sample_begin);
            if (prim == 0xffffffffu) {                               // __miss__ms (D6); AOVs stay 0 (raygen wrote them)
                if (INTEG == HJ_INTEGRATOR_PT || depth == 0) {
                    const float3 em = sky_radiance(S, f3(P.ray_d[slot]), F.ibl);
                    const float4 Lq = P.L[slot];
                    P.L[slot] = f4(f3(Lq) + f3(P.thr[slot]) * em, Lq.w);
                }
            } else {
                const Surface prd = make_surface(S, prim, hq.y, hq.z);
                if (first_sample) { P.aov_albedo[pix] = f4(prd.basecolor, 1.0f); P.aov_normal[pix] = f4(prd.normal, 1.0f); }   // rt.h:113-116
                if (prd.is_light) {
                    if (INTEG == HJ_INTEGRATOR_PT || depth == 0) {
                        const float4 Lq = P.L[slot];
                        P.L[slot] = f4(f3(Lq) + f3(P.thr[slot]) * prd.emission, Lq.w);
                    }
                } else {
                    P.sf0[slot] = f4(prd.position, prd.roughness); P.sf1[slot] = f4(prd.normal, prd.metallic);
                    P.sf2[slot] = f4(prd.basecolor, prd.sheen);
                    P.sf3[slot] = make_float4(prd.clearcoat, prd.ior, __uint_as_float((prd.is_specular ? 1u : 0u) | (prd.is_thinfilm ? 2u : 0u)), 0.0f);
                    cls = prd.is_specular ? HJ_CLS_GLASS : (prd.metallic > 0.5 ? HJ_CLS_GGX : HJ_CLS_DISNEY);   // BSDFs.h:995-1030
                }
            }
        }
        queue_push(&P.cnt->n_cls[HJ_CLS_DISNEY], P.q_cls[HJ_CLS_DISNEY], cls == HJ_CLS_DISNEY, slot);
        queue_push(&P.cnt->n_cls[HJ_CLS_GGX], P.q_cls[HJ_CLS_GGX], cls == HJ_CLS_GGX, slot);
        queue_push(&P.cnt->n_cls[HJ_CLS_GLASS], P.q_cls[HJ_CLS_GLASS], cls == HJ_CLS_GLASS, slot);
    }
}

// Light sampling + BSDF evaluation at a non-specular surface (rt.h:219-260 / 341-380): the first part of a loop iteration after the
// closest hit. Specular surfaces evaluate to zero (BSDFs.h:1006-1013), so their light sample contributes nothing and only moves the
// RNG counter. r02: this ran as a kernel of its own (k_nee) between k_surface and k_sample, re-reading the 64-B surface record, the ray
// and the throughput that k_sample reads again; it is now the head of k_sample (same draws, same arithmetic, one pass over the state).
template <int INTEG>
HJ_DI void nee_stage(const DevScene& S, const FrameK& F, const WfPtrs& P, uint32_t slot, const Surface& prd, const float3& thr, const float3& lwo,
                     const float3& t, const float3& nrm, const float3& b, Sampler& st, bool& push_shadow, bool& push_sky) {
    Disney bsdf; bsdf.init(prd);
    if (S.n_lights >= 1) {
        float lpdf; float3 lnrm, lem;
        const float3 lpos = sample_light(S, st, lpdf, lnrm, lem);
        float3 ld = lpos - prd.position;
        const float dist = length(ld);
        ld = normalize(ld);
        const float c1 = absdot(nrm, ld), c2 = absdot(lnrm, -ld);
        const float3 lwi = to_local(ld, t, nrm, b);
        const float3 f = bsdf.eval(S, lwo, lwi);
        const float G = c2 / (dist * dist);
        float3 contrib;
        if (INTEG == HJ_INTEGRATOR_MIS) {
            const float ppdf = bsdf.pdf(lwo, lwi) * G;
            const float w = lpdf / (lpdf + ppdf);
            contrib = thr * (f * G * c1 / lpdf) * w * lem;
        } else contrib = thr * (f * G * c1 / lpdf) * lem;
        if (!is_zero3(contrib)) {
            push_shadow = true;
            P.sh_o[slot] = f4(prd.position, dist - 0.001f); P.sh_d[slot] = f4(ld, 0.0f); P.sh_c[slot] = f4(contrib, 0.0f);
        }
    }
    if (INTEG == HJ_INTEGRATOR_MIS && F.sky) {
        // extension (sky_sampling): one direction drawn from the sky's luminance table, balance heuristic against the BSDF
        // pdf; the matching weight on BSDF-sampled rays that reach the sky is applied in finish_mis
        const float2 xi = draw2(st);
        float spdf;
        const float3 sd = sky_sample(S, xi, spdf);
        if (spdf > 0.0f) {
            const float c1 = absdot(nrm, sd);
            const float3 lwi = to_local(sd, t, nrm, b);
            const float3 f = bsdf.eval(S, lwo, lwi);
            const float w = spdf / (spdf + bsdf.pdf(lwo, lwi));
            const float3 contrib = thr * (f * c1 / spdf) * w * sky_radiance(S, sd, F.ibl);
            if (!is_zero3(contrib)) { push_sky = true; P.sk_d[slot] = f4(sd, 0.0f); P.sk_c[slot] = f4(contrib, 0.0f); }
        }
    }
}

// BSDF sampling of one surface class: [MIS only] a sample towards emitters (rt.h:383-420) -> MIS ray; the dead draw
// (rt.h:266 / 426); the continuation sample (rt.h:141 / 269 / 429), throughput update and the roulette of the next
// iteration. The two samples run through the same code (a 2-trip loop, not unrolled).
template <int CLS>
HJ_DI float3 sample_class(const DevScene& S, const Surface& prd, const float3& lwo, float3& lwi, float& pdf, Sampler& st) {
    if (CLS == HJ_CLS_GLASS) return metaglass_sample(f3(1.0f), prd.ior, lwo, lwi, pdf, st);
    if (CLS == HJ_CLS_GGX) { MultiScatterGGX g; g.init(prd.basecolor, prd.roughness); return g.sample(lwo, lwi, st, pdf); }
    Disney dz; dz.init(prd);
    return dz.sample(S, lwo, lwi, pdf, st);
}

// Tried in r02 and dropped: the two samples of a surface on two ADJACENT LANES (lane 2j the MIS sample, lane 2j+1 the continuation)
// so that a warp fetches each instruction of the 50 KB sampling code once for 16 surfaces x 2 samples — instruction-cache misses are
// this kernel's top stall. Same bits, but no gain: serial shading 124.5 -> 130 ms per C2 frame (twice the queue reads and surface-record
// loads, and the two lanes' epilogues diverge).
// HJ_SAMPLE_SYNC (r02): blocks of HJ_SAMPLE_THREADS threads whose warps pass through the three phases of an iteration (light sample /
// MIS sample / continuation sample) TOGETHER, separated by __syncthreads(), so that the 16 warps of a block fetch the same region of the
// 88 KB of code at the same time: instruction-cache misses were this kernel's top stall (ncu no_instruction 5.0 cycles per issue at 0.38
// IPC). Same arithmetic, same bits. Serial shading per C2 frame: 127.0 ms without -> 121.4 (128 threads) / 116.5 (256) / 113.6 (512) /
// 117.3 (1024); frame 361.4 -> 346.3 ms, C5 549 -> 536, C3 514 -> 512. The block size alone does nothing (256 threads, no barriers:
// 126.6); other block shapes with barriers: 512 x 1 block (128 registers, no spills) 125.4, 768 x 1 (80 registers) 122.6, 384 x 2 (80) 117.7.
// ncu after the change: no_instruction 5.0 -> 0.3-0.9 cycles per issue, barrier 1.9-4.1 (profiles/r02_ncu_shading_c2.txt). Also measured and dropped: one more barrier between the state load and the light sample (113.7), and a barrier between
// choosing the direction and evaluating the BSDF inside each sample (119.9: the extra live state and barrier waits cost more).
#ifndef HJ_SAMPLE_SYNC
#define HJ_SAMPLE_SYNC 1
#endif
#ifndef HJ_SAMPLE_THREADS
#define HJ_SAMPLE_THREADS 512
#endif
#ifndef HJ_SAMPLE_MINB
#define HJ_SAMPLE_MINB ((HJ_SHADE_MINB * WF_THREADS) / HJ_SAMPLE_THREADS)
#endif
#if HJ_SAMPLE_SYNC
template <int INTEG, int CLS>
__global__ void __launch_bounds__(HJ_SAMPLE_THREADS, HJ_SAMPLE_MINB) k_sample(DevScene S, FrameK F, WfPtrs P, int cur, int depth, uint32_t s0) {
    const uint32_t n = P.cnt->n_cls[CLS];
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; (i - threadIdx.x) < n; i += stride) {      // block-uniform trip count
        const uint32_t slot = (i < n) ? P.q_cls[CLS][i] : 0xffffffffu;
        const bool valid = slot != 0xffffffffu;
        bool push_mis = false, push_next = false, push_shadow = false, push_sky = false;
        Surface prd{}; float3 thr = f3(0.0f), t = f3(0.0f), b = f3(0.0f), nrm = f3(0.0f), lwo = f3(0.0f);
        Sampler st; st.n_spp = 0; st.scramble = F.frame; st.image_idx = 0; st.depth = 0;
        if (valid) {
            prd = load_surface(P, slot);
            const float3 d = f3(P.ray_d[slot]);
            thr = f3(P.thr[slot]);
            st.n_spp = s0 + slot / F.npix; st.image_idx = F.pix0 + slot % F.npix; st.depth = P.rng[slot];
            nrm = prd.normal;
            onb(nrm, t, b);
            lwo = to_local(-d, t, nrm, b);
            if (INTEG != HJ_INTEGRATOR_PT && CLS != HJ_CLS_GLASS) nee_stage<INTEG>(S, F, P, slot, prd, thr, lwo, t, nrm, b, st, push_shadow, push_sky);
            else {
                if (INTEG != HJ_INTEGRATOR_PT && S.n_lights >= 1) st.depth += 2;
                if (INTEG == HJ_INTEGRATOR_MIS && F.sky) st.depth += 1;
            }
            if (INTEG == HJ_INTEGRATOR_NEE) draw_skip(st);
        }
#pragma unroll 1
        for (int k = (INTEG == HJ_INTEGRATOR_MIS ? 0 : 1); k < 2; k++) {
            __syncthreads();
            if (valid) {
                float pdf = 1.0f;
                float3 lwi = f3(0.0f, 1.0f, 0.0f);
                const float3 f = sample_class<CLS>(S, prd, lwo, lwi, pdf, st);
                const float3 wi = to_world(lwi, t, nrm, b);
                if (k == 0) {
                    const float c1 = absdot(wi, nrm);
                    if (!is_zero3(f * c1)) {
                        push_mis = true;
                        P.mi_o[slot] = f4(prd.position, c1); P.mi_d[slot] = f4(wi, pdf);
                        P.mi_a[slot] = f4(f, prd.is_specular ? 1.0f : 0.0f); P.mi_t[slot] = f4(thr, 0.0f);
                    }
                    draw_skip(st);
                } else {
                    thr *= f * fabsf(dot(wi, nrm)) / pdf;
                    if (depth + 1 < F.max_depth && roulette(thr, st)) {
                        push_next = true;
                        P.ray_o[slot] = f4(prd.position, 0.0f); P.ray_d[slot] = f4(wi, 0.0f);
                        P.thr[slot] = f4(thr, 0.0f); P.rng[slot] = st.depth;
                    }
                }
            }
        }
        if (INTEG != HJ_INTEGRATOR_PT && CLS != HJ_CLS_GLASS) queue_push(&P.cnt->n_shadow, P.q_shadow, push_shadow, slot);
        if (INTEG == HJ_INTEGRATOR_MIS && CLS != HJ_CLS_GLASS && F.sky) queue_push(&P.cnt->n_sky, P.q_sky, push_sky, slot);
        if (INTEG == HJ_INTEGRATOR_MIS) queue_push(&P.cnt->n_mis, P.q_mis, push_mis, slot);
        queue_push(&P.cnt->n_extend[cur ^ 1], P.q_ext[cur ^ 1], push_next, slot);
        __syncthreads();
    }
}
#else
template <int INTEG, int CLS>
__global__ void __launch_bounds__(HJ_SAMPLE_THREADS, HJ_SAMPLE_MINB) k_sample(DevScene S, FrameK F, WfPtrs P, int cur, int depth, uint32_t s0) {
    const uint32_t n = P.cnt->n_cls[CLS];
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; (i & ~31u) < n; i += stride) {
        const uint32_t slot = (i < n) ? P.q_cls[CLS][i] : 0xffffffffu;
        bool push_mis = false, push_next = false, push_shadow = false, push_sky = false;
        if (slot != 0xffffffffu) {
            const Surface prd = load_surface(P, slot);
            const float3 d = f3(P.ray_d[slot]);
            float3 thr = f3(P.thr[slot]);
            Sampler st; st.n_spp = s0 + slot / F.npix; st.scramble = F.frame; st.image_idx = F.pix0 + slot % F.npix; st.depth = P.rng[slot];
            float3 t, b; const float3 nrm = prd.normal;
            onb(nrm, t, b);
            const float3 lwo = to_local(-d, t, nrm, b);
            if (INTEG != HJ_INTEGRATOR_PT && CLS != HJ_CLS_GLASS) nee_stage<INTEG>(S, F, P, slot, prd, thr, lwo, t, nrm, b, st, push_shadow, push_sky);
            else {                                                               // specular: the draws are made, their values never used
                if (INTEG != HJ_INTEGRATOR_PT && S.n_lights >= 1) st.depth += 2;   // light pick + barycentrics (light_sample.h:18,60)
                if (INTEG == HJ_INTEGRATOR_MIS && F.sky) st.depth += 1;             // the sky sample's draw
            }
            if (INTEG == HJ_INTEGRATOR_NEE) draw_skip(st);                      // dead draw, rt.h:266
#pragma unroll 1
            for (int k = (INTEG == HJ_INTEGRATOR_MIS ? 0 : 1); k < 2; k++) {
                float pdf = 1.0f;                   // defined (reference: uninitialised in MIS)
                float3 lwi = f3(0.0f, 1.0f, 0.0f);  // defined (reference: uninitialised in MIS)
                const float3 f = sample_class<CLS>(S, prd, lwo, lwi, pdf, st);
                const float3 wi = to_world(lwi, t, nrm, b);
                if (k == 0) {
                    const float c1 = absdot(wi, nrm);
                    if (!is_zero3(f * c1)) {
                        push_mis = true;
                        P.mi_o[slot] = f4(prd.position, c1); P.mi_d[slot] = f4(wi, pdf);
                        P.mi_a[slot] = f4(f, prd.is_specular ? 1.0f : 0.0f); P.mi_t[slot] = f4(thr, 0.0f);
                    }
                    draw_skip(st);                                               // dead draw, rt.h:426
                } else {
                    thr *= f * fabsf(dot(wi, nrm)) / pdf;
                    if (depth + 1 < F.max_depth && roulette(thr, st)) {          // opens iteration depth+1
                        push_next = true;
                        P.ray_o[slot] = f4(prd.position, 0.0f); P.ray_d[slot] = f4(wi, 0.0f);
                        P.thr[slot] = f4(thr, 0.0f); P.rng[slot] = st.depth;
                    }
                }
            }
        }
        if (INTEG != HJ_INTEGRATOR_PT && CLS != HJ_CLS_GLASS) queue_push(&P.cnt->n_shadow, P.q_shadow, push_shadow, slot);
        if (INTEG == HJ_INTEGRATOR_MIS && CLS != HJ_CLS_GLASS && F.sky) queue_push(&P.cnt->n_sky, P.q_sky, push_sky, slot);
        if (INTEG == HJ_INTEGRATOR_MIS) queue_push(&P.cnt->n_mis, P.q_mis, push_mis, slot);
        queue_push(&P.cnt->n_extend[cur ^ 1], P.q_ext[cur ^ 1], push_next, slot);
    }
}
#endif
// ---------------------------------------------------------------------------------------------- traversal kernels
// One persistent kernel per ray type (the reference's two optixTrace call sites, rt.h:53 RayTrace and rt.h:25
// TraceOcculution, plus the MIS emitter ray rt.h:389). Warps fetch rays from the queue dynamically: a lane whose ray
// has finished idles only until the warp has lost LOST_MAX lane-iterations, then the warp refills every idle lane
// with one warp-aggregated atomic — this keeps the SIMD lanes busy although ray lengths differ by 10x.
#ifndef HJ_MINB
#define HJ_MINB 8      // resident blocks per SM of the traversal kernels (64 registers)
#endif

HJ_DI void finish_extend(const FrameK& F, const WfPtrs& P, uint32_t slot, const Hit& h) {
    ST_STATE(&P.hit[slot], make_float4(h.t, h.b1, h.b2, __uint_as_float(h.prim)));
    if (F.debug) {                         // L.w = rays traced by this path: only the debug records read it
        float4 L = P.L[slot];
        uint2 g = P.dbg[slot];
        if (L.w == 0.0f) g.x = h.prim;     // first ray of the path
        g.y = (g.y ^ (h.prim + 1u)) * 0x01000193u;
        P.dbg[slot] = g;
        L.w += 1.0f;
        P.L[slot] = L;
    }
}
// TraceOcculution (rt.h:15-41): any hit in (1e-3, dist-1e-3) discards the light sample
HJ_DI void finish_shadow(const FrameK& F, const WfPtrs& P, uint32_t slot, const Hit& h) {
    if (h.prim != 0xffffffffu && !F.debug) return;         // occluded: nothing to add (the ray count is a debug record)
    float4 L = LD_STATE(&P.L[slot]);
    if (h.prim == 0xffffffffu) { const float4 c = LD_STATE(&P.sh_c[slot]); L.x += c.x; L.y += c.y; L.z += c.z; }
    L.w += 1.0f;
    ST_STATE(&P.L[slot], L);
}
// the extra BSDF-sampled ray of MIS (rt.h:383-420): emitter hit weighted by the balance heuristic, sky unweighted
HJ_DI void finish_mis(const DevScene& S, const FrameK& F, const WfPtrs& P, uint32_t slot, const Hit& h) {
    if (h.prim == HJ_HIT_BLOCKED && !F.debug) return;      // blocked by a non-emitter: no contribution, nothing to load
    const float4 oq = LD_STATE(&P.mi_o[slot]), dq = LD_STATE(&P.mi_d[slot]), aq = LD_STATE(&P.mi_a[slot]);
    const float3 org = f3(oq), wi = f3(dq), f = f3(aq), thr = f3(LD_STATE(&P.mi_t[slot]));
    const float c1 = oq.w, ppdf = dq.w;
    const float4 Lq = LD_STATE(&P.L[slot]);
    float3 L = f3(Lq);
    if (h.prim == HJ_HIT_BLOCKED) {
        // emitter-first query (MODE 3): something that is not a light is the closest hit — no contribution (rt.h:394)
    } else if (h.prim != 0xffffffffu) {
        const float4 a3 = S.tri_attr[4 * (size_t)h.prim + 3];
        const DevMaterial& m = S.materials[__float_as_uint(a3.w)];
        if (m.is_light) {
            const float4* Pp = S.tri_pos + 3 * (size_t)h.prim; const float4* A = S.tri_attr + 4 * (size_t)h.prim;
            const float b0 = 1.0f - h.b1 - h.b2;
            const float3 pos = b0 * f3(Pp[0]) + h.b1 * f3(Pp[1]) + h.b2 * f3(Pp[2]);
            const float3 ln = normalize(b0 * f3(A[0]) + h.b1 * f3(A[1]) + h.b2 * f3(A[2]));
            const float3 em = f3(m.emission[0], m.emission[1], m.emission[2]);
            const float c2 = absdot(-wi, ln);
            const float dist = length(pos - org);
            const float invG = dist * dist / c2;
            const float lpdf = (aq.w != 0.0f) ? 0.0f : light_pdf_area(S, h.prim) * invG;
            const float w = ppdf / (ppdf + lpdf);
            L += thr * w * c1 * em * f / ppdf;
        }
    } else {
        const float3 em = sky_radiance(S, wi, F.ibl);
        if (F.sky && aq.w == 0.0f) {                       // sky_sampling: balance heuristic against the sky's light-sampling pdf
            const float w = ppdf / (ppdf + sky_pdf(S, wi));
            L += thr * w * f * c1 * em / ppdf;
        } else L += thr * f * c1 * em / ppdf;              // reference: the sky through the BSDF branch, unweighted (rt.h:417-419)
    }
    if (F.debug) { uint2 g = P.dbg[slot]; g.y = (g.y ^ (h.prim + 1u)) * 0x01000193u; P.dbg[slot] = g; }
    ST_STATE(&P.L[slot], f4(L, Lq.w + 1.0f));
}

// the shadow ray of a sky sample (sky_sampling): unoccluded to infinity -> its contribution
HJ_DI void finish_sky(const FrameK& F, const WfPtrs& P, uint32_t slot, const Hit& h) {
    if (h.prim != 0xffffffffu && !F.debug) return;
    float4 L = P.L[slot];
    if (h.prim == 0xffffffffu) { const float4 c = P.sk_c[slot]; L.x += c.x; L.y += c.y; L.z += c.z; }
    L.w += 1.0f;
    P.L[slot] = L;
}

// MODE 4: shadow ray of a sky sample (any hit, origin = the surface position, no far limit).
// MODE 0: extend (closest hit), 1: shadow (any hit), 2: MIS emitter ray (closest hit), 3: MIS emitter ray answered by the
// emitter-first query (QUERY 2 of hj_trace.cuh: closest emitter, then any hit in front of it) — same contribution, and the
// walk through the whole scene may stop at the first blocker instead of finding the nearest one.
// The warp loop (dynamic ray fetch, node phases, pooled 32-wide triangle tests) is trace_warp_loop in hj_trace.cuh.
template <int MODE>
struct WfRayIO {
    const DevScene& S; const FrameK& F; const WfPtrs& P; const uint32_t* __restrict__ queue; uint32_t slot;
    HJ_DI bool load(uint32_t idx, float3& o, float3& d, float& tmin, float& tmax) {
        slot = LD_STATE(&queue[idx]);
        if (slot == 0xffffffffu) return false;
        const float4 oq = LD_STATE(MODE == 0 ? &P.ray_o[slot] : (MODE == 1 ? &P.sh_o[slot] : (MODE == 4 ? &P.sf0[slot] : &P.mi_o[slot])));
        const float4 dq = LD_STATE(MODE == 0 ? &P.ray_d[slot] : (MODE == 1 ? &P.sh_d[slot] : (MODE == 4 ? &P.sk_d[slot] : &P.mi_d[slot])));
        o = f3(oq); d = f3(dq); tmin = 0.001f; tmax = MODE == 1 ? oq.w : 1e16f;
        return true;
    }
    HJ_DI void finish(const Hit& h) {
        if (MODE == 0) finish_extend(F, P, slot, h);
        else if (MODE == 1) finish_shadow(F, P, slot, h);
        else if (MODE == 4) finish_sky(F, P, slot, h);
        else finish_mis(S, F, P, slot, h);
    }
};

template <int MODE, bool COUNT>
__global__ void __launch_bounds__(HJ_TRACE_THREADS, HJ_MINB) k_trace(DevScene S, FrameK F, WfPtrs P, int cur) {
    constexpr int QUERY = (MODE == 1 || MODE == 4) ? 1 : (MODE == 3 ? 2 : 0), CTR = MODE == 3 ? 2 : (MODE == 4 ? 3 : MODE);
    __shared__ TraceShared<QUERY> sh;
    uint2 l_stack[HJ_STACK];
    const uint32_t n = MODE == 0 ? P.cnt->n_extend[cur] : (MODE == 1 ? P.cnt->n_shadow : (MODE == 4 ? P.cnt->n_sky : P.cnt->n_mis));
    WfRayIO<MODE> io{S, F, P, MODE == 0 ? P.q_ext[cur] : (MODE == 1 ? P.q_shadow : (MODE == 4 ? P.q_sky : P.q_mis)), 0u};
    uint32_t nr = 0, nn = 0, nt = 0;
    trace_warp_loop<QUERY, COUNT>(S, io, n, &P.cnt->next[CTR], sh, l_stack, nr, nn, nt, &P.cnt->max_nodes, &P.cnt->max_tris, P.cnt->diag[CTR]);
    warp_add64((MODE == 1 || MODE == 4) ? &P.cnt->rays_shadow : &P.cnt->rays_extend, nr);
    if (COUNT) { warp_add64(&P.cnt->nodes, nn); warp_add64(&P.cnt->tris, nt); }
}

// ---------------------------------------------------------------------------------------------- accumulate / write-out
__global__ void k_reset_queues(WfCounters* c, uint32_t n0, int all) {
    if (all) { c->n_extend[0] = n0; c->n_extend[1] = 0; }
    c->n_shadow = 0; c->n_mis = 0; c->n_sky = 0; c->next[0] = c->next[1] = c->next[2] = c->next[3] = 0;
    c->n_cls[0] = c->n_cls[1] = c->n_cls[2] = c->n_ns = 0;
}
__global__ void k_swap_queues(WfCounters* c, int cur) {
    c->n_extend[cur] = 0; c->n_shadow = 0; c->n_mis = 0; c->n_sky = 0; c->next[0] = c->next[1] = c->next[2] = c->next[3] = 0;
    c->n_cls[0] = c->n_cls[1] = c->n_cls[2] = c->n_ns = 0;
}

// end of a pass: per pixel, add the pass's samples in sample order (same float addition order as the
// reference's spp loop), and export the per-path debug records
__global__ void k_accumulate(FrameK F, WfPtrs P, float4* accum, uint32_t s0, uint32_t n_samples, float4* path_rad, uint2* path_ids) {
    const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= F.npix) return;
    const uint32_t pix = F.pix0 + idx;
    float4 a = accum[pix];
    for (uint32_t s = 0; s < n_samples; s++) {
        const size_t slot = (size_t)s * F.npix + idx;
        const float4 L = P.L[slot];
        const bool skip = F.nan_repair && !(fabsf(L.x) <= FLT_MAX && fabsf(L.y) <= FLT_MAX && fabsf(L.z) <= FLT_MAX);
        if (!skip) { a.x += L.x; a.y += L.y; a.z += L.z; a.w += 1.0f; }
        if (F.debug) {
            const size_t o = (size_t)(s0 + s - F.sample_begin) * F.WH + pix;
            path_rad[o] = L; path_ids[o] = P.dbg[slot];
        }
    }
    accum[pix] = a;
}

// mean -> tonemap (kernel/color.h) -> float4 colour AOV; RGBA8 = float4ConvertColor (renderer.h:87-101)
__global__ void k_finalize(FrameK F, const float4* accum, float4* color, uchar4* rgba8) {
    const uint32_t pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= F.WH) return;
    const float4 a = accum[pix];
    // nan_repair: the mean over the samples that were finite (a.w counts them, summed over ranks); a.w == spp when none was dropped,
    // so a frame without non-finite samples is the same bits either way
    const float inv = 1.0f / ((F.nan_repair && a.w > 0.0f) ? a.w : (float)F.spp);
    float3 col = f3(a) * inv;
    if (F.tonemap == HJ_TONEMAP_ACES) col = tonemap_aces(col);
    else if (F.tonemap == HJ_TONEMAP_UCHIMURA) col = tonemap_uchimura(col);
    color[pix] = f4(col, 1.0f);
    const float3 sc = to_srgb(col);
    rgba8[pix] = make_uchar4(quantize8(sc.x), quantize8(sc.y), quantize8(sc.z), 255);
}

// ---------------------------------------------------------------------------------------------- batch trace (hj_trace)
// runs the same warp loop as the integrator's traversal kernels (persistent grid, dynamic fetch, pooled triangle tests)
template <bool ANY>
struct BatchRayIO {
    const float4* __restrict__ rays; uint32_t* prim; float* tuv; uint32_t i;
    HJ_DI bool load(uint32_t idx, float3& o, float3& d, float& tmin, float& tmax) {
        i = idx;
        const float4 oq = rays[2 * (size_t)idx], dq = rays[2 * (size_t)idx + 1];
        o = f3(oq); d = f3(dq); tmin = oq.w; tmax = dq.w;
        return true;
    }
    HJ_DI void finish(const Hit& h) {
        const bool hit = h.prim != 0xffffffffu;
        if (ANY) prim[i] = hit ? 1u : 0xffffffffu;
        else {
            prim[i] = h.prim;
            tuv[3 * (size_t)i] = hit ? h.t : 0.0f; tuv[3 * (size_t)i + 1] = hit ? h.b1 : 0.0f; tuv[3 * (size_t)i + 2] = hit ? h.b2 : 0.0f;
        }
    }
};
template <bool ANY>
__global__ void __launch_bounds__(HJ_TRACE_THREADS, HJ_MINB) k_trace_batch(DevScene S, const float4* rays, uint32_t n, uint32_t* prim, float* tuv, uint32_t* cursor) {
    __shared__ TraceShared<ANY ? 1 : 0> sh;
    uint2 l_stack[HJ_STACK];
    BatchRayIO<ANY> io{rays, prim, tuv, 0u};
    uint32_t nr = 0, nn = 0, nt = 0;
    trace_warp_loop<ANY ? 1 : 0, false>(S, io, n, cursor, sh, l_stack, nr, nn, nt, nullptr, nullptr);
}

int hj_trace_batch(hj_ctx* c, const float* rays_host, uint64_t n, int any_hit, uint32_t* prim_host, float* tuv_host) {
    if (n == 0) return HJ_OK;
    if (n > 0x7fffffffull) { c->err = "hj_trace: at most 2^31-1 rays per call"; return HJ_ERR_INVALID; }
    DevBuf &dr = c->trace_scratch[0], &dp = c->trace_scratch[1], &dt = c->trace_scratch[2], &dc = c->trace_scratch[3];   // kept across calls
    HJ_CUDA(c, dr.reserve(n * 32)); HJ_CUDA(c, dp.reserve(n * 4)); HJ_CUDA(c, dt.reserve(n * 12)); HJ_CUDA(c, dc.reserve(4));
    HJ_CUDA(c, cudaMemcpyAsync(dr.p, rays_host, n * 32, cudaMemcpyHostToDevice, c->stream));
    HJ_CUDA(c, cudaMemsetAsync(dc.p, 0, 4, c->stream));
    int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);
    const uint32_t grid = (uint32_t)std::min<uint64_t>((n + HJ_TRACE_THREADS - 1) / HJ_TRACE_THREADS, (uint64_t)sms * HJ_MINB);
    if (any_hit) k_trace_batch<true><<<grid, HJ_TRACE_THREADS, 0, c->stream>>>(c->S, dr.as<float4>(), (uint32_t)n, dp.as<uint32_t>(), dt.as<float>(), dc.as<uint32_t>());
    else k_trace_batch<false><<<grid, HJ_TRACE_THREADS, 0, c->stream>>>(c->S, dr.as<float4>(), (uint32_t)n, dp.as<uint32_t>(), dt.as<float>(), dc.as<uint32_t>());
    if (prim_host) HJ_CUDA(c, cudaMemcpyAsync(prim_host, dp.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
    if (tuv_host && !any_hit) HJ_CUDA(c, cudaMemcpyAsync(tuv_host, dt.p, n * 12, cudaMemcpyDeviceToHost, c->stream));
    HJ_CUDA(c, cudaStreamSynchronize(c->stream));
    HJ_CUDA(c, cudaGetLastError());
    return HJ_OK;
}

// ---------------------------------------------------------------------------------------------- hj_trace_noalloc
// hj_trace for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
// it). No counterpart in the reference. The scratch memory that hj_trace_batch keeps
// in DevBufs (device staging of the rays and of the results, and the work-fetch cursor of the persistent kernel) lives in ONE
// file-scope device array of a fixed, modest size; every GPU that loads the library holds it until the process exits.
//   layout: [0, 256) cursor (uint32, cleared before every launch) | rays 32 B x MAX | prim 4 B x MAX | (t,b1,b2) 12 B x MAX
// In HOST code the array's name is a host-side stand-in, not a device address: its device address is looked up once per device,
// with that device current, by cudaGetSymbolAddress, and that pointer is what the kernel gets. All calls share the array and it
// keeps the last call's contents: the cursor is cleared per call, every result slot below n is written by the kernel, and calls
// that overlap on one GPU (two contexts / streams / threads) race on it.
#define HJ_NOALLOC_BYTES (256 + (size_t)HJ_TRACE_NOALLOC_MAX_RAYS * (32 + 4 + 12))
static __device__ __align__(256) unsigned char g_trace_noalloc_scratch[HJ_NOALLOC_BYTES];
#define HJ_NOALLOC_MAX_DEVICES 64
static unsigned char* g_trace_noalloc_ptr[HJ_NOALLOC_MAX_DEVICES];      // device address of the array, per device (0 = not looked up yet)
static std::mutex g_trace_noalloc_mutex;

extern "C" int hj_trace_noalloc(hj_ctx* c, const float* rays_host, uint64_t n, int any_hit, uint32_t* prim_host, float* tuv_host) {
    if (!c || (n && !rays_host)) return HJ_ERR_INVALID;
    if (n > (uint64_t)HJ_TRACE_NOALLOC_MAX_RAYS) {
        char b[200]; snprintf(b, sizeof b, "hj_trace_noalloc: %llu rays exceed the fixed scratch of HJ_TRACE_NOALLOC_MAX_RAYS = %d rays per call; split the batch or use hj_trace",
                              (unsigned long long)n, (int)HJ_TRACE_NOALLOC_MAX_RAYS);
        c->err = b; return HJ_ERR_INVALID;
    }
    if (c->device < 0 || c->device >= HJ_NOALLOC_MAX_DEVICES) { c->err = "hj_trace_noalloc: device index beyond the per-device table"; return HJ_ERR_INVALID; }
    if (c->bvh.pending_set >= 0) { c->err = "hj_trace_noalloc: an asynchronous accel build is pending; call hj_accel_wait first (completing it may allocate)"; return HJ_ERR_NO_ACCEL; }
    if (!c->have_accel) { c->err = "hj_trace_noalloc: hj_accel_build has not been called"; return HJ_ERR_NO_ACCEL; }
    if (n == 0) return HJ_OK;
    HJ_CUDA(c, cudaSetDevice(c->device));
    unsigned char* base = nullptr;
    {   // set-up, once per device, with that device current
        std::lock_guard<std::mutex> lock(g_trace_noalloc_mutex);
        if (!g_trace_noalloc_ptr[c->device]) {
            void* p = nullptr;
            HJ_CUDA(c, cudaGetSymbolAddress(&p, g_trace_noalloc_scratch));
            g_trace_noalloc_ptr[c->device] = (unsigned char*)p;
        }
        base = g_trace_noalloc_ptr[c->device];
    }
    uint32_t* cursor = (uint32_t*)base;
    float4* d_rays = (float4*)(base + 256);
    uint32_t* d_prim = (uint32_t*)(base + 256 + (size_t)HJ_TRACE_NOALLOC_MAX_RAYS * 32);
    float* d_tuv = (float*)(base + 256 + (size_t)HJ_TRACE_NOALLOC_MAX_RAYS * 36);
    HJ_CUDA(c, cudaMemsetAsync(cursor, 0, 256, c->stream));             // the one thing that must start at zero
    HJ_CUDA(c, cudaMemcpyAsync(d_rays, rays_host, n * 32, cudaMemcpyHostToDevice, c->stream));
    int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);
    const uint32_t grid = (uint32_t)std::min<uint64_t>((n + HJ_TRACE_THREADS - 1) / HJ_TRACE_THREADS, (uint64_t)sms * HJ_MINB);
    if (any_hit) k_trace_batch<true><<<grid, HJ_TRACE_THREADS, 0, c->stream>>>(c->S, d_rays, (uint32_t)n, d_prim, d_tuv, cursor);
    else k_trace_batch<false><<<grid, HJ_TRACE_THREADS, 0, c->stream>>>(c->S, d_rays, (uint32_t)n, d_prim, d_tuv, cursor);
    if (prim_host) HJ_CUDA(c, cudaMemcpyAsync(prim_host, d_prim, n * 4, cudaMemcpyDeviceToHost, c->stream));
    if (tuv_host && !any_hit) HJ_CUDA(c, cudaMemcpyAsync(tuv_host, d_tuv, n * 12, cudaMemcpyDeviceToHost, c->stream));
    HJ_CUDA(c, cudaStreamSynchronize(c->stream));
    HJ_CUDA(c, cudaGetLastError());
    return HJ_OK;
}

// ---------------------------------------------------------------------------------------------- host driver
static inline uint32_t cdiv_u(uint64_t a, uint32_t b) { return (uint32_t)((a + b - 1) / b); }

static int wf_reserve(hj_ctx* c, WavefrontState& w, uint32_t cap, bool mis, bool debug, bool sky) {
    const size_t f4s = (size_t)cap * sizeof(float4);
    HJ_CUDA(c, w.ray_o.reserve(f4s)); HJ_CUDA(c, w.ray_d.reserve(f4s)); HJ_CUDA(c, w.hit.reserve(f4s));
    HJ_CUDA(c, w.thr.reserve(f4s)); HJ_CUDA(c, w.L.reserve(f4s)); HJ_CUDA(c, w.rng_depth.reserve((size_t)cap * 4));
    HJ_CUDA(c, w.dbg.reserve(debug ? (size_t)cap * 8 : 8));
    HJ_CUDA(c, w.q_extend[0].reserve((size_t)cap * 4)); HJ_CUDA(c, w.q_extend[1].reserve((size_t)cap * 4));
    HJ_CUDA(c, w.q_shadow.reserve((size_t)cap * 4)); HJ_CUDA(c, w.q_mis.reserve((size_t)cap * 4));
    for (int k = 0; k < 3; k++) HJ_CUDA(c, w.q_cls[k].reserve((size_t)cap * 4));
    for (int k = 0; k < 4; k++) HJ_CUDA(c, w.sf[k].reserve(f4s));
    HJ_CUDA(c, w.sh_o.reserve(f4s)); HJ_CUDA(c, w.sh_d.reserve(f4s)); HJ_CUDA(c, w.sh_c.reserve(f4s));
    if (mis) { HJ_CUDA(c, w.mi_o.reserve(f4s)); HJ_CUDA(c, w.mi_d.reserve(f4s)); HJ_CUDA(c, w.mi_a.reserve(f4s)); HJ_CUDA(c, w.mi_t.reserve(f4s)); }
    if (sky) { HJ_CUDA(c, w.q_sky.reserve((size_t)cap * 4)); HJ_CUDA(c, w.sk_d.reserve(f4s)); HJ_CUDA(c, w.sk_c.reserve(f4s)); }
    HJ_CUDA(c, w.counters.reserve(sizeof(WfCounters)));
    if (!w.stream) HJ_CUDA(c, cudaStreamCreateWithFlags(&w.stream, cudaStreamNonBlocking));
    if (!w.ev_acc) HJ_CUDA(c, cudaEventCreateWithFlags(&w.ev_acc, cudaEventDisableTiming));
    w.capacity = cap;
    return HJ_OK;
}

static WfPtrs wf_ptrs(hj_ctx* c, WavefrontState& w) {
    WfPtrs P;
    P.ray_o = w.ray_o.as<float4>(); P.ray_d = w.ray_d.as<float4>(); P.hit = w.hit.as<float4>(); P.thr = w.thr.as<float4>();
    P.L = w.L.as<float4>(); P.rng = w.rng_depth.as<uint32_t>(); P.dbg = w.dbg.as<uint2>();
    P.q_ext[0] = w.q_extend[0].as<uint32_t>(); P.q_ext[1] = w.q_extend[1].as<uint32_t>();
    P.q_shadow = w.q_shadow.as<uint32_t>(); P.q_mis = w.q_mis.as<uint32_t>();
    for (int k = 0; k < 3; k++) P.q_cls[k] = w.q_cls[k].as<uint32_t>();
    P.q_ns = w.q_ns.as<uint32_t>();
    P.sf0 = w.sf[0].as<float4>(); P.sf1 = w.sf[1].as<float4>(); P.sf2 = w.sf[2].as<float4>(); P.sf3 = w.sf[3].as<float4>();
    P.sh_o = w.sh_o.as<float4>(); P.sh_d = w.sh_d.as<float4>(); P.sh_c = w.sh_c.as<float4>();
    P.mi_o = w.mi_o.as<float4>(); P.mi_d = w.mi_d.as<float4>(); P.mi_a = w.mi_a.as<float4>(); P.mi_t = w.mi_t.as<float4>();
    P.q_sky = w.q_sky.as<uint32_t>(); P.sk_d = w.sk_d.as<float4>(); P.sk_c = w.sk_c.as<float4>();
    P.cnt = w.counters.as<WfCounters>();
    P.aov_albedo = c->aov_albedo.as<float4>(); P.aov_normal = c->aov_normal.as<float4>();
    return P;
}

// the shade stages of one bounce, in stream order
template <int INTEG>
static int launch_shade_t(hj_ctx* c, uint32_t grid, cudaStream_t ps, const FrameK& F, const WfPtrs& P, int cur, int depth, uint32_t s0) {
    int n = 4;
    k_surface<INTEG><<<grid, WF_THREADS, 0, ps>>>(c->S, F, P, cur, depth, s0);
    const uint32_t sgrid = std::max(1u, grid * WF_THREADS / HJ_SAMPLE_THREADS);      // same number of threads as the other shading kernels
    k_sample<INTEG, HJ_CLS_DISNEY><<<sgrid, HJ_SAMPLE_THREADS, 0, ps>>>(c->S, F, P, cur, depth, s0);
    k_sample<INTEG, HJ_CLS_GGX><<<sgrid, HJ_SAMPLE_THREADS, 0, ps>>>(c->S, F, P, cur, depth, s0);
    k_sample<INTEG, HJ_CLS_GLASS><<<sgrid, HJ_SAMPLE_THREADS, 0, ps>>>(c->S, F, P, cur, depth, s0);
    return n;
}
static int launch_shade(hj_ctx* c, int integrator, uint32_t grid, cudaStream_t ps, const FrameK& F, const WfPtrs& P, int cur, int depth, uint32_t s0) {
    if (integrator == HJ_INTEGRATOR_PT) return launch_shade_t<HJ_INTEGRATOR_PT>(c, grid, ps, F, P, cur, depth, s0);
    if (integrator == HJ_INTEGRATOR_NEE) return launch_shade_t<HJ_INTEGRATOR_NEE>(c, grid, ps, F, P, cur, depth, s0);
    return launch_shade_t<HJ_INTEGRATOR_MIS>(c, grid, ps, F, P, cur, depth, s0);
}

int hj_wavefront_render(hj_ctx* c, const hj_frame_params* fp) {
    const uint32_t W = fp->width, H = fp->height;
    const uint64_t WH = (uint64_t)W * H;
    if (W == 0 || H == 0 || fp->spp == 0 || fp->sample_end < fp->sample_begin || fp->max_depth < 0 || WH > 0x7fffffffull) {
        c->err = "hj_render: invalid frame parameters"; return HJ_ERR_INVALID;
    }
    if (fp->integrator < HJ_INTEGRATOR_PT || fp->integrator > HJ_INTEGRATOR_MIS) { c->err = "hj_render: unknown integrator"; return HJ_ERR_INVALID; }
    const uint32_t ns = fp->sample_end - fp->sample_begin;
    const bool debug = fp->debug_paths != 0;
    const bool timed = c->opt_time_stages != 0;
    // passes: whole samples of the whole frame; `np` pipelines keep that many passes in flight (1 when timing stages)
    // Pipelines in flight. r02 sweep (profiles/r02_summary.md): with the work-stealing traversal the tail of a pass costs little, and
    // what is left to gain from overlap is bought with smaller passes and 4x the launches. Two pipelines are best at every sample
    // count from 1 to 64 on C1 / C2 / C5 (C1 16 spp: 9.5 ms with 2, 11.8 with 8; C2 8 spp: 49.8 vs 53.8; 64 spp: 381 vs 386); only
    // the depth-32 glass scene, whose bounces 10-32 run on tiny queues, still prefers eight (525 vs 533 ms).
    int np = timed ? 1 : (c->opt_pipelines > 0 ? c->opt_pipelines : (fp->max_depth > 16 ? 8 : 2));
    // Pass size: the persistent traversal kernels keep ~150 k rays in flight and lose lanes while a queue drains, so a
    // pass should be as large as memory allows (r01-i, C2 64 spp: 2 Mi paths 1320, 16 Mi paths 1465 Mrays/s). Auto: the
    // frame split evenly over the pipelines, within half of the free device memory (~220 B of path state per slot).
    uint64_t target = (uint64_t)c->opt_paths_per_pass;
    if (c->opt_paths_per_pass <= 0) {
        size_t mfree = 0, mtotal = 0;
        cudaMemGetInfo(&mfree, &mtotal);
        size_t held = 0;                                   // what the pipelines already own counts as available
        for (int p = 0; p < HJ_MAX_PIPES; p++) for (int k = 0; c->wf[p].all[k]; k++) held += c->wf[p].all[k]->cap;
        const uint64_t budget_slots = (uint64_t)((mfree + held) / 2) / 240ull / (uint64_t)np;
        target = ((uint64_t)ns * WH + (uint64_t)np - 1) / (uint64_t)np;
        if (target > budget_slots) target = budget_slots;
        if (target > (64ull << 20)) target = 64ull << 20;
        target += WH - 1;                                  // round up to whole samples
    }
    uint32_t spass = (uint32_t)(target / WH); if (spass < 1) spass = 1; if (spass > ns && ns) spass = ns;
    if (spass * WH > 0xfffffff0ull) spass = (uint32_t)(0xfffffff0ull / WH);
    const uint32_t n_spasses = ns ? (ns + spass - 1) / spass : 0;
    // Few samples per frame (the real-time C4 case: 1-3 spp under the fps budget) leave pipelines idle, and every bounce of a
    // lone pass ends in the tail of its longest rays. Such frames are also cut into pixel bands so that all pipelines have
    // a pass: band = a contiguous pixel range of the pass's samples, at least 128 Ki paths.
    uint32_t bands = 1;
    if (!timed && n_spasses && n_spasses < (uint32_t)np) {
        bands = (uint32_t)np / n_spasses;
        const uint64_t most = std::max<uint64_t>(1, ((uint64_t)spass * WH) >> 17);
        if (bands > most) bands = (uint32_t)most;
    }
    if (c->opt_bands > 0) bands = (uint32_t)c->opt_bands;
    if (bands > WH) bands = (uint32_t)WH;
    const uint32_t band_pix = (uint32_t)((WH + bands - 1) / bands);
    bands = (uint32_t)((WH + band_pix - 1) / band_pix);
    const uint32_t n_passes = n_spasses * bands;
    if ((uint32_t)np > n_passes) np = (int)std::max(1u, n_passes);
    const uint32_t cap = (uint32_t)((uint64_t)spass * band_pix);
    const bool sky = fp->sky_sampling != 0 && fp->integrator == HJ_INTEGRATOR_MIS;
    for (int p = 0; p < np; p++) { int rc = wf_reserve(c, c->wf[p], cap, fp->integrator == HJ_INTEGRATOR_MIS, debug, sky); if (rc) return rc; }
    c->n_pipes_used = np;
    const size_t fb = (size_t)WH * sizeof(float4);
    HJ_CUDA(c, c->aov_color.reserve(fb)); HJ_CUDA(c, c->aov_albedo.reserve(fb)); HJ_CUDA(c, c->aov_normal.reserve(fb));
    HJ_CUDA(c, c->aov_accum.reserve(fb)); HJ_CUDA(c, c->aov_rgba8.reserve((size_t)WH * 4));
    if (debug) { HJ_CUDA(c, c->aov_path_rad.reserve((size_t)ns * fb)); HJ_CUDA(c, c->aov_path_ids.reserve((size_t)ns * WH * 8)); }
    if (!c->ev_frame) HJ_CUDA(c, cudaEventCreateWithFlags(&c->ev_frame, cudaEventDisableTiming));
    cudaStream_t st = c->stream;
    if (c->fb_w != W || c->fb_h != H) {
        HJ_CUDA(c, cudaMemsetAsync(c->aov_albedo.p, 0, fb, st)); HJ_CUDA(c, cudaMemsetAsync(c->aov_normal.p, 0, fb, st));
        c->fb_w = W; c->fb_h = H;
    }
    FrameK F;
    F.W = W; F.H = H; F.WH = (uint32_t)WH; F.spp = fp->spp; F.frame = fp->frame; F.sample_begin = fp->sample_begin;
    F.pix0 = 0; F.npix = (uint32_t)WH;
    F.cpos = make_float3(fp->camera_pos[0], fp->camera_pos[1], fp->camera_pos[2]);
    F.cdir = make_float3(fp->camera_dir[0], fp->camera_dir[1], fp->camera_dir[2]);
    F.cup = make_float3(fp->camera_up[0], fp->camera_up[1], fp->camera_up[2]);
    F.cright = make_float3(fp->camera_right[0], fp->camera_right[1], fp->camera_right[2]);
    F.cf = fp->camera_f; F.ibl = fp->ibl_intensity;
    F.integrator = fp->integrator; F.max_depth = fp->max_depth; F.tonemap = fp->tonemap; F.jitter = fp->pixel_jitter; F.debug = debug;
    F.sky = sky ? 1 : 0; F.nan_repair = c->opt_nan_repair;

    int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);
    const uint32_t grid = std::min(cdiv_u(cap, WF_THREADS), (uint32_t)sms * 16u);
    // persistent traversal grid: exactly the blocks that are resident at once (a multiple of the SM count)
    int occ = 8; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_trace<2, false>, HJ_TRACE_THREADS, 0);
    if (c->opt_trace_blocks_per_sm > 0) occ = c->opt_trace_blocks_per_sm;
    const uint32_t tgrid = std::min(cdiv_u(cap, HJ_TRACE_THREADS), (uint32_t)sms * (uint32_t)std::max(occ, 1));
    const bool count = c->opt_count_fetches != 0;
    uint64_t launches = 0;

    c->ev_kind.clear();
    auto mark_begin = [&](int kind, cudaStream_t s) {
        if (!timed) return;
        const size_t i = 2 * c->ev_kind.size();
        while (c->ev_pool.size() < i + 2) { cudaEvent_t e; cudaEventCreate(&e); c->ev_pool.push_back(e); }
        c->ev_kind.push_back(kind);
        cudaEventRecord(c->ev_pool[i], s);
    };
    auto mark_end = [&](cudaStream_t s) { if (timed) cudaEventRecord(c->ev_pool[2 * c->ev_kind.size() - 1], s); };

    // Tried in r02 and dropped: an L2 persisting window (cudaStreamAttributeAccessPolicyWindow) over the BVH nodes. With the carve-out at the
    // device maximum everything else thrashes (C2 380 -> 514 ms, C5 560 -> 739 ms); sized to the nodes alone (30 MB on C2) it changes nothing.
    HJ_CUDA(c, cudaEventRecord(c->ev0, st));
    HJ_CUDA(c, cudaMemsetAsync(c->aov_accum.p, 0, fb, st));
    for (int p = 0; p < np; p++) HJ_CUDA(c, cudaMemsetAsync(c->wf[p].counters.p, 0, sizeof(WfCounters), st));
    HJ_CUDA(c, cudaEventRecord(c->ev_frame, st));
    for (int p = 0; p < np; p++) HJ_CUDA(c, cudaStreamWaitEvent(c->wf[p].stream, c->ev_frame, 0));
    cudaEvent_t prev_acc = nullptr;
    uint32_t pass = 0;
    const FrameK F_frame = F;
    for (uint32_t s0 = fp->sample_begin; s0 < fp->sample_end; s0 += spass)
    for (uint32_t band = 0; band < bands; band++, pass++) {
        WavefrontState& w = c->wf[pass % np];
        FrameK F = F_frame;
        F.pix0 = band * band_pix; F.npix = std::min(band_pix, (uint32_t)WH - F.pix0);
        cudaStream_t ps = w.stream;
        const WfPtrs P = wf_ptrs(c, w);
        const uint32_t nsamp = std::min(spass, fp->sample_end - s0);
        const uint32_t nslots = (uint32_t)((uint64_t)nsamp * F.npix);
        k_reset_queues<<<1, 1, 0, ps>>>(P.cnt, nslots, 1);
        k_raygen<<<grid, WF_THREADS, 0, ps>>>(F, P, s0, nslots);
        launches += 2;
        int cur = 0;
        for (int depth = 0; depth < fp->max_depth; depth++) {
            mark_begin(0, ps);
            if (count) k_trace<0, true><<<tgrid, HJ_TRACE_THREADS, 0, ps>>>(c->S, F, P, cur);
            else k_trace<0, false><<<tgrid, HJ_TRACE_THREADS, 0, ps>>>(c->S, F, P, cur);
            mark_end(ps);
            mark_begin(1, ps);
            launches += launch_shade(c, fp->integrator, grid, ps, F, P, cur, depth, s0);
            mark_end(ps);
            launches += 1;
            if (fp->integrator != HJ_INTEGRATOR_PT) {
                mark_begin(0, ps);
                if (count) k_trace<1, true><<<tgrid, HJ_TRACE_THREADS, 0, ps>>>(c->S, F, P, cur);
                else k_trace<1, false><<<tgrid, HJ_TRACE_THREADS, 0, ps>>>(c->S, F, P, cur);
                mark_end(ps);
                launches++;
            }
            if (sky) {                   // shadow rays of the sky samples (extension, after the light samples' so the additions keep that order)
                mark_begin(0, ps);
                if (count) k_trace<4, true><<<tgrid, HJ_TRACE_THREADS, 0, ps>>>(c->S, F, P, cur);
                else k_trace<4, false><<<tgrid, HJ_TRACE_THREADS, 0, ps>>>(c->S, F, P, cur);
                mark_end(ps);
                launches++;
            }
            if (fp->integrator == HJ_INTEGRATOR_MIS) {
                mark_begin(0, ps);
                // the debug records hash the MIS ray's closest hit whatever it is: they need the plain closest-hit query
                const bool efirst = c->opt_mis_emitter_first && !debug;
                if (efirst) { if (count) k_trace<3, true><<<tgrid, HJ_TRACE_THREADS, 0, ps>>>(c->S, F, P, cur); else k_trace<3, false><<<tgrid, HJ_TRACE_THREADS, 0, ps>>>(c->S, F, P, cur); }
                else if (count) k_trace<2, true><<<tgrid, HJ_TRACE_THREADS, 0, ps>>>(c->S, F, P, cur);
                else k_trace<2, false><<<tgrid, HJ_TRACE_THREADS, 0, ps>>>(c->S, F, P, cur);
                mark_end(ps);
                launches++;
            }
            k_swap_queues<<<1, 1, 0, ps>>>(P.cnt, cur);
            launches++;
            cur ^= 1;
        }
        // accumulate in pass order: the float additions per pixel follow the sample order of the reference's spp loop
        if (prev_acc) HJ_CUDA(c, cudaStreamWaitEvent(ps, prev_acc, 0));
        k_accumulate<<<cdiv_u(F.npix, 256), 256, 0, ps>>>(F, P, c->aov_accum.as<float4>(), s0, nsamp, c->aov_path_rad.as<float4>(), c->aov_path_ids.as<uint2>());
        HJ_CUDA(c, cudaEventRecord(w.ev_acc, ps));
        prev_acc = w.ev_acc;
        launches++;
    }
    // join: every pipeline's last accumulate precedes the finalize on the context stream
    for (int p = 0; p < np; p++) { HJ_CUDA(c, cudaEventRecord(c->wf[p].ev_acc, c->wf[p].stream)); HJ_CUDA(c, cudaStreamWaitEvent(st, c->wf[p].ev_acc, 0)); }
    k_finalize<<<cdiv_u(WH, 256), 256, 0, st>>>(F, c->aov_accum.as<float4>(), c->aov_color.as<float4>(), c->aov_rgba8.as<uchar4>());
    launches++;
    HJ_CUDA(c, cudaEventRecord(c->ev1, st));
    if (c->accel[c->accel_cur].ev_last_read) HJ_CUDA(c, cudaEventRecord(c->accel[c->accel_cur].ev_last_read, st));   // a later build into this set waits for this frame
    HJ_CUDA(c, cudaGetLastError());
    c->fp = *fp; c->have_frame = true; c->have_output = false; c->frame_reduced = false;
    c->stats.kernel_launches = launches;
    c->stats.paths = (uint64_t)ns * WH;
    c->stats.passes = n_passes;
    return HJ_OK;
}

int hj_wavefront_finalize(hj_ctx* c) {
    const hj_frame_params& fp = c->fp;
    FrameK F{}; F.W = fp.width; F.H = fp.height; F.WH = fp.width * fp.height; F.spp = fp.spp; F.tonemap = fp.tonemap; F.nan_repair = c->opt_nan_repair;
    k_finalize<<<cdiv_u(F.WH, 256), 256, 0, c->stream>>>(F, c->aov_accum.as<float4>(), c->aov_color.as<float4>(), c->aov_rgba8.as<uchar4>());
    HJ_CUDA(c, cudaGetLastError());
    return HJ_OK;
}

// reads the device counters of the last frame into ctx->stats (after the stream has been synchronised)
int hj_wavefront_stats(hj_ctx* c) {
    WfCounters t{};
    for (int p = 0; p < c->n_pipes_used; p++) {
        WfCounters h;
        HJ_CUDA(c, cudaMemcpy(&h, c->wf[p].counters.p, sizeof h, cudaMemcpyDeviceToHost));
        t.rays_extend += h.rays_extend; t.rays_shadow += h.rays_shadow; t.nodes += h.nodes; t.tris += h.tris;
        t.max_nodes = std::max(t.max_nodes, h.max_nodes); t.max_tris = std::max(t.max_tris, h.max_tris);
        for (int m = 0; m < 4; m++) for (int k = 0; k < 8; k++) t.diag[m][k] += h.diag[m][k];
    }
    if (c->opt_count_fetches && getenv("HJ_TRACE_DIAG"))
        for (int m = 0; m < 3; m++) {
            const unsigned long long* d = t.diag[m]; const double ph = d[0] ? (double)d[0] : 1.0;
            fprintf(stderr, "[hj diag] ray type %d: node phases %llu, lanes/phase active %.2f walking %.2f waiting %.2f holding %.2f; "
                    "triangle batches %llu, lanes/batch %.2f\n", m, d[0], d[1] / ph, d[2] / ph, d[3] / ph, d[4] / ph, d[5], d[5] ? (double)d[6] / d[5] : 0.0);
        }
    c->stats.rays_extend = t.rays_extend; c->stats.rays_shadow = t.rays_shadow;
    c->stats.nodes_fetched = t.nodes; c->stats.tris_fetched = t.tris;
    c->stats.max_nodes_per_ray = t.max_nodes; c->stats.max_tris_per_ray = t.max_tris;
    float ms = 0; if (cudaEventElapsedTime(&ms, c->ev0, c->ev1) == cudaSuccess) c->stats.ms_render = ms;
    double tt = 0, ts = 0;
    for (size_t i = 0; i < c->ev_kind.size(); i++) {
        float m = 0; cudaEventElapsedTime(&m, c->ev_pool[2 * i], c->ev_pool[2 * i + 1]);
        (c->ev_kind[i] == 0 ? tt : ts) += m;
    }
    c->stats.ms_trace = (float)tt; c->stats.ms_shade = (float)ts;
    return HJ_OK;
}

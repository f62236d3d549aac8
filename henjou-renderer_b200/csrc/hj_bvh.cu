// hj_bvh.cu — GPU acceleration-structure build: the B200 replacement for the reference's optixAccelBuild call
// sites buildGAS / buildIAS / updateIASMatrix (/root/reference/include/renderer/renderer.h:319-396, 398-490,
// 257-291). The loader emits one instance per mesh node with no hierarchy (gltfloader.h:1508-1512), so the
// two-level GAS/IAS is flattened to world space each frame (semantics-preserving) and ONE
// BVH is built over all triangles:
//   k_flatten      instance transform -> world positions / normalised normals / uv / material per triangle
//   k_morton       63-bit Morton code of the triangle centroid inside the scene bounds
//   radix sort     hand-written LSD radix sort, 8 bits per pass, (u64 key, u32 value) pairs
//   k_karras       binary radix tree (Karras 2012)
//   k_fit          bottom-up AABB fit
//   k_collapse     level-by-level collapse to 8-wide nodes with 8-bit quantised child boxes (112 B / node)
#include "hj_ctx.h"
#include "hj_trace.cuh"
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

// results of a build, written by the device, mirrored into pinned host memory at the end of the enqueued work
struct BuildStatus {
    uint32_t n_nodes, n_tris, depth, error;        // error bits: HJ_BUILD_ERR_*
    uint32_t cur_n, node_ctr, pp, ploc_rounds;     // PLOC hand-over from the cooperative kernel to the tail
};
#define HJ_BUILD_ERR_NONFINITE 1u      // a world-space vertex is NaN / inf
#define HJ_BUILD_ERR_PLOC 2u           // a PLOC round merged nothing (only with non-finite boxes)
#define HJ_BUILD_ERR_COLLAPSE 4u       // the collapse did not terminate within the level bound

#define RS_THREADS 256
#define RS_ITEMS 16
#define RS_TILE (RS_THREADS * RS_ITEMS)

// ---------------------------------------------------------------------------------------------- flatten
// order-preserving float <-> int so atomicMin/atomicMax work on floats
HJ_DI int f2ord(float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
HJ_DI float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

// ≙ the vertex fetch + transform of __closesthit__ch (stale PTX :880-1287) hoisted out of the hit program:
// p_k = transform_position(transforms[inst], vertices[i_k]); n_k = normalize(transform_normal(inv[inst], normals[i_k]))
#define FLATTEN_THREADS 256
__global__ void __launch_bounds__(FLATTEN_THREADS) k_flatten(DevScene S, int* bounds, uint32_t* error) {
    // the 7 float4 a triangle produces are staged in shared memory and written out by the whole block: 48-B / 64-B records written by
    // one thread each are strided 16-B stores (r02: 1.70 ms for the 10.2 M triangles of C5 against 0.36 ms of pure traffic)
    __shared__ float4 s_pos[3 * FLATTEN_THREADS], s_attr[4 * FLATTEN_THREADS];
    const uint32_t t0 = blockIdx.x * FLATTEN_THREADS, t = t0 + threadIdx.x;
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (t < S.n_tris) {
        const uint32_t inst = S.tri_inst[t];
        const float4* m = S.xf + 3 * (size_t)inst; const float4* im = S.inv_xf + 3 * (size_t)inst;
        float3 p[3], n[3]; float2 uv[3];
        for (int k = 0; k < 3; k++) {
            const uint32_t vi = S.indices[3 * (size_t)t + k];
            p[k] = xform_point(m, ld3(S.vertices, vi));
            n[k] = normalize(xform_normal(im, ld3(S.normals, vi)));
            uv[k] = make_float2(S.texcoords[2 * (size_t)vi], S.texcoords[2 * (size_t)vi + 1]);
            if (!(fabsf(p[k].x) <= FLT_MAX && fabsf(p[k].y) <= FLT_MAX && fabsf(p[k].z) <= FLT_MAX)) atomicOr(error, HJ_BUILD_ERR_NONFINITE);   // a NaN box can never merge
            lo[0] = fminf(lo[0], p[k].x); lo[1] = fminf(lo[1], p[k].y); lo[2] = fminf(lo[2], p[k].z);
            hi[0] = fmaxf(hi[0], p[k].x); hi[1] = fmaxf(hi[1], p[k].y); hi[2] = fmaxf(hi[2], p[k].z);
        }
        float4* P = s_pos + 3 * threadIdx.x;
        P[0] = f4(p[0], 0.0f); P[1] = f4(p[1], 0.0f); P[2] = f4(p[2], 0.0f);
        float4* A = s_attr + 4 * threadIdx.x;
        A[0] = f4(n[0], uv[0].x); A[1] = f4(n[1], uv[0].y); A[2] = f4(n[2], uv[1].x);
        A[3] = make_float4(uv[1].y, uv[2].x, uv[2].y, __uint_as_float(S.material_ids[t]));
    }
    __syncthreads();
    const uint32_t n_here = min((uint32_t)FLATTEN_THREADS, S.n_tris > t0 ? S.n_tris - t0 : 0u);
    float4* Pg = S.tri_pos + 3 * (size_t)t0; float4* Ag = S.tri_attr + 4 * (size_t)t0;
    for (uint32_t i = threadIdx.x; i < 3 * n_here; i += FLATTEN_THREADS) Pg[i] = s_pos[i];
    for (uint32_t i = threadIdx.x; i < 4 * n_here; i += FLATTEN_THREADS) Ag[i] = s_attr[i];
    // block reduction of the scene bounds, one atomic per warp
    for (int a = 0; a < 3; a++) {
        float l = lo[a], h = hi[a];
        for (int o = 16; o; o >>= 1) { l = fminf(l, __shfl_xor_sync(0xffffffffu, l, o)); h = fmaxf(h, __shfl_xor_sync(0xffffffffu, h, o)); }
        if ((threadIdx.x & 31) == 0 && l <= h) { atomicMin(bounds + a, f2ord(l)); atomicMax(bounds + 3 + a, f2ord(h)); }
    }
}

HJ_DI uint64_t spread21(uint32_t v) {
    uint64_t x = v & 0x1fffffu;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

__global__ void k_morton(DevScene S, const int* bounds, uint64_t* keys, uint32_t* vals, int drop_bits) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= S.n_tris) return;
    const float3 lo = f3(ord2f(bounds[0]), ord2f(bounds[1]), ord2f(bounds[2]));
    const float3 hi = f3(ord2f(bounds[3]), ord2f(bounds[4]), ord2f(bounds[5]));
    const float4* P = S.tri_pos + 3 * (size_t)t;
    const float3 a = f3(P[0]), b = f3(P[1]), c = f3(P[2]);
    const float3 cen = f3(fminf(a.x, fminf(b.x, c.x)) + fmaxf(a.x, fmaxf(b.x, c.x)), fminf(a.y, fminf(b.y, c.y)) + fmaxf(a.y, fmaxf(b.y, c.y)),
                          fminf(a.z, fminf(b.z, c.z)) + fmaxf(a.z, fmaxf(b.z, c.z))) * 0.5f;
    const float ex = fmaxf(hi.x - lo.x, 1e-30f), ey = fmaxf(hi.y - lo.y, 1e-30f), ez = fmaxf(hi.z - lo.z, 1e-30f);
    const float e = fmaxf(ex, fmaxf(ey, ez));   // cubic cells: one scale for all axes
    const uint32_t qx = (uint32_t)fminf(fmaxf((cen.x - lo.x) / e * 2097152.0f, 0.0f), 2097151.0f);
    const uint32_t qy = (uint32_t)fminf(fmaxf((cen.y - lo.y) / e * 2097152.0f, 0.0f), 2097151.0f);
    const uint32_t qz = (uint32_t)fminf(fmaxf((cen.z - lo.z) / e * 2097152.0f, 0.0f), 2097151.0f);
    // option morton_bits: the low `drop_bits` of every axis are dropped, so the key has 3 x (21 - drop_bits) bits and the sort needs fewer passes
    keys[t] = (spread21(qx >> drop_bits) << 2) | (spread21(qy >> drop_bits) << 1) | spread21(qz >> drop_bits);
    vals[t] = t;
}

// ---------------------------------------------------------------------------------------------- radix sort
// One pass = histogram / scan / scatter. Warp w of a block owns a contiguous chunk of the tile, so a stable
// rank is (elements of the same digit in earlier chunks) + (same digit earlier in this warp's chunk).
__global__ void __launch_bounds__(RS_THREADS) k_rs_hist(const uint64_t* __restrict__ keys, uint32_t n, int shift, uint32_t* __restrict__ block_hist, uint32_t nb) {
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * RS_TILE;
    for (int i = 0; i < RS_ITEMS; i++) {
        const uint32_t idx = base + i * RS_THREADS + threadIdx.x;
        if (idx < n) atomicAdd(&h[(uint32_t)(keys[idx] >> shift) & 0xffu], 1u);
    }
    __syncthreads();
    block_hist[threadIdx.x * nb + blockIdx.x] = h[threadIdx.x];
}

// block_hist is [digit][tile]. One block per digit scans its row in place (exclusive) and writes the digit's total; the scatter
// kernel adds the exclusive scan of the 256 totals itself. (Round 1 scanned all 256 x tiles counters in ONE block.)
#define RS_SCAN_THREADS 256
__global__ void __launch_bounds__(RS_SCAN_THREADS) k_rs_scan_rows(uint32_t* hist, uint32_t nb, uint32_t* digit_total) {
    __shared__ uint32_t part[RS_SCAN_THREADS];
    uint32_t* a = hist + (size_t)blockIdx.x * nb;
    const uint32_t per = (nb + RS_SCAN_THREADS - 1) / RS_SCAN_THREADS;
    const uint32_t b = min(threadIdx.x * per, nb), e = min(b + per, nb);
    uint32_t s = 0;
    for (uint32_t i = b; i < e; i++) s += a[i];
    part[threadIdx.x] = s;
    __syncthreads();
    for (int o = 1; o < RS_SCAN_THREADS; o <<= 1) {
        uint32_t v = (threadIdx.x >= (uint32_t)o) ? part[threadIdx.x - o] : 0u;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    uint32_t run = part[threadIdx.x] - s;
    for (uint32_t i = b; i < e; i++) { uint32_t v = a[i]; a[i] = run; run += v; }
    if (threadIdx.x == RS_SCAN_THREADS - 1) digit_total[blockIdx.x] = part[RS_SCAN_THREADS - 1];
}

__global__ void __launch_bounds__(RS_THREADS) k_rs_scatter(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ vals, uint32_t n, int shift,
                                                           const uint32_t* __restrict__ block_off, const uint32_t* __restrict__ digit_total, uint32_t nb,
                                                           uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out) {
    __shared__ uint32_t wh[RS_THREADS / 32][256];
    __shared__ uint32_t dbase[256];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    {   // exclusive scan of the 256 digit totals (first write position of every digit)
        const uint32_t t = digit_total[threadIdx.x];
        uint32_t incl = t;
        for (int o = 1; o < 32; o <<= 1) { const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
        __shared__ uint32_t wsum[RS_THREADS / 32];
        if (lane == 31) wsum[w] = incl;
        __syncthreads();
        uint32_t off = 0;
        for (int ww = 0; ww < w; ww++) off += wsum[ww];
        dbase[threadIdx.x] = off + incl - t;
    }
    for (int i = threadIdx.x; i < (RS_THREADS / 32) * 256; i += RS_THREADS) (&wh[0][0])[i] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * RS_TILE + w * (32 * RS_ITEMS);
    uint64_t k[RS_ITEMS]; uint32_t rank[RS_ITEMS];
#pragma unroll
    for (int i = 0; i < RS_ITEMS; i++) {
        const uint32_t idx = base + i * 32 + lane;
        const bool ok = idx < n;
        k[i] = ok ? keys[idx] : ~0ull;
        const uint32_t d = ok ? ((uint32_t)(k[i] >> shift) & 0xffu) : 256u;
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const uint32_t before = __popc(peers & ((1u << lane) - 1u));
        uint32_t cnt = 0;
        if (ok) cnt = wh[w][d];
        __syncwarp();
        if (ok && before == 0) wh[w][d] = cnt + __popc(peers);
        __syncwarp();
        rank[i] = cnt + before;
    }
    __syncthreads();
    {   // digit = threadIdx.x: turn per-warp counts into global write offsets
        const int d = threadIdx.x;
        uint32_t run = dbase[d] + block_off[d * nb + blockIdx.x];
        for (int ww = 0; ww < RS_THREADS / 32; ww++) { uint32_t c = wh[ww][d]; wh[ww][d] = run; run += c; }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < RS_ITEMS; i++) {
        const uint32_t idx = base + i * 32 + lane;
        if (idx < n) {
            const uint32_t d = (uint32_t)(k[i] >> shift) & 0xffu;
            const uint32_t pos = wh[w][d] + rank[i];
            keys_out[pos] = k[i];
            vals_out[pos] = vals[idx];
        }
    }
}

// ---------------------------------------------------------------------------------------------- Karras 2012
// common-prefix length of sorted keys i and j; duplicates are disambiguated by the index
HJ_DI int delta(const uint64_t* __restrict__ k, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const uint64_t a = k[i], b = k[j];
    if (a == b) return 64 + __clz((uint32_t)i ^ (uint32_t)j);
    return __clzll((long long)(a ^ b));
}

// internal node i in [0,N-1); node ids: internal i -> i, leaf j -> N-1+j
__global__ void k_karras(const uint64_t* __restrict__ keys, int n, uint32_t* left, uint32_t* right, uint32_t* parent, uint32_t* cnt) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1) if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
        if (t == 1) break;
    }
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const uint32_t L = (lo == gamma) ? (uint32_t)(n - 1 + gamma) : (uint32_t)gamma;
    const uint32_t R = (hi == gamma + 1) ? (uint32_t)(n - 1 + gamma + 1) : (uint32_t)(gamma + 1);
    left[i] = L; right[i] = R; parent[L] = (uint32_t)i; parent[R] = (uint32_t)i;
    cnt[i] = (uint32_t)(hi - lo + 1);
    if (i == 0) parent[0] = 0xffffffffu;
}

// 1.0f when the triangle's material emits (the test finish_mis applies to a hit, hj_wavefront.cu): carried in the .w of the
// binary boxes (max over children) so that the collapse can mark the wide-node children that contain emitters
HJ_DI float tri_emits(const DevScene& S, uint32_t prim) {
    return S.materials[__float_as_uint(S.tri_attr[4 * (size_t)prim + 3].w)].is_light ? 1.0f : 0.0f;
}

__global__ void k_fit(const DevScene S, const uint32_t* __restrict__ sorted, int n, const uint32_t* __restrict__ left, const uint32_t* __restrict__ right,
                      const uint32_t* __restrict__ parent, uint32_t* flags, float4* blo, float4* bhi) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const float4* P = S.tri_pos + 3 * (size_t)sorted[j];
    const float3 a = f3(P[0]), b = f3(P[1]), c = f3(P[2]);
    float3 lo = f3(fminf(a.x, fminf(b.x, c.x)), fminf(a.y, fminf(b.y, c.y)), fminf(a.z, fminf(b.z, c.z)));
    float3 hi = f3(fmaxf(a.x, fmaxf(b.x, c.x)), fmaxf(a.y, fmaxf(b.y, c.y)), fmaxf(a.z, fmaxf(b.z, c.z)));
    uint32_t id = (uint32_t)(n - 1 + j);
    float em = tri_emits(S, sorted[j]);
    blo[id] = f4(lo, em); bhi[id] = f4(hi, 0.0f);
    if (n == 1) return;
    uint32_t p = parent[id];
    while (p != 0xffffffffu) {
        __threadfence();
        if (atomicAdd(&flags[p], 1u) == 0u) return;     // the sibling subtree is not finished: it will carry on
        const uint32_t l = left[p], r = right[p];
        const float4 ll = __ldcg(blo + l), lh = __ldcg(bhi + l), rl = __ldcg(blo + r), rh = __ldcg(bhi + r);   // L2 loads: written by other SMs
        lo = f3(fminf(ll.x, rl.x), fminf(ll.y, rl.y), fminf(ll.z, rl.z));
        hi = f3(fmaxf(lh.x, rh.x), fmaxf(lh.y, rh.y), fmaxf(lh.z, rh.z));
        em = fmaxf(ll.w, rl.w);
        blo[p] = f4(lo, em); bhi[p] = f4(hi, 0.0f);
        p = parent[p];
    }
}

// ---------------------------------------------------------------------------------------------- PLOC
// Parallel locally-ordered clustering (Meister & Bittner 2018): bottom-up agglomeration of the Morton-ordered
// clusters. Every round each cluster looks PLOC_RADIUS neighbours to either side for the partner that minimises the
// surface area of the merged box; mutual nearest neighbours merge. Builds a much better tree than the Karras radix
// tree over the same order (fewer node visits per ray, no pathological long rays through flat tessellated surfaces).
//
// r02: the whole agglomeration is ONE cooperative kernel (grid-wide barriers between the phases of a round) followed by a
// single-block tail for the last <= 1024 clusters. Round 1 ran every round as four launches and read the cluster count back
// to the host after each (one cudaStreamSynchronize per round, ~20 rounds): that made hj_accel_build a blocking call and
// kept it from running under the previous frame's shading. Internal node ids are now assigned by prefix
// sums instead of a global atomic, so the same input gives the same tree layout in every run.
#ifndef PLOC_RADIUS
#define PLOC_RADIUS 8      // r01-u sweep on C2 (2..100): 6-8 give the fewest node visits and the shortest build
#endif
#define PLOC_THREADS 256
#define PLOC_TAIL 1024


__global__ void k_ploc_init(const DevScene S, const uint32_t* __restrict__ sorted, int n, uint32_t* cid, float4* clo, float4* chi, float4* blo, float4* bhi) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const float4* P = S.tri_pos + 3 * (size_t)sorted[j];
    const float3 a = f3(P[0]), b = f3(P[1]), c = f3(P[2]);
    const float4 lo = make_float4(fminf(a.x, fminf(b.x, c.x)), fminf(a.y, fminf(b.y, c.y)), fminf(a.z, fminf(b.z, c.z)), tri_emits(S, sorted[j]));
    const float4 hi = make_float4(fmaxf(a.x, fmaxf(b.x, c.x)), fmaxf(a.y, fmaxf(b.y, c.y)), fmaxf(a.z, fmaxf(b.z, c.z)), 0.0f);
    const uint32_t id = (uint32_t)(n - 1 + j);
    cid[j] = id; clo[j] = lo; chi[j] = hi; blo[id] = lo; bhi[id] = hi;
}

HJ_DI float merged_area(const float4& alo, const float4& ahi, const float4& blo, const float4& bhi) {
    const float dx = fmaxf(ahi.x, bhi.x) - fminf(alo.x, blo.x), dy = fmaxf(ahi.y, bhi.y) - fminf(alo.y, blo.y),
                dz = fmaxf(ahi.z, bhi.z) - fminf(alo.z, blo.z);
    return dx * dy + dy * dz + dz * dx;
}

// nearest neighbour of cluster i inside the window; ties resolved by the (min index, max index) pair so that "nearest" is a
// strict total order on pairs and the globally closest pair is always mutual (guarantees progress on finite boxes)
HJ_DI int ploc_nearest(int i, int cur, const float4* clo, const float4* chi) {   // no __restrict__: other blocks wrote these arrays in the previous round (no ld.global.nc)
    const float4 lo = clo[i], hi = chi[i];
    float best = FLT_MAX; int bj = -1;
    const int j0 = max(0, i - PLOC_RADIUS), j1 = min(cur - 1, i + PLOC_RADIUS);
    for (int j = j0; j <= j1; j++) {
        if (j == i) continue;
        const float d = merged_area(lo, hi, clo[j], chi[j]);
        bool better = d < best;
        if (d == best && bj >= 0) {
            const int a0 = min(i, j), a1 = max(i, j), b0 = min(i, bj), b1 = max(i, bj);
            better = a0 < b0 || (a0 == b0 && a1 < b1);
        }
        if (better) { best = d; bj = j; }
    }
    return bj;
}

struct PlocArgs {
    int n, tail_at;
    uint32_t* cid[2]; float4* clo[2]; float4* chi[2];
    int* nn; uint32_t *bkeep, *bmerge;             // per-block keep / merge counts of the current round
    uint32_t *left, *right, *parent, *cnt; float4 *blo, *bhi;
    BuildStatus* st;
};

// the merge of clusters (a, b) into the new internal node `id`
HJ_DI void ploc_make_node(const PlocArgs& A, uint32_t id, uint32_t a, uint32_t b, const float4& alo, const float4& ahi, const float4& xlo, const float4& xhi,
                          float4& lo, float4& hi) {
    lo = make_float4(fminf(alo.x, xlo.x), fminf(alo.y, xlo.y), fminf(alo.z, xlo.z), fmaxf(alo.w, xlo.w));
    hi = make_float4(fmaxf(ahi.x, xhi.x), fmaxf(ahi.y, xhi.y), fmaxf(ahi.z, xhi.z), 0.0f);
    const uint32_t first_leaf = (uint32_t)(A.n - 1);
    A.left[id] = a; A.right[id] = b; A.parent[a] = id; A.parent[b] = id;
    A.cnt[id] = (a >= first_leaf ? 1u : A.cnt[a]) + (b >= first_leaf ? 1u : A.cnt[b]);
    A.blo[id] = lo; A.bhi[id] = hi;
}

// block-wide sum of two counters (every thread gets both totals)
HJ_DI void block_sum2(uint32_t& x, uint32_t& y, uint32_t* s /* [2 * warps + 2] */) {
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    const uint32_t sx = __reduce_add_sync(0xffffffffu, x), sy = __reduce_add_sync(0xffffffffu, y);
    __syncthreads();
    if (lane == 0) { s[w] = sx; s[nw + w] = sy; }
    __syncthreads();
    uint32_t tx = 0, ty = 0;
    for (int k = 0; k < nw; k++) { tx += s[k]; ty += s[nw + k]; }
    x = tx; y = ty;
}

// Every block owns a contiguous chunk of the cluster array, so a stable compaction needs only the counts of the blocks before it.
__global__ void __launch_bounds__(PLOC_THREADS) k_ploc_coop(PlocArgs A) {
    cg::grid_group grid = cg::this_grid();
    __shared__ uint32_t s_red[2 * (PLOC_THREADS / 32) + 2];
    __shared__ uint32_t s_wk[PLOC_THREADS / 32], s_wm[PLOC_THREADS / 32];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, w = tid >> 5;
    const uint32_t gtid = blockIdx.x * PLOC_THREADS + tid, gsize = gridDim.x * PLOC_THREADS;
    uint32_t cur = (uint32_t)A.n, node_base = 0, rounds = 0, err = 0; int pp = 0;
    while (cur > (uint32_t)A.tail_at && cur > 1u) {
        const float4* clo = A.clo[pp]; const float4* chi = A.chi[pp]; const uint32_t* cid = A.cid[pp];
        for (uint32_t i = gtid; i < cur; i += gsize) A.nn[i] = ploc_nearest((int)i, (int)cur, clo, chi);
        grid.sync();
        uint32_t chunk = (cur + gridDim.x - 1) / gridDim.x;
        chunk = (chunk + PLOC_THREADS - 1) / PLOC_THREADS * PLOC_THREADS;
        const uint32_t c0 = min(blockIdx.x * chunk, cur), c1 = min(c0 + chunk, cur);
        uint32_t nk = 0, nm = 0;                          // clusters this block keeps / merges it leads
        for (uint32_t i = c0 + tid; i < c1; i += PLOC_THREADS) {
            const int j = A.nn[i];
            const bool mutual = j >= 0 && A.nn[j] == (int)i;
            nk += (mutual && (int)i > j) ? 0u : 1u; nm += (mutual && (int)i < j) ? 1u : 0u;
        }
        block_sum2(nk, nm, s_red);
        if (tid == 0) { A.bkeep[blockIdx.x] = nk; A.bmerge[blockIdx.x] = nm; }
        grid.sync();
        uint32_t koff = 0, moff = 0, ktot = 0, mtot = 0;  // counts of the blocks before this one, and of all blocks
        for (uint32_t b = tid; b < gridDim.x; b += PLOC_THREADS) {
            const uint32_t k = A.bkeep[b], m = A.bmerge[b];
            ktot += k; mtot += m; if (b < blockIdx.x) { koff += k; moff += m; }
        }
        block_sum2(ktot, mtot, s_red); block_sum2(koff, moff, s_red);
        uint32_t* cid2 = A.cid[pp ^ 1]; float4* clo2 = A.clo[pp ^ 1]; float4* chi2 = A.chi[pp ^ 1];
        for (uint32_t t0 = c0; t0 < c1; t0 += PLOC_THREADS) {
            const uint32_t i = t0 + tid;
            bool keep = false, lead = false; int j = -1;
            if (i < c1) { j = A.nn[i]; const bool mutual = j >= 0 && A.nn[j] == (int)i; keep = !(mutual && (int)i > j); lead = mutual && (int)i < j; }
            const uint32_t mk = __ballot_sync(0xffffffffu, keep), mm = __ballot_sync(0xffffffffu, lead);
            __syncthreads();
            if (lane == 0) { s_wk[w] = __popc(mk); s_wm[w] = __popc(mm); }
            __syncthreads();
            uint32_t bk = 0, bm = 0, tk = 0, tm = 0;
            for (uint32_t ww = 0; ww < PLOC_THREADS / 32; ww++) { if (ww < w) { bk += s_wk[ww]; bm += s_wm[ww]; } tk += s_wk[ww]; tm += s_wm[ww]; }
            if (keep) {
                const uint32_t pos = koff + bk + __popc(mk & ((1u << lane) - 1u));
                uint32_t id = cid[i]; float4 lo = clo[i], hi = chi[i];
                if (lead) {
                    const uint32_t nid = node_base + moff + bm + __popc(mm & ((1u << lane) - 1u));
                    ploc_make_node(A, nid, id, cid[j], lo, hi, clo[j], chi[j], lo, hi);
                    id = nid;
                }
                cid2[pos] = id; clo2[pos] = lo; chi2[pos] = hi;
            }
            koff += tk; moff += tm;
        }
        if (mtot == 0u) { err = HJ_BUILD_ERR_PLOC; break; }      // uniform over the grid: no pair was mutual (non-finite boxes)
        cur = ktot; node_base += mtot; pp ^= 1; rounds++;
        grid.sync();
    }
    if (grid.thread_rank() == 0) {
        A.st->cur_n = cur; A.st->node_ctr = node_base; A.st->pp = (uint32_t)pp; A.st->ploc_rounds = rounds;
        if (err) atomicOr(&A.st->error, err);
    }
}

// The last rounds of PLOC work on a handful of clusters: once they fit one block they run in shared memory, same rules, same
// prefix-sum node ids. Also sets the root's parent link. Launched unconditionally after k_ploc_coop; reads its state from *st.
__global__ void __launch_bounds__(PLOC_TAIL) k_ploc_tail(PlocArgs A) {
    __shared__ float4 s_lo[PLOC_TAIL], s_hi[PLOC_TAIL];
    __shared__ uint32_t s_cid[PLOC_TAIL];
    __shared__ int s_nn[PLOC_TAIL];
    __shared__ uint32_t s_wk[PLOC_TAIL / 32], s_wm[PLOC_TAIL / 32];
    const int i = threadIdx.x, w = i >> 5, lane = i & 31;
    if (A.st->error) return;
    int cur = (int)A.st->cur_n; const int pp = (int)A.st->pp; uint32_t node_base = A.st->node_ctr, rounds = 0;
    if (cur > PLOC_TAIL) return;                         // (ploc_tail off and more than one cluster left cannot happen: the loop above runs to 1)
    if (i < cur) { s_cid[i] = A.cid[pp][i]; s_lo[i] = A.clo[pp][i]; s_hi[i] = A.chi[pp][i]; }
    __syncthreads();
    while (cur > 1) {
        if (i < cur) s_nn[i] = ploc_nearest(i, cur, s_lo, s_hi);
        __syncthreads();
        bool keep = false, lead = false;
        uint32_t my_cid = 0, other = 0; float4 my_lo = make_float4(0, 0, 0, 0), my_hi = my_lo, xlo = my_lo, xhi = my_lo;
        if (i < cur) {
            const int j = s_nn[i];
            const bool mutual = j >= 0 && s_nn[j] == i;
            keep = !(mutual && i > j); lead = mutual && i < j;
            my_cid = s_cid[i]; my_lo = s_lo[i]; my_hi = s_hi[i];
            if (lead) { other = s_cid[j]; xlo = s_lo[j]; xhi = s_hi[j]; }
        }
        const uint32_t mk = __ballot_sync(0xffffffffu, keep), mm = __ballot_sync(0xffffffffu, lead);
        if (lane == 0) { s_wk[w] = (uint32_t)__popc(mk); s_wm[w] = (uint32_t)__popc(mm); }
        __syncthreads();                                 // also: every read of s_* of this round is done
        uint32_t bk = 0, bm = 0, tk = 0, tm = 0;
        for (int ww = 0; ww < PLOC_TAIL / 32; ww++) { if (ww < w) { bk += s_wk[ww]; bm += s_wm[ww]; } tk += s_wk[ww]; tm += s_wm[ww]; }
        if (keep) {
            if (lead) {
                const uint32_t nid = node_base + bm + (uint32_t)__popc(mm & ((1u << lane) - 1u));
                ploc_make_node(A, nid, my_cid, other, my_lo, my_hi, xlo, xhi, my_lo, my_hi);
                my_cid = nid;
            }
            const uint32_t pos = bk + (uint32_t)__popc(mk & ((1u << lane) - 1u));
            s_cid[pos] = my_cid; s_lo[pos] = my_lo; s_hi[pos] = my_hi;
        }
        __syncthreads();
        if (tm == 0u) { if (i == 0) atomicOr(&A.st->error, HJ_BUILD_ERR_PLOC); return; }     // no mutual pair: non-finite boxes; never spin
        cur = (int)tk; node_base += tm; rounds++;
    }
    if (i == 0) {
        A.parent[A.n - 2] = 0xffffffffu;                 // ids are handed out in order: the last merge made the root
        A.st->cur_n = 1u; A.st->node_ctr = node_base; A.st->ploc_rounds += rounds;
    }
}

// ---------------------------------------------------------------------------------------------- collapse to 8-wide
struct CollapseArgs {
    int n; uint32_t leaf_max; int leaf_split;
    const uint32_t *left, *right, *cnt, *sorted;
    const float4 *blo, *bhi;
    const float4* tri_pos;
    uint4* nodes; float4* btris;
    uint32_t* counters;     // [0] nodes allocated, [1] triangles allocated, [2] next-level queue length
};

HJ_DI uint32_t bn_count(const CollapseArgs& a, uint32_t id) { return id >= (uint32_t)(a.n - 1) ? 1u : a.cnt[id]; }
// the (<= 3) leaves below a binary node, left to right, as indices into the Morton-sorted triangle order. Written out
// case by case so that the result lives in three registers (no local-memory stack, see k_collapse).
HJ_DI uint32_t bn_leaves3(const CollapseArgs& a, uint32_t id, uint32_t& l0, uint32_t& l1, uint32_t& l2) {
    const uint32_t first = (uint32_t)(a.n - 1);
    l0 = l1 = l2 = 0u;
    if (id >= first) { l0 = id - first; return 1u; }
    const uint32_t l = a.left[id], r = a.right[id];
    if (l < first) { l0 = a.left[l] - first; l1 = a.right[l] - first; l2 = r - first; return 3u; }      // ((x x) x)
    if (r < first) { l0 = l - first; l1 = a.left[r] - first; l2 = a.right[r] - first; return 3u; }      // (x (x x))
    l0 = l - first; l1 = r - first; return 2u;
}
HJ_DI float bn_area(const CollapseArgs& a, uint32_t id) {
    const float4 l = a.blo[id], h = a.bhi[id];
    const float dx = h.x - l.x, dy = h.y - l.y, dz = h.z - l.z;
    return dx * dy + dy * dz + dz * dx;
}
// grid exponent of one axis: the smallest power-of-two step with 255 steps covering [lo, hi]
HJ_DI uint32_t grid_exponent(float lo, float hi) {
    const float ext = hi - lo;
    uint32_t e = 1;
    if (ext > 0.0f) {
        const uint32_t bits = __float_as_uint(ext / 255.0f);
        e = (bits >> 23) & 0xffu;
        if (bits & 0x7fffffu) e++;
        if (e < 1u) e = 1u;
    }
    while (e < 254u && lo + 255.0f * __uint_as_float(e << 23) < hi) e++;
    return e;
}
// conservative 8-bit grid coordinates of [clo, chi] on the grid (lo, scale)
HJ_DI uint32_t grid_floor(float clo, float lo, float scale) {
    int q = (int)floorf((clo - lo) / scale);
    q = max(0, min(255, q));
    while (q > 0 && lo + (float)q * scale > clo) q--;
    return (uint32_t)q;
}
HJ_DI uint32_t grid_ceil(float chi, float lo, float scale) {
    int q = (int)ceilf((chi - lo) / scale);
    q = max(0, min(255, q));
    while (q < 255 && lo + (float)q * scale < chi) q++;
    return (uint32_t)q;
}

// One thread turns one binary subtree root into one wide node: repeatedly opens the child with the largest
// surface area until 8 children (or nothing left to open), orders the children into octant slots, quantises
// their boxes conservatively on the node's 8-bit grid and copies the triangles of its leaf children.
//
// Per-thread state that is indexed dynamically (the 8 children and their boxes) lives in SHARED memory, everything else
// in scalar registers and bit fields. r01-w: the earlier form kept ~600 B of dynamically indexed local arrays; a small
// source change made nvcc 12.9 produce a kernel whose child boxes were overwritten while still live (children's
// quantised boxes no longer contained their triangles -> missed hits). k_verify (option verify_accel) checks the result.
#define COLLAPSE_THREADS 128
struct CollapseShared {
    uint32_t ch[8][COLLAPSE_THREADS];
    float lo[3][8][COLLAPSE_THREADS], hi[3][8][COLLAPSE_THREADS];
};
HJ_DI void collapse_one(const CollapseArgs& a, const uint2 work, uint2* __restrict__ qout, uint32_t* qlen_out, CollapseShared& sh) {
    uint32_t (&s_ch)[8][COLLAPSE_THREADS] = sh.ch;
    float (&s_lo)[3][8][COLLAPSE_THREADS] = sh.lo; float (&s_hi)[3][8][COLLAPSE_THREADS] = sh.hi;
    const uint32_t tid = threadIdx.x;
    const uint32_t first_leaf = (uint32_t)(a.n - 1);
    // Phase 0 opens the child with the largest surface area among those holding more than leaf_max triangles. Phase 1
    // (leaf_split): slots left over -> split multi-triangle leaf children too, largest box first: the node test looks at
    // all 8 slots anyway, so tighter per-triangle boxes cost nothing and save triangle tests.
    int nc = 1; s_ch[0][tid] = work.x;
    bool split = false;
    while (nc < 8) {
        int best = -1; float ba = -1.0f;
        for (int k = 0; k < nc; k++) {
            const uint32_t id = s_ch[k][tid];
            const bool open = id < first_leaf && (split || bn_count(a, id) > a.leaf_max || nc == 1);
            if (open) { const float ar = bn_area(a, id); if (ar > ba) { ba = ar; best = k; } }
        }
        if (best < 0) {
            if (split || !a.leaf_split) break;
            split = true; continue;
        }
        const uint32_t id = s_ch[best][tid];
        s_ch[best][tid] = a.left[id]; s_ch[nc][tid] = a.right[id]; nc++;
    }
    // child boxes and the node box
    float lox = FLT_MAX, loy = FLT_MAX, loz = FLT_MAX, hix = -FLT_MAX, hiy = -FLT_MAX, hiz = -FLT_MAX;
    for (int k = 0; k < nc; k++) {
        const uint32_t id = s_ch[k][tid];
        const float4 l = a.blo[id], h = a.bhi[id];
        s_lo[0][k][tid] = l.x; s_lo[1][k][tid] = l.y; s_lo[2][k][tid] = l.z;
        s_hi[0][k][tid] = h.x; s_hi[1][k][tid] = h.y; s_hi[2][k][tid] = h.z;
        lox = fminf(lox, l.x); loy = fminf(loy, l.y); loz = fminf(loz, l.z);
        hix = fmaxf(hix, h.x); hiy = fmaxf(hiy, h.y); hiz = fmaxf(hiz, h.z);
    }
    // octant slots: greedy assignment maximising (child centre - node centre) . (+-1,+-1,+-1).
    // slot_of: nibble k = slot of child k; child_in: nibble s = child in slot s (0xf: empty)
    uint32_t slot_of = 0u, slots_used = 0u, done = 0u;
    const float ncx = (lox + hix) * 0.5f, ncy = (loy + hiy) * 0.5f, ncz = (loz + hiz) * 0.5f;
    for (int round = 0; round < nc; round++) {
        float bc = -FLT_MAX; int bk = -1, bs = -1;
        for (int k = 0; k < nc; k++) {
            if (done & (1u << k)) continue;
            const float c0 = (s_lo[0][k][tid] + s_hi[0][k][tid]) * 0.5f - ncx, c1 = (s_lo[1][k][tid] + s_hi[1][k][tid]) * 0.5f - ncy,
                        c2 = (s_lo[2][k][tid] + s_hi[2][k][tid]) * 0.5f - ncz;
            for (int sl = 0; sl < 8; sl++) {
                if (slots_used & (1u << sl)) continue;
                const float cost = ((sl & 1) ? c0 : -c0) + ((sl & 2) ? c1 : -c1) + ((sl & 4) ? c2 : -c2);
                if (cost > bc) { bc = cost; bk = k; bs = sl; }
            }
        }
        if (bk < 0) {                                    // only with non-finite boxes: first free child into the first free slot
            bk = __ffs((int)(~done & ((1u << nc) - 1u))) - 1; bs = __ffs((int)(~slots_used & 0xffu)) - 1;
        }
        slot_of |= (uint32_t)bs << (4 * bk); slots_used |= 1u << bs; done |= 1u << bk;
    }
    uint32_t child_in = 0xffffffffu;
    for (int k = 0; k < nc; k++) {
        const uint32_t sl = (slot_of >> (4 * k)) & 7u;
        child_in = (child_in & ~(0xfu << (4u * sl))) | ((uint32_t)k << (4u * sl));
    }
    // grid: per axis the smallest power-of-two step with 255 steps covering the node
    const uint32_t ebx = grid_exponent(lox, hix), eby = grid_exponent(loy, hiy), ebz = grid_exponent(loz, hiz);
    const float scx = __uint_as_float(ebx << 23), scy = __uint_as_float(eby << 23), scz = __uint_as_float(ebz << 23);
    uint4* N = a.nodes + HJ_NODE_QUADS * (size_t)work.y;
    uint32_t* Nw = (uint32_t*)N;
    uint32_t imask = 0, n_int = 0, n_tri = 0, em_word = 0;
    unsigned long long qlx = 0, qly = 0, qlz = 0, qhx = 0, qhy = 0, qhz = 0;       // byte s = grid coordinate of slot s
    for (int sl = 0; sl < 8; sl++) {
        const uint32_t k = (child_in >> (4 * sl)) & 0xfu;
        // child mask of slot sl: where its hit lands in the traversal's 32-bit hit word (bits 0..23 triangles of this node,
        // bit 24+s inner child in slot s; the ray-octant permutation of the inner bits is applied by the traversal)
        uint32_t cm = 0u, q[6] = {255u, 255u, 255u, 0u, 0u, 0u};
        if (k != 0xfu) {
            const uint32_t id = s_ch[k][tid];
            const uint32_t cnt = bn_count(a, id);
            const bool emits = a.blo[id].w > 0.0f;
            if (id < first_leaf && cnt > a.leaf_max) {
                imask |= 1u << sl; n_int++; cm = 1u << (24 + sl);
                if (emits) em_word |= cm;
            } else {
                cm = ((1u << cnt) - 1u) << n_tri;
                if (emits) {                              // which of its (<= 3) triangles emit
                    uint32_t l0, l1, l2;
                    const uint32_t m = bn_leaves3(a, id, l0, l1, l2);
                    if (a.blo[first_leaf + l0].w > 0.0f) em_word |= 1u << n_tri;
                    if (m > 1u && a.blo[first_leaf + l1].w > 0.0f) em_word |= 1u << (n_tri + 1u);
                    if (m > 2u && a.blo[first_leaf + l2].w > 0.0f) em_word |= 1u << (n_tri + 2u);
                }
                n_tri += cnt;
            }
            q[0] = grid_floor(s_lo[0][k][tid], lox, scx); q[1] = grid_floor(s_lo[1][k][tid], loy, scy); q[2] = grid_floor(s_lo[2][k][tid], loz, scz);
            q[3] = grid_ceil(s_hi[0][k][tid], lox, scx); q[4] = grid_ceil(s_hi[1][k][tid], loy, scy); q[5] = grid_ceil(s_hi[2][k][tid], loz, scz);
        }
        const int sh = 8 * sl;
        qlx |= (unsigned long long)q[0] << sh; qly |= (unsigned long long)q[1] << sh; qlz |= (unsigned long long)q[2] << sh;
        qhx |= (unsigned long long)q[3] << sh; qhy |= (unsigned long long)q[4] << sh; qhz |= (unsigned long long)q[5] << sh;
        Nw[sl < 2 ? 6 + sl : 18 + sl] = cm;               // cm[0..1] = q1.zw, cm[2..5] = q5, cm[6..7] = q6.xy
    }
    const uint32_t child_base = n_int ? atomicAdd(&a.counters[0], n_int) : 0u;
    const uint32_t tri_base = n_tri ? atomicAdd(&a.counters[1], n_tri) : 0u;
    N[0] = make_uint4(__float_as_uint(lox), __float_as_uint(loy), __float_as_uint(loz), ebx | (eby << 8) | (ebz << 16) | (imask << 24));
    Nw[4] = child_base; Nw[5] = tri_base;
    N[2] = make_uint4((uint32_t)qlx, (uint32_t)(qlx >> 32), (uint32_t)qly, (uint32_t)(qly >> 32));
    N[3] = make_uint4((uint32_t)qlz, (uint32_t)(qlz >> 32), (uint32_t)qhx, (uint32_t)(qhx >> 32));
    N[4] = make_uint4((uint32_t)qhy, (uint32_t)(qhy >> 32), (uint32_t)qhz, (uint32_t)(qhz >> 32));
    Nw[26] = em_word; Nw[27] = 0u;
    // children: queue the internal ones (contiguous, in slot order), copy the triangles of the leaves (same order as above)
    const uint32_t qbase = n_int ? atomicAdd(qlen_out, n_int) : 0u;
    uint32_t r = 0, off = 0;
    for (int sl = 0; sl < 8; sl++) {
        const uint32_t k = (child_in >> (4 * sl)) & 0xfu;
        if (k == 0xfu) continue;
        const uint32_t id = s_ch[k][tid];
        if (imask & (1u << sl)) { qout[qbase + r] = make_uint2(id, child_base + r); r++; }
        else {
            uint32_t l0, l1, l2;
            const uint32_t cnt = bn_leaves3(a, id, l0, l1, l2);
            for (uint32_t j = 0; j < cnt; j++) {
                const uint32_t prim = a.sorted[j == 0u ? l0 : (j == 1u ? l1 : l2)];
                const float4* P = a.tri_pos + 3 * (size_t)prim;
                float4* T = a.btris + 3 * (size_t)(tri_base + off + j);
                float4 v0 = P[0]; v0.w = __uint_as_float(prim);
                T[0] = v0; T[1] = P[1]; T[2] = P[2];
            }
            off += cnt;
        }
    }
}

// All levels in ONE cooperative kernel: level L reads queue L & 1 (length qlen[L % 3]), appends the internal children to the other
// queue (length qlen[(L + 1) % 3]) and clears qlen[(L + 2) % 3] for the level after; a grid-wide barrier separates the levels.
// Round 1 launched one kernel per level and read the next level's length back to the host (a stream synchronisation per level).
// counters: [0] nodes allocated, [1] triangles allocated, [2..4] queue lengths.
#define COLLAPSE_MAX_LEVELS 256
__global__ void __launch_bounds__(COLLAPSE_THREADS) k_collapse_coop(CollapseArgs a, uint2* q0, uint2* q1, BuildStatus* st) {
    cg::grid_group grid = cg::this_grid();
    __shared__ CollapseShared sh;
    uint32_t* qlen = a.counters + 2;
    const uint32_t gtid = blockIdx.x * COLLAPSE_THREADS + threadIdx.x, gsize = gridDim.x * COLLAPSE_THREADS;
    if (st->error) return;                               // uniform: written before this kernel started
    uint32_t level = 0;
    for (;; level++) {
        const uint32_t n_in = *(volatile uint32_t*)&qlen[level % 3];
        if (n_in == 0u || level >= COLLAPSE_MAX_LEVELS) break;
        if (gtid == 0) qlen[(level + 2) % 3] = 0u;
        const uint2* qin = (level & 1u) ? q1 : q0; uint2* qout = (level & 1u) ? q0 : q1;
        for (uint32_t i = gtid; i < n_in; i += gsize) collapse_one(a, qin[i], qout, &qlen[(level + 1) % 3], sh);
        grid.sync();
    }
    if (gtid == 0) {
        st->n_nodes = a.counters[0]; st->n_tris = a.counters[1]; st->depth = level;
        if (level >= COLLAPSE_MAX_LEVELS) atomicOr(&st->error, HJ_BUILD_ERR_COLLAPSE);
    }
}

// option verify_accel: every leaf triangle inside its slot's decoded box, every internal child's origin inside too, child masks
// disjoint, every primitive in exactly one leaf (error bits in err[0], count of primitives not seen once in err[1])
__global__ void k_verify(const uint4* __restrict__ nodes, const float4* __restrict__ btris, uint32_t n_nodes, uint32_t n_tris, uint32_t* seen, uint32_t* err) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    const uint4* N = nodes + HJ_NODE_QUADS * (size_t)i;
    const uint4 q0 = N[0], q1 = N[1], q2 = N[2], q3 = N[3], q4 = N[4], q5 = N[5], q6 = N[6];
    const float lo[3] = {__uint_as_float(q0.x), __uint_as_float(q0.y), __uint_as_float(q0.z)};
    const float sc[3] = {__uint_as_float((q0.w & 0xffu) << 23), __uint_as_float(((q0.w >> 8) & 0xffu) << 23), __uint_as_float(((q0.w >> 16) & 0xffu) << 23)};
    const uint32_t imask = q0.w >> 24;
    const uint32_t cm[8] = {q1.z, q1.w, q5.x, q5.y, q5.z, q5.w, q6.x, q6.y};
    const uint32_t ql[3][2] = {{q2.x, q2.y}, {q2.z, q2.w}, {q3.x, q3.y}}, qh[3][2] = {{q3.z, q3.w}, {q4.x, q4.y}, {q4.z, q4.w}};
    uint32_t tri_union = 0;
    for (int s = 0; s < 8; s++) {
        float blo[3], bhi[3];
        for (int ax = 0; ax < 3; ax++) {
            blo[ax] = lo[ax] + (float)((ql[ax][s >> 2] >> (8 * (s & 3))) & 0xffu) * sc[ax];
            bhi[ax] = lo[ax] + (float)((qh[ax][s >> 2] >> (8 * (s & 3))) & 0xffu) * sc[ax];
        }
        if (imask & (1u << s)) {
            if (cm[s] != (1u << (24 + s))) atomicOr(err, 16u);
            const uint32_t child = q1.x + __popc(imask & ((1u << s) - 1u));
            if (child >= n_nodes) { atomicOr(err, 32u); continue; }
            const uint4 c0 = nodes[HJ_NODE_QUADS * (size_t)child];
            const float co[3] = {__uint_as_float(c0.x), __uint_as_float(c0.y), __uint_as_float(c0.z)};
            for (int ax = 0; ax < 3; ax++) if (co[ax] < blo[ax] || co[ax] > bhi[ax]) atomicOr(err, 64u);
        } else if (cm[s]) {
            if (cm[s] & tri_union) atomicOr(err, 128u);
            tri_union |= cm[s];
            for (uint32_t m = cm[s]; m; m &= m - 1u) {
                const uint32_t t = q1.y + (uint32_t)__ffs((int)m) - 1u;
                if (t >= n_tris) { atomicOr(err, 256u); continue; }
                const float4 v0 = btris[3 * (size_t)t], v1 = btris[3 * (size_t)t + 1], v2 = btris[3 * (size_t)t + 2];
                const uint32_t prim = __float_as_uint(v0.w);
                if (prim >= n_tris) { atomicOr(err, 512u); continue; }
                atomicAdd(&seen[prim], 1u);
                const float p[3][3] = {{v0.x, v0.y, v0.z}, {v1.x, v1.y, v1.z}, {v2.x, v2.y, v2.z}};
                for (int v = 0; v < 3; v++) for (int ax = 0; ax < 3; ax++) if (p[v][ax] < blo[ax] || p[v][ax] > bhi[ax]) atomicOr(err, 1024u);
            }
        }
    }
}
__global__ void k_verify_seen(const uint32_t* seen, uint32_t n, uint32_t* err) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && seen[i] != 1u) { atomicOr(err, 2048u); atomicAdd(err + 1, 1u); }
}

// ---------------------------------------------------------------------------------------------- host driver
static inline uint32_t cdiv(uint64_t a, uint32_t b) { return (uint32_t)((a + b - 1) / b); }

__global__ void k_build_begin(BuildStatus* st, int* bounds) {
    if (threadIdx.x == 0) { st->n_nodes = st->n_tris = st->depth = st->error = 0u; st->cur_n = st->node_ctr = st->pp = st->ploc_rounds = 0u; }
    if (threadIdx.x < 3) bounds[threadIdx.x] = 0x7fffffff;
    else if (threadIdx.x < 6) bounds[threadIdx.x] = (int)0x80000000;
}
__global__ void k_collapse_begin(uint32_t* counters, uint2* q0, uint32_t root) {
    counters[0] = 1u; counters[1] = 0u; counters[2] = 1u; counters[3] = 0u; counters[4] = 0u;
    q0[0] = make_uint2(root, 0u);
}

static int coop_grid(hj_ctx* c, const void* kernel, int threads, int& cache) {
    if (cache > 0) return cache;
    int occ = 0, sms = 148;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, threads, 0);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);
    cache = std::max(1, occ) * sms;
    return cache;
}

// General 3x4 affine inverse by cofactors in float (the reference uses glm::inverse on the 4x4, renderer.h:272-284)
static void invert_affine(const float* m, float* r) {
    const float a = m[0], b = m[1], cc = m[2], d = m[4], e = m[5], f = m[6], g = m[8], h = m[9], i = m[10];
    const float A = e * i - f * h, B = -(d * i - f * g), C = d * h - e * g;
    const float det = a * A + b * B + cc * C;
    const float id = 1.0f / det;
    r[0] = A * id; r[1] = -(b * i - cc * h) * id; r[2] = (b * f - cc * e) * id;
    r[4] = B * id; r[5] = (a * i - cc * g) * id; r[6] = -(a * f - cc * d) * id;
    r[8] = C * id; r[9] = -(a * h - b * g) * id; r[10] = (a * e - b * d) * id;
    const float tx = m[3], ty = m[7], tz = m[11];
    r[3] = -(r[0] * tx + r[1] * ty + r[2] * tz);
    r[7] = -(r[4] * tx + r[5] * ty + r[6] * tz);
    r[11] = -(r[8] * tx + r[9] * ty + r[10] * tz);
}

// Enqueues ≙ updateIASMatrix (+ buildGAS/buildIAS in mode BUILD) for the set the renderer is NOT reading, on the build stream:
// matrices H2D, flatten, [Morton, sort, PLOC | Karras + fit] or [refit: fit], collapse, status D2H. Not one host synchronisation:
// loop control (PLOC rounds, collapse levels) lives in two cooperative kernels, results come back through a pinned status word.
int hj_bvh_enqueue(hj_ctx* c, const float* xf, const float* inv_xf, int mode) {
    const uint32_t n = c->S.n_tris, n_inst = c->S.n_instances;
    if (n == 0) { c->err = "hj_accel_build: scene has no triangles"; return HJ_ERR_INVALID; }
    BvhState& B = c->bvh;
    if (!B.build_stream) {
        int lo = 0, hi = 0; cudaDeviceGetStreamPriorityRange(&lo, &hi);          // lowest priority: a build yields to a frame in flight
        HJ_CUDA(c, cudaStreamCreateWithPriority(&B.build_stream, cudaStreamNonBlocking, lo));
        HJ_CUDA(c, cudaEventCreate(&B.ev_b0)); HJ_CUDA(c, cudaEventCreate(&B.ev_b1));
        HJ_CUDA(c, cudaHostAlloc((void**)&B.h_status, sizeof(BuildStatus), cudaHostAllocDefault));
    }
    const int set = c->have_accel ? (c->accel_cur ^ 1) : c->accel_cur;
    AccelSet& A = c->accel[set];
    cudaStream_t st = B.build_stream;
    // pinned staging of the matrices (the caller's arrays are borrowed for the call only)
    if (B.h_xf_cap < 24 * (size_t)n_inst) {
        if (B.h_xf) cudaFreeHost(B.h_xf);
        B.h_xf = nullptr; B.h_xf_cap = 0;
        HJ_CUDA(c, cudaHostAlloc((void**)&B.h_xf, 24 * (size_t)n_inst * sizeof(float), cudaHostAllocDefault));
        B.h_xf_cap = 24 * (size_t)n_inst;
    }
    float* m = B.h_xf; float* im = B.h_xf + 12 * (size_t)n_inst;
    for (uint32_t i = 0; i < n_inst; i++) {
        float* p = m + 12 * (size_t)i; float* q = im + 12 * (size_t)i;
        if (xf) memcpy(p, xf + 12 * (size_t)i, 48);
        else { memset(p, 0, 48); p[0] = p[5] = p[10] = 1.0f; }
        if (inv_xf) memcpy(q, inv_xf + 12 * (size_t)i, 48);
        else if (xf) invert_affine(p, q);
        else memcpy(q, p, 48);
    }
    HJ_CUDA(c, A.xf.reserve(48 * (size_t)n_inst)); HJ_CUDA(c, A.inv_xf.reserve(48 * (size_t)n_inst));
    HJ_CUDA(c, A.tri_pos.reserve(3 * (size_t)n * sizeof(float4))); HJ_CUDA(c, A.tri_attr.reserve(4 * (size_t)n * sizeof(float4)));
    HJ_CUDA(c, A.nodes.reserve((size_t)n * 16 * HJ_NODE_QUADS)); HJ_CUDA(c, A.btris.reserve(3 * (size_t)n * sizeof(float4)));
    HJ_CUDA(c, B.bounds.reserve(8 * sizeof(int))); HJ_CUDA(c, B.status.reserve(sizeof(BuildStatus)));
    if (!A.ev_last_read) HJ_CUDA(c, cudaEventCreateWithFlags(&A.ev_last_read, cudaEventDisableTiming));
    else HJ_CUDA(c, cudaStreamWaitEvent(st, A.ev_last_read, 0));                 // frames that still read this set finish first
    HJ_CUDA(c, cudaStreamWaitEvent(st, c->ev_scene, 0));                         // the scene upload on the context stream
    DevScene S = c->S;                                                           // the scene as the build kernels see it: target set
    S.xf = A.xf.as<float4>(); S.inv_xf = A.inv_xf.as<float4>(); S.tri_pos = A.tri_pos.as<float4>(); S.tri_attr = A.tri_attr.as<float4>();
    BuildStatus* dst = B.status.as<BuildStatus>();
    HJ_CUDA(c, cudaEventRecord(B.ev_b0, st));
    HJ_CUDA(c, cudaMemcpyAsync(A.xf.p, m, 48 * (size_t)n_inst, cudaMemcpyHostToDevice, st));
    HJ_CUDA(c, cudaMemcpyAsync(A.inv_xf.p, im, 48 * (size_t)n_inst, cudaMemcpyHostToDevice, st));
    k_build_begin<<<1, 32, 0, st>>>(dst, B.bounds.as<int>());
    k_flatten<<<cdiv(n, FLATTEN_THREADS), FLATTEN_THREADS, 0, st>>>(S, B.bounds.as<int>(), &dst->error);
    const bool topo = (mode == HJ_ACCEL_BUILD) || !B.have_topology;
    uint32_t root = 0;
    if (topo) {
        const uint32_t nb = cdiv(n, RS_TILE);
        HJ_CUDA(c, B.keys.reserve((size_t)n * 8)); HJ_CUDA(c, B.keys_alt.reserve((size_t)n * 8));
        HJ_CUDA(c, B.vals.reserve((size_t)n * 4)); HJ_CUDA(c, B.vals_alt.reserve((size_t)n * 4));
        HJ_CUDA(c, B.block_hist.reserve((size_t)256 * nb * 4)); HJ_CUDA(c, B.digit_total.reserve(256 * 4));
        HJ_CUDA(c, B.left.reserve((size_t)n * 4)); HJ_CUDA(c, B.right.reserve((size_t)n * 4));
        HJ_CUDA(c, B.parent.reserve(2 * (size_t)n * 4));
        const int bpa = c->opt_morton_bits, key_bits = 3 * bpa, n_passes = (key_bits + 7) / 8;
        k_morton<<<cdiv(n, 256), 256, 0, st>>>(S, B.bounds.as<int>(), B.keys.as<uint64_t>(), B.vals.as<uint32_t>(), 21 - bpa);
        uint64_t *k0 = B.keys.as<uint64_t>(), *k1 = B.keys_alt.as<uint64_t>();
        uint32_t *v0 = B.vals.as<uint32_t>(), *v1 = B.vals_alt.as<uint32_t>();
        for (int pass = 0; pass < n_passes; pass++) {      // 8 bits per pass over the populated key bits
            k_rs_hist<<<nb, RS_THREADS, 0, st>>>(k0, n, 8 * pass, B.block_hist.as<uint32_t>(), nb);
            k_rs_scan_rows<<<256, RS_SCAN_THREADS, 0, st>>>(B.block_hist.as<uint32_t>(), nb, B.digit_total.as<uint32_t>());
            k_rs_scatter<<<nb, RS_THREADS, 0, st>>>(k0, v0, n, 8 * pass, B.block_hist.as<uint32_t>(), B.digit_total.as<uint32_t>(), nb, k1, v1);
            std::swap(k0, k1); std::swap(v0, v1);
        }
        if (n_passes & 1) {         // an odd number of passes leaves the result in the alternate buffers: the tree builders read B.keys / B.vals
            HJ_CUDA(c, cudaMemcpyAsync(B.keys.p, B.keys_alt.p, (size_t)n * 8, cudaMemcpyDeviceToDevice, st));
            HJ_CUDA(c, cudaMemcpyAsync(B.vals.p, B.vals_alt.p, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
        }
        HJ_CUDA(c, B.cnt.reserve((size_t)n * 4));
        HJ_CUDA(c, B.blo.reserve(2 * (size_t)n * sizeof(float4))); HJ_CUDA(c, B.bhi.reserve(2 * (size_t)n * sizeof(float4)));
        if (n > 1 && c->opt_builder == 0) {
            k_karras<<<cdiv(n - 1, 256), 256, 0, st>>>(B.keys.as<uint64_t>(), (int)n, B.left.as<uint32_t>(), B.right.as<uint32_t>(),
                                                       B.parent.as<uint32_t>(), B.cnt.as<uint32_t>());
        } else if (n > 1) {
            const int grid = std::min<int>(coop_grid(c, (const void*)k_ploc_coop, PLOC_THREADS, B.coop_blocks_ploc), (int)cdiv(n, PLOC_THREADS));
            HJ_CUDA(c, B.p_cid[0].reserve((size_t)n * 4)); HJ_CUDA(c, B.p_cid[1].reserve((size_t)n * 4));
            for (int k = 0; k < 2; k++) { HJ_CUDA(c, B.p_lo[k].reserve((size_t)n * 16)); HJ_CUDA(c, B.p_hi[k].reserve((size_t)n * 16)); }
            HJ_CUDA(c, B.p_nn.reserve((size_t)n * 4));
            HJ_CUDA(c, B.p_keep.reserve((size_t)B.coop_blocks_ploc * 4)); HJ_CUDA(c, B.p_bcount.reserve((size_t)B.coop_blocks_ploc * 4));
            k_ploc_init<<<cdiv(n, 256), 256, 0, st>>>(S, B.vals.as<uint32_t>(), (int)n, B.p_cid[0].as<uint32_t>(), B.p_lo[0].as<float4>(), B.p_hi[0].as<float4>(),
                                                      B.blo.as<float4>(), B.bhi.as<float4>());
            PlocArgs P;
            P.n = (int)n; P.tail_at = c->opt_ploc_tail ? PLOC_TAIL : 1;
            for (int k = 0; k < 2; k++) { P.cid[k] = B.p_cid[k].as<uint32_t>(); P.clo[k] = B.p_lo[k].as<float4>(); P.chi[k] = B.p_hi[k].as<float4>(); }
            P.nn = B.p_nn.as<int>(); P.bkeep = B.p_keep.as<uint32_t>(); P.bmerge = B.p_bcount.as<uint32_t>();
            P.left = B.left.as<uint32_t>(); P.right = B.right.as<uint32_t>(); P.parent = B.parent.as<uint32_t>(); P.cnt = B.cnt.as<uint32_t>();
            P.blo = B.blo.as<float4>(); P.bhi = B.bhi.as<float4>(); P.st = dst;
            void* args[] = {&P};
            HJ_CUDA(c, cudaLaunchCooperativeKernel((const void*)k_ploc_coop, dim3((unsigned)grid), dim3(PLOC_THREADS), args, 0, st));
            k_ploc_tail<<<1, PLOC_TAIL, 0, st>>>(P);
            root = n - 2;        // ids are handed out in order: the last merge made the root
        }
    } else root = (n > 1 && c->opt_builder == 1) ? n - 2 : 0;
    if (!topo || c->opt_builder == 0 || n == 1) {   // PLOC computes the boxes while it merges; a refit recomputes them bottom-up
        HJ_CUDA(c, B.flags.reserve((size_t)n * 4));
        HJ_CUDA(c, cudaMemsetAsync(B.flags.p, 0, (size_t)n * 4, st));
        k_fit<<<cdiv(n, 256), 256, 0, st>>>(S, B.vals.as<uint32_t>(), (int)n, B.left.as<uint32_t>(), B.right.as<uint32_t>(), B.parent.as<uint32_t>(),
                                            B.flags.as<uint32_t>(), B.blo.as<float4>(), B.bhi.as<float4>());
    }
    // collapse: all levels in one cooperative kernel
    HJ_CUDA(c, B.queue_a.reserve((size_t)n * sizeof(uint2))); HJ_CUDA(c, B.queue_b.reserve((size_t)n * sizeof(uint2)));
    HJ_CUDA(c, B.counters.reserve(8 * sizeof(uint32_t)));
    k_collapse_begin<<<1, 1, 0, st>>>(B.counters.as<uint32_t>(), B.queue_a.as<uint2>(), root);
    CollapseArgs a;
    a.n = (int)n; a.leaf_max = (uint32_t)c->opt_leaf_max; a.leaf_split = c->opt_leaf_split;
    a.left = B.left.as<uint32_t>(); a.right = B.right.as<uint32_t>(); a.cnt = B.cnt.as<uint32_t>();
    a.sorted = B.vals.as<uint32_t>(); a.blo = B.blo.as<float4>(); a.bhi = B.bhi.as<float4>(); a.tri_pos = S.tri_pos;
    a.nodes = A.nodes.as<uint4>(); a.btris = A.btris.as<float4>(); a.counters = B.counters.as<uint32_t>();
    {
        const int grid = std::min<int>(coop_grid(c, (const void*)k_collapse_coop, COLLAPSE_THREADS, B.coop_blocks_collapse), (int)cdiv(n, COLLAPSE_THREADS));
        uint2* q0 = B.queue_a.as<uint2>(); uint2* q1 = B.queue_b.as<uint2>();
        void* args[] = {&a, &q0, &q1, &dst};
        HJ_CUDA(c, cudaLaunchCooperativeKernel((const void*)k_collapse_coop, dim3((unsigned)grid), dim3(COLLAPSE_THREADS), args, 0, st));
    }
    HJ_CUDA(c, cudaMemcpyAsync(B.h_status, dst, sizeof(BuildStatus), cudaMemcpyDeviceToHost, st));
    HJ_CUDA(c, cudaEventRecord(B.ev_b1, st));
    HJ_CUDA(c, cudaGetLastError());
    B.pending_set = set; B.pending_mode = mode; B.pending_topo = topo;
    return HJ_OK;
}

// Waits for the build in flight (if any), validates what the device reported and makes its buffers the ones frames read.
int hj_bvh_finish(hj_ctx* c) {
    BvhState& B = c->bvh;
    if (B.pending_set < 0) return HJ_OK;
    const int set = B.pending_set; B.pending_set = -1;
    AccelSet& A = c->accel[set];
    const uint32_t n = c->S.n_tris;
    HJ_CUDA(c, cudaEventSynchronize(B.ev_b1));
    HJ_CUDA(c, cudaGetLastError());
    const BuildStatus h = *B.h_status;
    float ms = 0; cudaEventElapsedTime(&ms, B.ev_b0, B.ev_b1);
    auto fail = [&](const std::string& msg) { if (B.pending_topo) B.have_topology = false; c->err = msg; return HJ_ERR_INVALID; };
    if (h.error & HJ_BUILD_ERR_NONFINITE) return fail("hj_accel_build: a vertex is not finite after the instance transform (NaN / inf in the vertices or matrices)");
    if (h.error & HJ_BUILD_ERR_PLOC) return fail("hj_accel_build: PLOC made no progress (non-finite bounding boxes)");
    if (h.error & HJ_BUILD_ERR_COLLAPSE) return fail("hj_accel_build: collapse did not terminate");
    if (h.n_tris != n) return fail("hj_accel_build: internal error, leaf triangles != scene triangles");
    if (2 * h.depth > HJ_STACK) {
        char b[160]; snprintf(b, sizeof b, "hj_accel_build: wide BVH depth %u exceeds the traversal stack (%d)", h.depth, HJ_STACK);
        return fail(b);
    }
    if (B.pending_topo) { B.have_topology = true; B.ploc_rounds = (int)h.ploc_rounds; }
    if (c->opt_verify_accel) {      // structural check of the wide BVH (k_verify); the tests switch it on
        cudaStream_t st = B.build_stream;
        uint32_t e[2] = {0, 0};
        HJ_CUDA(c, B.verr.reserve(8)); HJ_CUDA(c, B.flags.reserve((size_t)n * 4));
        HJ_CUDA(c, cudaMemsetAsync(B.flags.p, 0, (size_t)n * 4, st)); HJ_CUDA(c, cudaMemsetAsync(B.verr.p, 0, 8, st));
        k_verify<<<cdiv(h.n_nodes, 128), 128, 0, st>>>(A.nodes.as<uint4>(), A.btris.as<float4>(), h.n_nodes, n, B.flags.as<uint32_t>(), B.verr.as<uint32_t>());
        k_verify_seen<<<cdiv(n, 256), 256, 0, st>>>(B.flags.as<uint32_t>(), n, B.verr.as<uint32_t>());
        HJ_CUDA(c, cudaMemcpyAsync(e, B.verr.p, 8, cudaMemcpyDeviceToHost, st));
        HJ_CUDA(c, cudaStreamSynchronize(st));
        if (e[0]) {
            char b[200]; snprintf(b, sizeof b, "hj_accel_build: verify_accel found an inconsistent BVH (error bits 0x%x, %u primitives not in exactly one leaf)", e[0], e[1]);
            return fail(b);
        }
    }
    A.n_nodes = h.n_nodes; A.depth = h.depth;
    c->accel_cur = set;
    c->S.xf = A.xf.as<float4>(); c->S.inv_xf = A.inv_xf.as<float4>(); c->S.tri_pos = A.tri_pos.as<float4>(); c->S.tri_attr = A.tri_attr.as<float4>();
    c->S.nodes = A.nodes.as<uint4>(); c->S.btris = A.btris.as<float4>();
    c->have_accel = true;
    c->stats.ms_accel = ms; c->stats.bvh_nodes = A.n_nodes; c->stats.bvh_node_bytes = 16 * HJ_NODE_QUADS; c->stats.bvh_depth = A.depth;
    return HJ_OK;
}

void hj_bvh_release(hj_ctx* c) {
    BvhState& B = c->bvh;
    if (B.build_stream) { cudaStreamSynchronize(B.build_stream); cudaStreamDestroy(B.build_stream); B.build_stream = nullptr; }
    if (B.ev_b0) cudaEventDestroy(B.ev_b0);
    if (B.ev_b1) cudaEventDestroy(B.ev_b1);
    if (B.h_status) cudaFreeHost(B.h_status);
    if (B.h_xf) cudaFreeHost(B.h_xf);
    for (AccelSet& A : c->accel) {
        DevBuf* bufs[] = {&A.xf, &A.inv_xf, &A.tri_pos, &A.tri_attr, &A.nodes, &A.btris};
        for (DevBuf* b : bufs) b->release();
        if (A.ev_last_read) cudaEventDestroy(A.ev_last_read);
    }
}

// hj_trace.cuh — ray traversal of the compressed 8-wide BVH: the B200 replacement for the reference's
// optixTrace calls (/root/reference/include/kernel/rt.h:25 TraceOcculution, :53 RayTrace), which run on RT
// cores inside the closed-source OptiX runtime. B200 has no RT cores: this is plain sm_100a CUDA.
//
// Node (112 B = seven 16-byte words; built by hj_bvh.cu):
//   q0: origin.xyz (float bits) | ex | ey<<8 | ez<<16 | imask<<24      (e* = biased float exponents of the grid step)
//   q1: child_base (first internal child node), tri_base (first triangle of this node's leaves), cm[0], cm[1]
//   q2: qlo.x[0..7] qlo.y[0..7]   q3: qlo.z[0..7] qhi.x[0..7]   q4: qhi.y[0..7] qhi.z[0..7]      (8-bit grid coordinates)
//   q5: cm[2..5]   q6: cm[6], cm[7], em, 0
//       cm[s] = child mask of slot s: ((1 << count) - 1) << offset  leaf with 1..3 triangles at tri_base + offset
//                                     1 << (24 + s)                  internal child (node = child_base + popc(imask below s))
//                                     0                              empty slot
//       so a hit ORs cm[s] into a 32-bit word: bits 0..23 = triangles of this node, bits 24..31 = internal children
//       (the hit word of Ylitie, Karras, Laine 2017, precomputed per child instead of shifted together per visit).
//       em = the bits of that hit word whose child contains / whose triangle is an emitter (material.is_light): the
//       emitter phase of the MIS query (QUERY 2 below) ANDs its hit word with it.
// Child slots are assigned at build time by octant (slot bit k set = child on the + side of axis k), so
// `slot ^ ray_octant` orders children front to back without sorting (Ylitie et al. 2017).
//
// The triangle test is the watertight test of Woop/Benthin/Wald 2013 without FMA contraction: a pure
// function of (ray, triangle), so the closest hit — ties on t broken by the lower global primitive id — does
// not depend on traversal order and is bit-identical to a brute-force loop over all triangles.
#pragma once
#include "hj_scene.cuh"

#define HJ_STACK 96          // local-memory part of the traversal stack
// Shared-memory part of the stack (entries per thread). What the 8 resident blocks do not take stays L1, in steps of the
// carve-out (ncu launch__shared_mem_config_size): the 20.5 KB layout of r01-v (8 entries) got 196 KiB, i.e. 60 KB of L1;
// 8 x (15.4 KB + 1 KB reserved) = 131 KB fits the 132 KiB step and leaves 124 KB: L1 hit 51 -> 63 %, C2 -1.2 %, C5 -0.9 %.
// 2 entries lose it again to local-memory traffic. The queries without the key0 array have room for a fourth entry.
#ifndef HJ_SSTACK_Q2
#define HJ_SSTACK_Q2 3
#endif
#ifndef HJ_SSTACK
#define HJ_SSTACK 4
#endif
#define HJ_TRACE_THREADS 128 // block size of every kernel that traverses
#define HJ_NODE_QUADS 7      // 16-byte words per node (112 B)

struct RayPrep {
    float3 o;
    int kx, ky, kz; float Sx, Sy, Sz;     // watertight shear
    float3 idir; uint32_t oct;            // safe reciprocal direction, sign bits
};

HJ_DI RayPrep prep_ray(const float3& o, const float3& d) {
    RayPrep r; r.o = o;
    float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    r.kz = (ax > ay) ? ((ax > az) ? 0 : 2) : ((ay > az) ? 1 : 2);
    r.kx = r.kz + 1; if (r.kx == 3) r.kx = 0;
    r.ky = r.kx + 1; if (r.ky == 3) r.ky = 0;
    if (comp(d, r.kz) < 0.0f) { int t = r.kx; r.kx = r.ky; r.ky = t; }
    float dz = comp(d, r.kz);
    r.Sx = comp(d, r.kx) / dz; r.Sy = comp(d, r.ky) / dz; r.Sz = 1.0f / dz;
    // box test only: a zero component gets a huge finite reciprocal so that 0*inf never appears
    r.idir.x = (ax > 1e-30f) ? 1.0f / d.x : copysignf(1e30f, d.x);
    r.idir.y = (ay > 1e-30f) ? 1.0f / d.y : copysignf(1e30f, d.y);
    r.idir.z = (az > 1e-30f) ? 1.0f / d.z : copysignf(1e30f, d.z);
    r.oct = (r.idir.x < 0.0f ? 1u : 0u) | (r.idir.y < 0.0f ? 2u : 0u) | (r.idir.z < 0.0f ? 4u : 0u);   // from idir: -0.0 counts as negative
    return r;
}

HJ_DI bool tri_test(const RayPrep& r, const float3& p0, const float3& p1, const float3& p2,
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

struct Hit { float t, b1, b2; uint32_t prim; };   // prim == 0xffffffff: miss

// k4b = bits of 2^23, which must sit in a REGISTER: PRMT takes one immediate, and that has to be the byte selector
// (r01-g SASS: with a literal 2^23 ptxas kept the four selectors in registers instead, 45 extra moves per node).
static __constant__ uint32_t c_two23 = 0x4B000000u;
HJ_DI uint32_t load_two23() { return *(const volatile uint32_t*)&c_two23; }

// Intersects the 8 quantised child boxes of one 112-byte node with a ray. A hit of slot s ORs the node's precomputed
// child mask cm[s] into a 32-bit word: bits 0..23 triangles (relative to tri_base), bit 24+s inner child s. The inner
// bits are then permuted by the ray octant (bit 24 + (s ^ oct_inv), so that the highest bit is the nearest child) with
// one lookup in a 2 KB shared-memory table. r01-p SASS had the masks assembled from 8 meta bytes with ~53 ALU-pipe
// instructions (extract, shift, select) per node; the ALU pipe (PRMT, FMNMX, SHF, LOP3, SEL at one warp instruction
// per 2 cycles) is what bounds the node test, so those moved into the node: 80 -> 112 B, -40 ALU instructions.
template <bool EM>
HJ_DI void node_boxes(const uint4* __restrict__ nodes, uint32_t node, const float3& o, const float3& idir, uint32_t oct,
                      float tmin, float tcur, uint32_t k4b, const uint8_t* perm_row, bool em, uint2& ng, uint2& tg) {
    const uint4* N = nodes + HJ_NODE_QUADS * (size_t)node;
    const uint4 q0 = __ldg(N), q1 = __ldg(N + 1), q2 = __ldg(N + 2), q3 = __ldg(N + 3), q4 = __ldg(N + 4), q5 = __ldg(N + 5);
    uint4 q6;
    if (EM) q6 = __ldg(N + 6); else { const uint2 t = __ldg((const uint2*)(N + 6)); q6 = make_uint4(t.x, t.y, 0u, 0u); }
    const float sx = __uint_as_float((q0.w & 0xffu) << 23), sy = __uint_as_float(((q0.w >> 8) & 0xffu) << 23),
                sz = __uint_as_float(((q0.w >> 16) & 0xffu) << 23);
    const float ax = sx * idir.x, ay = sy * idir.y, az = sz * idir.z;
    const float ox = (__uint_as_float(q0.x) - o.x) * idir.x, oy = (__uint_as_float(q0.y) - o.y) * idir.y,
                oz = (__uint_as_float(q0.z) - o.z) * idir.z;
    // Plane distance t = q*a + o with q an 8-bit grid coordinate. q enters as the float 2^23+q (one PRMT, no
    // int->float conversion) and the 2^23*a is folded into o: t = fma(2^23+q, a, o - 2^23 a). That costs half a grid
    // cell of accuracy, so every slab is widened by one cell |a| plus a slack proportional to the magnitudes involved:
    // rounding here can never cull a triangle that the (exact) triangle test accepts.
    const float slack = 2e-6f * (fabsf(ox) + fabsf(oy) + fabsf(oz) + 255.0f * (fabsf(ax) + fabsf(ay) + fabsf(az)));
    const float bx = 8388608.0f * ax, by = 8388608.0f * ay, bz = 8388608.0f * az;
    const float oxn = (ox - slack - fabsf(ax)) - bx, oyn = (oy - slack - fabsf(ay)) - by, ozn = (oz - slack - fabsf(az)) - bz;
    const float oxf = (ox + slack + fabsf(ax)) - bx, oyf = (oy + slack + fabsf(ay)) - by, ozf = (oz + slack + fabsf(az)) - bz;
    const bool nx = (oct & 1u) != 0, ny = (oct & 2u) != 0, nz = (oct & 4u) != 0;
    const uint32_t cm[8] = {q1.z, q1.w, q5.x, q5.y, q5.z, q5.w, q6.x, q6.y};
    uint32_t hitmask = 0;
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const uint32_t lox = h ? q2.y : q2.x, loy = h ? q2.w : q2.z, loz = h ? q3.y : q3.x;
        const uint32_t hix = h ? q3.w : q3.z, hiy = h ? q4.y : q4.x, hiz = h ? q4.w : q4.z;
        const uint32_t nearx = nx ? hix : lox, farx = nx ? lox : hix;
        const uint32_t neary = ny ? hiy : loy, fary = ny ? loy : hiy;
        const uint32_t nearz = nz ? hiz : loz, farz = nz ? loz : hiz;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint32_t sel = 0x7650u | (uint32_t)j;
            const float tnx = __fmaf_rn(__uint_as_float(__byte_perm(nearx, k4b, sel)), ax, oxn);
            const float tny = __fmaf_rn(__uint_as_float(__byte_perm(neary, k4b, sel)), ay, oyn);
            const float tnz = __fmaf_rn(__uint_as_float(__byte_perm(nearz, k4b, sel)), az, ozn);
            const float tfx = __fmaf_rn(__uint_as_float(__byte_perm(farx, k4b, sel)), ax, oxf);
            const float tfy = __fmaf_rn(__uint_as_float(__byte_perm(fary, k4b, sel)), ay, oyf);
            const float tfz = __fmaf_rn(__uint_as_float(__byte_perm(farz, k4b, sel)), az, ozf);
            const float tn = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, tmin));
            const float tf = fminf(fminf(tfx, tfy), fminf(tfz, tcur));
            if (tn <= tf) hitmask |= cm[4 * h + j];      // empty slots have cm = 0
        }
    }
    if (EM && em) hitmask &= q6.z;                 // emitter phase: only children that contain / triangles that are emitters
    ng = make_uint2(q1.x, ((uint32_t)perm_row[hitmask >> 24] << 24) | (q0.w >> 24));
    tg = make_uint2(q1.y, hitmask & 0x00ffffffu);
}

// perm[o][x]: bit s of x moved to bit s ^ o (o = ray octant ^ 7); filled by every traversal kernel at its start
HJ_DI void fill_octant_perm(uint8_t (*perm)[256]) {
    for (uint32_t i = threadIdx.x; i < 8u * 256u; i += blockDim.x) {
        const uint32_t o = i >> 8, x = i & 255u;
        uint32_t y = 0;
#pragma unroll
        for (uint32_t s = 0; s < 8u; s++) if (x & (1u << s)) y |= 1u << (s ^ o);
        perm[o][x] = (uint8_t)y;
    }
    __syncthreads();
}

// ------------------------------------------------------------------------------------------------------------------
// Warp-cooperative traversal (the loop every traversal kernel runs).
//
// Lanes own rays and walk the node hierarchy; the triangle tests they generate are NOT run by the owning lane: every
// lane with leaf hits puts one record into a per-warp ring in shared memory and the warp expands the records and tests
// 32 (owner lane, triangle) pairs at a time, one per lane, whoever owns them. The owner's ray (origin + watertight shear) and its current best hit live in shared memory;
// a tester merges a hit with one 64-bit atomicMin on the key (t bits << 32 | global primitive id) — exactly the
// closest-hit rule (smaller t, ties -> lower id), so the result is still a pure function of (ray, scene). Lanes never
// wait inside a triangle phase; r01-e measured 13-15 of 32 lanes per instruction with per-lane triangle tests.
//   * a lane is finished when its node walk is done AND every pair it queued has been tested (ticket <= head);
//   * rays are fetched dynamically: once REFILL_MIN lanes are idle the warp takes that many queue entries with one
//     warp-aggregated atomicAdd.
#define HJ_RCAP 64           // records per warp: at most 32 new per node phase + fewer than 32 left over
#ifndef REFILL_MIN
#define REFILL_MIN 8         // idle lanes before a warp fetches new rays
#endif
#ifndef TRACE_BURST
#define TRACE_BURST 4        // node phases between refill checks
#endif
#ifndef WAIT_MAX
#define WAIT_MAX 4           // lanes waiting for their last triangle tests that force a partial batch
#endif
#ifndef HJ_STEAL
#define HJ_STEAL 1           // intra-warp work stealing at the tail of the queue
#endif
#ifndef STEAL_ALWAYS
#define STEAL_ALWAYS 0        // 1: also while the queue still has rays (lanes waiting for the next refill)
#endif
#ifndef STEAL_MIN_IDLE
#define STEAL_MIN_IDLE 1     // idle lanes before a warp steals (r01-w sweep 1/16/24: 1 is best on C2, C3 and C5)
#endif
#ifndef TRI_K
#define TRI_K 1              // a batch runs once pending pairs >= min(32, TRI_K x lanes doing node work). r02 sweep: 1 beats 2 by 0.9 % (C2, C3), 1.3 % (C5)
#endif

template <int QUERY>
struct TraceShared {
    static constexpr int SS = QUERY == 2 ? HJ_SSTACK_Q2 : HJ_SSTACK;
    uint2 stack[(SS ? SS : 1) * HJ_TRACE_THREADS];
    uint32_t ray[9][HJ_TRACE_THREADS];              // ox oy oz Sx Sy Sz (kx | ky<<2 | kz<<4) tmin tmax
    unsigned long long key[HJ_TRACE_THREADS];       // best hit: t bits << 32 | prim (miss: tmax bits << 32 | 0xffffffff)
    float2 bary[HJ_TRACE_THREADS];                  // barycentrics of the best hit
    uint8_t perm[8][256];                           // octant permutation of the inner-child hit bits (fill_octant_perm)
    uint2 rring[HJ_TRACE_THREADS / 32][HJ_RCAP];    // pending leaf records: first triangle (leaf order), hit bits | owner lane << 24
    unsigned long long key0[QUERY == 2 ? HJ_TRACE_THREADS : 1];   // QUERY 2: the emitter phase's result = the key the occlusion phase starts from
    uint32_t helpers[HJ_TRACE_THREADS];             // lanes currently walking a stolen subtree of this thread's ray (HJ_STEAL)
};

// QUERY 0: closest hit. 1: any hit (occlusion). 2: "is the closest hit an emitter, and which" — the MIS emitter ray
// (rt.h:383-420 uses its closest hit only if that is a light; anything else contributes nothing). Answered in two walks
// of the same lane: (a) closest hit among EMITTERS only (hit words ANDed with the nodes' em bits: a few node visits),
// giving the key (t_e, prim_e) or (tmax, none); (b) an any-hit walk for a key below it — smaller t, or the same t and a
// lower primitive id, exactly the closest-hit order — which ends at the first such triangle. Result: Hit of the emitter /
// miss (nothing at all: sky) / prim = HJ_HIT_BLOCKED. Same answer as QUERY 0 whenever that one's hit is an emitter or a miss.
#define HJ_HIT_BLOCKED 0xfffffffeu

// IO: uint32_t count; bool load(idx, o, d, tmin, tmax) (false: empty entry); void finish(const Hit&)
template <int QUERY, bool COUNT, class IO>
HJ_DI void trace_warp_loop(const DevScene& S, IO& io, uint32_t n, uint32_t* next, TraceShared<QUERY>& sh, uint2* l_stack,
                           uint32_t& n_rays, uint32_t& n_nodes, uint32_t& n_tris, uint32_t* max_nodes, uint32_t* max_tris,
                           unsigned long long* diag = nullptr) {
    const uint32_t FULL = 0xffffffffu;
    constexpr int SS = TraceShared<QUERY>::SS;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, wbase = tid & ~31u;
    uint2* rring = sh.rring[tid >> 5];
    uint32_t npairs = 0;                            // warp-uniform: triangle tests pending in the ring
    uint2* sstack = sh.stack + tid;
    uint32_t head = 0, tail = 0;                    // warp-uniform absolute ring positions
    uint32_t ticket = 0;                            // ring position after this lane's last queued pair
    float3 o = f3(0.0f), idir = f3(0.0f); uint32_t oct = 0; float tmin = 0.0f;
    uint2 ng = make_uint2(0u, 0u), tg = make_uint2(0u, 0u); int sp = 0;
    bool active = false, exhausted = false;
    bool emph = false;                              // QUERY 2: this lane is in its emitter phase
    uint32_t own = tid;                             // thread whose ray this lane walks (itself unless it helps, HJ_STEAL)
    constexpr bool ANY = QUERY == 1;
    uint32_t nn0 = 0, own_tris = 0;
    const uint32_t k4b = load_two23();
    fill_octant_perm(sh.perm);
    uint32_t d_ph = 0, d_act = 0, d_walk = 0, d_wait = 0, d_hold = 0, d_bat = 0, d_blanes = 0;
    while (true) {
        // ---- lanes whose walk is over and whose queued pairs are all tested hand their result over (once per burst,
        //      so that the write-out runs several lanes wide)
        if (active && own != tid && ng.y <= 0x00ffffffu && sp == 0 && tg.y == 0u && (int32_t)(head - ticket) >= 0) {
            atomicSub(&sh.helpers[own], 1u);        // a helper's part of its owner's ray is done (its triangle tests included)
            active = false;
        }
        else if (active && ng.y <= 0x00ffffffu && sp == 0 && tg.y == 0u && (int32_t)(head - ticket) >= 0 &&
                 (!HJ_STEAL || *(volatile uint32_t*)&sh.helpers[tid] == 0u)) {
            const unsigned long long k = sh.key[tid];
            if (QUERY == 2 && emph) {               // emitter phase over: walk again from the root for anything in front of it
                emph = false; sh.key0[tid] = k;
                ng = make_uint2(0u, 0x80000000u);
            } else {
                const float2 b = sh.bary[tid];
                Hit h; h.t = __uint_as_float((uint32_t)(k >> 32)); h.prim = (uint32_t)k; h.b1 = b.x; h.b2 = b.y;
                if (QUERY == 2 && k != sh.key0[tid]) h.prim = HJ_HIT_BLOCKED;
                io.finish(h);
                n_rays++; active = false;
                if (COUNT) { atomicMax(max_nodes, n_nodes - nn0); atomicMax(max_tris, own_tris); }
            }
        }
        const uint32_t idle = __ballot_sync(FULL, !active);
        if (!exhausted && __popc(idle) >= REFILL_MIN) {
            const int leader = __ffs(idle) - 1;
            uint32_t base = 0;
            if ((int)lane == leader) base = atomicAdd(next, (uint32_t)__popc(idle));
            base = __shfl_sync(FULL, base, leader);
            if (!active) {
                const uint32_t idx = base + __popc(idle & ((1u << lane) - 1u));
                float3 d; float tmax;
                if (idx < n && io.load(idx, o, d, tmin, tmax)) {
                    const RayPrep r = prep_ray(o, d);
                    idir = r.idir; oct = r.oct;
                    sh.ray[0][tid] = __float_as_uint(o.x); sh.ray[1][tid] = __float_as_uint(o.y); sh.ray[2][tid] = __float_as_uint(o.z);
                    sh.ray[3][tid] = __float_as_uint(r.Sx); sh.ray[4][tid] = __float_as_uint(r.Sy); sh.ray[5][tid] = __float_as_uint(r.Sz);
                    sh.ray[6][tid] = (uint32_t)r.kx | ((uint32_t)r.ky << 2) | ((uint32_t)r.kz << 4);
                    sh.ray[7][tid] = __float_as_uint(tmin); sh.ray[8][tid] = __float_as_uint(tmax);
                    sh.key[tid] = ((unsigned long long)__float_as_uint(tmax) << 32) | 0xffffffffull;
                    sh.bary[tid] = make_float2(0.0f, 0.0f);
                    ng = make_uint2(0u, 0x80000000u);     // the root, as the only inner hit of a virtual parent
                    tg = make_uint2(0u, 0u); sp = 0; ticket = head;
                    active = true; nn0 = n_nodes; own_tris = 0; emph = QUERY == 2; own = tid; sh.helpers[tid] = 0u;
                }
            }
            if (base + (uint32_t)__popc(idle) >= n) exhausted = true;
        }
#if HJ_STEAL
        // ---- tail of the queue: idle lanes take over pending subtrees of the lanes that are still walking. The closest hit
        //      is a minimum over triangles (atomicMin on the owner's key), so it does not matter which lane walks which
        //      subtree: a helper walks with the owner's ray, prunes with the owner's current best and queues leaf records in
        //      the owner's name. One helper per victim and round; helpers can be robbed in turn. The longest rays
        //      (thousands of node visits against a mean of 15) otherwise keep a warp at 1 of 32 lanes for milliseconds.
        if (exhausted || STEAL_ALWAYS) {
            const uint32_t idle2 = __ballot_sync(FULL, !active);
            const uint32_t vict = __ballot_sync(FULL, active && sp > 0);
            if (__popc(idle2) >= STEAL_MIN_IDLE && vict) {
                const uint32_t npair = min(__popc(idle2), __popc(vict));
                const uint32_t irank = __popc(idle2 & ((1u << lane) - 1u)), vrank = __popc(vict & ((1u << lane) - 1u));
                const bool thief = !active && irank < npair, robbed = active && sp > 0 && vrank < npair;
                const uint32_t src = thief ? __fns(vict, 0, (int)irank + 1) : lane;
                const float sox = __shfl_sync(FULL, o.x, src), soy = __shfl_sync(FULL, o.y, src), soz = __shfl_sync(FULL, o.z, src);
                const float six = __shfl_sync(FULL, idir.x, src), siy = __shfl_sync(FULL, idir.y, src), siz = __shfl_sync(FULL, idir.z, src);
                const uint32_t soct = __shfl_sync(FULL, oct, src), sown = __shfl_sync(FULL, own, src);
                const float stmin = __shfl_sync(FULL, tmin, src);
                const uint32_t semph = __shfl_sync(FULL, (uint32_t)emph, src);
                uint2 loot = make_uint2(0u, 0u);
                if (thief) loot = sh.stack[wbase + src];                    // the victim's oldest entry: the largest pending subtrees
                __syncwarp();
                if (robbed) {                                               // the newest entry takes its place
                    sp--;
                    if (sp > 0) sstack[0] = sp < SS ? sstack[sp * HJ_TRACE_THREADS] : l_stack[sp - SS];
                }
                if (thief) {
                    o = f3(sox, soy, soz); idir = f3(six, siy, siz); oct = soct; tmin = stmin; emph = semph != 0u; own = sown;
                    ng = loot; tg = make_uint2(0u, 0u); sp = 0; ticket = head; active = true; nn0 = n_nodes; own_tris = 0;
                    atomicAdd(&sh.helpers[own], 1u);
                }
                __syncwarp();
            }
        }
#endif
        if (!__any_sync(FULL, active)) { if (exhausted) break; else continue; }
#pragma unroll 1
        for (int it = 0; it < TRACE_BURST; it++) {
            // ---- node phase: every lane without un-queued triangles visits its nearest pending inner child
            bool did = false;
            if (active && tg.y == 0u) {
                if (ng.y <= 0x00ffffffu && sp > 0) { sp--; ng = sp < SS ? sstack[sp * HJ_TRACE_THREADS] : l_stack[sp - SS]; }
                if (ng.y > 0x00ffffffu) {
                    const uint32_t k = 31u - (uint32_t)__clz(ng.y);      // highest bit = nearest (slots permuted by the octant)
                    ng.y &= ~(1u << k);
                    const uint32_t slot = (k - 24u) ^ (oct ^ 7u);
                    const uint32_t node = ng.x + __popc(ng.y & 0xffu & ((1u << slot) - 1u));
                    if (ng.y > 0x00ffffffu) { if (sp < SS) sstack[sp * HJ_TRACE_THREADS] = ng; else l_stack[sp - SS] = ng; sp++; }
                    if (COUNT) n_nodes++;
                    const float tcur = __uint_as_float((uint32_t)(sh.key[own] >> 32));
                    node_boxes<QUERY == 2>(S.nodes, node, o, idir, oct, tmin, tcur, k4b, sh.perm[oct ^ 7u], emph, ng, tg);
                    did = true;
                }
            }
            // ---- queue the leaf hits: one record (first triangle, hit bits | owner lane << 24) per lane and node visit,
            //      placed with a ballot prefix; the warp expands records into (owner, triangle) pairs when it tests them
            const uint32_t cnt = active ? (uint32_t)__popc(tg.y) : 0u;
            const uint32_t has = __ballot_sync(FULL, cnt != 0u);
            if (has) {
                if (cnt != 0u) {
                    const uint32_t pos = tail + (uint32_t)__popc(has & ((1u << lane) - 1u));
                    rring[pos & (HJ_RCAP - 1)] = make_uint2(tg.x, tg.y | ((own & 31u) << 24));
                    tg.y = 0u; ticket = pos + 1u; if (COUNT) own_tris += cnt;
                }
                tail += (uint32_t)__popc(has);
                npairs += __reduce_add_sync(FULL, cnt);
                __syncwarp();
            }
            // ---- triangle phase: 32 pairs at a time, one per lane
            const bool walked = active && ng.y <= 0x00ffffffu && sp == 0;
            const uint32_t waiting = __ballot_sync(FULL, walked && (int32_t)(head - ticket) < 0);
            const uint32_t nwork = (uint32_t)__popc(__ballot_sync(FULL, did));
            if (COUNT && diag) {       // lane-state census per node phase: phases, active, walking, waiting
                const uint32_t na = __popc(__ballot_sync(FULL, active));
                if (lane == 0) { d_ph++; d_act += na; d_walk += nwork; d_wait += __popc(waiting); }
            }
            uint32_t thr = nwork * TRI_K; if (thr > 32u) thr = 32u; if (thr < 1u) thr = 1u;
            if (__popc(waiting) >= WAIT_MAX) thr = 1u;
            while (npairs >= thr) {
                // the first (up to) 32 records, one per lane; an exclusive prefix sum of their triangle counts tells which
                // records are consumed whole (incl <= 32) and which one in part
                const uint32_t nrec = tail - head;
                uint2 rec = make_uint2(0u, 0u);
                if (lane < nrec) rec = rring[(head + lane) & (HJ_RCAP - 1)];
                const uint32_t c = (uint32_t)__popc(rec.y & 0x00ffffffu);
                uint32_t incl = c;
#pragma unroll
                for (int dlt = 1; dlt < 32; dlt <<= 1) { const uint32_t v = __shfl_up_sync(FULL, incl, dlt); if ((int)lane >= dlt) incl += v; }
                const uint32_t excl = incl - c;
                const uint32_t starts = __reduce_or_sync(FULL, (c != 0u && excl < 32u) ? (1u << excl) : 0u);
                uint32_t nb = __shfl_sync(FULL, incl, 31); if (nb > 32u) nb = 32u;
                // pair number `lane` belongs to the record whose start flag is the last one at or below bit `lane`
                const uint32_t k = (uint32_t)__popc(starts & (0xffffffffu >> (31u - lane))) - 1u;
                const uint32_t kbase = __shfl_sync(FULL, rec.x, (int)(k & 31u)), kmask = __shfl_sync(FULL, rec.y, (int)(k & 31u)),
                               kexcl = __shfl_sync(FULL, excl, (int)(k & 31u));
                bool hit = false; unsigned long long key = 0; uint32_t ow = 0; float b1 = 0.0f, b2 = 0.0f;
                if (lane < nb) {
                    uint32_t m = kmask & 0x00ffffffu;
                    for (uint32_t i = lane - kexcl; i > 0u; i--) m &= m - 1u;      // records are consumed from their lowest bit
                    const uint32_t tri = kbase + (uint32_t)__ffs((int)m) - 1u;
                    ow = wbase + (kmask >> 24);
                    if (!ANY || (uint32_t)sh.key[ow] == 0xffffffffu) {
                        const float4* T = S.btris + 3 * (size_t)tri;
                        const float4 v0 = __ldg(T), v1 = __ldg(T + 1), v2 = __ldg(T + 2);
                        if (COUNT) n_tris++;
                        RayPrep r;
                        r.o = f3(__uint_as_float(sh.ray[0][ow]), __uint_as_float(sh.ray[1][ow]), __uint_as_float(sh.ray[2][ow]));
                        r.Sx = __uint_as_float(sh.ray[3][ow]); r.Sy = __uint_as_float(sh.ray[4][ow]); r.Sz = __uint_as_float(sh.ray[5][ow]);
                        const uint32_t kp = sh.ray[6][ow];
                        r.kx = (int)(kp & 3u); r.ky = (int)((kp >> 2) & 3u); r.kz = (int)(kp >> 4);
                        float t;
                        if (tri_test(r, f3(v0), f3(v1), f3(v2), __uint_as_float(sh.ray[7][ow]), __uint_as_float(sh.ray[8][ow]), t, b1, b2)) {
                            hit = true;
                            key = ((unsigned long long)__float_as_uint(t) << 32) | (unsigned long long)__float_as_uint(v0.w);
                            atomicMin(&sh.key[ow], key);
                        }
                    }
                }
                // the record cut by the batch keeps its untested (higher) bits; whole records leave the ring
                if (lane < nrec && excl < 32u && incl > 32u) {
                    uint32_t m = rec.y & 0x00ffffffu;
                    for (uint32_t i = 32u - excl; i > 0u; i--) m &= m - 1u;
                    rring[(head + lane) & (HJ_RCAP - 1)].y = m | (rec.y & 0xff000000u);
                }
                head += (uint32_t)__popc(__ballot_sync(FULL, lane < nrec && incl <= 32u));
                npairs -= nb;
                __syncwarp();
                if (hit && sh.key[ow] == key) sh.bary[ow] = make_float2(b1, b2);
                __syncwarp();
                if (COUNT && diag && lane == 0) { d_bat++; d_blanes += nb; }
                thr = 32u;                                         // after a forced partial batch only full ones
            }
            // ---- any hit ends an occlusion ray
            if (ANY && active && (uint32_t)sh.key[own] != 0xffffffffu) { ng.y = 0u; tg.y = 0u; sp = 0; }
            if (QUERY == 2 && active && !emph && sh.key[own] != sh.key0[own]) { ng.y = 0u; tg.y = 0u; sp = 0; }
        }
    }
    if (COUNT && diag && lane == 0) {
        atomicAdd(diag + 0, (unsigned long long)d_ph); atomicAdd(diag + 1, (unsigned long long)d_act); atomicAdd(diag + 2, (unsigned long long)d_walk);
        atomicAdd(diag + 3, (unsigned long long)d_wait); atomicAdd(diag + 4, (unsigned long long)d_hold);
        atomicAdd(diag + 5, (unsigned long long)d_bat); atomicAdd(diag + 6, (unsigned long long)d_blanes);
    }
}

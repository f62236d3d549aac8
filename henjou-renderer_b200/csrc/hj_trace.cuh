// hj_trace.cuh — ray traversal of the compressed 8-wide BVH: the B200 replacement for the reference's
// optixTrace calls (/root/reference/include/kernel/rt.h:25 TraceOcculution, :53 RayTrace), which run on RT
// cores inside the closed-source OptiX runtime. B200 has no RT cores: this is plain sm_100a CUDA.
//
// Node (80 B = five 16-byte loads; built by hj_bvh.cu):
//   q0: origin.xyz (float bits) | ex | ey<<8 | ez<<16 | imask<<24      (e* = biased float exponents of the grid step)
//   q1: child_base (first internal child node), tri_base (first triangle of this node's leaves), meta[0..3], meta[4..7]
//       meta[s] = (count<<5) | offset for a leaf slot (count 1..4 triangles at tri_base+offset), 0 otherwise
//   q2: qlo.x[0..7] qlo.y[0..7]   q3: qlo.z[0..7] qhi.x[0..7]   q4: qhi.y[0..7] qhi.z[0..7]      (8-bit grid coordinates)
// Child slots are assigned at build time by octant (slot bit k set = child on the + side of axis k), so
// `slot ^ ray_octant` orders children front to back without sorting (Ylitie et al. 2017).
//
// The triangle test is the watertight test of Woop/Benthin/Wald 2013 without FMA contraction: a pure
// function of (ray, triangle), so the closest hit — ties on t broken by the lower global primitive id — does
// not depend on traversal order and is bit-identical to a brute-force loop over all triangles.
#pragma once
#include "hj_scene.cuh"

#define HJ_STACK 64

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

// bit s of `m` moves to bit s^oct: three conditional swaps
HJ_DI uint32_t perm_mask(uint32_t m, uint32_t oct) {
    if (oct & 1u) m = ((m & 0x55u) << 1) | ((m >> 1) & 0x55u);
    if (oct & 2u) m = ((m & 0x33u) << 2) | ((m >> 2) & 0x33u);
    if (oct & 4u) m = ((m & 0x0fu) << 4) | ((m >> 4) & 0x0fu);
    return m;
}

// Intersects the 8 quantised child boxes of one node with the ray; returns the slot mask of children whose
// box overlaps [tmin, tcur]. Conservative: the slab bounds are widened by a slack proportional to the
// magnitudes involved, so rounding in this test can never cull a triangle the (exact) triangle test accepts.
HJ_DI uint32_t node_children_hit(const uint4 q0, const uint4 q2, const uint4 q3, const uint4 q4, uint32_t valid,
                                 const RayPrep& r, float tmin, float tcur) {
    const float sx = __uint_as_float((q0.w & 0xffu) << 23), sy = __uint_as_float(((q0.w >> 8) & 0xffu) << 23),
                sz = __uint_as_float(((q0.w >> 16) & 0xffu) << 23);
    const float ax = sx * r.idir.x, ay = sy * r.idir.y, az = sz * r.idir.z;
    const float ox = (__uint_as_float(q0.x) - r.o.x) * r.idir.x, oy = (__uint_as_float(q0.y) - r.o.y) * r.idir.y,
                oz = (__uint_as_float(q0.z) - r.o.z) * r.idir.z;
    const float slack = 2e-6f * (fabsf(ox) + fabsf(oy) + fabsf(oz) + 255.0f * (fabsf(ax) + fabsf(ay) + fabsf(az)));
    // near/far planes per axis by ray sign
    const uint32_t lox[2] = {q2.x, q2.y}, loy[2] = {q2.z, q2.w}, loz[2] = {q3.x, q3.y};
    const uint32_t hix[2] = {q3.z, q3.w}, hiy[2] = {q4.x, q4.y}, hiz[2] = {q4.z, q4.w};
    const bool nx = (r.oct & 1u) != 0, ny = (r.oct & 2u) != 0, nz = (r.oct & 4u) != 0;
    uint32_t mask = 0;
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const uint32_t nearx = nx ? hix[h] : lox[h], farx = nx ? lox[h] : hix[h];
        const uint32_t neary = ny ? hiy[h] : loy[h], fary = ny ? loy[h] : hiy[h];
        const uint32_t nearz = nz ? hiz[h] : loz[h], farz = nz ? loz[h] : hiz[h];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const float tnx = (float)((nearx >> (8 * j)) & 0xffu) * ax + ox, tfx = (float)((farx >> (8 * j)) & 0xffu) * ax + ox;
            const float tny = (float)((neary >> (8 * j)) & 0xffu) * ay + oy, tfy = (float)((fary >> (8 * j)) & 0xffu) * ay + oy;
            const float tnz = (float)((nearz >> (8 * j)) & 0xffu) * az + oz, tfz = (float)((farz >> (8 * j)) & 0xffu) * az + oz;
            const float tn = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, tmin)) - slack;
            const float tf = fminf(fminf(tfx, tfy), fminf(tfz, tcur)) + slack;
            if (tn <= tf) mask |= 1u << (4 * h + j);
        }
    }
    return mask & valid;
}

HJ_DI uint32_t meta_of(const uint4& q1, int s) { return ((s < 4 ? q1.z : q1.w) >> (8 * (s & 3))) & 0xffu; }

// Closest hit in (tmin,tmax); ties on t -> lower global primitive id. COUNT: fetch counters for the roofline.
template <bool ANY, bool COUNT>
HJ_DI Hit traverse(const DevScene& S, const float3& o, const float3& d, float tmin, float tmax, uint32_t& n_nodes, uint32_t& n_tris) {
    Hit best; best.t = tmax; best.b1 = best.b2 = 0.0f; best.prim = 0xffffffffu;
    const RayPrep r = prep_ray(o, d);
    uint2 stack[HJ_STACK];
    int sp = 0;
    uint32_t node = 0;
    while (true) {
        const uint4* N = S.nodes + 5 * (size_t)node;
        const uint4 q0 = __ldg(N), q1 = __ldg(N + 1), q2 = __ldg(N + 2), q3 = __ldg(N + 3), q4 = __ldg(N + 4);
        if (COUNT) n_nodes++;
        const uint32_t imask = q0.w >> 24;
        uint32_t leafmask = 0;
#pragma unroll
        for (int s = 0; s < 8; s++) if (meta_of(q1, s)) leafmask |= 1u << s;
        const float tcur = (best.prim != 0xffffffffu) ? best.t : tmax;
        const uint32_t hits = node_children_hit(q0, q2, q3, q4, imask | leafmask, r, tmin, tcur);
        uint32_t lh = hits & leafmask;
        while (lh) {
            const int s = __ffs(lh) - 1; lh &= lh - 1;
            const uint32_t m = meta_of(q1, s);
            const uint32_t first = q1.y + (m & 31u), cnt = m >> 5;
            for (uint32_t k = 0; k < cnt; k++) {
                const float4* T = S.btris + 3 * (size_t)(first + k);
                const float4 v0 = __ldg(T), v1 = __ldg(T + 1), v2 = __ldg(T + 2);
                if (COUNT) n_tris++;
                float t, b1, b2;
                if (tri_test(r, f3(v0), f3(v1), f3(v2), tmin, tmax, t, b1, b2)) {
                    const uint32_t p = __float_as_uint(v0.w);
                    if (ANY) { best.t = t; best.b1 = b1; best.b2 = b2; best.prim = p; return best; }
                    if (best.prim == 0xffffffffu || t < best.t || (t == best.t && p < best.prim)) { best.t = t; best.b1 = b1; best.b2 = b2; best.prim = p; }
                }
            }
        }
        uint32_t ih = hits & imask, base = q1.x, im = imask;
        if (!ih) {
            if (sp == 0) break;
            const uint2 e = stack[--sp];
            base = e.x; im = e.y >> 8; ih = e.y & 0xffu;
        }
        // nearest remaining child of the group: lowest set bit after permuting slots by the ray octant
        const uint32_t s = (uint32_t)(__ffs(perm_mask(ih, r.oct)) - 1) ^ r.oct;
        ih &= ~(1u << s);
        if (ih && sp < HJ_STACK) stack[sp++] = make_uint2(base, (im << 8) | ih);
        node = base + __popc(im & ((1u << s) - 1u));
    }
    return best;
}

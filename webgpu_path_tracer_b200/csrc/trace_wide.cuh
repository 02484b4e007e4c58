// trace_wide.cuh -- closest hit over the 8-wide quantised BVH (wide_bvh.cuh).
//
// The wide traversal is a SUPERSET search: child boxes are conservative (quantised outwards, and
// the slab test carries a 2^-19 relative slack), so every triangle the reference's own walk
// (hitRay.wgsl:42-110) would test is tested here too, with the reference's exact triangle
// arithmetic (tri_test, pinned).  The closest hit is then order-independent EXCEPT when two
// triangles report exactly the same distance (the reference keeps the last one it visited,
// common.wgsl:219 accepts dst == tmax); such rays -- and rays whose traversal stack would
// overflow -- are flagged and re-traced with the reference-order binary walk (trace_binary.cuh).
#pragma once
#include <cuda_fp16.h>
#include "intersect.cuh"
#include "trace_binary.cuh"
#include "wide_bvh.cuh"

namespace wg {

#ifndef WIDE_STACK
#define WIDE_STACK 24
#endif

struct WideRay {
    float idx, idy, idz;      // 1/d with tiny components clamped (box tests only); traversal rays: times the ray scale
    uint32_t octinv;          // bit 2/1/0 set when the (clamped) d.x/y/z is positive: ALWAYS the sign of idx/idy/idz
    float far_scale;          // traversal rays: 1.000002 * ray scale (closest distance -> scaled far cap)
};

__device__ __forceinline__ float rcp_approx(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

__device__ __forceinline__ WideRay make_wide_ray(const Ray& r) {
    WideRay w;
    const float eps = 1e-12f;
    const float dx = fabsf(r.d.x) < eps ? copysignf(eps, r.d.x) : r.d.x;
    const float dy = fabsf(r.d.y) < eps ? copysignf(eps, r.d.y) : r.d.y;
    const float dz = fabsf(r.d.z) < eps ? copysignf(eps, r.d.z) : r.d.z;
    // box tests only (conservative, 2^-19 slack): the 1-ulp MUFU.RCP is enough, |d| is in [1e-12, ~1]
    w.idx = rcp_approx(dx); w.idy = rcp_approx(dy); w.idz = rcp_approx(dz);
    // near / far planes are chosen from the sign BIT of the clamped component, i.e. of the reciprocal itself: a
    // -0.0 component (shootRay's centre row, axis-aligned mirrors) clamps to -eps and must pair with the far planes
    // (`d >= 0` is true for -0.0 and paired a negative reciprocal with the near planes: every child missed)
    w.octinv = ((__float_as_uint(dx) >> 31) ? 0u : 4u) | ((__float_as_uint(dy) >> 31) ? 0u : 2u) | ((__float_as_uint(dz) >> 31) ? 0u : 1u);
    w.far_scale = 1.000002f;
    return w;
}

// The ray of a wide TRAVERSAL.  The slab test multiplies plane bytes by the cell size along the ray, cell_t =
// 2^(e-127) / d, held as a binary16 operand of the mixed-precision FMA (wide_node_hits), so all ray distances of one
// traversal are scaled by a per-ray power of two s (exact in f32; every comparison is between scaled values, so
// nothing changes but the exponent range): s puts 2^24 x (the largest cell any node can have along the axis of the
// smallest |d|) at 2^14..2^15, the top of binary16's range (the 2^24 undoes the 2^-24 of the plane operand, a
// binary16 subnormal).  Cells only shrink below the root, so nothing overflows; cells that fall below binary16's
// normal range lose precision, which the slab test's absolute slack covers.
// scale_k = 371 - (largest plane exponent byte of the root node), from the host (WideBvh::scale_k).
// `ok` = false, the reference walk decides: the scale does not fit f32 (caller-supplied rays with absurd directions),
// or one direction component is below 2^-20 of another -- the ray runs (almost) inside a plane of the grid, its
// reciprocal there is 2^20 times the others, the common scale pushes every other axis into the absolute slack and the
// walk would visit every node that plane cuts (measured: an exactly horizontal ray in the atrium's floor, 4,000 nodes,
// 10 ms); about 6 rays in a million (cosine sampling returns an exact 0 component with probability 2^-25 per draw).
__device__ __forceinline__ WideRay make_wide_ray_scaled(const Ray& r, uint32_t scale_k, bool& ok) {
    WideRay w = make_wide_ray(r);
    const float ax = fabsf(w.idx), ay = fabsf(w.idy), az = fabsf(w.idz);
    const float m = fmaxf(ax, fmaxf(ay, az)), mn = fminf(ax, fminf(ay, az));
    const int se = (int)scale_k - (int)(__float_as_uint(m) >> 23);
    ok = se >= 1 && se <= 254 && m <= mn * 1048576.0f;
    const float s = __uint_as_float((uint32_t)(ok ? se : 127) << 23);
    w.idx *= s; w.idy *= s; w.idz *= s;
    w.far_scale = 1.000002f * s;
    return w;
}

// Octant permutation table: lut[c * 256 + H] has bit (s ^ c) set for every bit s of H, so a hit
// byte in slot order becomes a byte in visit-priority order (highest bit = nearest octant first)
// with one shared-memory load.  Every kernel that traverses fills its own copy once per block.
#define WIDE_LUT_BYTES 2048
__device__ __forceinline__ void wide_lut_init(unsigned char* lut) {
    for (uint32_t i = threadIdx.x; i < WIDE_LUT_BYTES; i += blockDim.x) {
        const uint32_t c = i >> 8, H = i & 255u;
        uint32_t r = 0;
        #pragma unroll
        for (uint32_t b = 0; b < 8; b++) r |= ((H >> b) & 1u) << (b ^ c);
        lut[i] = (unsigned char)r;
    }
    __syncthreads();
}

// Plane bytes meet the ray in ONE instruction per plane: PRMT turns two bytes b of a plane word into the two
// binary16 SUBNORMALS 0x00bb = b * 2^-24 (exact), and the mixed-precision FMA of sm_100a (PTX fma.rn.f32.f16, SASS
// FHFMA, which reads either half of a register) computes  t = (b 2^-24) * cell_h + origin_t  in f32, where cell_h is
// 2^24 x the cell size along the scaled ray, rounded to binary16.
// Before: PRMT -> subnormal half -> widening HADD2.F32 -> FFMA, i.e. 48 more FMA-pipe instructions per node visit
// (I2F.U8 runs on the XU pipe at 8 issue cycles per warp on sm_100a: never an option).
// Error budget, in cells of the node grid: |b (cell_h - cell)| <= 255 * 2^-12 = 0.06226 (binary16 rounding of the
// cell) + f32 rounding of origin_t and of the result.  K2 quantises every child box outwards by 1/16 cell beyond
// floor / ceil (wide_bvh.cu), which covers the cell part; the relative part is the 2^-19 slack below.  A cell_h below
// 2^-14 is a binary16 subnormal with absolute error 2^-25: times 255 * 2^-24 that is < 2^-41, covered by
// WIDE_ABS_SLACK on each side of the interval (applied at the comparison).
#define WIDE_ABS_SLACK 9.094947017729282e-13f     /* 2^-40, in scaled ray distance */
// halves are picked INSIDE the asm block (mov.b32 {lo, hi}, r): ptxas sees 32-bit registers and emits R.H0 / R.H1
#define WGPT_FHFMA(NAME, A, B)                                                                                  \
    __device__ __forceinline__ float NAME(uint32_t a, uint32_t b, float c) {                                      \
        float d;                                                                                                  \
        asm("{\n\t.reg .b16 al, ah, bl, bh;\n\tmov.b32 {al, ah}, %1;\n\tmov.b32 {bl, bh}, %2;\n\t"                \
            "fma.rn.f32.f16 %0, " A ", " B ", %3;\n\t}" : "=f"(d) : "r"(a), "r"(b), "f"(c));                      \
        return d;                                                                                                 \
    }
WGPT_FHFMA(fhfma_ll, "al", "bl")      // a.lo * b.lo + c
WGPT_FHFMA(fhfma_lh, "al", "bh")      // a.lo * b.hi + c
WGPT_FHFMA(fhfma_hl, "ah", "bl")
WGPT_FHFMA(fhfma_hh, "ah", "bh")
__device__ __forceinline__ uint32_t pack_h2(float lo, float hi) {
    const __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&h);
}

// Intersects the 8 children of one wide node (its five 16-byte records n0..n4); returns the hit byte in SLOT order
// (bit s = the box in slot s is hit over (0, closest]); empty slots never report.  `wr` must come from
// make_wide_ray_scaled; only the origin of `ray` is read.
__device__ __forceinline__ uint32_t wide_node_test(const uint4 n0, const uint4 n1, const uint4 n2, const uint4 n3, const uint4 n4,
                                                   const Ray& ray, const WideRay& wr, float closest, uint32_t& child_base,
                                                   uint32_t& tri_base, uint32_t& imask, uint32_t& lmask) {
    child_base = n1.x; tri_base = n1.y; imask = n0.w >> 24; lmask = n1.z & 0xffu;
    // 2^24 x cell size along the (scaled) ray, per axis, as binary16: x in the low half of cxy, y in the high half
    const uint32_t cxy = pack_h2(__uint_as_float(((n0.w & 0xffu) + 24u) << 23) * wr.idx, __uint_as_float((((n0.w >> 8) & 0xffu) + 24u) << 23) * wr.idy);
    const uint32_t cz = pack_h2(__uint_as_float((((n0.w >> 16) & 0xffu) + 24u) << 23) * wr.idz, 0.0f);
    // distance of the grid origin along each axis
    const float cx = (__uint_as_float(n0.x) - ray.o.x) * wr.idx;
    const float cy = (__uint_as_float(n0.y) - ray.o.y) * wr.idy;
    const float cz_ = (__uint_as_float(n0.z) - ray.o.z) * wr.idz;
    // near / far plane words by direction sign; [0] = slots 0..3, [1] = slots 4..7
    const bool px = (wr.octinv & 4u) != 0, py = (wr.octinv & 2u) != 0, pz = (wr.octinv & 1u) != 0;
    const uint32_t nx[2] = {px ? n2.x : n3.z, px ? n2.y : n3.w}, fx[2] = {px ? n3.z : n2.x, px ? n3.w : n2.y};
    const uint32_t ny[2] = {py ? n2.z : n4.x, py ? n2.w : n4.y}, fy[2] = {py ? n4.x : n2.z, py ? n4.y : n2.w};
    const uint32_t nz[2] = {pz ? n3.x : n4.z, pz ? n3.y : n4.w}, fz[2] = {pz ? n4.z : n3.x, pz ? n4.w : n3.y};
    const float far_cap = fmaf(closest, wr.far_scale, WIDE_ABS_SLACK);
    // Straight-line code for all 8 slots (no divergence inside a node).  The sign of
    // the three differences below is the MISS flag; a funnel shift collects the flags, slot 7 first, so
    // slot s ends up in bit s.
    uint32_t miss = 0;
    #pragma unroll
    for (int pr = 3; pr >= 0; pr--) {
        const int w = pr >> 1;
        const uint32_t sel = (pr & 1) ? 0x4342u : 0x4140u;
        // two slots per PRMT: low half = slot 2 pr, high half = slot 2 pr + 1, each 0x00bb = byte * 2^-24
        const uint32_t hnx = __byte_perm(nx[w], 0u, sel), hfx = __byte_perm(fx[w], 0u, sel);
        const uint32_t hny = __byte_perm(ny[w], 0u, sel), hfy = __byte_perm(fy[w], 0u, sel);
        const uint32_t hnz = __byte_perm(nz[w], 0u, sel), hfz = __byte_perm(fz[w], 0u, sel);
        #pragma unroll
        for (int k = 1; k >= 0; k--) {
            // hit <=> tlo3 - slack <= (thi3 + slack) * (1 + 2^-19)  and  tlo3 - slack <= far  and  thi3 + slack >= 0 (the
            // box is not behind the origin); each test is the sign of one FMA-pipe result, OR-ed in a single LOP3
            const float tlo3 = k ? fmaxf(fmaxf(fhfma_hl(hnx, cxy, cx), fhfma_hh(hny, cxy, cy)), fhfma_hl(hnz, cz, cz_))
                                 : fmaxf(fmaxf(fhfma_ll(hnx, cxy, cx), fhfma_lh(hny, cxy, cy)), fhfma_ll(hnz, cz, cz_));
            const float thi3 = k ? fminf(fminf(fhfma_hl(hfx, cxy, cx), fhfma_hh(hfy, cxy, cy)), fhfma_hl(hfz, cz, cz_))
                                 : fminf(fminf(fhfma_ll(hfx, cxy, cx), fhfma_lh(hfy, cxy, cy)), fhfma_ll(hfz, cz, cz_));
            const float hi_s = fmaf(thi3, 1.000002f, 3.0f * WIDE_ABS_SLACK);     // 2 slacks for the interval ends, 1 so that a thi3 just below 0 cannot flip d1
            const float d1 = hi_s - tlo3;
            const float d2 = far_cap - tlo3;
            const uint32_t d = __float_as_uint(d1) | __float_as_uint(d2) | __float_as_uint(hi_s);
            miss = __funnelshift_l(d, miss, 1);
        }
    }
    return ~miss & (imask | lmask);
}

__device__ __forceinline__ uint32_t wide_node_hits(const uint4* __restrict__ nodes, uint32_t ni, const Ray& ray,
                                                   const WideRay& wr, float closest, uint32_t& child_base,
                                                   uint32_t& tri_base, uint32_t& imask, uint32_t& lmask) {
    const uint4* n = &nodes[5 * (size_t)ni];
    return wide_node_test(__ldg(&n[0]), __ldg(&n[1]), __ldg(&n[2]), __ldg(&n[3]), __ldg(&n[4]), ray, wr, closest,
                          child_base, tri_base, imask, lmask);
}

// Triangle part of hitScene over the wide BVH.  `closest`/`h` carry in the sphere/quad result.
// Returns true if a triangle was accepted; sets `ambiguous` when the result must be re-derived
// with the reference-order walk.  `lut` = the block's octant permutation table (wide_lut_init).
__device__ __forceinline__ bool hit_bvh_wide(const DeviceScene& S, const WideBvh& WB, const unsigned char* lut, const Ray& ray,
                                             Counts* cnt, ObjRayCache& oc, float& closest, HitLite& h, bool& ambiguous) {
    ambiguous = false;
    if (WB.n_nodes == 0) return false;
    if (WB.force_reference_walk) { ambiguous = true; return false; }
    bool ray_ok;
    const WideRay wr = make_wide_ray_scaled(ray, WB.scale_k, ray_ok);
    if (!ray_ok) { ambiguous = true; return false; }
    const unsigned char* lut_r = lut + (wr.octinv << 8);
    bool hit_anything = false;
    uint2 stack[WIDE_STACK];
    int sp = 0;
    uint2 G = make_uint2(0u, 0x80000000u);
    unsigned n_nodes = 0, n_tris = 0;

    while (true) {
        if (G.y & 0xff000000u) {
            const uint32_t bit = 31u - (uint32_t)__clz((int)(G.y & 0xff000000u));
            const uint32_t hits_imask = G.y;
            G.y &= ~(1u << bit);
            if (G.y & 0xff000000u) {
                if (sp >= WIDE_STACK) { ambiguous = true; break; }
                stack[sp++] = G;
            }
            const uint32_t slot = (bit - 24u) ^ wr.octinv;
            const uint32_t rel = __popc(hits_imask & ~(0xffffffffu << slot) & 0xffu);
            const uint32_t ni = G.x + rel;
            uint32_t child_base, tri_base, imask, lmask;
            n_nodes++;
            const uint32_t hits = wide_node_hits(WB.nodes, ni, ray, wr, closest, child_base, tri_base, imask, lmask);
            G = make_uint2(child_base, ((uint32_t)lut_r[hits & imask] << 24) | imask);
            uint32_t tm = hits & lmask;
            while (tm) {
                const int b = __ffs((int)tm) - 1;
                tm &= tm - 1u;
                const uint32_t ti = tri_base + (uint32_t)__popc(lmask & ~(0xffffffffu << b));
                const float4* t = &WB.tris[3 * (size_t)ti];
                const float4 A = __ldg(&t[0]), B = __ldg(&t[1]), C = __ldg(&t[2]);
                oc.fetch(S, __float_as_int(B.w), ray);
                float dst, uu, vv;
                n_tris++;
                if (tri_test(V3(A.x, A.y, A.z), V3(B.x, B.y, B.z), V3(C.x, C.y, C.z), oc.oray, WGPT_RAY_TMIN, closest, dst, uu, vv)) {
                    if (hit_anything && dst == closest) ambiguous = true;       // equal-distance triangles: visit order decides
                    hit_anything = true; closest = dst;
                    h.t = dst; h.kind = HIT_TRI; h.index = __float_as_int(A.w); h.uu = uu; h.vv = vv;
                    h.mat_id = oc.m.material_id;
                }
            }
        } else {
            if (sp == 0) break;
            G = stack[--sp];
        }
    }
    if (cnt) { cnt->nodes += n_nodes; cnt->tris += n_tris; }
    return hit_anything;
}

// hitScene for the fast path: analytic primitives, wide BVH, reference-order fallback.
__device__ __forceinline__ bool hitScene_wide(const DeviceScene& S, const WideBvh& WB, const unsigned char* lut, const Ray& ray, uint32_t& rng,
                                              Counts* cnt, ObjRayCache& oc, HitLite& h, unsigned* fallbacks) {
    float closest = WGPT_MAX_FLOAT;
    if (cnt) cnt->rays++;
    oc.reset();
    bool hit = hit_analytic(S, ray, rng, cnt, closest, h);
    const HitLite h0 = h; const float c0 = closest;
    bool ambiguous;
    bool hit_t = hit_bvh_wide(S, WB, lut, ray, cnt, oc, closest, h, ambiguous);
    if (ambiguous) {
        h = h0; closest = c0;
        hit_t = hit_bvh_binary(S, ray, nullptr, oc, closest, h);
        if (fallbacks) (*fallbacks)++;
    }
    return hit | hit_t;
}

}  // namespace wg

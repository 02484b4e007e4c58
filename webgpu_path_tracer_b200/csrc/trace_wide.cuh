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
    float idx, idy, idz;      // 1/d with tiny components clamped (box tests only)
    uint32_t octinv;          // bit 2/1/0 set when d.x/y/z >= 0
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
    w.octinv = (r.d.x >= 0.0f ? 4u : 0u) | (r.d.y >= 0.0f ? 2u : 0u) | (r.d.z >= 0.0f ? 1u : 0u);
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

// Two quantised plane bytes -> two floats (byte * 2^-24) without the conversion unit: PRMT builds
// the half2 whose lanes are the SUBNORMAL halves 0x00bb (= b * 2^-24 exactly) and the widening
// HADD2.F32 runs on the FMA pipe (I2F.U8 costs 8 issue cycles per warp on sm_100a, measured; 48
// of them per node visit made the conversion unit the bottleneck of the whole traversal).  The
// 2^24 is folded into the per-node scale.
__device__ __forceinline__ void cvt2(uint32_t w, uint32_t sel, float& f0, float& f1) {
    const uint32_t p = __byte_perm(w, 0u, sel);
    const __half2 h = *reinterpret_cast<const __half2*>(&p);
    f0 = __low2float(h); f1 = __high2float(h);
}

// Intersects the 8 children of one wide node; returns the hit byte in SLOT order (bit s = the box
// in slot s is hit over (0, closest]); empty slots never report.
__device__ __forceinline__ uint32_t wide_node_hits(const uint4* __restrict__ nodes, uint32_t ni, const Ray& ray,
                                                   const WideRay& wr, float closest, uint32_t& child_base,
                                                   uint32_t& tri_base, uint32_t& imask, uint32_t& lmask) {
    const uint4 n0 = __ldg(&nodes[5 * (size_t)ni]);
    const uint4 n1 = __ldg(&nodes[5 * (size_t)ni + 1]);
    const uint4 n2 = __ldg(&nodes[5 * (size_t)ni + 2]);
    const uint4 n3 = __ldg(&nodes[5 * (size_t)ni + 3]);
    const uint4 n4 = __ldg(&nodes[5 * (size_t)ni + 4]);
    child_base = n1.x; tri_base = n1.y; imask = n0.w >> 24; lmask = n1.z & 0xffu;
    // scale = 2^(e-127) * 2^24 (the subnormal-half conversion yields byte * 2^-24)
    const float adjx = __uint_as_float(((n0.w & 0xffu) + 24u) << 23) * wr.idx;
    const float adjy = __uint_as_float((((n0.w >> 8) & 0xffu) + 24u) << 23) * wr.idy;
    const float adjz = __uint_as_float((((n0.w >> 16) & 0xffu) + 24u) << 23) * wr.idz;
    const float ox = (__uint_as_float(n0.x) - ray.o.x) * wr.idx;
    const float oy = (__uint_as_float(n0.y) - ray.o.y) * wr.idy;
    const float oz = (__uint_as_float(n0.z) - ray.o.z) * wr.idz;
    // near / far plane words by direction sign; [0] = slots 0..3, [1] = slots 4..7
    const bool px = (wr.octinv & 4u) != 0, py = (wr.octinv & 2u) != 0, pz = (wr.octinv & 1u) != 0;
    const uint32_t nx[2] = {px ? n2.x : n3.z, px ? n2.y : n3.w}, fx[2] = {px ? n3.z : n2.x, px ? n3.w : n2.y};
    const uint32_t ny[2] = {py ? n2.z : n4.x, py ? n2.w : n4.y}, fy[2] = {py ? n4.x : n2.z, py ? n4.y : n2.w};
    const uint32_t nz[2] = {pz ? n3.x : n4.z, pz ? n3.y : n4.w}, fz[2] = {pz ? n4.z : n3.x, pz ? n4.w : n3.y};
    const float far_cap = closest * 1.000002f;
    // Straight-line code for all 8 slots (no divergence inside a node).  The sign of
    // the three differences below is the MISS flag; a funnel shift collects the flags, slot 7 first, so
    // slot s ends up in bit s.
    uint32_t miss = 0;
    #pragma unroll
    for (int pr = 3; pr >= 0; pr--) {
        const int w = pr >> 1;
        const uint32_t sel = (pr & 1) ? 0x4342u : 0x4140u;
        float tlx[2], thx[2], tly[2], thy[2], tlz[2], thz[2];
        cvt2(nx[w], sel, tlx[0], tlx[1]); cvt2(fx[w], sel, thx[0], thx[1]);
        cvt2(ny[w], sel, tly[0], tly[1]); cvt2(fy[w], sel, thy[0], thy[1]);
        cvt2(nz[w], sel, tlz[0], tlz[1]); cvt2(fz[w], sel, thz[0], thz[1]);
        #pragma unroll
        for (int k = 1; k >= 0; k--) {
            // hit <=> tlo3 <= thi3 * (1 + 2^-19)  and  tlo3 <= far  and  thi3 >= 0  (the box is not behind the
            // origin); each test is the sign of one FMA-pipe result, OR-ed in a single LOP3
            const float tlo3 = fmaxf(fmaxf(fmaf(tlx[k], adjx, ox), fmaf(tly[k], adjy, oy)), fmaf(tlz[k], adjz, oz));
            const float thi3 = fminf(fminf(fmaf(thx[k], adjx, ox), fmaf(thy[k], adjy, oy)), fmaf(thz[k], adjz, oz));
            const float d1 = fmaf(thi3, 1.000002f, -tlo3);
            const float d2 = far_cap - tlo3;
            const uint32_t d = __float_as_uint(d1) | __float_as_uint(d2) | __float_as_uint(thi3);
            miss = __funnelshift_l(d, miss, 1);
        }
    }
    return ~miss & (imask | lmask);
}

// Triangle part of hitScene over the wide BVH.  `closest`/`h` carry in the sphere/quad result.
// Returns true if a triangle was accepted; sets `ambiguous` when the result must be re-derived
// with the reference-order walk.  `lut` = the block's octant permutation table (wide_lut_init).
__device__ __forceinline__ bool hit_bvh_wide(const DeviceScene& S, const WideBvh& WB, const unsigned char* lut, const Ray& ray,
                                             Counts* cnt, ObjRayCache& oc, float& closest, HitLite& h, bool& ambiguous) {
    ambiguous = false;
    if (WB.n_nodes == 0) return false;
    if (WB.force_reference_walk) { ambiguous = true; return false; }
    const WideRay wr = make_wide_ray(ray);
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

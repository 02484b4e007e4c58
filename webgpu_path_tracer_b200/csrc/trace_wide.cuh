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
#include "intersect.cuh"
#include "trace_binary.cuh"
#include "wide_bvh.cuh"

namespace wg {

#define WIDE_STACK 24

struct WideRay {
    float idx, idy, idz;      // 1/d with tiny components clamped (box tests only)
    uint32_t octinv;          // bit 2/1/0 set when d.x/y/z >= 0
};

__device__ __forceinline__ WideRay make_wide_ray(const Ray& r) {
    WideRay w;
    const float eps = 1e-12f;
    const float dx = fabsf(r.d.x) < eps ? copysignf(eps, r.d.x) : r.d.x;
    const float dy = fabsf(r.d.y) < eps ? copysignf(eps, r.d.y) : r.d.y;
    const float dz = fabsf(r.d.z) < eps ? copysignf(eps, r.d.z) : r.d.z;
    w.idx = 1.0f / dx; w.idy = 1.0f / dy; w.idz = 1.0f / dz;
    w.octinv = (r.d.x >= 0.0f ? 4u : 0u) | (r.d.y >= 0.0f ? 2u : 0u) | (r.d.z >= 0.0f ? 1u : 0u);
    return w;
}

__device__ __forceinline__ uint32_t byte_of(uint32_t lo, uint32_t hi, int s) {
    return ((s < 4 ? lo : hi) >> ((s & 3) * 8)) & 0xffu;
}

// Intersects the 8 children of one wide node; returns the hit mask (bits 24..31: inner children
// in octant order, bits 0..23: triangles of the leaf children) and the node's q1 words.
__device__ __forceinline__ uint32_t wide_node_hits(const uint4* __restrict__ nodes, uint32_t ni, const Ray& ray,
                                                   const WideRay& wr, float closest, uint32_t& child_base,
                                                   uint32_t& tri_base, uint32_t& imask) {
    const uint4 n0 = __ldg(&nodes[5 * (size_t)ni]);
    const uint4 n1 = __ldg(&nodes[5 * (size_t)ni + 1]);
    const uint4 n2 = __ldg(&nodes[5 * (size_t)ni + 2]);
    const uint4 n3 = __ldg(&nodes[5 * (size_t)ni + 3]);
    const uint4 n4 = __ldg(&nodes[5 * (size_t)ni + 4]);
    child_base = n1.x; tri_base = n1.y; imask = n0.w >> 24;
    const float adjx = __uint_as_float((n0.w & 0xffu) << 23) * wr.idx;
    const float adjy = __uint_as_float(((n0.w >> 8) & 0xffu) << 23) * wr.idy;
    const float adjz = __uint_as_float(((n0.w >> 16) & 0xffu) << 23) * wr.idz;
    const float ox = (__uint_as_float(n0.x) - ray.o.x) * wr.idx;
    const float oy = (__uint_as_float(n0.y) - ray.o.y) * wr.idy;
    const float oz = (__uint_as_float(n0.z) - ray.o.z) * wr.idz;
    // near / far planes by direction sign
    const bool px = (wr.octinv & 4u) != 0, py = (wr.octinv & 2u) != 0, pz = (wr.octinv & 1u) != 0;
    const uint32_t nxa = px ? n2.x : n3.z, nxb = px ? n2.y : n3.w, fxa = px ? n3.z : n2.x, fxb = px ? n3.w : n2.y;
    const uint32_t nya = py ? n2.z : n4.x, nyb = py ? n2.w : n4.y, fya = py ? n4.x : n2.z, fyb = py ? n4.y : n2.w;
    const uint32_t nza = pz ? n3.x : n4.z, nzb = pz ? n3.y : n4.w, fza = pz ? n4.z : n3.x, fzb = pz ? n4.w : n3.y;
    const float far_cap = closest * 1.000002f;
    uint32_t hitmask = 0;
    #pragma unroll
    for (int s = 0; s < 8; s++) {
        const uint32_t m = byte_of(n1.z, n1.w, s);
        const uint32_t child_bits = m >> 5;
        if (child_bits == 0) continue;
        const float tlx = fmaf((float)byte_of(nxa, nxb, s), adjx, ox), thx = fmaf((float)byte_of(fxa, fxb, s), adjx, ox);
        const float tly = fmaf((float)byte_of(nya, nyb, s), adjy, oy), thy = fmaf((float)byte_of(fya, fyb, s), adjy, oy);
        const float tlz = fmaf((float)byte_of(nza, nzb, s), adjz, oz), thz = fmaf((float)byte_of(fza, fzb, s), adjz, oz);
        const float tlo = fmaxf(fmaxf(tlx, tly), fmaxf(tlz, 0.0f));
        const float thi = fminf(fminf(thx, thy), thz) * 1.000002f;
        if (tlo <= fminf(thi, far_cap)) {
            const bool inner = (m & 0x18u) == 0x18u;
            const uint32_t bit = inner ? (24u + ((m & 7u) ^ wr.octinv)) : (m & 31u);
            hitmask |= child_bits << bit;
        }
    }
    return hitmask;
}

// Triangle part of hitScene over the wide BVH.  `closest`/`h` carry in the sphere/quad result.
// Returns true if a triangle was accepted; sets `ambiguous` when the result must be re-derived
// with the reference-order walk.
__device__ __forceinline__ bool hit_bvh_wide(const DeviceScene& S, const WideBvh& WB, const Ray& ray, Counts* cnt,
                                             ObjRayCache& oc, float& closest, HitLite& h, bool& ambiguous) {
    ambiguous = false;
    if (WB.n_nodes == 0) return false;
    const WideRay wr = make_wide_ray(ray);
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
            uint32_t child_base, tri_base, imask;
            n_nodes++;
            const uint32_t hm = wide_node_hits(WB.nodes, ni, ray, wr, closest, child_base, tri_base, imask);
            G = make_uint2(child_base, (hm & 0xff000000u) | imask);
            uint32_t tm = hm & 0x00ffffffu;
            while (tm) {
                const int b = __ffs((int)tm) - 1;
                tm &= tm - 1u;
                const float4* t = &WB.tris[3 * (size_t)(tri_base + (uint32_t)b)];
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
__device__ __forceinline__ bool hitScene_wide(const DeviceScene& S, const WideBvh& WB, const Ray& ray, uint32_t& rng,
                                              Counts* cnt, ObjRayCache& oc, HitLite& h, unsigned* fallbacks) {
    float closest = WGPT_MAX_FLOAT;
    if (cnt) cnt->rays++;
    oc.reset();
    bool hit = hit_analytic(S, ray, rng, cnt, closest, h);
    const HitLite h0 = h; const float c0 = closest;
    bool ambiguous;
    bool hit_t = hit_bvh_wide(S, WB, ray, cnt, oc, closest, h, ambiguous);
    if (ambiguous) {
        h = h0; closest = c0;
        hit_t = hit_bvh_binary(S, ray, nullptr, oc, closest, h);
        if (fallbacks) (*fallbacks)++;
    }
    return hit | hit_t;
}

}  // namespace wg

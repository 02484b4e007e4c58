// trace_binary.cuh -- closest hit with the reference's own traversal: brute-force spheres and
// quads, then the stack-based front-to-back walk of the binary BVH, same visit order, same
// slab arithmetic, same stack guard (/root/reference/shaders/hitRay.wgsl:1-113).
// This is the arbiter of "what the reference would have hit"; the wide-BVH traversal falls back
// to it for rays it flags as ambiguous.
#pragma once
#include "intersect.cuh"

namespace wg {

#define WGPT_MAX_STACK 64

// spheres + quads part (hitRay.wgsl:6-40), shared by both traversals.  SPH = false: the caller knows the scene has
// no spheres (binding 1 empty) and the sphere / volume code is not even compiled in.
template <bool SPH = true>
__device__ __forceinline__ bool hit_analytic(const DeviceScene& S, const Ray& ray, uint32_t& rng,
                                             Counts* cnt, float& closest, HitLite& h) {
    bool hit_anything = false;
    unsigned draws = 0;
    const int n_spheres = SPH ? S.NUM_SPHERES : 0;
    for (int i = 0; i < n_spheres; i++) {
        const Sphere s = S.spheres[i];
        const float medium = __ldg(&S.materials[(int)s.material_id].material_type);
        if (medium < 3.0f) {
            if (hit_sphere(s, i, WGPT_RAY_TMIN, closest, ray, h)) { hit_anything = true; closest = h.t; }
        } else {
            if (hit_volume(s, i, WGPT_RAY_TMIN, closest, ray, S.materials, rng, draws, h)) { hit_anything = true; closest = h.t; }
        }
    }
    for (int i = 0; i < S.NUM_QUADS; i++) {
        if (hit_quad(S.quads[i], i, WGPT_RAY_TMIN, closest, ray, h)) { hit_anything = true; closest = h.t; }
    }
    if (cnt) { cnt->spheres += S.NUM_SPHERES; cnt->quads += S.NUM_QUADS; cnt->rng += draws; }
    return hit_anything;
}

// BVH part (hitRay.wgsl:42-110).  `closest` carries in the sphere/quad result.
// `box_tmin` is the lower end of the ray interval for the SLAB tests only (triangles always use
// the reference's tmin); the reference's value is WGPT_RAY_TMIN.  The tie resolver of the fast path
// raises it to just below the known closest distance, which skips subtrees that cannot hold a
// triangle at that distance without changing the visit order of the ones that can.
__device__ __forceinline__ bool hit_bvh_binary(const DeviceScene& S, const Ray& ray, Counts* cnt,
                                               ObjRayCache& oc, float& closest, HitLite& h,
                                               float box_tmin = WGPT_RAY_TMIN) {
    if (S.NUM_AABB <= 0) return false;
    bool hit_anything = false;
    const v3 invDir = 1.0f / ray.d;
    int toVisitOffset = 0, cur = 0;
    int stack[WGPT_MAX_STACK];
    const int STACK_SIZE = S.K.STACK_SIZE < WGPT_MAX_STACK ? S.K.STACK_SIZE : WGPT_MAX_STACK;
    const float4* nodes = reinterpret_cast<const float4*>(S.bvh);
    unsigned n_nodes = 0, n_tris = 0;

    while (true) {
        const float4 a = __ldg(&nodes[3 * cur]);         // min.xyz, right_offset
        const float4 b = __ldg(&nodes[3 * cur + 1]);     // max.xyz, prim_type
        n_nodes++;
        if (hit_aabb(V3(a.x, a.y, a.z), V3(b.x, b.y, b.z), box_tmin, closest, ray.o, invDir)) {
            if ((int)b.w == 2) {                         // leaf
                const float4 c = __ldg(&nodes[3 * cur + 2]);   // prim_id, prim_count, skip_link, axis
                const int start = (int)c.x, count = (int)c.y;
                for (int j = 0; j < count; j++) {
                    const float4* t = reinterpret_cast<const float4*>(&S.triangles[start + j]);
                    const float4 A = __ldg(&t[0]), B = __ldg(&t[1]), C = __ldg(&t[2]);
                    const float mesh_id = __ldg(&S.triangles[start + j].mesh_id);
                    oc.fetch(S, (int)mesh_id, ray);
                    float dst, uu, vv;
                    n_tris++;
                    if (tri_test(V3(A.x, A.y, A.z), V3(B.x, B.y, B.z), V3(C.x, C.y, C.z), oc.oray,
                                 WGPT_RAY_TMIN, closest, dst, uu, vv)) {
                        hit_anything = true; closest = dst;
                        h.t = dst; h.kind = HIT_TRI; h.index = start + j; h.uu = uu; h.vv = vv;
                        h.mat_id = oc.m.material_id;
                    }
                }
                if (toVisitOffset == 0) break;
                cur = stack[--toVisitOffset];
            } else {
                const float axis = __ldg(&S.bvh[cur].axis);
                if (comp(ray.d, (int)axis) < 0.0f) { stack[toVisitOffset++] = cur + 1; cur = (int)a.w; }
                else { stack[toVisitOffset++] = (int)a.w; cur = cur + 1; }
            }
        } else {
            if (toVisitOffset == 0) break;
            cur = stack[--toVisitOffset];
        }
        if (toVisitOffset >= STACK_SIZE) { if (cnt) cnt->aborts++; break; }
    }
    if (cnt) { cnt->nodes += n_nodes; cnt->tris += n_tris; }
    return hit_anything;
}

// hitScene (hitRay.wgsl:1-113)
__device__ __forceinline__ bool hitScene_binary(const DeviceScene& S, const Ray& ray, uint32_t& rng,
                                                Counts* cnt, ObjRayCache& oc, HitLite& h) {
    float closest = WGPT_MAX_FLOAT;
    if (cnt) cnt->rays++;
    oc.reset();
    bool hit = hit_analytic(S, ray, rng, cnt, closest, h);
    hit |= hit_bvh_binary(S, ray, cnt, oc, closest, h);
    return hit;
}

}  // namespace wg

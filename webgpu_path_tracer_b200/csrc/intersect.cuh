// intersect.cuh -- RNG and primitive intersectors of the hot path, pinned arithmetic.
//
// Restates (not ports) /root/reference/shaders/common.wgsl:7-12 (PCG hash), :29-146 (sphere,
// constant-density volume), :148-187 (quad), :191-242 (triangle), :245-256 (slab test).
// Differences of FORM that keep every bit: the hit record is kept "lite" (t, kind, index,
// barycentrics) and finalised once per closest-hit query instead of at every accept; the
// object-space ray of hit_triangle is computed once per (ray, mesh) and reused (same inputs,
// same operations => same bits as recomputing it per triangle, common.wgsl:193-197).
#pragma once
#include "pvec.cuh"
#include "scene.cuh"
#include "../../include/wgpt_pinned_math.h"

namespace wg {

#define WGPT_MAX_FLOAT 999999999.999f       /* header.wgsl:3, rounds to 1e9f; MAX_FLOAT+1 rounds to the same */
#define WGPT_MIN_FLOAT 0.0001f
#define WGPT_RAY_TMIN 0.000001f             /* header.wgsl:33 */
#define WGPT_PI_F 3.14159265358979323846f
#define WGPT_TWO_PI_F 6.28318530717958647692f

enum HitKind { HIT_NONE = -1, HIT_SPHERE = 0, HIT_QUAD = 1, HIT_TRI = 2, HIT_VOLUME = 3 };

struct Ray { v3 o, d; };

// closest-hit result before finalisation
struct HitLite {
    float t;
    int kind;        // HitKind
    int index;       // sphere / quad index, or POSITION in the triangle buffer
    float uu, vv;    // triangle barycentrics as hit_triangle computes them
    int mat_id;      // hitRec.material as an index (sticky: see hit_volume)
};

struct Counts { unsigned rays, nodes, tris, spheres, quads, rng, samples, aborts; };

__device__ __forceinline__ v3 ld3(const float* p) { return V3(p[0], p[1], p[2]); }
__device__ __forceinline__ v3 at(const Ray& r, float t) { return r.o + t * r.d; }

// common.wgsl:7-12; the divisor 4294967295 rounds to 2^32 in f32, and dividing by a power of two
// is the same correctly rounded value as multiplying by its (exact) reciprocal
__device__ __forceinline__ float rand2D(uint32_t& s) {
    s = s * 747796405u + 2891336453u;
    const uint32_t word = ((s >> ((s >> 28u) + 4u)) ^ s) * 277803737u;
    return fm((float)((word >> 22u) ^ word), 2.3283064365386962890625e-10f);
}

// roots of common.wgsl:33-52 / :78-97; returns false when no root lies in (tmin, tmax)
__device__ __forceinline__ bool sphere_root(const Sphere& s, float tmin, float tmax, const Ray& ray, float& root) {
    const v3 oc = ray.o - V3(s.cx, s.cy, s.cz);
    const float a = dot(ray.d, ray.d);
    const float half_b = dot(ray.d, oc);
    const float c = fs(dot(oc, oc), fm(s.r, s.r));
    const float disc = fs(fm(half_b, half_b), fm(a, c));
    if (disc < 0.0f) return false;
    const float sqrtd = fsq(disc);
    root = fd(fs(-half_b, sqrtd), a);
    if (root <= tmin || root >= tmax) {
        root = fd(fa(-half_b, sqrtd), a);
        if (root <= tmin || root >= tmax) return false;
    }
    return true;
}

// common.wgsl:29-73 (accept part deferred to finalize_hit)
__device__ __forceinline__ bool hit_sphere(const Sphere& s, int idx, float tmin, float tmax, const Ray& ray, HitLite& h) {
    float root;
    if (!sphere_root(s, tmin, tmax, ray, root)) return false;
    h.t = root; h.kind = HIT_SPHERE; h.index = idx; h.mat_id = (int)s.material_id;
    return true;
}

// common.wgsl:102-146.  hit_sphere_local's sentinel MAX_FLOAT+1 == 1e9f can also be a genuine
// root value only if root == 1e9f, which (tmin,tmax) exclusion already rules out.
// The material is stored BEFORE the accept test (:130 vs :136) and one number is drawn (:134).
__device__ __forceinline__ bool hit_volume(const Sphere& s, int idx, float tmin, float tmax, const Ray& ray,
                                           const Material* __restrict__ materials, uint32_t& rng, unsigned& draws, HitLite& h) {
    float rec1, rec2;
    if (!sphere_root(s, -WGPT_MAX_FLOAT, WGPT_MAX_FLOAT, ray, rec1)) return false;
    if (!sphere_root(s, fa(rec1, 0.0001f), WGPT_MAX_FLOAT, ray, rec2)) return false;
    if (rec1 < tmin) rec1 = tmin;
    if (rec2 > tmax) rec2 = tmax;
    if (rec1 >= rec2) return false;
    if (rec1 < 0.0f) rec1 = 0.0f;

    h.mat_id = (int)s.material_id;
    const float ray_length = length(ray.d);
    const float dist_inside = fm(fs(rec2, rec1), ray_length);
    draws++;
    const float hit_dist = fm(__ldg(&materials[h.mat_id].roughness), wgpt_pm::logf_(rand2D(rng)));
    if (hit_dist > dist_inside) return false;

    h.t = fa(rec1, fd(hit_dist, ray_length));
    h.kind = HIT_VOLUME; h.index = idx;
    return true;
}

// common.wgsl:148-187
__device__ __forceinline__ bool hit_quad(const Quad& q, int idx, float tmin, float tmax, const Ray& ray, HitLite& h) {
    const v3 qn = ld3(q.normal);
    if (dot(ray.d, qn) > 0.0f) return false;
    const float denom = dot(qn, ray.d);
    if (fabsf(denom) < 1e-8f) return false;
    const float t = fd(fs(q.D, dot(qn, ray.o)), denom);
    if (t <= tmin || t >= tmax) return false;
    const v3 planar = at(ray, t) - ld3(q.Q);
    const float alpha = dot(ld3(q.w), cross(planar, ld3(q.v)));
    const float beta = dot(ld3(q.w), cross(ld3(q.u), planar));
    if (alpha < 0.0f || 1.0f < alpha || beta < 0.0f || 1.0f < beta) return false;
    h.t = t; h.kind = HIT_QUAD; h.index = idx; h.mat_id = (int)q.material_id;
    return true;
}

// common.wgsl:199-222 on an already object-space ray; A, B, C object-space vertices.
// `eps` is the reference's tmin doubling as determinant / barycentric epsilon; accepts dst == tmax.
__device__ __forceinline__ bool tri_test(v3 A, v3 B, v3 C, const Ray& oray, float eps, float tmax,
                                         float& dst, float& uu, float& vv) {
    const v3 AB = B - A, AC = C - A;
    const v3 normal = cross(AB, AC);
    const float det = -dot(oray.d, normal);
    if (fabsf(det) < eps) return false;
    const v3 ao = oray.o - A;
    const v3 dao = cross(ao, oray.d);
    const float invDet = fd(1.0f, det);
    dst = fm(dot(ao, normal), invDet);
    uu = fm(dot(AC, dao), invDet);
    vv = fm(-dot(AB, dao), invDet);
    const float ww = fs(fs(1.0f, uu), vv);
    return !(dst < eps || dst > tmax || uu < eps || vv < eps || ww < eps);
}

// common.wgsl:245-256
__device__ __forceinline__ bool hit_aabb(v3 mn, v3 mx, float tmin, float tmax, v3 o, v3 invDir) {
    const v3 t0 = (mn - o) * invDir, t1 = (mx - o) * invDir;
    const float sx = bmin(t0.x, t1.x), sy = bmin(t0.y, t1.y), sz = bmin(t0.z, t1.z);
    const float bx = bmax(t0.x, t1.x), by = bmax(t0.y, t1.y), bz = bmax(t0.z, t1.z);
    const float t_min = bmax(tmin, bmax(sx, bmax(sy, sz)));
    const float t_max = bmin(tmax, bmin(bx, bmin(by, bz)));
    return t_max > t_min;
}

// Object-space ray cache for hit_triangle (common.wgsl:193-197).
struct ObjRayCache {
    int mesh;            // index into meshes[] the cached ray belongs to, -1 = none
    Mesh m;
    Ray oray;
    __device__ __forceinline__ void reset() { mesh = -1; }
    __device__ __forceinline__ void fetch(const DeviceScene& S, int mesh_idx, const Ray& ray) {
        if (mesh_idx == mesh) return;
        mesh = mesh_idx;
        m = S.meshes[mesh_idx];
        const float* inv = S.transforms[m.global_id].invModelMatrix;
        oray.o = mul4_xyz(inv, ray.o, 1.0f);
        oray.d = mul4_xyz(inv, ray.d, 0.0f);
    }
};

}  // namespace wg

/*
 * oracle/oracle_shader.cpp -- TEST INFRASTRUCTURE (CPU oracle).  Not part of the product path.
 * Restatement of the reference WGSL; see oracle_shader.h.  PARITY UNPINNED (no reference tests).
 * Build: g++ -O2 -ffp-contract=off -fno-fast-math (oracle/Makefile).
 */
#include "oracle_shader.h"

namespace orc {

using wgpt_pm::sinf_; using wgpt_pm::cosf_; using wgpt_pm::acosf_; using wgpt_pm::logf_; using wgpt_pm::powf_;

/* common.wgsl:29-73 */
bool Invocation::hit_sphere(int idx, float tmin, float tmax, const Ray& ray) {
    cnt.sphere_tests++;
    const Sphere& sphere = S.spheres[idx];
    const vec3 center = V3(sphere.cx, sphere.cy, sphere.cz);
    const vec3 oc = ray.origin - center;
    const float a = dot(ray.dir, ray.dir);
    const float half_b = dot(ray.dir, oc);
    const float c = dot(oc, oc) - sphere.r * sphere.r;
    const float discriminant = half_b * half_b - a * c;
    if (discriminant < 0) return false;

    const float sqrtd = sqrtf(discriminant);
    float root = (-half_b - sqrtd) / a;
    if (root <= tmin || root >= tmax) {
        root = (-half_b + sqrtd) / a;
        if (root <= tmin || root >= tmax) return false;
    }
    hit_t = root;
    hit_p = at(ray, root);
    hit_normal = normalize((hit_p - center) / sphere.r);
    hit_front_face = dot(ray.dir, hit_normal) < 0;
    if (hit_front_face == false) hit_normal = -hit_normal;
    hit_material = (int)sphere.material_id;
    hit_type = 0; hit_index = idx;
    return true;
}

/* common.wgsl:75-100 */
float Invocation::hit_sphere_local(const Sphere& sphere, float tmin, float tmax, const Ray& ray) {
    const vec3 center = V3(sphere.cx, sphere.cy, sphere.cz);
    const vec3 oc = ray.origin - center;
    const float a = dot(ray.dir, ray.dir);
    const float half_b = dot(ray.dir, oc);
    const float c = dot(oc, oc) - sphere.r * sphere.r;
    const float discriminant = half_b * half_b - a * c;
    if (discriminant < 0) return MAX_FLOAT_P1;
    const float sqrtd = sqrtf(discriminant);
    float root = (-half_b - sqrtd) / a;
    if (root <= tmin || root >= tmax) {
        root = (-half_b + sqrtd) / a;
        if (root <= tmin || root >= tmax) return MAX_FLOAT_P1;
    }
    return root;
}

/* common.wgsl:102-146 -- constant-density medium; note the material store at :130 happens
 * BEFORE the accept test at :136 (quirk Q2) and draws exactly one random number at :134. */
bool Invocation::hit_volume(int idx, float tmin, float tmax, const Ray& ray) {
    cnt.sphere_tests++;
    const Sphere& sphere = S.spheres[idx];
    float rec1 = hit_sphere_local(sphere, -MAX_FLOAT, MAX_FLOAT, ray);
    if (rec1 == MAX_FLOAT_P1) return false;
    float rec2 = hit_sphere_local(sphere, rec1 + 0.0001f, MAX_FLOAT, ray);
    if (rec2 == MAX_FLOAT_P1) return false;
    if (rec1 < tmin) rec1 = tmin;
    if (rec2 > tmax) rec2 = tmax;
    if (rec1 >= rec2) return false;
    if (rec1 < 0) rec1 = 0;

    hit_material = (int)sphere.material_id;
    const float ray_length = length(ray.dir);
    const float dist_inside = (rec2 - rec1) * ray_length;
    const float hit_dist = mat().roughness * logf_(rand2D());
    if (hit_dist > dist_inside) return false;

    hit_t = rec1 + (hit_dist / ray_length);
    hit_p = at(ray, hit_t);
    hit_normal = normalize(hit_p - V3(sphere.cx, sphere.cy, sphere.cz));
    hit_front_face = true;
    hit_type = 0; hit_index = idx;
    return true;
}

/* common.wgsl:148-187 */
bool Invocation::hit_quad(int idx, float tmin, float tmax, const Ray& ray) {
    cnt.quad_tests++;
    const Quad& quad = S.quads[idx];
    const vec3 qn = v3(quad.normal);
    if (dot(ray.dir, qn) > 0) return false;
    const float denom = dot(qn, ray.dir);
    if (fabsf(denom) < 1e-8f) return false;
    const float t = (quad.D - dot(qn, ray.origin)) / denom;
    if (t <= tmin || t >= tmax) return false;

    const vec3 intersection = at(ray, t);
    const vec3 planar_hitpt_vector = intersection - v3(quad.Q);
    const float alpha = dot(v3(quad.w), cross(planar_hitpt_vector, v3(quad.v)));
    const float beta = dot(v3(quad.w), cross(v3(quad.u), planar_hitpt_vector));
    if (alpha < 0 || 1 < alpha || beta < 0 || 1 < beta) return false;

    hit_t = t;
    hit_p = intersection;
    hit_normal = normalize(qn);
    hit_front_face = dot(ray.dir, hit_normal) < 0;
    if (hit_front_face == false) hit_normal = -hit_normal;
    hit_material = (int)quad.material_id;
    hit_type = 1; hit_index = idx;
    return true;
}

/* common.wgsl:191-242 -- object-space test; `tmin` doubles as the determinant / barycentric
 * epsilon; accepts dst == tmax. */
bool Invocation::hit_triangle(int idx, float tmin, float tmax, const Ray& incidentRay) {
    cnt.tri_tests++;
    const Triangle& tri = S.triangles[idx];
    const Mesh& mesh = S.meshes[(int)tri.mesh_id];
    const mat4& invModelMatrix = S.transforms[mesh.global_id].invModelMatrix;

    Ray ray;
    ray.origin = xyz(mul(invModelMatrix, V4(incidentRay.origin, 1.0f)));
    ray.dir = xyz(mul(invModelMatrix, V4(incidentRay.dir, 0.0f)));

    const vec3 A = v3(tri.A);
    const vec3 AB = v3(tri.B) - A;
    const vec3 AC = v3(tri.C) - A;
    const vec3 normal = cross(AB, AC);
    const float determinant = -dot(ray.dir, normal);
    if (fabsf(determinant) < tmin) return false;

    const vec3 ao = ray.origin - A;
    const vec3 dao = cross(ao, ray.dir);
    const float invDet = 1.0f / determinant;
    const float dst = dot(ao, normal) * invDet;
    const float uu = dot(AC, dao) * invDet;
    const float vv = -dot(AB, dao) * invDet;
    const float ww = (1.0f - uu) - vv;
    if (dst < tmin || dst > tmax || uu < tmin || vv < tmin || ww < tmin) return false;

    hit_t = dst;
    hit_p = at(incidentRay, dst);
    vec3 n = (v3(tri.nA) * ww + v3(tri.nB) * uu) + v3(tri.nC) * vv;
    hit_normal = normalize(xyz(mulT(invModelMatrix, V4(n, 0.0f))));
    hit_front_face = dot(incidentRay.dir, hit_normal) < 0;
    if (hit_front_face == false) hit_normal = -hit_normal;
    hit_material = mesh.material_id;
    hit_type = 2; hit_index = (int)tri.local_id;
    return true;
}

/* common.wgsl:245-256 */
bool Invocation::hit_aabb(const BvhNode& box, float tmin, float tmax, const Ray& ray, vec3 invDir) {
    cnt.node_visits++;
    const vec3 t0s = (v3(box.mn) - ray.origin) * invDir;
    const vec3 t1s = (v3(box.mx) - ray.origin) * invDir;
    const vec3 tsmaller = vmin(t0s, t1s);
    const vec3 tbigger = vmax(t0s, t1s);
    const float t_min = fmax_(tmin, fmax_(tsmaller.x, fmax_(tsmaller.y, tsmaller.z)));
    const float t_max = fmin_(tmax, fmin_(tbigger.x, fmin_(tbigger.y, tbigger.z)));
    return t_max > t_min;
}

/* common.wgsl:258-269 */
void Invocation::get_lights() {
    for (int i = 0; i < S.NUM_QUADS; i++) {
        const Material m = mat_at((int)S.quads[i].material_id);
        if (m.emissionColor[0] > 0.0f) { lights = S.quads[i]; break; }
    }
}

/* hitRay.wgsl:1-113 */
bool Invocation::hitScene(const Ray& ray) {
    cnt.rays++;
    float closest_so_far = MAX_FLOAT;
    bool hit_anything = false;

    for (int i = 0; i < S.NUM_SPHERES; i++) {
        const float medium = mat_at((int)S.spheres[i].material_id).material_type;
        if (medium < ISOTROPIC) {
            if (hit_sphere(i, ray_tmin, closest_so_far, ray)) { hit_anything = true; closest_so_far = hit_t; }
        } else {
            if (hit_volume(i, ray_tmin, closest_so_far, ray)) { hit_anything = true; closest_so_far = hit_t; }
        }
    }
    for (int i = 0; i < S.NUM_QUADS; i++) {
        if (hit_quad(i, ray_tmin, closest_so_far, ray)) { hit_anything = true; closest_so_far = hit_t; }
    }

    if (S.NUM_AABB <= 0) return hit_anything;      /* the reference always binds >= 1 node */

    const int leafNode = 2;
    const vec3 invDir = 1.0f / ray.dir;
    int toVisitOffset = 0;
    int curNodeIdx = 0;
    int stack[64];
    const int STACK_SIZE = K.STACK_SIZE < 64 ? K.STACK_SIZE : 64;

    while (true) {
        const BvhNode& node = S.bvh[curNodeIdx];
        if (hit_aabb(node, ray_tmin, closest_so_far, ray, invDir)) {
            if ((int)node.prim_type == leafNode) {
                const int startPrim = (int)node.prim_id;
                const int countPrim = (int)node.prim_count;
                for (int j = 0; j < countPrim; j++) {
                    if (hit_triangle(startPrim + j, ray_tmin, closest_so_far, ray)) {
                        hit_anything = true; closest_so_far = hit_t;
                    }
                }
                if (toVisitOffset == 0) break;
                toVisitOffset--;
                curNodeIdx = stack[toVisitOffset];
            } else {
                if (comp(ray.dir, (int)node.axis) < 0) {
                    stack[toVisitOffset] = curNodeIdx + 1;
                    toVisitOffset++;
                    curNodeIdx = (int)node.right_offset;
                } else {
                    stack[toVisitOffset] = (int)node.right_offset;
                    toVisitOffset++;
                    curNodeIdx++;
                }
            }
        } else {
            if (toVisitOffset == 0) break;
            toVisitOffset--;
            curNodeIdx = stack[toVisitOffset];
        }
        if (toVisitOffset >= STACK_SIZE) { cnt.stack_aborts++; break; }
    }
    return hit_anything;
}

/* The reference's own commented-out specification of closest hit without the BVH
 * (hitRay.wgsl:188-221): spheres, quads, then every triangle in array order. */
bool Invocation::hitScene_bruteForce(const Ray& ray) {
    float closest_so_far = MAX_FLOAT;
    bool hit_anything = false;
    for (int i = 0; i < S.NUM_SPHERES; i++) {
        const float medium = mat_at((int)S.spheres[i].material_id).material_type;
        if (medium < ISOTROPIC) {
            if (hit_sphere(i, ray_tmin, closest_so_far, ray)) { hit_anything = true; closest_so_far = hit_t; }
        } else {
            if (hit_volume(i, ray_tmin, closest_so_far, ray)) { hit_anything = true; closest_so_far = hit_t; }
        }
    }
    for (int i = 0; i < S.NUM_QUADS; i++)
        if (hit_quad(i, ray_tmin, closest_so_far, ray)) { hit_anything = true; closest_so_far = hit_t; }
    for (int i = 0; i < S.NUM_TRIANGLES; i++)
        if (hit_triangle(i, ray_tmin, closest_so_far, ray)) { hit_anything = true; closest_so_far = hit_t; }
    return hit_anything;
}

}  /* namespace orc */

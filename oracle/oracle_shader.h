/*
 * oracle/oracle_shader.h -- TEST INFRASTRUCTURE (CPU oracle).  Not part of the product path.
 *
 * Line-by-line CPU restatement of the reference's WGSL compute path
 * (/root/reference/shaders/{header,common,importanceSampling,scatterRay,traceRay,hitRay,
 * shootRay,main}.wgsl and shaders/fragment.js).  Each function cites the lines it follows.
 * The reference's per-invocation `var<private>` globals live in `Invocation` so their sticky
 * behaviour (hitRec.material overwritten by hit_volume before its accept test, scatterRec.skip_pdf,
 * the ONB u/v/unit_w, doSpecular) is preserved.
 *
 * PARITY UNPINNED: the reference has no tests, golden vectors or runnable implementation on
 * this box (WGSL needs a WebGPU stack; no JS engine), so this oracle is pinned only by the
 * independent checks in tests/ (RNG known answers, closed-form intersections, brute-force
 * equivalence, furnace test).
 */
#ifndef WGPT_ORACLE_SHADER_H
#define WGPT_ORACLE_SHADER_H

#include <stdint.h>
#include "ovec.h"
#include "../include/wgpt_pinned_math.h"

namespace orc {

/* header.wgsl:1-8 */
static const float PI_F = 3.14159265358979323846f;
static const float TWO_PI_F = 6.28318530717958647692f;     /* `2 * PI` folds in abstract float */
static const float MIN_FLOAT = 0.0001f;
static const float MAX_FLOAT = 999999999.999f;             /* rounds to 1e9f */
static const float MAX_FLOAT_P1 = 1000000000.999f;         /* `MAX_FLOAT + 1` folds, rounds to 1e9f */
static const float LAMBERTIAN = 0.0f, MIRROR = 1.0f, GLASS = 2.0f, ISOTROPIC = 3.0f;

/* header.wgsl:9-13 -- compile-time constants of the shader, run-time here. */
struct Consts {
    int NUM_SAMPLES = 1;
    int MAX_BOUNCES = 100;
    int STRATIFY = 0;
    int IMPORTANCE_SAMPLING = 0;
    int STACK_SIZE = 20;
};

/* Byte layouts of the storage buffers: SURVEY.md §8(a) T-0..T-9 (header.wgsl:41-117). */
struct Uniforms { float screenW, screenH, frameNum, resetBuffer; mat4 viewMatrix; };      /* 80 B */
struct Sphere   { float cx, cy, cz, r, global_id, local_id, material_id, pad; };          /* 32 B */
struct Quad     { float Q[3], pad0, u[3], local_id, v[3], global_id, normal[3], D, w[3], material_id; }; /* 80 B */
struct Triangle { float A[3], p0, B[3], p1, C[3], p2, nA[3], p3, nB[3], local_id, nC[3], mesh_id; };     /* 96 B */
struct Mesh     { int32_t num_triangles, offset, global_id, material_id; };               /* 16 B */
struct ModelTransform { mat4 modelMatrix, invModelMatrix; };                              /* 128 B */
struct Material { float color[3], p0, specularColor[3], p1, emissionColor[3], specularStrength,
                  roughness, eta, material_type, p2; };                                   /* 64 B */
struct BvhNode  { float mn[3], right_offset, mx[3], prim_type, prim_id, prim_count, skip_link, axis; }; /* 48 B */

static_assert(sizeof(Uniforms) == 80 && sizeof(Sphere) == 32 && sizeof(Quad) == 80, "layout");
static_assert(sizeof(Triangle) == 96 && sizeof(Mesh) == 16 && sizeof(ModelTransform) == 128, "layout");
static_assert(sizeof(Material) == 64 && sizeof(BvhNode) == 48, "layout");

struct SceneView {
    Uniforms uniforms;
    const Sphere* spheres; int NUM_SPHERES;
    const Quad* quads; int NUM_QUADS;
    const Triangle* triangles; int NUM_TRIANGLES;
    const Mesh* meshes; int NUM_MESHES;
    const ModelTransform* transforms; int NUM_TRANSFORMS;
    const Material* materials; int NUM_MATERIALS;
    const BvhNode* bvh; int NUM_AABB;
};

struct Ray { vec3 origin, dir; };

struct Counters {
    uint64_t rays = 0, node_visits = 0, tri_tests = 0, sphere_tests = 0, quad_tests = 0,
             rng_draws = 0, samples = 0, stack_aborts = 0;
    void add(const Counters& o) {
        rays += o.rays; node_visits += o.node_visits; tri_tests += o.tri_tests;
        sphere_tests += o.sphere_tests; quad_tests += o.quad_tests; rng_draws += o.rng_draws;
        samples += o.samples; stack_aborts += o.stack_aborts;
    }
};

static inline vec3 v3(const float* p) { return V3(p[0], p[1], p[2]); }

/* One shader invocation: the `var<private>` state of header.wgsl:25-39, shootRay.wgsl:51-52,
 * scatterRay.wgsl:1, importanceSampling.wgsl:56-58. */
struct Invocation {
    const SceneView& S;
    const Consts& K;
    Counters cnt;

    uint32_t randState = 0;
    vec3 pixelCoords = {0, 0, 0};
    /* HitRecord (header.wgsl:119-125); material kept as an index: the table is read-only. */
    vec3 hit_p = {0, 0, 0}; float hit_t = 0; vec3 hit_normal = {0, 0, 0}; bool hit_front_face = false;
    int hit_material = -1;
    /* primitive id of the accepted hit (the reference exposes none; SURVEY.md §8(c) convention:
     * type 0 sphere / 1 quad / 2 triangle (Triangle.local_id) / -1 miss) */
    int hit_type = -1, hit_index = -1;
    /* ScatterRecord (header.wgsl:127-131) */
    bool skip_pdf = false; Ray skip_pdf_ray = {{0, 0, 0}, {0, 0, 0}};
    Quad lights;
    float ray_tmin = 0.000001f, ray_tmax = MAX_FLOAT;
    float fovFactor = 0; vec3 cam_origin = {0, 0, 0};
    float doSpecular = 0;
    vec3 unit_w = {0, 0, 0}, u = {0, 0, 0}, v = {0, 0, 0};
    /* first hitScene() result of the first sample, recorded for the parity hooks */
    bool primary_done = false; int primary_type = -1, primary_index = -1; float primary_t = 0;

    Invocation(const SceneView& s, const Consts& k) : S(s), K(k) {
        for (unsigned i = 0; i < sizeof(Quad) / 4; i++) ((float*)&lights)[i] = 0.0f;
    }

    Material zero_material() const { Material m; for (int i = 0; i < 16; i++) ((float*)&m)[i] = 0.0f; return m; }
    Material mat() const { return (hit_material >= 0 && hit_material < S.NUM_MATERIALS) ? S.materials[hit_material] : zero_material(); }
    Material mat_at(int i) const { return (i >= 0 && i < S.NUM_MATERIALS) ? S.materials[i] : zero_material(); }

    /* common.wgsl:1-3 */
    static vec3 at(const Ray& ray, float t) { return ray.origin + t * ray.dir; }

    /* common.wgsl:7-12 (PCG hash; divisor 4294967295 rounds to 2^32 in f32) */
    float rand2D() {
        cnt.rng_draws++;
        randState = randState * 747796405u + 2891336453u;
        uint32_t word = ((randState >> ((randState >> 28u) + 4u)) ^ randState) * 277803737u;
        return (float)((word >> 22u) ^ word) / 4294967296.0f;
    }

    bool hit_sphere(int idx, float tmin, float tmax, const Ray& ray);
    float hit_sphere_local(const Sphere& sphere, float tmin, float tmax, const Ray& ray);
    bool hit_volume(int idx, float tmin, float tmax, const Ray& ray);
    bool hit_quad(int idx, float tmin, float tmax, const Ray& ray);
    bool hit_triangle(int idx, float tmin, float tmax, const Ray& incidentRay);
    bool hit_aabb(const BvhNode& box, float tmin, float tmax, const Ray& ray, vec3 invDir);
    void get_lights();
    bool hitScene(const Ray& ray);
    bool hitScene_bruteForce(const Ray& ray);

    float reflectance(float cosine, float ref_idx);
    vec3 uniform_random_in_unit_sphere();
    vec3 cosine_sampling_wrt_Z();
    void onb_build_from_w(vec3 w);
    vec3 onb_get_local(vec3 a) const { return (u * a.x + v * a.y) + unit_w * a.z; }
    float onb_lambertian_scattering_pdf(const Ray& scattered);
    Ray get_random_on_quad(const Quad& q, vec3 origin);
    float light_pdf(const Ray& ray, const Quad& quad);
    Ray material_scatter(const Ray& ray_in);
    vec3 ray_color(const Ray& incidentRay);
    Ray getCameraRay(float s, float t);
    vec3 pathTrace();
    vec3 computeFrameBuffer(uint32_t pixelIndex);
};

}  /* namespace orc */
#endif

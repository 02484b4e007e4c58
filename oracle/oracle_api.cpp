/*
 * oracle/oracle_api.cpp -- TEST INFRASTRUCTURE (CPU oracle).  Not part of the product path.
 *
 * C entry points around the WGSL restatement, shaped like include/wgpt.h so the parity tests
 * drive both sides with the same calls.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this library.  PARITY UNPINNED (see
 * oracle_shader.h).  The dispatch / framebuffer semantics follow main.wgsl:20-27,
 * fragment.js:22-36 and renderer.js:163-191 of the reference.
 */
#include <stdlib.h>
#include <string.h>
#include <vector>
#include <omp.h>
#include "oracle_shader.h"

using namespace orc;

struct OracleCtx {
    int W, H;
    Consts K;
    std::vector<unsigned char> buf[10];
    std::vector<float> framebuffer;   /* binding 3: vec4 per pixel */
    Counters counters;
    int threads;
};

static SceneView make_view(OracleCtx* c) {
    SceneView S;
    memset(&S, 0, sizeof(S));
    if (c->buf[0].size() >= sizeof(Uniforms)) memcpy(&S.uniforms, c->buf[0].data(), sizeof(Uniforms));
    S.spheres = (const Sphere*)c->buf[1].data();     S.NUM_SPHERES = (int)(c->buf[1].size() / sizeof(Sphere));
    S.quads = (const Quad*)c->buf[2].data();         S.NUM_QUADS = (int)(c->buf[2].size() / sizeof(Quad));
    S.triangles = (const Triangle*)c->buf[5].data(); S.NUM_TRIANGLES = (int)(c->buf[5].size() / sizeof(Triangle));
    S.meshes = (const Mesh*)c->buf[6].data();        S.NUM_MESHES = (int)(c->buf[6].size() / sizeof(Mesh));
    S.transforms = (const ModelTransform*)c->buf[7].data(); S.NUM_TRANSFORMS = (int)(c->buf[7].size() / sizeof(ModelTransform));
    S.materials = (const Material*)c->buf[8].data(); S.NUM_MATERIALS = (int)(c->buf[8].size() / sizeof(Material));
    S.bvh = (const BvhNode*)c->buf[9].data();        S.NUM_AABB = (int)(c->buf[9].size() / sizeof(BvhNode));
    return S;
}

extern "C" {

void* oracle_create(int W, int H, int num_samples, int max_bounces, int stratify,
                    int importance_sampling, int stack_size, int threads) {
    OracleCtx* c = new OracleCtx();
    c->W = W; c->H = H;
    c->K.NUM_SAMPLES = num_samples; c->K.MAX_BOUNCES = max_bounces; c->K.STRATIFY = stratify;
    c->K.IMPORTANCE_SAMPLING = importance_sampling; c->K.STACK_SIZE = stack_size;
    c->framebuffer.assign((size_t)W * H * 4, 0.0f);
    c->threads = threads > 0 ? threads : omp_get_max_threads();
    return c;
}

void oracle_destroy(void* h) { delete (OracleCtx*)h; }

int oracle_threads(void* h) { return ((OracleCtx*)h)->threads; }

int oracle_write_buffer(void* h, int binding, const void* data, size_t bytes) {
    OracleCtx* c = (OracleCtx*)h;
    if (binding < 0 || binding > 9 || binding == 4) return -1;
    if (binding == 3) {
        if (bytes != c->framebuffer.size() * 4) return -2;
        memcpy(c->framebuffer.data(), data, bytes);
        return 0;
    }
    c->buf[binding].assign((const unsigned char*)data, (const unsigned char*)data + bytes);
    return 0;
}

int oracle_read_framebuffer(void* h, float* out) {
    OracleCtx* c = (OracleCtx*)h;
    memcpy(out, c->framebuffer.data(), c->framebuffer.size() * 4);
    return 0;
}

/* One computeFrameBuffer dispatch with the uniforms currently in binding 0 (main.wgsl:1-28).
 * `radiance_out` (optional, 3 floats/pixel) receives pathTrace()'s value per pixel;
 * `prim_*` (optional) the first closest-hit query of each pixel.  If `accumulate` is 0 the
 * framebuffer is left untouched (parity hooks). */
int oracle_dispatch_ex(void* h, int accumulate, float* radiance_out, int* prim_type, int* prim_index, float* prim_t) {
    OracleCtx* c = (OracleCtx*)h;
    const SceneView S = make_view(c);
    const int npix = c->W * c->H;
    const bool reset = S.uniforms.resetBuffer != 0.0f;
    Counters total;
    #pragma omp parallel num_threads(c->threads)
    {
        Counters local;
        #pragma omp for schedule(dynamic, 256)
        for (int p = 0; p < npix; p++) {
            Invocation inv(S, c->K);
            const vec3 col = inv.computeFrameBuffer((uint32_t)p);
            local.add(inv.cnt);
            if (radiance_out) { radiance_out[3 * p] = col.x; radiance_out[3 * p + 1] = col.y; radiance_out[3 * p + 2] = col.z; }
            if (prim_type) prim_type[p] = inv.primary_type;
            if (prim_index) prim_index[p] = inv.primary_index;
            if (prim_t) prim_t[p] = inv.primary_t;
            if (accumulate) {
                float* fb = &c->framebuffer[(size_t)p * 4];
                if (!reset) { fb[0] = fb[0] + col.x; fb[1] = fb[1] + col.y; fb[2] = fb[2] + col.z; }
                else { fb[0] = col.x; fb[1] = col.y; fb[2] = col.z; }
                fb[3] = 1.0f;
            }
        }
        #pragma omp critical
        total.add(local);
    }
    c->counters.add(total);
    return 0;
}

int oracle_dispatch(void* h) { return oracle_dispatch_ex(h, 1, 0, 0, 0, 0); }

/* Frames first_frame .. first_frame+n-1 with resetBuffer = 0 (renderer.js:163-188 loop). */
int oracle_render(void* h, int first_frame, int n_frames) {
    OracleCtx* c = (OracleCtx*)h;
    if (c->buf[0].size() < sizeof(Uniforms)) return -3;
    Uniforms* U = (Uniforms*)c->buf[0].data();
    for (int f = 0; f < n_frames; f++) {
        U->frameNum = (float)(first_frame + f);
        U->resetBuffer = 0.0f;
        oracle_dispatch(h);
    }
    return 0;
}

/* fs of fragment.js:22-36: sum/frameNum -> ACES (common.wgsl:273-282) -> pow 1/2.2; clears the
 * framebuffer when resetBuffer == 1.  rgba_out: 4 floats per pixel. */
int oracle_present(void* h, float* rgba_out) {
    OracleCtx* c = (OracleCtx*)h;
    const SceneView S = make_view(c);
    const int npix = c->W * c->H;
    const float frameNum = S.uniforms.frameNum;
    for (int p = 0; p < npix; p++) {
        float* fb = &c->framebuffer[(size_t)p * 4];
        for (int k = 0; k < 3; k++) {
            float col = fb[k] / frameNum;
            const float v1 = col * 0.6f;
            float r = (v1 * (2.51f * v1 + 0.03f)) / (v1 * (2.43f * v1 + 0.59f) + 0.14f);
            r = fmin_(fmax_(r, 0.0f), 1.0f);
            rgba_out[4 * p + k] = wgpt_pm::powf_(r, 0.45454545454545453f);
        }
        rgba_out[4 * p + 3] = 1.0f;
        if (S.uniforms.resetBuffer == 1.0f) { fb[0] = fb[1] = fb[2] = fb[3] = 0.0f; }
    }
    return 0;
}

/* Closest-hit query for caller-supplied rays (6 floats each: origin, dir); seeds[i] is the RNG
 * state on entry (hit_volume draws).  brute_force != 0 uses the triangle loop of
 * hitRay.wgsl:188-221 instead of the BVH. */
int oracle_trace_rays(void* h, int n, const float* rays, const uint32_t* seeds, int brute_force,
                      int* out_type, int* out_index, float* out_t) {
    OracleCtx* c = (OracleCtx*)h;
    const SceneView S = make_view(c);
    Counters total;
    #pragma omp parallel num_threads(c->threads)
    {
        Counters local;
        #pragma omp for schedule(dynamic, 256)
        for (int i = 0; i < n; i++) {
            Invocation inv(S, c->K);
            inv.randState = seeds ? seeds[i] : 0u;
            Ray r = {V3(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]), V3(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5])};
            const bool hit = brute_force ? inv.hitScene_bruteForce(r) : inv.hitScene(r);
            out_type[i] = hit ? inv.hit_type : -1;
            out_index[i] = hit ? inv.hit_index : -1;
            out_t[i] = hit ? inv.hit_t : 0.0f;
            local.add(inv.cnt);
        }
        #pragma omp critical
        total.add(local);
    }
    c->counters.add(total);
    return 0;
}

/* out[8]: rays, node_visits, tri_tests, sphere_tests, quad_tests, rng_draws, samples, stack_aborts */
int oracle_counters(void* h, uint64_t* out, int reset) {
    OracleCtx* c = (OracleCtx*)h;
    out[0] = c->counters.rays; out[1] = c->counters.node_visits; out[2] = c->counters.tri_tests;
    out[3] = c->counters.sphere_tests; out[4] = c->counters.quad_tests; out[5] = c->counters.rng_draws;
    out[6] = c->counters.samples; out[7] = c->counters.stack_aborts;
    if (reset) c->counters = Counters();
    return 0;
}

/* direct access to the RNG and the pinned math for known-answer tests */
void oracle_rand_sequence(uint32_t seed, int n, uint32_t* states, float* values) {
    SceneView S; memset(&S, 0, sizeof(S)); Consts K;
    Invocation inv(S, K);
    inv.randState = seed;
    for (int i = 0; i < n; i++) { values[i] = inv.rand2D(); states[i] = inv.randState; }
}
void oracle_pinned_math(int fn, int n, const float* x, const float* y, float* out) {
    for (int i = 0; i < n; i++) {
        switch (fn) {
            case 0: out[i] = wgpt_pm::sinf_(x[i]); break;
            case 1: out[i] = wgpt_pm::cosf_(x[i]); break;
            case 2: out[i] = wgpt_pm::acosf_(x[i]); break;
            case 3: out[i] = wgpt_pm::logf_(x[i]); break;
            default: out[i] = wgpt_pm::powf_(x[i], y[i]); break;
        }
    }
}

}  /* extern "C" */

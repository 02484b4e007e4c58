// exact.cu -- K1, the exact-mode integrator: one thread per (pixel, frame), the reference's own
// binary-BVH traversal and operation order, pinned IEEE arithmetic.  Bit-identical to the CPU
// oracle by construction; it is the arbiter the wavefront path is checked against.
// Restates /root/reference/shaders/main.wgsl:1-28, shootRay.wgsl:5-60, traceRay.wgsl:3-83.
#include "kernels.cuh"
#include "shade.cuh"
#include "camera.cuh"
#include "trace_binary.cuh"

namespace wg {

struct Primary { int type, index; float t; bool done; };

// traceRay.wgsl:3-83
__device__ __forceinline__ v3 ray_color(const DeviceScene& S, Ray ray, uint32_t& rng, Counts* cnt,
                                        ShadeState& st, Primary& prim) {
    v3 acc = V3s(0.0f), throughput = V3s(1.0f);
    ObjRayCache oc;
    HitLite h; h.mat_id = -1; h.uu = h.vv = 0.0f;
    for (int i = 0; i < S.K.MAX_BOUNCES; i++) {
        const bool hit = hitScene_binary(S, ray, rng, cnt, oc, h);
        if (!prim.done) {
            prim.done = true;
            prim.type = hit ? (h.kind == HIT_VOLUME ? 0 : h.kind) : -1;
            prim.index = hit ? (h.kind == HIT_TRI ? (int)S.triangles[h.index].local_id : h.index) : -1;
            prim.t = hit ? h.t : 0.0f;
        }
        if (!hit) { acc = acc + V3(0.0f, 1.0f, 1.0f) * throughput; break; }
        TriNormals tn; tn.a = tn.b = tn.c = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (h.kind == HIT_TRI) tn = load_tri_normals(S, h.index);
        finalize_hit(S, ray, h, tn, st);
        const Material m = S.materials[h.mat_id];
        unsigned draws = 0;
        bool early;
        const bool alive = bounce_update(S, m, i, ray, st, rng, draws, acc, throughput, early);
        if (cnt) cnt->rng += draws;
        if (!alive) break;
    }
    return acc;
}

// shootRay.wgsl:5-49 + main.wgsl:3-18
__device__ __forceinline__ v3 compute_pixel(const DeviceScene& S, uint32_t pixelIndex, Counts* cnt, Primary& prim) {
    const PixelSetup ps = pixel_setup(S, pixelIndex);
    uint32_t rng = pixel_seed(pixelIndex, S.uniforms.frameNum);

    ShadeState st;
    st.p = st.n = st.unit_w = st.u = st.v = V3s(0.0f);
    st.front_face = false; st.skip_pdf = false; st.doSpecular = 0.0f;
    st.skip_ray.o = st.skip_ray.d = V3s(0.0f);
    prim.done = false; prim.type = prim.index = -1; prim.t = 0.0f;

    v3 pix = V3s(0.0f);
    if (S.K.STRATIFY) {
        const int side = strat_side(S);
        float numSamples = 0.0f;
        for (int i = 0; i < side; i++) {
            for (int j = 0; j < side; j++) {
                const Ray r = sample_ray(S, ps, i, j, rng);
                if (cnt) { cnt->rng += 2; cnt->samples++; }
                pix = pix + ray_color(S, r, rng, cnt, st, prim);
                numSamples += 1.0f;
            }
        }
        pix = pix / numSamples;
    } else {
        for (int i = 0; i < S.K.NUM_SAMPLES; i++) {
            const Ray r = sample_ray(S, ps, 0, 0, rng);
            if (cnt) { cnt->rng += 2; cnt->samples++; }
            pix = pix + ray_color(S, r, rng, cnt, st, prim);
        }
        pix = pix / (float)S.K.NUM_SAMPLES;
    }
    return pix;
}

__device__ __forceinline__ void flush_counts(const DeviceScene& S, const Counts& c) {
    // warp-aggregate, one atomic per counter per warp
    unsigned v[8] = {c.rays, c.nodes, c.tris, c.spheres, c.quads, c.rng, c.samples, c.aborts};
    #pragma unroll
    for (int k = 0; k < 8; k++) {
        unsigned long long x = v[k];
        for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
        if ((threadIdx.x & 31) == 0 && x) atomicAdd(&S.counters[k], x);
    }
}

// One computeFrameBuffer dispatch.  fb == nullptr => probe only (binding 3 untouched).
template <bool COUNT>
__global__ void __launch_bounds__(128) k_exact_frame(DeviceScene S, float4* __restrict__ fb,
                                                     float* __restrict__ radiance_out, int* __restrict__ prim_type,
                                                     int* __restrict__ prim_index, float* __restrict__ prim_t) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t npix = (uint32_t)S.W * (uint32_t)S.H;
    if (S.shard.pixel_limit < npix) npix = S.shard.pixel_limit;
    uint32_t p = npix;
    if (i < S.shard.n_pixels) p = shard_pixel(S.shard, S.W, i);
    Counts c = {0, 0, 0, 0, 0, 0, 0, 0};
    if (p < npix) {
        Primary prim;
        const v3 col = compute_pixel(S, p, COUNT ? &c : nullptr, prim);
        if (fb) {
            float4 out;
            if (S.uniforms.resetBuffer == 0.0f) {
                const float4 old = fb[p];
                out = make_float4(fa(old.x, col.x), fa(old.y, col.y), fa(old.z, col.z), 1.0f);
            } else {
                out = make_float4(col.x, col.y, col.z, 1.0f);
            }
            fb[p] = out;
        }
        if (radiance_out) { radiance_out[3 * p] = col.x; radiance_out[3 * p + 1] = col.y; radiance_out[3 * p + 2] = col.z; }
        if (prim_type) prim_type[p] = prim.type;
        if (prim_index) prim_index[p] = prim.index;
        if (prim_t) prim_t[p] = prim.t;
    }
    if (COUNT) flush_counts(S, c);
}

template <bool COUNT>
__global__ void __launch_bounds__(128) k_trace_rays_binary(DeviceScene S, int n, const float* __restrict__ rays,
                                                           const uint32_t* __restrict__ seeds, int* __restrict__ out_type,
                                                           int* __restrict__ out_index, float* __restrict__ out_t) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    Counts c = {0, 0, 0, 0, 0, 0, 0, 0};
    if (i < n) {
        Ray r; r.o = V3(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]); r.d = V3(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]);
        uint32_t rng = seeds ? seeds[i] : 0u;
        ObjRayCache oc; HitLite h; h.mat_id = -1; h.uu = h.vv = 0.0f;
        const bool hit = hitScene_binary(S, r, rng, COUNT ? &c : nullptr, oc, h);
        out_type[i] = hit ? (h.kind == HIT_VOLUME ? 0 : h.kind) : -1;
        out_index[i] = hit ? (h.kind == HIT_TRI ? (int)S.triangles[h.index].local_id : h.index) : -1;
        out_t[i] = hit ? h.t : 0.0f;
    }
    if (COUNT) flush_counts(S, c);
}

__global__ void k_pinned_math(int fn, int n, const float* x, const float* y, float* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    switch (fn) {
        case 0: out[i] = wgpt_pm::sinf_(x[i]); break;
        case 1: out[i] = wgpt_pm::cosf_(x[i]); break;
        case 2: out[i] = wgpt_pm::acosf_(x[i]); break;
        case 3: out[i] = wgpt_pm::logf_(x[i]); break;
        case 4: out[i] = wgpt_pm::powf_(x[i], y[i]); break;
        // the device-only forms of the shading code, checked against the plain ones by the tests:
        case 5: out[i] = (V3(x[i], -x[i], 0.0f) / y[i]).x; break;                    // shared-reciprocal vector division
        case 6: { float sn, cs; wgpt_pm::sincosf_(x[i], &sn, &cs); out[i] = sn; break; }
        default: { float sn, cs; wgpt_pm::sincosf_(x[i], &sn, &cs); out[i] = cs; break; }
    }
}

// ---- host launchers (kernels.cuh) ----
void launch_exact_frame(const DeviceScene& S, float4* fb, float* radiance_out, int* prim_type, int* prim_index,
                        float* prim_t, cudaStream_t stream) {
    const unsigned threads = 128;
    const unsigned blocks = (S.shard.n_pixels + threads - 1) / threads;
    if (blocks == 0) return;
    if (S.counters) k_exact_frame<true><<<blocks, threads, 0, stream>>>(S, fb, radiance_out, prim_type, prim_index, prim_t);
    else k_exact_frame<false><<<blocks, threads, 0, stream>>>(S, fb, radiance_out, prim_type, prim_index, prim_t);
}

void launch_trace_rays_binary(const DeviceScene& S, int n, const float* rays, const uint32_t* seeds, int* out_type,
                              int* out_index, float* out_t, cudaStream_t stream) {
    const unsigned threads = 128, blocks = (n + threads - 1) / threads;
    if (blocks == 0) return;
    if (S.counters) k_trace_rays_binary<true><<<blocks, threads, 0, stream>>>(S, n, rays, seeds, out_type, out_index, out_t);
    else k_trace_rays_binary<false><<<blocks, threads, 0, stream>>>(S, n, rays, seeds, out_type, out_index, out_t);
}

void launch_pinned_math(int fn, int n, const float* x, const float* y, float* out, cudaStream_t stream) {
    if (n <= 0) return;
    k_pinned_math<<<(n + 255) / 256, 256, 0, stream>>>(fn, n, x, y, out);
}

}  // namespace wg

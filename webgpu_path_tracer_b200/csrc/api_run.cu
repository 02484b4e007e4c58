// api_run.cu -- the per-frame half of the C ABI: dispatch / render / present and the parity hooks.
// Mirrors /root/reference/renderer.js:163-191 and webgpu-utils.js:125-148.
#include <vector>
#include <string.h>
#include "context.h"
#include "kernels.cuh"

using namespace wg;

namespace {

// renders frame(s) in exact mode: one launch per frame on the context's stream
int exact_render(wgpt_ctx* c, int first_frame, int n_frames, uint32_t pixel_limit) {
    for (int f = 0; f < n_frames; f++) {
        const int frame = first_frame + f;
        if (c->cfg.strata_world > 1 && (frame % c->cfg.strata_world) != c->cfg.strata_rank) continue;
        c->uniforms.frameNum = (float)frame;
        const DeviceScene S = make_device_scene(c, pixel_limit);
        launch_exact_frame(S, (float4*)c->dbuf[3], nullptr, nullptr, nullptr, nullptr, c->stream);
        c->last_launches++;
    }
    return cuda_err(c, cudaGetLastError(), "exact_render");
}

struct HostScratch {
    void* p = nullptr;
    ~HostScratch() { if (p) cudaFree(p); }
};

}  // namespace

extern "C" {

int wgpt_read_buffer(wgpt_ctx* c, int binding, void* out, size_t bytes) {
    if (!c || !out) return WGPT_E_INVALID;
    cudaSetDevice(c->cfg.device);
    if (binding == 0) {
        if (bytes != 80 || !c->have_uniforms) return set_err(c, WGPT_E_INVALID, "uniforms: 80 bytes, after a write");
        memcpy(out, &c->uniforms, 80);
        return WGPT_OK;
    }
    if (binding < 1 || binding > 9 || binding == 4) return set_err(c, WGPT_E_INVALID, "unknown binding");
    if (bytes != c->dbytes[binding]) return set_err(c, WGPT_E_INVALID, "size does not match the bound buffer");
    int rc = fast_flush(c);                        // frames dispatched so far are in binding 3 before it is read
    if (rc) return rc;
    cudaError_t e = cudaMemcpyAsync(out, c->dbuf[binding], bytes, cudaMemcpyDeviceToHost, c->stream);
    if (e != cudaSuccess) return cuda_err(c, e, "read_buffer");
    return cuda_err(c, cudaStreamSynchronize(c->stream), "read_buffer");
}

int wgpt_device_ptr(wgpt_ctx* c, int binding, void** dptr, size_t* bytes) {
    if (!c || binding < 1 || binding > 9 || binding == 4) return WGPT_E_INVALID;
    if (dptr) *dptr = c->dbuf[binding];
    if (bytes) *bytes = c->dbytes[binding];
    return WGPT_OK;
}

int wgpt_dispatch(wgpt_ctx* c, uint32_t workgroups) {
    int rc = ready(c);
    if (rc) return rc;
    cudaSetDevice(c->cfg.device);
    const uint64_t inv = (uint64_t)workgroups * 64u;
    const uint32_t limit = inv > 0xffffffffull ? 0xffffffffu : (uint32_t)inv;
    cudaEventRecord(c->ev0, c->stream);
    c->last_launches = 0;
    if (c->cfg.mode != WGPT_MODE_EXACT && c->cfg.max_bounces >= 1 && limit >= (uint32_t)c->cfg.width * c->cfg.height) {
        // the frame joins the living wavefront and the call returns (strata belong to wgpt_render: a dispatch is
        // one whole frame in both modes)
        rc = fast_dispatch(c, (int)c->uniforms.frameNum, c->uniforms.resetBuffer != 0.0f);
    } else {
        if ((rc = fast_close(c)) != WGPT_OK) return rc;
        // (a partial dispatch -- fewer workgroups than pixels -- always takes the one-thread-per-pixel path)
        const DeviceScene S = make_device_scene(c, limit);
        launch_exact_frame(S, (float4*)c->dbuf[3], nullptr, nullptr, nullptr, nullptr, c->stream);
        c->last_launches = 1;
        rc = cuda_err(c, cudaGetLastError(), "dispatch");
    }
    cudaEventRecord(c->ev1, c->stream);
    c->timing_valid = true;
    return rc;
}

int wgpt_render(wgpt_ctx* c, int first_frame, int n_frames) {
    int rc = ready(c);
    if (rc) return rc;
    if (n_frames < 0 || first_frame < 0) return set_err(c, WGPT_E_INVALID, "first_frame, n_frames must be >= 0");
    cudaSetDevice(c->cfg.device);
    c->uniforms.resetBuffer = 0.0f;
    cudaEventRecord(c->ev0, c->stream);
    c->last_launches = 0;
    if (c->cfg.mode != WGPT_MODE_EXACT && c->cfg.max_bounces >= 1) rc = fast_render(c, first_frame, n_frames, 0);
    else { rc = fast_close(c); if (!rc) rc = exact_render(c, first_frame, n_frames, 0xffffffffu); }
    if (n_frames > 0) c->uniforms.frameNum = (float)(first_frame + n_frames - 1);
    cudaEventRecord(c->ev1, c->stream);
    c->timing_valid = true;
    return rc;
}

int wgpt_present(wgpt_ctx* c, float* rgba_out, uint8_t* rgba8_out) {
    if (!c) return WGPT_E_INVALID;
    if (!c->have_uniforms) return set_err(c, WGPT_E_STATE, "uniforms (binding 0) not written");
    cudaSetDevice(c->cfg.device);
    {   // `fs` runs after the compute passes submitted before it (webgpu-utils.js:133,147: one in-order queue)
        const int rc = fast_flush(c);
        if (rc) return rc;
    }
    const uint32_t npix = (uint32_t)c->cfg.width * c->cfg.height;
    // device staging lives in the context (the image size is fixed at wgpt_create): the per-frame loop of
    // renderer.js:172-184 calls this every frame, and a cudaMalloc / cudaFree pair per call cost ~3 ms of its 9
    cudaError_t e;
    if (rgba_out && !c->d_present[0] && (e = cudaMalloc(&c->d_present[0], (size_t)npix * 16)) != cudaSuccess) return cuda_err(c, e, "present");
    if (rgba8_out && !c->d_present[1] && (e = cudaMalloc(&c->d_present[1], (size_t)npix * 4)) != cudaSuccess) return cuda_err(c, e, "present");
    float4* d_rgba = rgba_out ? (float4*)c->d_present[0] : nullptr;
    uchar4* d_rgba8 = rgba8_out ? (uchar4*)c->d_present[1] : nullptr;
    const int clear = c->uniforms.resetBuffer == 1.0f;
    launch_present((const float4*)c->dbuf[3], (float4*)c->dbuf[3], d_rgba, d_rgba8, npix, c->uniforms.frameNum, clear, c->stream);
    if (rgba_out) cudaMemcpyAsync(rgba_out, d_rgba, (size_t)npix * 16, cudaMemcpyDeviceToHost, c->stream);
    if (rgba8_out) cudaMemcpyAsync(rgba8_out, d_rgba8, (size_t)npix * 4, cudaMemcpyDeviceToHost, c->stream);
    return cuda_err(c, cudaStreamSynchronize(c->stream), "present");
}

int wgpt_probe(wgpt_ctx* c, float* radiance_out, int32_t* prim_type, int32_t* prim_index, float* prim_t) {
    int rc = ready(c);
    if (rc) return rc;
    cudaSetDevice(c->cfg.device);
    if ((rc = fast_close(c)) != WGPT_OK) return rc;
    const size_t npix = (size_t)c->cfg.width * c->cfg.height;
    cudaMemsetAsync(c->d_radiance, 0, npix * 12, c->stream);
    cudaMemsetAsync(c->d_ptype, 0xff, npix * 4, c->stream);
    cudaMemsetAsync(c->d_pindex, 0xff, npix * 4, c->stream);
    cudaMemsetAsync(c->d_pt, 0, npix * 4, c->stream);
    if (c->cfg.mode != WGPT_MODE_EXACT && c->cfg.max_bounces >= 1) {
        // through the wavefront: render the frame into a scratch copy of binding 3 (reset semantics:
        // the scratch then holds exactly this frame's per-pixel radiance), binding 3 itself untouched
        HostScratch d_fb;
        cudaError_t e = cudaMalloc(&d_fb.p, npix * 16);
        if (e != cudaSuccess) return cuda_err(c, e, "probe");
        cudaMemsetAsync(d_fb.p, 0, npix * 16, c->stream);
        void* real_fb = c->dbuf[3];
        c->dbuf[3] = d_fb.p;
        rc = fast_render_probe(c, (int)c->uniforms.frameNum, 1, 1, c->d_ptype, c->d_pindex, c->d_pt);
        c->dbuf[3] = real_fb;
        fast_close(c);                             // the scratch buffer goes away: nothing may append to this session
        if (rc) return rc;
        launch_fb_to_rgb((const float4*)d_fb.p, c->d_radiance, (uint32_t)npix, c->stream);
        if (radiance_out) cudaMemcpyAsync(radiance_out, c->d_radiance, npix * 12, cudaMemcpyDeviceToHost, c->stream);
        if (prim_type) cudaMemcpyAsync(prim_type, c->d_ptype, npix * 4, cudaMemcpyDeviceToHost, c->stream);
        if (prim_index) cudaMemcpyAsync(prim_index, c->d_pindex, npix * 4, cudaMemcpyDeviceToHost, c->stream);
        if (prim_t) cudaMemcpyAsync(prim_t, c->d_pt, npix * 4, cudaMemcpyDeviceToHost, c->stream);
        return cuda_err(c, cudaStreamSynchronize(c->stream), "probe");
    }
    const DeviceScene S = make_device_scene(c, 0xffffffffu);
    launch_exact_frame(S, nullptr, c->d_radiance, c->d_ptype, c->d_pindex, c->d_pt, c->stream);
    if (radiance_out) cudaMemcpyAsync(radiance_out, c->d_radiance, npix * 12, cudaMemcpyDeviceToHost, c->stream);
    if (prim_type) cudaMemcpyAsync(prim_type, c->d_ptype, npix * 4, cudaMemcpyDeviceToHost, c->stream);
    if (prim_index) cudaMemcpyAsync(prim_index, c->d_pindex, npix * 4, cudaMemcpyDeviceToHost, c->stream);
    if (prim_t) cudaMemcpyAsync(prim_t, c->d_pt, npix * 4, cudaMemcpyDeviceToHost, c->stream);
    return cuda_err(c, cudaStreamSynchronize(c->stream), "probe");
}

// Path signatures of the last wgpt_probe in a fast mode: the RNG state of every pixel's item when it finished.  The first
// call (out == NULL allowed) switches the recording on for the following probes.
int wgpt_probe_rng(wgpt_ctx* c, uint32_t* out) {
    if (!c) return WGPT_E_INVALID;
    if (c->cfg.mode == WGPT_MODE_EXACT) return set_err(c, WGPT_E_STATE, "wgpt_probe_rng: fast modes only");
    cudaSetDevice(c->cfg.device);
    const size_t npix = (size_t)c->cfg.width * c->cfg.height;
    if (!c->d_prng) {
        cudaError_t e = cudaMalloc((void**)&c->d_prng, npix * 4);
        if (e != cudaSuccess) return cuda_err(c, e, "probe_rng");
        cudaMemsetAsync(c->d_prng, 0, npix * 4, c->stream);
    }
    if (!out) return WGPT_OK;
    cudaMemcpyAsync(out, c->d_prng, npix * 4, cudaMemcpyDeviceToHost, c->stream);
    return cuda_err(c, cudaStreamSynchronize(c->stream), "probe_rng");
}

int wgpt_trace_rays(wgpt_ctx* c, int n, const float* rays, const uint32_t* seeds, int32_t* prim_type,
                    int32_t* prim_index, float* prim_t) {
    if (!c || n < 0 || (n && (!rays || !prim_type || !prim_index || !prim_t))) return WGPT_E_INVALID;
    int rc = validate_scene(c);
    if (rc) return rc;
    if (n == 0) return WGPT_OK;
    cudaSetDevice(c->cfg.device);
    HostScratch d_rays, d_seeds, d_type, d_index, d_t;
    cudaError_t e;
    if ((e = cudaMalloc(&d_rays.p, (size_t)n * 24)) != cudaSuccess || (e = cudaMalloc(&d_seeds.p, (size_t)n * 4)) != cudaSuccess ||
        (e = cudaMalloc(&d_type.p, (size_t)n * 4)) != cudaSuccess || (e = cudaMalloc(&d_index.p, (size_t)n * 4)) != cudaSuccess ||
        (e = cudaMalloc(&d_t.p, (size_t)n * 4)) != cudaSuccess) return cuda_err(c, e, "trace_rays");
    cudaMemcpyAsync(d_rays.p, rays, (size_t)n * 24, cudaMemcpyHostToDevice, c->stream);
    if (seeds) cudaMemcpyAsync(d_seeds.p, seeds, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream);
    else cudaMemsetAsync(d_seeds.p, 0, (size_t)n * 4, c->stream);
    if (c->cfg.mode != WGPT_MODE_EXACT) {
        rc = fast_trace_rays(c, n, (const float*)d_rays.p, (const uint32_t*)d_seeds.p, (int*)d_type.p, (int*)d_index.p, (float*)d_t.p);
        if (rc) return rc;
    } else {
        if (!c->have_uniforms) memset(&c->uniforms, 0, sizeof(c->uniforms));
        const DeviceScene S = make_device_scene(c, 0xffffffffu);
        launch_trace_rays_binary(S, n, (const float*)d_rays.p, (const uint32_t*)d_seeds.p, (int*)d_type.p, (int*)d_index.p, (float*)d_t.p, c->stream);
    }
    cudaMemcpyAsync(prim_type, d_type.p, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream);
    cudaMemcpyAsync(prim_index, d_index.p, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream);
    cudaMemcpyAsync(prim_t, d_t.p, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream);
    return cuda_err(c, cudaStreamSynchronize(c->stream), "trace_rays");
}

int wgpt_set_counters(wgpt_ctx* c, int enabled) {
    if (!c) return WGPT_E_INVALID;
    c->counters_on = enabled != 0;
    return WGPT_OK;
}

int wgpt_counters(wgpt_ctx* c, wgpt_counter_block* out, int reset) {
    if (!c || !out) return WGPT_E_INVALID;
    cudaSetDevice(c->cfg.device);
    {
        const int rc = fast_flush(c);
        if (rc) return rc;
    }
    unsigned long long h[8];
    cudaError_t e = cudaMemcpyAsync(h, c->d_counters, 64, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess && reset) e = cudaMemsetAsync(c->d_counters, 0, 64, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) return cuda_err(c, e, "counters");
    out->rays = h[0]; out->node_visits = h[1]; out->tri_tests = h[2]; out->sphere_tests = h[3];
    out->quad_tests = h[4]; out->rng_draws = h[5]; out->samples = h[6]; out->stack_aborts = h[7];
    return WGPT_OK;
}

int wgpt_fast_statistics(wgpt_ctx* c, wgpt_fast_stats* out) {
    if (!c || !out) return WGPT_E_INVALID;
    cudaSetDevice(c->cfg.device);
    uint32_t v[8];
    fast_stats(c, v);
    out->wide_nodes = v[0]; out->wide_tris = v[1]; out->wide_levels = v[2]; out->pool_slots = v[3]; out->ring_frames = v[4];
    out->fallback_rays = v[5]; out->iterations = v[6]; out->trace_fallback_rays = v[7];
    return WGPT_OK;
}

int wgpt_set_stage_profiling(wgpt_ctx* c, int enabled) {
    if (!c) return WGPT_E_INVALID;
    c->stage_profiling = enabled != 0;
    return WGPT_OK;
}

int wgpt_stage_timings(wgpt_ctx* c, float ms[4], uint64_t launches[4]) {
    if (!c || !ms || !launches) return WGPT_E_INVALID;
    cudaSetDevice(c->cfg.device);
    return fast_stage_timings(c, ms, launches);
}

int wgpt_extend_history(wgpt_ctx* c, uint32_t out[64]) {
    if (!c || !out) return WGPT_E_INVALID;
    cudaSetDevice(c->cfg.device);
    return fast_extend_history(c, out);
}

int wgpt_synchronize(wgpt_ctx* c) {
    if (!c) return WGPT_E_INVALID;
    cudaSetDevice(c->cfg.device);
    const int rc = fast_flush(c);
    if (rc) return rc;
    return cuda_err(c, cudaStreamSynchronize(c->stream), "synchronize");
}

int wgpt_last_timing(wgpt_ctx* c, float* ms, uint64_t* launches) {
    if (!c) return WGPT_E_INVALID;
    if (!c->timing_valid) return set_err(c, WGPT_E_STATE, "no dispatch/render timed yet");
    cudaSetDevice(c->cfg.device);
    cudaError_t e = cudaEventSynchronize(c->ev1);
    if (e != cudaSuccess) return cuda_err(c, e, "last_timing");
    float t = 0.0f;
    if ((e = cudaEventElapsedTime(&t, c->ev0, c->ev1)) != cudaSuccess) return cuda_err(c, e, "last_timing");
    if (ms) *ms = t;
    if (launches) *launches = c->last_launches;
    return WGPT_OK;
}

// Device-side stopwatch on the context's own stream (bench.py: a step is render + reduce, both enqueued there, and
// torch.cuda.Event only sees torch's stream): mark(0) ... mark(1), then elapsed() waits for mark 1.
int wgpt_timer_mark(wgpt_ctx* c, int which) {
    if (!c || which < 0 || which > 1) return WGPT_E_INVALID;
    cudaSetDevice(c->cfg.device);
    cudaError_t e = cudaSuccess;
    if (!c->tm_ev[0] && ((e = cudaEventCreate(&c->tm_ev[0])) != cudaSuccess || (e = cudaEventCreate(&c->tm_ev[1])) != cudaSuccess))
        return cuda_err(c, e, "timer_mark");
    return cuda_err(c, cudaEventRecord(c->tm_ev[which], c->stream), "timer_mark");
}

int wgpt_timer_elapsed(wgpt_ctx* c, float* ms) {
    if (!c || !ms || !c->tm_ev[0]) return WGPT_E_INVALID;
    cudaSetDevice(c->cfg.device);
    cudaError_t e = cudaEventSynchronize(c->tm_ev[1]);
    if (e == cudaSuccess) e = cudaEventElapsedTime(ms, c->tm_ev[0], c->tm_ev[1]);
    return cuda_err(c, e, "timer_elapsed");
}

int wgpt_pinned_math(wgpt_ctx* c, int fn, int n, const float* x, const float* y, float* out) {
    if (!c || n < 0 || (n && (!x || !out)) || fn < 0 || fn > 7) return WGPT_E_INVALID;
    if (n == 0) return WGPT_OK;
    cudaSetDevice(c->cfg.device);
    HostScratch dx, dy, dout;
    cudaError_t e;
    if ((e = cudaMalloc(&dx.p, (size_t)n * 4)) != cudaSuccess || (e = cudaMalloc(&dy.p, (size_t)n * 4)) != cudaSuccess ||
        (e = cudaMalloc(&dout.p, (size_t)n * 4)) != cudaSuccess) return cuda_err(c, e, "pinned_math");
    cudaMemcpyAsync(dx.p, x, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream);
    if (y) cudaMemcpyAsync(dy.p, y, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream);
    else cudaMemsetAsync(dy.p, 0, (size_t)n * 4, c->stream);
    launch_pinned_math(fn, n, (const float*)dx.p, (const float*)dy.p, (float*)dout.p, c->stream);
    cudaMemcpyAsync(out, dout.p, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream);
    return cuda_err(c, cudaStreamSynchronize(c->stream), "pinned_math");
}

int wgpt_k4_timing(wgpt_ctx* c, int kernel, int reps, float* ms_per_launch, uint64_t* bytes_per_launch) {
    if (!c || kernel < 0 || kernel > 2 || reps < 1 || !ms_per_launch) return WGPT_E_INVALID;
    cudaSetDevice(c->cfg.device);
    const uint32_t npix = (uint32_t)c->cfg.width * c->cfg.height;
    HostScratch d_fb, d_frame, d_out;
    cudaError_t e;
    if ((e = cudaMalloc(&d_fb.p, (size_t)npix * 16)) != cudaSuccess || (e = cudaMalloc(&d_frame.p, (size_t)npix * 16)) != cudaSuccess ||
        (e = cudaMalloc(&d_out.p, (size_t)npix * 16)) != cudaSuccess) return cuda_err(c, e, "k4_timing");
    cudaMemsetAsync(d_fb.p, 0, (size_t)npix * 16, c->stream);
    {   // a frame of varied radiance sums (LCG, 0 .. 16 x frameNum), w = 1: every pixel is "rendered"
        std::vector<float> h((size_t)npix * 4);
        uint32_t x = 12345u;
        for (size_t i = 0; i < h.size(); i++) { x = x * 1664525u + 1013904223u; h[i] = (i & 3) == 3 ? 1.0f : (float)(x >> 8) * (112.0f / 16777216.0f); }
        cudaMemcpyAsync(d_frame.p, h.data(), h.size() * 4, cudaMemcpyHostToDevice, c->stream);
        cudaStreamSynchronize(c->stream);
    }
    cudaEvent_t a = nullptr, b = nullptr;
    if ((e = cudaEventCreate(&a)) != cudaSuccess || (e = cudaEventCreate(&b)) != cudaSuccess) {
        if (a) cudaEventDestroy(a);
        return cuda_err(c, e, "k4_timing");
    }
    for (int pass = 0; pass < 2; pass++) {                                   // pass 0 = warm-up
        if (pass) cudaEventRecord(a, c->stream);
        for (int r = 0; r < (pass ? reps : 3); r++) {
            if (kernel == 0) launch_accumulate((float4*)d_fb.p, (const float4*)d_frame.p, npix, 0, c->stream);
            else launch_present((const float4*)d_frame.p, nullptr, kernel == 2 ? (float4*)d_out.p : nullptr,
                                kernel == 1 ? (uchar4*)d_out.p : nullptr, npix, 7.0f, 0, c->stream);
        }
        if (pass) cudaEventRecord(b, c->stream);
    }
    e = cudaStreamSynchronize(c->stream);
    float t = 0.0f;
    if (e == cudaSuccess) e = cudaEventElapsedTime(&t, a, b);
    cudaEventDestroy(a); cudaEventDestroy(b);
    if (e != cudaSuccess) return cuda_err(c, e, "k4_timing");
    *ms_per_launch = t / (float)reps;
    if (bytes_per_launch) *bytes_per_launch = (uint64_t)npix * (kernel == 0 ? 48u : kernel == 1 ? 20u : 32u);
    return WGPT_OK;
}

int wgpt_present_quant_check(wgpt_ctx* c, uint32_t first_bits, uint32_t count, uint64_t* mismatches, uint64_t* exact_path) {
    if (!c || !mismatches) return WGPT_E_INVALID;
    cudaSetDevice(c->cfg.device);
    HostScratch d;
    cudaError_t e = cudaMalloc(&d.p, 16);
    if (e != cudaSuccess) return cuda_err(c, e, "present_quant_check");
    cudaMemsetAsync(d.p, 0, 16, c->stream);
    launch_quant_check(first_bits, count, (unsigned long long*)d.p, c->stream);
    unsigned long long h[2] = {0, 0};
    cudaMemcpyAsync(h, d.p, 16, cudaMemcpyDeviceToHost, c->stream);
    if ((e = cudaStreamSynchronize(c->stream)) != cudaSuccess) return cuda_err(c, e, "present_quant_check");
    *mismatches = h[0];
    if (exact_path) *exact_path = h[1];
    return WGPT_OK;
}

}  // extern "C"

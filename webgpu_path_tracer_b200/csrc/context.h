// context.h -- the state behind a wgpt_ctx: device copies of the bind-group buffers, the
// stream, scratch for the parity hooks and (fast mode) the derived wide BVH + wavefront pool.
#pragma once
#include <string>
#include <vector>
#include <cuda_runtime.h>
#include "../../include/wgpt.h"
#include "scene.cuh"

namespace wg { struct FastState; }

struct wgpt_ctx {
    wgpt_config cfg;
    cudaStream_t stream = nullptr;
    void* dbuf[10] = {nullptr};            // device copy of each binding (0 unused: uniforms travel by value)
    size_t dbytes[10] = {0};
    std::vector<unsigned char> hbuf[10];   // host mirrors of the small tables (lights lookup, validation)
    wg::Uniforms uniforms;
    bool have_uniforms = false;
    wg::Quad lights;
    bool lights_valid = false;
    unsigned long long* d_counters = nullptr;
    bool counters_on = false;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t tm_ev[2] = {nullptr, nullptr};       // wgpt_timer_mark
    bool timing_valid = false;
    uint64_t last_launches = 0;
    uint32_t shard_pixels = 0;
    // probe scratch
    float* d_radiance = nullptr; int* d_ptype = nullptr; int* d_pindex = nullptr; float* d_pt = nullptr;
    void* d_present[2] = {nullptr, nullptr};    // wgpt_present's device staging (f32 image, RGBA8 image): kept between calls
    uint32_t* d_prng = nullptr;            // final RNG state per pixel of the last fast-mode probe (allocated by wgpt_probe_rng_enable)
    wg::FastState* fast = nullptr;         // wide BVH + wavefront pool (fast.cu)
    bool wide_dirty = true;
    bool stage_profiling = false;
    // summary of the big device-only tables, gathered at upload time for validation
    int tri_mesh_max = -1; bool tri_bad = false;
    long long bvh_prim_end_max = 0; bool bvh_bad = false;
    int bvh_depth = 0;                     // deepest leaf of the binary BVH (root = 0)
    // multi-GPU (mgpu.cu): this context's NCCL communicator, if wgpt_comm_init / wgpt_group_create made one
    void* nccl_comm = nullptr; int nccl_rank = 0, nccl_world = 1;
    std::string err;
};

namespace wg {
// assembles the by-value descriptor for a launch; pixel_limit = invocations dispatched
DeviceScene make_device_scene(wgpt_ctx* c, uint32_t pixel_limit);
int set_err(wgpt_ctx* c, int code, const std::string& msg);
int cuda_err(wgpt_ctx* c, cudaError_t e, const char* what);
int ready(wgpt_ctx* c);
int validate_scene(wgpt_ctx* c);

// fast.cu
int fast_render(wgpt_ctx* c, int first_frame, int n_frames, int reset_first);
int fast_dispatch(wgpt_ctx* c, int frame, int reset_first);      // non-blocking: the frame joins the open wavefront session
int fast_flush(wgpt_ctx* c);                                      // run the open session until its frames are in binding 3
int fast_close(wgpt_ctx* c);                                      // flush, and nothing may append to the session afterwards
int fast_render_probe(wgpt_ctx* c, int first_frame, int n_frames, int reset_first, int* d_ptype, int* d_pindex, float* d_pt);
int fast_stats(wgpt_ctx* c, uint32_t out[8]);
int fast_stage_timings(wgpt_ctx* c, float ms[4], uint64_t launches[4]);
int fast_extend_history(wgpt_ctx* c, uint32_t out[64]);
int fast_trace_rays(wgpt_ctx* c, int n, const float* d_rays, const uint32_t* d_seeds, int* d_type, int* d_index, float* d_t);
void fast_destroy(wgpt_ctx* c);
// mgpu.cu
void mgpu_destroy(wgpt_ctx* c);
}

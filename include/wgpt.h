/*
 * wgpt.h -- C ABI of libwgpt.so: the B200 (sm_100a CUDA) replacement for the reference's
 * compute-shader path-tracing hot path.
 *
 * The boundary mirrors what /root/reference/renderer.js does to WebGPU:
 *   - bind group 0 of `computeFrameBuffer`, bindings 0,1,2,3,5,6,7,8,9
 *     (renderer.js:109-123, shaders/header.wgsl:15-23)            -> wgpt_write_buffer
 *   - queue.writeBuffer(uniforms) + dispatchWorkgroups per frame
 *     (renderer.js:182-188, webgpu-utils.js:125-134)              -> wgpt_dispatch / wgpt_render
 *   - the present pass `fs` reading binding 3
 *     (renderer.js:130-136,191, shaders/fragment.js:22-36)        -> wgpt_present
 * Plain pointers and sizes only.  Every call returns 0 on success or a negative WGPT_E_* code;
 * wgpt_last_error() gives the message.  One context = one CUDA device = one caller thread at a
 * time.  Calls enqueue on the context's stream and return; read_buffer/present/counters/
 * primary_hits/sample_radiance/trace_rays synchronise.  There is NO CPU fallback: without a
 * CUDA device wgpt_create fails with WGPT_E_CUDA.
 */
#ifndef WGPT_H
#define WGPT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WGPT_VERSION 1

/* error codes */
#define WGPT_OK            0
#define WGPT_E_INVALID    -1   /* bad argument / binding / size not a multiple of the stride */
#define WGPT_E_CUDA       -2   /* CUDA runtime error (or no device) */
#define WGPT_E_STATE      -3   /* call order: scene or uniforms missing */
#define WGPT_E_NOMEM      -4

/* binding numbers of shaders/header.wgsl:15-23 and their element strides in bytes */
#define WGPT_BINDING_UNIFORMS    0   /* 80 B  Uniforms       header.wgsl:41-46  */
#define WGPT_BINDING_SPHERES     1   /* 32 B  Sphere         header.wgsl:68-74  */
#define WGPT_BINDING_QUADS       2   /* 80 B  Quad           header.wgsl:76-86  */
#define WGPT_BINDING_FRAMEBUFFER 3   /* 16 B  vec4f / pixel  header.wgsl:18     */
#define WGPT_BINDING_TRIANGLES   5   /* 96 B  Triangle       header.wgsl:88-98  */
#define WGPT_BINDING_MESHES      6   /* 16 B  Mesh (4 x i32) header.wgsl:100-105 */
#define WGPT_BINDING_TRANSFORMS  7   /* 128 B modelTransform header.wgsl:63-66  */
#define WGPT_BINDING_MATERIALS   8   /* 64 B  Material       header.wgsl:53-61  */
#define WGPT_BINDING_BVH         9   /* 48 B  AABB node      header.wgsl:107-117 */

/* integrator implementations */
#define WGPT_MODE_EXACT 0   /* one thread per pixel, the reference's binary BVH and operation order;
                               bit-identical to the oracle (pinned IEEE arithmetic, no FMA)          */
#define WGPT_MODE_FAST  1   /* persistent-thread wavefront over SoA ray queues on the 8-wide quantised
                               BVH; same RNG stream and intersection arithmetic, fast shading math   */

#define WGPT_MODE_FAST_TOL 2 /* OPT-IN tolerance mode: WGPT_MODE_FAST's pipeline with the SHADING of a hit (normal, BSDF
                               sample, pdfs, throughput) on fast arithmetic -- contracted FMAs, MUFU reciprocals + one
                               Newton step, f32 libm -- while the camera ray, every intersection and the RNG stream stay
                               pinned.  Primary hits (type, index, t) remain bit-identical to the reference arithmetic;
                               per-sample radiance agrees to <= 1e-4 relative except on the paths that diverge (a
                               discrete decision flips on a last-bit difference: measured 1e-4 .. 1e-3 of the samples,
                               tests/test_gpu_tol.py, bench.py key `tol_mode`).  NOT bit-comparable with the oracle.  */

/* The WGSL compile-time constants (shaders/header.wgsl:9-13) are run-time parameters here. */
typedef struct wgpt_config {
    uint32_t struct_size;          /* sizeof(wgpt_config), for forward compatibility */
    int32_t  width, height;        /* canvas size; uniforms.screenDims must match */
    int32_t  device;               /* CUDA device ordinal */
    int32_t  mode;                 /* WGPT_MODE_* */
    int32_t  num_samples;          /* NUM_SAMPLES          (reference HEAD: 1); fast mode: <= 16383 */
    int32_t  max_bounces;          /* MAX_BOUNCES          (100)                 */
    int32_t  stratify;             /* STRATIFY             (false)               */
    int32_t  importance_sampling;  /* IMPORTANCE_SAMPLING  (false)               */
    int32_t  stack_size;           /* STACK_SIZE           (20), <= 64           */
    /* multi-GPU sharding; world = 1 renders everything.
     * Tile split: rows are grouped in blocks of `tile_rows`; this context renders block b iff
     * b % tile_world == tile_rank.  Pixels outside the shard are never written (they stay 0), so
     * a sum-reduce of the shards' framebuffers reproduces the single-GPU buffer bit for bit.
     * Frame-strata split: wgpt_render renders frame f iff f % strata_world == strata_rank.   */
    int32_t  tile_rank, tile_world, tile_rows;
    int32_t  strata_rank, strata_world;
} wgpt_config;

typedef struct wgpt_ctx wgpt_ctx;

/* counters reported by wgpt_counters (same order as the oracle's) */
typedef struct wgpt_counter_block {
    uint64_t rays;          /* closest-hit queries (hitScene calls, hitRay.wgsl:1)      */
    uint64_t node_visits;   /* hit_aabb calls on the reference's binary BVH (exact mode)
                               or wide-node visits (fast mode)                          */
    uint64_t tri_tests;     /* hit_triangle calls                                       */
    uint64_t sphere_tests, quad_tests, rng_draws, samples, stack_aborts;
} wgpt_counter_block;

void wgpt_default_config(wgpt_config* cfg);
int  wgpt_create(const wgpt_config* cfg, wgpt_ctx** out);
void wgpt_destroy(wgpt_ctx* ctx);
const char* wgpt_last_error(const wgpt_ctx* ctx);     /* ctx may be NULL: last create error */

/* createBuffer + queue.writeBuffer (webgpu-utils.js:15-55): copies `bytes` from host memory at
 * call time; element count = bytes / stride (arrayLength(), main.wgsl:10-14).  Writing binding
 * 5 or 9 invalidates the derived wide BVH (rebuilt on the next dispatch in fast mode). */
int wgpt_write_buffer(wgpt_ctx* ctx, int binding, const void* data, size_t bytes);
/* device -> host copy of a binding (the reference never reads back; needed headless). */
int wgpt_read_buffer(wgpt_ctx* ctx, int binding, void* out, size_t bytes);
/* raw device pointer + size of a binding (zero-copy views for the host layer; call wgpt_synchronize first: frames
 * submitted with wgpt_dispatch may still be in flight). */
int wgpt_device_ptr(wgpt_ctx* ctx, int binding, void** dptr, size_t* bytes);

/* One `computePass(...)`: one frame with the uniforms currently in binding 0
 * (main.wgsl:1-28).  `workgroups` is what the caller would pass to dispatchWorkgroups
 * (renderer.js:187-188: W*H/64 + 1); invocations with index >= W*H are masked.
 * Enqueues and returns, like webgpu-utils.js:125-134 (submit): in fast mode the frame JOINS the living wavefront
 * (consecutive frame numbers with the same view and resetBuffer = 0 share one session; anything else -- another
 * view, a gap, a reset, a buffer write -- first finishes the frames already submitted, which keep the uniforms they
 * were dispatched with).  present / read_buffer / counters / synchronize wait for the submitted frames.  A dispatch
 * is one whole frame whatever the strata shard (strata belong to wgpt_render). */
int wgpt_dispatch(wgpt_ctx* ctx, uint32_t workgroups);
/* Batched form of the renderAnimation loop (renderer.js:163-188) without camera motion:
 * frames first_frame .. first_frame+n_frames-1, uniforms.frameNum advanced internally,
 * resetBuffer = 0; leaves uniforms.frameNum at the last frame rendered. */
int wgpt_render(wgpt_ctx* ctx, int first_frame, int n_frames);
/* `fs` (fragment.js:22-36): rgba_out[4*i..] = (pow(aces(fb[i].xyz / frameNum), 1/2.2), 1);
 * clears binding 3 when uniforms.resetBuffer == 1.  rgba_out: host, 4 floats per pixel;
 * rgba8_out (optional, may be NULL): host, 4 bytes per pixel, round(c*255). */
int wgpt_present(wgpt_ctx* ctx, float* rgba_out, uint8_t* rgba8_out);
/* Bytes-only present that ALLOCATES NOTHING.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same kernel and the same bytes as wgpt_present(ctx, NULL, rgba8_out) -- wgpt_present cudaMallocs its
 * device staging on a context's first call -- but the RGBA8 image is staged in a static device array of
 * WGPT_STATIC_PRESENT_MAX_PIXELS x 4 bytes (8 MiB) that every GPU that loads the library holds until exit.  Images with
 * more pixels are rejected with WGPT_E_INVALID.  Every call on a device shares that array (it keeps the last call's
 * bytes; each call overwrites what it reads back), so calls that OVERLAP on the GPU -- two contexts of one device, two
 * host threads -- race: serialise them.  Calls no cudaMalloc / cudaFree / allocating library routine. */
#define WGPT_STATIC_PRESENT_MAX_PIXELS (2u << 20)
int wgpt_present_rgba8_static(wgpt_ctx* ctx, uint8_t* rgba8_out);

/* ---- parity hooks (no reference counterpart) ---- */
/* Re-runs one dispatch with the current uniforms WITHOUT touching binding 3 and returns, per
 * pixel, the first closest-hit query of the invocation: type 0 sphere / 1 quad / 2 triangle /
 * -1 miss, index (sphere/quad index, Triangle.local_id), t; and pathTrace()'s radiance (3 f32). */
int wgpt_probe(wgpt_ctx* ctx, float* radiance_out, int32_t* prim_type, int32_t* prim_index, float* prim_t);
/* Fast modes: path signatures of the last wgpt_probe -- per pixel, the RNG state of the (pixel, frame) item when it finished xor a hash
 * of its hit sequence (kind and index of every closest hit): equal <=> the path drew the same numbers and hit the same things.  Call once before the probes of interest (out may be NULL)
 * to switch the recording on.  tests/test_gpu_tol.py uses it to tell diverged paths from non-diverged ones. */
int wgpt_probe_rng(wgpt_ctx* ctx, uint32_t* out);
/* Closest hit of caller-supplied rays (6 f32 each: origin, dir); seeds[i] = RNG state on entry. */
int wgpt_trace_rays(wgpt_ctx* ctx, int n, const float* rays, const uint32_t* seeds,
                    int32_t* prim_type, int32_t* prim_index, float* prim_t);
int wgpt_counters(wgpt_ctx* ctx, wgpt_counter_block* out, int reset);
/* enable/disable counter collection (adds atomics to the kernels; off by default) */
int wgpt_set_counters(wgpt_ctx* ctx, int enabled);

/* ---- device-resident timing of the last wgpt_render/wgpt_dispatch (CUDA events on the
 * context's stream): total milliseconds and number of kernel launches issued. ---- */
int wgpt_last_timing(wgpt_ctx* ctx, float* ms, uint64_t* launches);
int wgpt_synchronize(wgpt_ctx* ctx);
/* stopwatch on the context's stream: wgpt_timer_mark(ctx, 0) ... wgpt_timer_mark(ctx, 1); wgpt_timer_elapsed waits for
 * mark 1 and returns the device time between the marks (renders AND reduces enqueued in between). */
int wgpt_timer_mark(wgpt_ctx* ctx, int which);
int wgpt_timer_elapsed(wgpt_ctx* ctx, float* ms);

/* fast-mode introspection: the derived wide BVH, the wavefront pool, and how many rays had to be
 * re-traced with the reference-order binary walk (equal-distance triangles / stack overflow). */
typedef struct wgpt_fast_stats {
    uint32_t wide_nodes, wide_tris, wide_levels;   /* K2 output */
    uint32_t pool_slots, ring_frames;              /* K3 pool capacity, frames in flight */
    uint32_t fallback_rays, iterations;            /* of the last wgpt_render / wgpt_dispatch / wgpt_probe */
    uint32_t trace_fallback_rays;                  /* of wgpt_trace_rays since creation */
} wgpt_fast_stats;
int wgpt_fast_statistics(wgpt_ctx* ctx, wgpt_fast_stats* out);

/* Per-stage device time of the fast path, for the roofline line of bench.py.  When enabled, every
 * extend / shade launch is bracketed by CUDA events on the context's stream (about 1 us each);
 * wgpt_stage_timings synchronises and returns the sums since it was last called:
 * ms[0] extend (K3 traversal), ms[1] shade, ms[2] control + accumulate (K4), ms[3] hierarchy build (K5 BVH build + K2 wide flatten);
 * launches[i] = kernel launches counted in ms[i]. */
/* measurement: rays the extend stage traced in iteration i (i < 64) of the last render -- what an ncu capture of launch i
 * processed (bench.py's same-launch traffic / instructions-per-ray figures, tools/evidence_collect.py) */
int wgpt_extend_history(wgpt_ctx* ctx, uint32_t out[64]);
int wgpt_set_stage_profiling(wgpt_ctx* ctx, int enabled);
int wgpt_stage_timings(wgpt_ctx* ctx, float ms[4], uint64_t launches[4]);

/* pinned-math probe for tests: fn 0 sin, 1 cos, 2 acos, 3 log, 4 pow(x,y), evaluated ON THE GPU;
 * 5 x/y through the shading code's shared-reciprocal vector division, 6 / 7 sin / cos through its
 * one-reduction sincos (device-only forms that must equal the plain ones bit for bit) */
int wgpt_pinned_math(wgpt_ctx* ctx, int fn, int n, const float* x, const float* y, float* out);

/* K4 stream timing for the roofline of the accumulate / present passes ("a pure
 * stream"): runs `reps` launches of kernel 0 = accumulate (16 B read staging + 16 B read + 16 B
 * write of binding 3 per pixel), 1 = present to RGBA8 (16 B read + 4 B write), 2 = present to f32
 * RGBA (16 B + 16 B) on device scratch of the context's image size, WITHOUT touching binding 3, and
 * returns the mean device time per launch (CUDA events) and the bytes one launch moves. */
int wgpt_k4_timing(wgpt_ctx* ctx, int kernel, int reps, float* ms_per_launch, uint64_t* bytes_per_launch);

/* The bytes-only present (rgba_out == NULL) decides round(255 pow(c, 1/2.2)) from an SFU estimate
 * when that is provably enough and from the pinned pow otherwise.  This hook compares the two for
 * every float whose bit pattern lies in [first_bits, first_bits + count) ON THE GPU and returns the
 * number of differing bytes (must be 0) and how many inputs took the pinned path. */
int wgpt_present_quant_check(wgpt_ctx* ctx, uint32_t first_bits, uint32_t count, uint64_t* mismatches, uint64_t* exact_path);

/* ---- multi-GPU (csrc/mgpu.cu).  The scene is replicated; wgpt_config's tile_* / strata_* fields say
 * which pixels / frames a context renders; the partial accumulation buffers are combined by ONE ncclReduce(sum) of
 * binding 3 that the library issues on the context's stream, behind the render kernels (no host synchronisation).
 * Tile shards have disjoint support, so their sum equals the single-GPU buffer bit for bit; frame strata change the
 * per-pixel summation order (f32 rounding).  NCCL is resolved with dlopen at the first of these calls.
 *
 * One process per GPU: rank 0 calls wgpt_nccl_unique_id, the caller ships the 128 bytes to every rank (MPI,
 * torch.distributed, a file ...), every rank calls wgpt_comm_init on its context (collective), then
 * wgpt_reduce_framebuffer(ctx, root) after its wgpt_render calls (collective; root = -1: every rank gets the sum).
 * The call enqueues and returns; wgpt_synchronize / wgpt_read_buffer / wgpt_present wait for it. */
int wgpt_nccl_unique_id(void* id128);
int wgpt_comm_init(wgpt_ctx* ctx, const void* id128, int rank, int world);
int wgpt_reduce_framebuffer(wgpt_ctx* ctx, int root);

/* One process, N devices: a group owns one context per device (rank r: tile shard r % tile_groups of strata shard
 * r / tile_groups; cfg->device / tile_* / strata_* are overwritten per rank; devices == NULL: 0 .. n-1) and one host
 * thread per context for the calls that block.  wgpt_group_reduce leaves the sum in rank `root`'s binding 3 and
 * returns when it is there; wgpt_group_ctx(g, root) is then an ordinary context to read / present. */
typedef struct wgpt_group wgpt_group;
int  wgpt_group_create(const wgpt_config* cfg, int n_devices, const int* devices, int tile_groups, wgpt_group** out);
void wgpt_group_destroy(wgpt_group* g);
const char* wgpt_group_last_error(const wgpt_group* g);
int  wgpt_group_size(const wgpt_group* g);
wgpt_ctx* wgpt_group_ctx(wgpt_group* g, int rank);
int  wgpt_group_write_buffer(wgpt_group* g, int binding, const void* data, size_t bytes);
int  wgpt_group_render(wgpt_group* g, int first_frame, int n_frames);
int  wgpt_group_reduce(wgpt_group* g, int root);

#ifdef __cplusplus
}
#endif
#endif

// fast.cu -- host orchestration of the fast path: builds the 8-wide BVH when bindings 5/9 change
// (K2), owns the wavefront pool, and drives the stage kernels (K3/K4) in batches without
// stalling on the device: the only host<->device traffic per batch is one 4-byte "finished" flag
// read back asynchronously, one batch behind.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <vector>
#include "context.h"
#include "kernels.cuh"
#include "wavefront.cuh"
#include "bvh_build.cuh"

namespace wg {

// the open wavefront session (see fast_submit)
struct Session {
    bool open = false, pending = false, probe = false, counters_on = false;
    int frame0 = 0, stride = 1, reset_first = 0;
    uint32_t frames = 0;                   // frame ordinals submitted so far
    Uniforms uniforms;                     // the view / screen size every frame of the session is rendered with
    void* fb = nullptr;                    // the buffer the frames fold into (binding 3, or the probe's scratch)
    WfParams P;
};

struct FastState {
    Session sess;
    unsigned acc_phase = 0;
    unsigned extend_blocks[2] = {0, 0};    // persistent extend grid on this context's device, without / with counters
    // wide BVH
    uint4* wide_nodes = nullptr; float4* wide_tris = nullptr; uint32_t* wide_scratch = nullptr;
    uint32_t wide_cap = 0, wide_tri_cap = 0;
    // device-built hierarchy (K5, bvh_build.cu) that feeds K2 instead of binding 9
    BvhNode* own_bvh = nullptr; void* build_scratch = nullptr; uint32_t own_cap = 0;
    bool own_used = false;
    WideBuildStats wstats = {0, 0, 0, 0, 127};
    // wavefront
    WfParams P;
    void* pool_mem = nullptr;
    uint32_t capacity = 0, ring = 0;
    uint32_t* h_flag = nullptr;            // pinned: finished flag copies
    cudaEvent_t flag_ev[2] = {nullptr, nullptr};
    uint32_t* d_fallbacks = nullptr;
    uint32_t* d_path_sig = nullptr;        // per-slot hit-sequence hash, probes with signatures only
    bool pool_ready = false;
    // stage profiling: a ring of event quadruples, folded into the sums whenever it fills up
    std::vector<cudaEvent_t> prof_ev;
    size_t prof_used = 0;
    double prof_ms[4] = {0, 0, 0, 0};
    uint64_t prof_launches[4] = {0, 0, 0, 0};
    cudaEvent_t wide_ev[2] = {nullptr, nullptr};
    bool wide_timed = false;
    int wide_launches = 0;              // kernels of the last K2 run
};

static const size_t PROF_QUADS = 2048;
#ifndef WF_ACC_EVERY
#define WF_ACC_EVERY 4
#endif

static void prof_fold(FastState* F) {
    if (!F->prof_used) return;
    cudaEventSynchronize(F->prof_ev[4 * (F->prof_used - 1) + 3]);
    // developer aid: WGPT_PROF_DUMP=file appends one line per iteration (extend, shade, control + accumulate ms)
    const char* dump_path = getenv("WGPT_PROF_DUMP");
    FILE* dump = dump_path && *dump_path ? fopen(dump_path, "a") : nullptr;
    for (size_t i = 0; i < F->prof_used; i++) {
        float a = 0, b = 0, c2 = 0;
        cudaEventElapsedTime(&a, F->prof_ev[4 * i], F->prof_ev[4 * i + 1]);
        cudaEventElapsedTime(&b, F->prof_ev[4 * i + 1], F->prof_ev[4 * i + 2]);
        cudaEventElapsedTime(&c2, F->prof_ev[4 * i + 2], F->prof_ev[4 * i + 3]);
        F->prof_ms[0] += a; F->prof_ms[1] += b; F->prof_ms[2] += c2;
        if (dump) fprintf(dump, "%.4f %.4f %.4f\n", a, b, c2);
        F->prof_launches[0] += 1; F->prof_launches[1] += 1; F->prof_launches[2] += 1;
    }
    if (dump) { fprintf(dump, "\n"); fclose(dump); }
    F->prof_used = 0;
}

static cudaEvent_t* prof_next(FastState* F) {
    if (F->prof_ev.empty()) {
        F->prof_ev.resize(4 * PROF_QUADS);
        for (auto& e : F->prof_ev) cudaEventCreate(&e);
    }
    if (F->prof_used == PROF_QUADS) prof_fold(F);
    return &F->prof_ev[4 * F->prof_used++];
}

static int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}

static bool transforms_consistent(const wgpt_ctx* c);

static int ensure_wide(wgpt_ctx* c) {
    if (!c->fast) c->fast = new FastState();
    FastState* F = c->fast;
    if (!c->wide_dirty) return WGPT_OK;
    const int n_bvh = (int)(c->dbytes[9] / 48), n_tri = (int)(c->dbytes[5] / 96);
    cudaError_t e;
    if (n_bvh == 0) { F->wstats = {0, 0, 0, 0, 127}; c->wide_dirty = false; return WGPT_OK; }
    const uint32_t cap = (uint32_t)std::max(16, n_bvh / 2 + 16);
    if (cap > F->wide_cap || (uint32_t)n_tri > F->wide_tri_cap) {
        if (F->wide_nodes) cudaFree(F->wide_nodes);
        if (F->wide_tris) cudaFree(F->wide_tris);
        if (F->wide_scratch) cudaFree(F->wide_scratch);
        F->wide_nodes = nullptr; F->wide_tris = nullptr; F->wide_scratch = nullptr;
        if ((e = cudaMalloc((void**)&F->wide_nodes, (size_t)cap * 80)) != cudaSuccess ||
            (e = cudaMalloc((void**)&F->wide_tris, (size_t)std::max(1, n_tri) * 48)) != cudaSuccess ||
            (e = cudaMalloc((void**)&F->wide_scratch, wide_scratch_bytes(cap))) != cudaSuccess) return cuda_err(c, e, "wide bvh alloc");
        F->wide_cap = cap; F->wide_tri_cap = (uint32_t)n_tri;
    }
    if (c->stage_profiling) {
        if (!F->wide_ev[0]) { cudaEventCreate(&F->wide_ev[0]); cudaEventCreate(&F->wide_ev[1]); }
        cudaEventRecord(F->wide_ev[0], c->stream);
    }
    // Which hierarchy K2 flattens: by default one built here on the device from the triangles of
    // binding 5 (K5: better surface-area quality than the reference's median split); WGPT_BVH=reference
    // flattens binding 9 itself.  Binding 9 always remains the tree of exact mode and of the tie resolver.
    const char* which = getenv("WGPT_BVH");
    bool own = !(which && which[0] == 'r') && n_tri > 0 && n_bvh == 2 * n_tri - 1 && c->dbytes[6] && c->dbytes[7] && transforms_consistent(c);
    for (int attempt = 0; attempt < 2; attempt++) {
        const BvhNode* src = (const BvhNode*)c->dbuf[9];
        if (own) {
            if ((uint32_t)n_tri > F->own_cap) {
                if (F->own_bvh) cudaFree(F->own_bvh);
                if (F->build_scratch) cudaFree(F->build_scratch);
                F->own_bvh = nullptr; F->build_scratch = nullptr; F->own_cap = 0;
                if ((e = cudaMalloc((void**)&F->own_bvh, (size_t)(2 * n_tri) * sizeof(BvhNode))) != cudaSuccess ||
                    (e = cudaMalloc(&F->build_scratch, bvh_build_scratch_bytes(n_tri))) != cudaSuccess) return cuda_err(c, e, "bvh build alloc");
                F->own_cap = (uint32_t)n_tri;
            }
            launch_bvh_build((const Triangle*)c->dbuf[5], n_tri, (const Mesh*)c->dbuf[6], (const ModelTransform*)c->dbuf[7],
                             F->own_bvh, F->build_scratch, c->stream);
            c->last_launches += BVH_BUILD_LAUNCHES;
            src = F->own_bvh;
        }
        F->wide_launches = launch_wide_build(src, n_bvh, (const Triangle*)c->dbuf[5], n_tri, F->wide_nodes, F->wide_tris,
                                             F->wide_cap, F->wide_scratch, c->stream);
        c->last_launches += F->wide_launches;
        F->wstats = wide_build_stats(F->wide_scratch, c->stream);
        if ((e = cudaGetLastError()) != cudaSuccess) return cuda_err(c, e, "wide bvh build");
        F->own_used = own;
        if (!(F->wstats.overflow || F->wstats.n_tris != (uint32_t)n_tri) || !own) break;
        own = false;                        // a degenerate clustering (too deep for K2): flatten the reference's tree instead
    }
    if (c->stage_profiling) { cudaEventRecord(F->wide_ev[1], c->stream); F->wide_timed = true; }
    if (F->wstats.overflow || F->wstats.n_tris != (uint32_t)n_tri)
        return set_err(c, WGPT_E_INVALID, "binary BVH does not meet the flatten contract (pre-order, 1-triangle leaves, depth <= 48 wide levels)");
    c->wide_dirty = false;
    return WGPT_OK;
}

static WideBvh wide_view(const wgpt_ctx* c, FastState* F) {
    WideBvh w; w.nodes = F->wide_nodes; w.tris = F->wide_tris; w.n_nodes = F->wstats.n_nodes; w.n_tris = F->wstats.n_tris;
    w.force_reference_walk = c->bvh_depth >= std::min(c->cfg.stack_size, 64) ? 1u : 0u;
    // per-ray distance scale of the slab test (trace_wide.cuh): root cells of 2^-40 .. 2^40 world units are representable
    const int er = (int)F->wstats.root_emax;
    if (er < 127 - 40 || er > 127 + 40) w.force_reference_walk = 1u;
    w.scale_k = (uint32_t)(371 - er);
    return w;
}

// M * M^-1 == I to f32 accuracy for every transform of binding 7?  The reference's shaders never read modelMatrix
// (common.wgsl:195 loads it, dead): hits are decided by invModelMatrix and the boxes of binding 9.  K5 takes its world
// boxes from modelMatrix, so a caller whose modelMatrix is stale / identity / ill-conditioned must not get a device-built
// tree: binding 9 is flattened instead.
static bool transforms_consistent(const wgpt_ctx* c) {
    const ModelTransform* t = (const ModelTransform*)c->hbuf[7].data();
    const size_t n = c->hbuf[7].size() / sizeof(ModelTransform);
    for (size_t i = 0; i < n; i++) {
        const float* M = t[i].modelMatrix; const float* V = t[i].invModelMatrix;
        for (int r = 0; r < 4; r++)
            for (int col = 0; col < 4; col++) {
                double acc = 0.0, mag = 0.0;
                for (int k = 0; k < 4; k++) { const double p = (double)M[k * 4 + r] * (double)V[col * 4 + k]; acc += p; mag += p < 0 ? -p : p; }
                const double err = acc - (r == col ? 1.0 : 0.0);
                if (!((err < 0 ? -err : err) <= 4e-6 * (1.0 + mag))) return false;          // also false for NaN
            }
    }
    return true;
}

static void free_pool(FastState* F) {
    if (F->pool_mem) cudaFree(F->pool_mem);
    if (F->h_flag) cudaFreeHost(F->h_flag);
    if (F->flag_ev[0]) cudaEventDestroy(F->flag_ev[0]);
    if (F->flag_ev[1]) cudaEventDestroy(F->flag_ev[1]);
    F->pool_mem = nullptr; F->h_flag = nullptr; F->flag_ev[0] = F->flag_ev[1] = nullptr;
    F->pool_ready = false;
}

static int ensure_pool(wgpt_ctx* c) {
    FastState* F = c->fast;
    if (F->pool_ready) return WGPT_OK;
    free_pool(F);                                      // leftovers of a failed attempt
    const uint32_t npix = (uint32_t)c->cfg.width * c->cfg.height;
    uint32_t cap = (uint32_t)env_int("WGPT_POOL", 0);
    // 4 frames of the WHOLE image in flight, at most 8 Mi slots; a tile shard gets the same pool as the unsharded image (it
    // just holds more frames of its fewer pixels: measured at 8 GPUs, quarter-image tiles with a quarter-size pool lost 11 %)
    if (!cap) cap = std::min<uint64_t>(1u << 23, (uint64_t)npix * 4);
    cap = std::max(1024u, (cap + 1023u) & ~1023u);
    // Frames in flight: a frame is folded into binding 3 only when its slowest path has ended (up to MAX_BOUNCES
    // iterations later), so the window must be long enough to keep the slots fed meanwhile: 8 GiB of staging,
    // between 64 and 256 frames (1080p: 256 frames; 4K: 64).
    uint32_t ring = (uint32_t)env_int("WGPT_RING", 0);
    if (!ring) ring = (uint32_t)std::min<uint64_t>(WF_MAX_RING, std::max<uint64_t>(64, (8ull << 30) / ((uint64_t)npix * 16)));
    ring = std::max(2u, std::min<uint32_t>(ring, WF_MAX_RING));
    // one allocation, carved into 16-byte-aligned arrays; when the device cannot give that much (a smaller GPU, several
    // contexts or ranks on one GPU) the staging ring is halved down to 2 frames before giving up
    const size_t per_slot = WF_BYTES_PER_SLOT;
    size_t bytes = 0;
    cudaError_t e = cudaSuccess;
    for (;; ring = std::max(2u, ring / 2)) {
        bytes = (size_t)cap * per_slot + (size_t)ring * npix * 16 + (size_t)cap / WF_SHADE_THREADS / 8 + WF_FRONT_BLOCKS * 4 + WF_WORK_SHARDS * 128 + 56 * 256;
        e = cudaMalloc(&F->pool_mem, bytes);
        if (e != cudaErrorMemoryAllocation || ring == 2) break;
        cudaGetLastError();
    }
    if (e != cudaSuccess) { F->pool_mem = nullptr; return cuda_err(c, e, "wavefront pool alloc"); }
    if ((e = cudaMemsetAsync(F->pool_mem, 0, bytes, c->stream)) != cudaSuccess) { free_pool(F); return cuda_err(c, e, "wavefront pool clear"); }
    char* p = (char*)F->pool_mem;
    auto carve = [&](size_t n) { char* r = p; p += (n + 255) & ~(size_t)255; return r; };
    WfParams& P = F->P;
    memset(&P, 0, sizeof(P));
    P.ctl = (WfControl*)carve(sizeof(WfControl));
    P.pool.e0 = (float4*)carve((size_t)cap * 16); P.pool.e1 = (float4*)carve((size_t)cap * 16);
    P.pool.s0 = (float4*)carve((size_t)cap * 16); P.pool.s1 = (float4*)carve((size_t)cap * 16);
    P.pool.x = (float4*)carve((size_t)cap * 16);  P.pool.pix = (float4*)carve((size_t)cap * 16);
    P.pool.capacity = cap;
    P.q_extend = (uint32_t*)carve((size_t)cap * 4);
    P.q_live[0] = (uint32_t*)carve((size_t)cap * 4); P.q_live[1] = (uint32_t*)carve((size_t)cap * 4);
    P.work_shards = (unsigned long long*)carve((size_t)WF_WORK_SHARDS * 128);
    P.tie_bitmap = (uint32_t*)carve(((size_t)cap / WF_SHADE_THREADS / 32 + 2) * 4);
    P.tie_blocks = (uint32_t*)carve((size_t)WF_FRONT_BLOCKS * 4);
    P.stage = (float4*)carve((size_t)ring * npix * 16);
    P.ring = ring;
    P.items_per_frame = c->shard_pixels;
    P.inv_items_per_frame = 1.0 / (double)(c->shard_pixels ? c->shard_pixels : 1u);
    if ((size_t)(p - (char*)F->pool_mem) > bytes) { free_pool(F); return set_err(c, WGPT_E_NOMEM, "pool carve overflow"); }
    F->capacity = cap; F->ring = ring;
    if ((e = cudaMallocHost((void**)&F->h_flag, 64)) != cudaSuccess) { free_pool(F); return cuda_err(c, e, "pinned flag"); }
    memset(F->h_flag, 0, 64);
    if ((e = cudaEventCreateWithFlags(&F->flag_ev[0], cudaEventDisableTiming)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&F->flag_ev[1], cudaEventDisableTiming)) != cudaSuccess) { free_pool(F); return cuda_err(c, e, "wavefront pool"); }
    // the persistent extend grid of THIS context's device (a process may drive several devices)
    F->extend_blocks[0] = wf_extend_grid(false);
    F->extend_blocks[1] = wf_extend_grid(true);
    F->pool_ready = true;
    return WGPT_OK;
}

static int check_materials(wgpt_ctx* c) {
    const Material* m = (const Material*)c->hbuf[8].data();
    for (size_t i = 0; i < c->hbuf[8].size() / 64; i++) {
        const float t = m[i].material_type;
        if (!(t == 0.0f || t == 1.0f || t == 2.0f || t == 3.0f))
            return set_err(c, WGPT_E_INVALID, "fast mode shades material_type 0..3 only (header.wgsl:4-8); use exact mode for others");
    }
    return WGPT_OK;
}

// ---- the wavefront SESSION --------------------------------------------------------------------------------------
// A session is one living wavefront: the frames submitted so far (frame0, frame0 + stride, ...), the uniforms they
// are rendered with, the framebuffer they fold into.  wgpt_render opens one, pumps iterations until the device reports
// that every frame has been folded, and leaves it open; wgpt_dispatch APPENDS its frame to the open session when it
// continues it (next frame number, same view, resetBuffer = 0) and only enqueues a few iterations: the call returns
// while the GPU works (webgpu-utils.js:125-134 submits and returns), later frames overlap the tails of earlier ones,
// and whoever needs the pixels (present, read_buffer, counters, a buffer write, ...) flushes first.
static bool same_view(const Uniforms& a, const Uniforms& b) {
    return a.screenW == b.screenW && a.screenH == b.screenH && memcmp(a.viewMatrix, b.viewMatrix, sizeof(a.viewMatrix)) == 0;
}

int fast_pump(wgpt_ctx* c, int iterations) {
    FastState* F = c->fast;
    if (!F || !F->sess.open) return WGPT_OK;
    Session& s = F->sess;
    DeviceScene S = make_device_scene(c, 0xffffffffu);
    S.uniforms = s.uniforms;                               // the view the session's frames were submitted with
    S.counters = s.counters_on ? c->d_counters : nullptr;
    const WideBvh WB = wide_view(c, F);
    // the accumulate pass runs every WF_ACC_EVERY-th iteration and at the end of every pump: frames that complete in
    // between wait in the staging ring (its window follows the frames FOLDED, so nothing is overwritten)
    for (int i = 0; i < iterations; i++) {
        const bool do_acc = i + 1 == iterations || (++F->acc_phase % WF_ACC_EVERY) == 0;
        c->last_launches += (uint64_t)wf_launch_iteration(S, WB, s.P, (float4*)s.fb, s.reset_first, F->extend_blocks[S.counters ? 1 : 0],
                                                          do_acc, c->stream, c->stage_profiling ? prof_next(F) : nullptr);
    }
    return cuda_err(c, cudaGetLastError(), "wavefront iteration");
}

// runs the open session until every submitted frame is in the framebuffer (the device stays one batch ahead of the
// host: the only traffic is one 4-byte "finished" flag per batch, read back asynchronously)
int fast_flush(wgpt_ctx* c) {
    FastState* F = c->fast;
    if (!F || !F->sess.open || !F->sess.pending) return WGPT_OK;
    const int batch = std::max(1, env_int("WGPT_BATCH", 8));
    F->h_flag[0] = F->h_flag[1] = 0;
    for (int b = 0;; b++) {
        int rc = fast_pump(c, batch);
        if (rc) return rc;
        cudaMemcpyAsync(&F->h_flag[b & 1], &F->sess.P.ctl->finished, 4, cudaMemcpyDeviceToHost, c->stream);
        cudaEventRecord(F->flag_ev[b & 1], c->stream);
        if (b >= 1) {                          // look at the previous batch's flag: the device stays one batch ahead
            cudaError_t e = cudaEventSynchronize(F->flag_ev[(b - 1) & 1]);
            if (e != cudaSuccess) return cuda_err(c, e, "wavefront batch");
            if (F->h_flag[(b - 1) & 1]) break;
        }
        if (b > 4000000) return set_err(c, WGPT_E_STATE, "wavefront did not converge");
    }
    F->sess.pending = false;
    return cuda_err(c, cudaGetLastError(), "fast_flush");
}

// finishes and forgets the session (buffer writes, probes, mode changes): nothing may append to it afterwards
int fast_close(wgpt_ctx* c) {
    if (!c->fast) return WGPT_OK;
    const int rc = fast_flush(c);
    c->fast->sess.open = false; c->fast->sess.pending = false;
    return rc;
}

// Adds frames first_frame .. first_frame + n_frames - 1 (those of this context's strata shard when `strata`) to the
// wavefront: appended to the open session when they continue it, else a new session after the old one has drained.
// Enqueues no iteration by itself.
int fast_submit(wgpt_ctx* c, int first_frame, int n_frames, bool strata, int reset_first, int* d_ptype, int* d_pindex, float* d_pt) {
    int rc = check_materials(c);
    if (rc) return rc;
    if (c->cfg.max_bounces > WF_MAX_BOUNCE_FIELD || c->cfg.num_samples > WF_MAX_SAMPLE_FIELD)
        return set_err(c, WGPT_E_INVALID, "fast mode packs the bounce / sample counters: max_bounces <= 1023, num_samples <= 16383");
    if ((rc = ensure_wide(c)) != WGPT_OK) return rc;
    if ((rc = ensure_pool(c)) != WGPT_OK) return rc;
    FastState* F = c->fast;
    // frames of this strata shard: f in [first, first+n) with f % world == rank
    const int world = strata ? c->cfg.strata_world : 1, rank = strata ? c->cfg.strata_rank : 0;
    int f0 = first_frame;
    while (world > 1 && (f0 % world) != rank) f0++;
    const int last = first_frame + n_frames;
    const uint32_t my_frames = f0 < last ? (uint32_t)((last - 1 - f0) / world + 1) : 0u;
    if (my_frames == 0) return WGPT_OK;
    Session& s = F->sess;
    const bool probe = d_ptype || d_pindex || d_pt;
    const bool append = s.open && !reset_first && !probe && !s.probe && s.stride == world &&
                        (long long)f0 == (long long)s.frame0 + (long long)s.frames * world && same_view(s.uniforms, c->uniforms) &&
                        s.fb == c->dbuf[3] && s.counters_on == c->counters_on && s.frames + my_frames < (1u << 30);
    if (append) {
        wf_launch_append(s.P, my_frames, c->stream);
        s.frames += my_frames;
    } else {
        if ((rc = fast_close(c)) != WGPT_OK) return rc;
        s.P = F->P;
        s.P.frame0 = f0; s.P.frame_stride = world;
        s.P.bvh_depth = c->bvh_depth;
        s.P.tol = c->cfg.mode == WGPT_MODE_FAST_TOL ? 1 : 0;
        s.P.prim_type = d_ptype; s.P.prim_index = d_pindex; s.P.prim_t = d_pt;
        s.P.prim_rng = probe ? c->d_prng : nullptr;
        s.P.path_sig = nullptr;
        if (s.P.prim_rng) {                                 // signatures requested (wgpt_probe_rng): the per-slot hit-sequence hash
            if (!F->d_path_sig) {
                cudaError_t e = cudaMalloc((void**)&F->d_path_sig, (size_t)F->capacity * 4);
                if (e != cudaSuccess) return cuda_err(c, e, "path signature array");
            }
            cudaMemsetAsync(F->d_path_sig, 0, (size_t)F->capacity * 4, c->stream);
            s.P.path_sig = F->d_path_sig;
        }
        s.frame0 = f0; s.stride = world; s.frames = my_frames;
        s.uniforms = c->uniforms; s.fb = c->dbuf[3]; s.reset_first = reset_first; s.probe = probe;
        s.counters_on = c->counters_on;
        wf_launch_init(s.P, my_frames, c->stream);
        s.open = true;
    }
    c->last_launches += 1;
    s.pending = true;
    return cuda_err(c, cudaGetLastError(), "fast_submit");
}

int fast_render_probe(wgpt_ctx* c, int first_frame, int n_frames, int reset_first, int* d_ptype, int* d_pindex, float* d_pt) {
    int rc = fast_submit(c, first_frame, n_frames, true, reset_first, d_ptype, d_pindex, d_pt);
    if (rc) return rc;
    return fast_flush(c);
}

int fast_render(wgpt_ctx* c, int first_frame, int n_frames, int reset_first) {
    return fast_render_probe(c, first_frame, n_frames, reset_first, nullptr, nullptr, nullptr);
}

// wgpt_dispatch in fast mode: the frame joins the wavefront and the call returns after enqueueing a few iterations
int fast_dispatch(wgpt_ctx* c, int frame, int reset_first) {
    int rc = fast_submit(c, frame, 1, false, reset_first, nullptr, nullptr, nullptr);
    if (rc) return rc;
    return fast_pump(c, std::max(1, env_int("WGPT_DISPATCH_ITERS", 4)));
}

int fast_trace_rays(wgpt_ctx* c, int n, const float* d_rays, const uint32_t* d_seeds, int* d_type, int* d_index, float* d_t) {
    int rc = ensure_wide(c);
    if (rc) return rc;
    FastState* F = c->fast;
    if (!F->d_fallbacks) {
        cudaError_t e = cudaMalloc((void**)&F->d_fallbacks, 4);
        if (e != cudaSuccess) return cuda_err(c, e, "fallback counter");
        cudaMemsetAsync(F->d_fallbacks, 0, 4, c->stream);
    }
    if (!c->have_uniforms) memset(&c->uniforms, 0, sizeof(c->uniforms));
    const DeviceScene S = make_device_scene(c, 0xffffffffu);
    launch_trace_rays_wide(S, wide_view(c, F), n, d_rays, d_seeds, d_type, d_index, d_t, F->d_fallbacks, c->stream);
    return cuda_err(c, cudaGetLastError(), "fast_trace_rays");
}

// fallbacks (rays re-traced in reference order) of the last render + of trace_rays since creation; wide BVH stats
int fast_stats(wgpt_ctx* c, uint32_t out[8]) {
    memset(out, 0, 32);
    FastState* F = c->fast;
    if (!F) return WGPT_OK;
    out[0] = F->wstats.n_nodes; out[1] = F->wstats.n_tris; out[2] = F->wstats.n_levels; out[3] = F->capacity; out[4] = F->ring;
    fast_flush(c);
    cudaStreamSynchronize(c->stream);
    if (F->pool_ready) {
        WfControl h;
        cudaMemcpy(&h, F->P.ctl, sizeof(h), cudaMemcpyDeviceToHost);
        out[5] = h.fallbacks; out[6] = h.iterations;
    }
    if (F->d_fallbacks) cudaMemcpy(&out[7], F->d_fallbacks, 4, cudaMemcpyDeviceToHost);
    return WGPT_OK;
}

// rays the extend stage traced in iteration i (i < 64) of the last session: what an ncu capture of launch i processed
int fast_extend_history(wgpt_ctx* c, uint32_t out[64]) {
    memset(out, 0, 64 * sizeof(uint32_t));
    FastState* F = c->fast;
    if (!F || !F->pool_ready) return WGPT_OK;
    int rc = fast_flush(c);
    if (rc) return rc;
    cudaStreamSynchronize(c->stream);
    return cuda_err(c, cudaMemcpy(out, F->P.ctl->hist_n_extend, 64 * sizeof(uint32_t), cudaMemcpyDeviceToHost), "extend_history");
}

int fast_stage_timings(wgpt_ctx* c, float ms[4], uint64_t launches[4]) {
    for (int i = 0; i < 4; i++) { ms[i] = 0.0f; launches[i] = 0; }
    FastState* F = c->fast;
    if (!F) return WGPT_OK;
    prof_fold(F);
    if (F->wide_timed) {
        float t = 0;
        cudaEventSynchronize(F->wide_ev[1]);
        cudaEventElapsedTime(&t, F->wide_ev[0], F->wide_ev[1]);
        F->prof_ms[3] += t; F->prof_launches[3] += F->wide_launches + (F->own_used ? BVH_BUILD_LAUNCHES : 0); F->wide_timed = false;
    }
    for (int i = 0; i < 4; i++) { ms[i] = (float)F->prof_ms[i]; launches[i] = F->prof_launches[i]; F->prof_ms[i] = 0; F->prof_launches[i] = 0; }
    return WGPT_OK;
}

void fast_destroy(wgpt_ctx* c) {
    FastState* F = c->fast;
    if (!F) return;
    for (auto& e : F->prof_ev) cudaEventDestroy(e);
    if (F->wide_ev[0]) { cudaEventDestroy(F->wide_ev[0]); cudaEventDestroy(F->wide_ev[1]); }
    if (F->wide_nodes) cudaFree(F->wide_nodes);
    if (F->wide_tris) cudaFree(F->wide_tris);
    if (F->wide_scratch) cudaFree(F->wide_scratch);
    if (F->own_bvh) cudaFree(F->own_bvh);
    if (F->build_scratch) cudaFree(F->build_scratch);
    free_pool(F);
    if (F->d_fallbacks) cudaFree(F->d_fallbacks);
    if (F->d_path_sig) cudaFree(F->d_path_sig);
    delete F;
    c->fast = nullptr;
}

}  // namespace wg

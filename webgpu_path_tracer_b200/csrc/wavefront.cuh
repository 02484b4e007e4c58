// wavefront.cuh -- state of the persistent-thread wavefront integrator (K3): SoA path pool, ray /
// shade queues, the device-resident control block, and the frame staging ring that keeps the
// per-pixel float sum in FRAME ORDER (main.wgsl:24-27 adds one frame at a time) while many frames
// are in flight.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>
#include "scene.cuh"
#include "wide_bvh.cuh"

namespace wg {

#define WF_MAX_RING 256
#ifndef WF_SHADE_THREADS
#define WF_SHADE_THREADS 256     // slots per shade block (the smaller of the two block sizes: sizes bitmaps)
#endif
#ifndef WF_FRONT_BLOCKS
#define WF_FRONT_BLOCKS 512      // leading shade blocks that take over the blocks holding marked slots
#endif
// slot status / shade class, kept in hid.w: 0..3 = material_type of the hit to shade, 4 = miss to
// shade, 5 = idle (no path; waiting for the ring window to release work), 6 = dead (no path, no
// work left), 7 = in the extend queue
#define WF_CLS_MISS 4
#define WF_CLS_IDLE 5
#define WF_CLS_DEAD 6
#define WF_CLS_TRACE 7      // ray queued for the extend stage (which overwrites the class)
#define WF_CLS_TIE 8        // extend found >= 2 triangles at exactly the closest distance: shade re-walks in reference order
#define WF_CLS_OVERFLOW 9   // the wide stack overflowed: shade re-walks the whole reference tree
#define WF_CLASSES 10

// One path per slot, structure-of-arrays, 16-byte records (LDG.128 / STG.128), 80 bytes per slot
// + the 16-byte pixel sum that is only touched when a sample starts or ends.  The shade stage is
// HBM-bound on exactly these bytes (it reads and rewrites every live slot once per bounce), so
// the state is packed: no padding lanes, ids in one word.
//   x  = (tag, hit index, hit u, hit v)               tag = (material id << 8) | (shade class << 4) | (hit kind + 1);
//        the record the shade stage reads FIRST: it holds the sort key and everything needed to address
//        the hit's data (triangle normals, material), so all further loads can be issued at once.
//        For WF_CLS_TIE the u lane carries the closest distance the extend stage found.
//   e0 = (ray origin.xyz, hit t)
//   e1 = (ray direction.xyz, counters)                 counters = (ring slot << 24) | (bounce << 14) | sample
//   s0 = (throughput.xyz, radiance.x)
//   s1 = (radiance.y, radiance.z, rng state, shard-local pixel)
// The extend stage reads e0, e1 and x.tag and writes x (and the t lane of e0) when a triangle wins.
struct PathPool {
    float4* x; float4* e0; float4* e1; float4* s0; float4* s1;
    float4* pix;     // pixColor of the (pixel, frame) so far .xyz, .w = samples taken (stratified divisor)
    uint32_t capacity;
};
#define WF_BYTES_PER_SLOT (6 * 16 + 4 + 8)  /* 6 records + one ray-queue entry + two live-list entries */
#define WF_MAX_BOUNCE_FIELD 1023
#define WF_MAX_SAMPLE_FIELD 16383
#define WF_CTR_FIRST_QUERY_MASK 0x00ffffffu   /* bounce == 0 and sample == 0: the primary ray of the item */

__host__ __device__ inline uint32_t wf_tag(int kind, int cls, int mat_id) { return ((uint32_t)mat_id << 8) | ((uint32_t)cls << 4) | (uint32_t)(kind + 1); }
__host__ __device__ inline int wf_tag_kind(uint32_t tag) { return (int)(tag & 15u) - 1; }
__host__ __device__ inline int wf_tag_cls(uint32_t tag) { return (int)((tag >> 4) & 15u); }
__host__ __device__ inline int wf_tag_mat(uint32_t tag) { return (int)(tag >> 8) == 0xffffff ? -1 : (int)(tag >> 8); }
__host__ __device__ inline uint32_t wf_ctr(uint32_t ring_slot, uint32_t bounce, uint32_t scnt) { return (ring_slot << 24) | (bounce << 14) | (scnt & 0x3fffu); }

// Work items (frame ordinal, pixel) are numbered g = 0 .. total_work - 1, frame-major.  One global counter handing them
// out is a single L2 address taking an atomic from every shade warp of every launch (~0.5 per ns: its round trip was
// 8-15 % of the shade stage's stall samples).  The item space is therefore dealt out in chunks of WF_WORK_CHUNK
// consecutive items to WF_WORK_SHARDS counters, one cache line each (chunk q belongs to shard q % SHARDS); a shade block
// claims from shard (block % SHARDS).  Inside a shard the local index l counts that shard's items in increasing g.
#ifndef WF_WORK_SHARDS
#define WF_WORK_SHARDS 16
#endif
#define WF_WORK_CHUNK_LOG2 8
// number of items g < n that belong to shard c
__host__ __device__ inline unsigned long long wf_shard_count(unsigned long long n, uint32_t c) {
    const unsigned long long q = n >> WF_WORK_CHUNK_LOG2, r = n & ((1ull << WF_WORK_CHUNK_LOG2) - 1ull);
    const unsigned long long full = q > c ? (q - c + WF_WORK_SHARDS - 1) / WF_WORK_SHARDS : 0ull;
    return (full << WF_WORK_CHUNK_LOG2) + ((q % WF_WORK_SHARDS) == c ? r : 0ull);
}
// global item number of local index l of shard c
__host__ __device__ inline unsigned long long wf_shard_item(unsigned long long l, uint32_t c) {
    const unsigned long long q = (l >> WF_WORK_CHUNK_LOG2) * WF_WORK_SHARDS + c;
    return (q << WF_WORK_CHUNK_LOG2) + (l & ((1ull << WF_WORK_CHUNK_LOG2) - 1ull));
}

struct WfControl {
    unsigned long long next_work;      // (unused since the counter is sharded: WfParams::work_shards)
    unsigned long long work_limit;     // items below this may start (ring window)
    unsigned long long total_work;
    uint32_t n_extend;                 // rays queued for the extend stage of this iteration
    uint32_t extend_cursor;            // persistent-warp fetch cursor
    uint32_t n_next;                   // rays compacted by shade for the next iteration
    uint32_t done[WF_MAX_RING];        // finished items per ring slot
    uint32_t acc_k;                    // frames whose every item has finished (complete in the staging ring)
    uint32_t acc_done;                 // frames folded into binding 3 (<= acc_k: the accumulate pass runs every few iterations);
                                       // the ring window and `finished` follow THIS counter: a staging slot is free once folded
    uint32_t ready_first, n_ready;     // frames the accumulate pass launched behind this control step folds (0: none launched)
    uint32_t n_frames;                 // frame ordinals to render in this call
    uint32_t finished;
    uint32_t fallbacks;                // rays re-traced with the reference-order walk
    uint32_t iterations;
    // Drain: once every work item has been handed out no slot can start a new path, and the live slots thin out
    // over many iterations (a few long paths keep a render going).  Shade then lists the slots that are still
    // alive, and once that list is short the next shade launches run over the LIST instead of over every slot.
    uint32_t drain;                    // all work handed out (set by k_wf_ctl_end; shade lists its live slots from then on)
    uint32_t sparse;                   // this iteration's shade runs over q_live[live_read][0 .. n_live)
    uint32_t live_read;                // which of the two live lists is read (the other one is being written)
    uint32_t n_live, n_live_next;
    unsigned long long shard_lim[WF_WORK_SHARDS];   // wf_shard_count(work_limit, c), refreshed with work_limit
    uint32_t n_tie_blocks;             // shade blocks that hold a slot the extend stage marked (tie / stack overflow) this iteration
    uint32_t hist_n_extend[64];        // rays queued for the extend stage of iteration i (i < 64) of the session: measurement only
};

struct WfParams {
    PathPool pool;
    WfControl* ctl;
    uint32_t* q_extend;                // slot ids of live rays, compacted by shade (capacity entries)
    uint32_t* q_live[2];               // slot ids of live paths (drain phase), double-buffered
    // The reference-order walk of a marked slot is one long chain of dependent loads (tens of microseconds); a shade
    // launch cannot end before its slowest one.  The extend stage therefore notes WHICH shade blocks hold marked
    // slots, and the shade launch runs exactly those blocks first (WF_FRONT_BLOCKS leading blocks take them, the
    // block at the natural position leaves), so the walks overlap the rest of the launch instead of trailing it.
    unsigned long long* work_shards;   // WF_WORK_SHARDS counters, 16 words (128 B) apart: next local index of each shard
    uint32_t* tie_bitmap;              // one bit per shade block (capacity / WF_SHADE_THREADS)
    uint32_t* tie_blocks;              // WF_FRONT_BLOCKS entries
    float4* stage;                     // ring x W*H: (pixColor, 1)
    uint32_t ring;                     // frames in flight
    uint32_t items_per_frame;          // shard.n_pixels
    double inv_items_per_frame;        // 1.0 / items_per_frame (quotient estimate, fixed up with integers)
    int bvh_depth;                     // deepest leaf of the binary BVH (bounds the reference's stack use)
    int frame0, frame_stride;          // frame number of ordinal k = frame0 + k * frame_stride
    // Slots per shade block of THIS launch: 256, or 512 for scenes with several shade classes in flight (spheres in
    // binding 1, or IMPORTANCE_SAMPLING on): the block-local material sort then leaves fewer mixed warps (shade C1 -12 %,
    // C3 -9 %); single-material scenes gain nothing from the sort and pay the heavier block (C4 +11 %), so they keep 256.
    // The extend stage marks tie / overflow slots per shade block, so it needs the same number.
    uint32_t shade_threads;
    int tol;                           // WGPT_MODE_FAST_TOL: shading of a hit on fast arithmetic (shade.cuh wg::tol); intersections, camera, RNG stay pinned
    int do_acc;                        // the accumulate pass is launched right after this iteration's control step (set per launch)
    // optional probe outputs (first closest-hit query of every pixel), indexed by pixel
    int* prim_type; int* prim_index; float* prim_t;
    uint32_t* path_sig;                // optional, per SLOT: hash of the hit sequence of the path in the slot (probes with signatures only)
    uint32_t* prim_rng;                // optional: path signature of every (pixel, frame) item when it FINISHED: its RNG state (the draws it
                                       // made) xor the hash of its hit sequence; indexed by pixel
};


// wavefront.cu
void wf_launch_init(const WfParams& P, uint32_t n_frames, cudaStream_t stream);
void wf_launch_append(const WfParams& P, uint32_t more_frames, cudaStream_t stream);
unsigned wf_extend_grid(bool counters);      // CTAs of the persistent extend grid on the current device
// ev: optional 4 events recorded before extend, after extend, after shade, after control (+ accumulate)
// do_acc: also launch the accumulate pass (frames complete since the last one are folded); returns the launches issued
int wf_launch_iteration(const DeviceScene& S, const WideBvh& WB, const WfParams& P, float4* fb, int reset_first,
                        unsigned extend_blocks, bool do_acc, cudaStream_t stream, cudaEvent_t* ev = nullptr);
void launch_trace_rays_wide(const DeviceScene& S, const WideBvh& WB, int n, const float* rays, const uint32_t* seeds,
                            int* out_type, int* out_index, float* out_t, uint32_t* fallbacks, cudaStream_t stream);

}  // namespace wg

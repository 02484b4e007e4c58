// wavefront.cu -- K3: the persistent-thread wavefront integrator.
//
// Path state lives in a structure-of-arrays pool (one slot per path); one iteration advances every
// live path by one bounce with two kernels sized from a device-resident control block (no host
// round trip):
//   extend   persistent warps.  Every lane owns one ray; lanes that finish pull the next slot id
//            from the compacted ray queue with one warp-aggregated atomic (ballot + popc), so a
//            warp never idles behind its longest traversal (dynamic ray fetch).  Closest hit =
//            spheres/quads + 8-wide BVH (trace_wide.cuh); result and shade class go back to the slot.
//   shade    threads map to slots in slot order (coalesced SoA); each block first sorts its 256
//            slots by material class in shared memory so warps shade one material at a time;
//            miss -> background; hit -> emission, BSDF sample, mixture pdf against the light quad
//            (the reference's "shadow" term is the analytic light_pdf, importanceSampling.wgsl:88-125:
//            it casts no occlusion ray), Russian roulette.  Finished samples fold into the pixel,
//            finished (pixel, frame) items go to the staging ring, and the slot REGENERATES from
//            the next work item (generate stage).  Live rays are compacted (block-level
//            ballot/prefix, one atomic per block) into the next ray queue.
//   accumulate  K4: frames whose every item finished are added to binding 3 in frame order.
// The RNG stream of a (pixel, frame) item lives in its slot, so the draw order equals the
// reference's whatever the schedule; all arithmetic is pinned.
#include <stdio.h>
#include "wavefront.cuh"
#include "shade.cuh"
#include "camera.cuh"
#include "trace_wide.cuh"

namespace wg {

#ifndef WF_SPARSE_BELOW
#define WF_SPARSE_BELOW 1        // shade runs over the live list once it holds at most capacity / this many slots (1, 2, 4, 8 measured: 1 is best by 0.3 %)
#endif
#ifndef WF_REFILL_BELOW
#define WF_REFILL_BELOW 22      // refill the warp when fewer lanes than this still traverse (finished lanes write back together then)
#endif

__device__ __forceinline__ uint32_t lanemask_lt() { uint32_t m; asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m)); return m; }

__global__ void k_wf_init(WfParams P, uint32_t n_frames) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < P.pool.capacity) P.pool.x[i] = make_float4(__uint_as_float(wf_tag(HIT_NONE, WF_CLS_IDLE, -1)), 0.0f, 0.0f, 0.0f);
    if (i == 0) {
        WfControl* c = P.ctl;
        c->next_work = 0;
        c->total_work = (unsigned long long)n_frames * P.items_per_frame;
        const unsigned long long win = (unsigned long long)P.ring * P.items_per_frame;
        c->work_limit = win < c->total_work ? win : c->total_work;
        for (uint32_t k = 0; k < WF_WORK_SHARDS; k++) { P.work_shards[16 * k] = 0ull; c->shard_lim[k] = wf_shard_count(c->work_limit, k); }
        c->n_extend = 0; c->extend_cursor = 0; c->n_next = 0;
        for (int k = 0; k < WF_MAX_RING; k++) c->done[k] = 0;
        c->acc_k = 0; c->acc_done = 0; c->ready_first = 0; c->n_ready = 0; c->n_frames = n_frames;
        c->finished = n_frames == 0 ? 1u : 0u; c->fallbacks = 0; c->iterations = 0;
        c->drain = 0; c->sparse = 0; c->live_read = 0; c->n_live = 0; c->n_live_next = 0; c->n_tie_blocks = 0;
        for (int k = 0; k < 64; k++) c->hist_n_extend[k] = 0;
    }
}

// More frames join a wavefront that is still alive (or has just run dry): wgpt_dispatch appends one frame per call
// to the open session instead of filling and draining the pool for every frame.  Slots that found no work left
// (DEAD) poll the work counters again; the window, the totals and the drain bookkeeping restart from the new total.
__global__ void k_wf_append(WfParams P, uint32_t more_frames) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < P.pool.capacity) {
        uint32_t* tag = reinterpret_cast<uint32_t*>(&P.pool.x[i]);
        if (wf_tag_cls(*tag) == WF_CLS_DEAD) *tag = wf_tag(HIT_NONE, WF_CLS_IDLE, -1);
    }
    if (i == 0) {
        WfControl* c = P.ctl;
        c->n_frames += more_frames;
        c->total_work += (unsigned long long)more_frames * P.items_per_frame;
        const unsigned long long lim = (unsigned long long)(c->acc_done + P.ring) * P.items_per_frame;
        c->work_limit = lim < c->total_work ? lim : c->total_work;
        for (uint32_t k = 0; k < WF_WORK_SHARDS; k++) c->shard_lim[k] = wf_shard_count(c->work_limit, k);
        c->finished = 0; c->drain = 0; c->sparse = 0; c->n_live = 0; c->n_live_next = 0;
    }
}

// first closest-hit query of a (pixel, frame) item -> parity probe arrays
__device__ __forceinline__ void write_probe(const DeviceScene& S, const WfParams& P, uint32_t pix_i, bool hit, const HitLite& h) {
    const uint32_t p = shard_pixel(S.shard, S.W, pix_i);
    P.prim_type[p] = hit ? (h.kind == HIT_VOLUME ? 0 : h.kind) : -1;
    P.prim_index[p] = hit ? (h.kind == HIT_TRI ? (int)S.triangles[h.index].local_id : h.index) : -1;
    P.prim_t[p] = hit ? h.t : 0.0f;
}

// Conservative test of a ray against the root box of the BVH over (0, closest]: rays that fail it
// cannot hit any triangle and never enter the extend queue.
__device__ __forceinline__ bool root_box_hit(const DeviceScene& S, const Ray& ray, float closest) {
    const WideRay wr = make_wide_ray(ray);
    const float4 a = __ldg(reinterpret_cast<const float4*>(S.bvh)), b = __ldg(reinterpret_cast<const float4*>(S.bvh) + 1);
    const float x0 = (a.x - ray.o.x) * wr.idx, x1 = (b.x - ray.o.x) * wr.idx;
    const float y0 = (a.y - ray.o.y) * wr.idy, y1 = (b.y - ray.o.y) * wr.idy;
    const float z0 = (a.z - ray.o.z) * wr.idz, z1 = (b.z - ray.o.z) * wr.idz;
    const float tlo = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), 0.0f));
    const float thi = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), closest));
    return tlo * 0.999998f <= thi * 1.000002f;
}

// ---------------------------------------------------------------- extend
// The sphere/quad part of hitScene (hitRay.wgsl:6-40) was already evaluated by the shade stage
// when it produced the ray (all lanes in lock step there); this kernel walks the wide BVH.
// Node visits and triangle tests run as separate warp phases: a lane that found candidate
// triangles parks them and goes on with its next node; the triangle phase runs once TRI_BATCH
// lanes have parked triangles
#ifndef WF_TRI_BATCH
#define WF_TRI_BATCH 16
#endif
#ifndef WF_TRI_ADAPT
#define WF_TRI_ADAPT 2       // ... or as soon as ADAPT x (lanes with parked triangles) >= lanes with node work: measured sweep, DESIGN.md 4
#endif

#ifndef WF_EXTEND_MINB
#define WF_EXTEND_MINB 7
#endif
#ifndef WF_SSTACK
#define WF_SSTACK 12        // traversal-stack levels kept in shared memory (8 B x 128 threads each)
#endif
template <bool COUNT>
__global__ void __launch_bounds__(128, WF_EXTEND_MINB) k_wf_extend(DeviceScene S, WideBvh WB, WfParams P) {
    __shared__ unsigned char s_lut[WIDE_LUT_BYTES];
    WfControl* ctl = P.ctl;
    const uint32_t n = ctl->n_extend;
    if (n == 0) return;                                    // nothing queued (e.g. the iterations after the last frame finished)
    wide_lut_init(s_lut);
    const uint32_t lane = threadIdx.x & 31u;
    Counts c = {0, 0, 0, 0, 0, 0, 0, 0};

    // Per-lane state is kept to what the walk itself needs (the kernel sits at its register cap): the ray direction is
    // only used to set up the object-space ray (re-read from the slot in the rare multi-mesh case), the slot's tag and
    // its sphere/quad result stay in the slot and are re-read at write-back, and of the triangle hit only
    // (index, u, v, material) are carried -- its distance IS `closest`.
    bool has = false, exhausted = (n == 0);
    uint32_t slot = 0;
    v3 ro = V3s(0.0f);
    WideRay wr; wr.idx = wr.idy = wr.idz = 0.0f; wr.octinv = 0; wr.far_scale = 1.0f;
    ObjRayCache oc; oc.reset();
    int h_index = -1, h_mat = -1;
    float h_uu = 0.0f, h_vv = 0.0f;
    float closest = WGPT_MAX_FLOAT;
    bool hit_t = false, ambiguous = false, tie = false;
    uint2 G = make_uint2(0u, 0u);
    uint32_t tm = 0, tri_base = 0, tri_lmask = 0;   // parked candidate triangles of the last node visited: hit leaf slots, first
                                                    // packed triangle, leaf-slot mask (triangle of slot b = tri_base + popc(lmask below b))
    // A second, PARKED set of candidate triangles: a lane whose node visit found leaf hits parks them and goes
    // on with its next node instead of idling until enough lanes have triangles for a triangle phase; it only
    // blocks when a later visit finds leaf hits while the parked set is still waiting.
    uint32_t pk = 0, ptri_base = 0;                 // parked: hit leaf slots (bits 0-7) | leaf-slot mask << 8; first packed triangle
    // Traversal stack: the first WF_SSTACK levels live in SHARED memory ([level][thread]: conflict-free), deeper ones in
    // local memory.  With the whole stack in local memory the pop was the largest single source of long-scoreboard
    // stalls of the kernel (24 % of them, ncu source view: 28 warps x 32 lanes x 192 B of stack per SM do not stay in
    // L1 next to the node and triangle lines); a wide-node walk pushes at most one entry per level, so 12 levels
    // cover practically every ray of an 8-wide tree.  (Measured and dropped: copying the NEXT node global -> shared
    // with cp.async right after each visit, so its latency runs under the votes and the triangle phase: the extra
    // divergent advance step cost 14 % more warp instructions and the smaller L1 10 points of hit rate, extend +24 %.)
    __shared__ uint2 s_stack[WF_SSTACK][128];
    uint2 stack[WIDE_STACK - WF_SSTACK];
    int sp = 0;
    unsigned dbg_nodes = 0;

    while (true) {
        // ---- lanes whose walk is over write their result back and free the lane ----
        if (has && tm == 0u && (pk & 0xffu) == 0u && !(G.y & 0xff000000u) && sp == 0) {
#ifdef WF_DEBUG_LONG_WALKS
            if (COUNT && dbg_nodes > WF_DEBUG_LONG_WALKS) {
                const float4 d4 = P.pool.e1[slot];
                printf("long walk: %u nodes o=(%g %g %g) d=(%g %g %g) id=(%g %g %g) closest=%g hit_t=%d amb=%d\n", dbg_nodes, ro.x, ro.y, ro.z,
                       d4.x, d4.y, d4.z, wr.idx, wr.idy, wr.idz, closest, (int)hit_t, (int)ambiguous);
            }
#endif
            const uint32_t tag0 = __float_as_uint(P.pool.x[slot].x);      // the slot as shade left it (kind / material of the sphere-quad hit)
            if (ambiguous | tie) {
                // equal-distance triangles or a stack overflow: the reference-order walk decides.  The slot
                // still holds the ray and its sphere/quad result; it is only MARKED here and the shade stage
                // resolves it (inline, one such ray stalled the other 31 lanes of this warp for the length
                // of a latency-bound binary walk; in shade the other warps of the SM hide that latency).
                // A tie carries the closest distance along, an overflow of the wide stack does not.
                if (!ambiguous) reinterpret_cast<float*>(&P.pool.x[slot])[2] = closest;
                reinterpret_cast<uint32_t*>(&P.pool.x[slot])[0] = (tag0 & ~0xf0u) | ((uint32_t)(ambiguous ? WF_CLS_OVERFLOW : WF_CLS_TIE) << 4);
                {   // note the shade block of this slot (first marker of a block appends it to the front list)
                    const uint32_t blk = slot / P.shade_threads, bit = 1u << (blk & 31u);
                    if (!(atomicOr(&P.tie_bitmap[blk >> 5], bit) & bit)) {
                        const uint32_t at = atomicAdd(&ctl->n_tie_blocks, 1u);
                        if (at < WF_FRONT_BLOCKS) P.tie_blocks[at] = blk;
                    }
                }
            } else {
                const bool hit_a = wf_tag_kind(tag0) != HIT_NONE;
                const bool hit = hit_a | hit_t;
                const int mat = hit_t ? h_mat : wf_tag_mat(tag0);
                const int cls = hit ? (int)__ldg(&S.materials[mat].material_type) : WF_CLS_MISS;
                if (hit_t) {
                    reinterpret_cast<float*>(&P.pool.e0[slot])[3] = closest;
                    P.pool.x[slot] = make_float4(__uint_as_float(wf_tag(HIT_TRI, cls, mat)), __int_as_float(h_index), h_uu, h_vv);
                } else {
                    reinterpret_cast<uint32_t*>(&P.pool.x[slot])[0] = (tag0 & ~0xf0u) | ((uint32_t)cls << 4);   // the sphere/quad hit (or the miss) stands
                }
                if (P.prim_type) {
                    const float4 e1v = P.pool.e1[slot], s1v = P.pool.s1[slot];
                    if ((__float_as_uint(e1v.w) & WF_CTR_FIRST_QUERY_MASK) == 0u) {                // first query of the item: the primary ray
                        HitLite h;
                        h.kind = hit_t ? (int)HIT_TRI : wf_tag_kind(tag0); h.mat_id = mat; h.uu = h_uu; h.vv = h_vv;
                        h.index = hit_t ? h_index : __float_as_int(P.pool.x[slot].y);
                        h.t = hit_t ? closest : P.pool.e0[slot].w;
                        write_probe(S, P, __float_as_uint(s1v.w), hit, h);
                    }
                }
            }
            has = false;
        }
        // ---- dynamic ray fetch: every lane without a ray takes the next queue entry ----
        if (!exhausted) {
            const uint32_t m = __ballot_sync(0xffffffffu, !has);
            if (m) {
                const int leader = __ffs((int)m) - 1;
                const uint32_t want = (uint32_t)__popc(m);
                uint32_t base = 0;
                if ((int)lane == leader) base = atomicAdd(&ctl->extend_cursor, want);
                base = __shfl_sync(0xffffffffu, base, leader);
                if (!has) {
                    const uint32_t idx = base + (uint32_t)__popc(m & lanemask_lt());
                    if (idx < n) {
                        slot = P.q_extend[idx];
                        const float4 o4 = P.pool.e0[slot], d4 = P.pool.e1[slot];
                        const uint32_t tag0 = __float_as_uint(P.pool.x[slot].x);
                        Ray ray; ray.o = V3(o4.x, o4.y, o4.z); ray.d = V3(d4.x, d4.y, d4.z);
                        ro = ray.o;
                        // sphere/quad result carried in by shade: only its distance matters here
                        closest = wf_tag_kind(tag0) != HIT_NONE ? o4.w : WGPT_MAX_FLOAT;
                        oc.reset();
#ifndef WF_NO_OBJRAY_AT_FETCH
                        // One mesh in the scene: its object-space ray (common.wgsl:193-197) is set up here, where at least
                        // a third of the warp fetches together; done lazily at the first triangle test it ran ~4 lanes wide.
                        if (S.NUM_MESHES == 1) oc.fetch(S, 0, ray);
#endif
                        hit_t = false; ambiguous = false; tie = false; sp = 0; tm = 0; pk = 0; dbg_nodes = 0;
                        bool ray_ok;
                        wr = make_wide_ray_scaled(ray, WB.scale_k, ray_ok);
                        G = make_uint2(0u, 0x80000000u);
                        if (WB.force_reference_walk || !ray_ok) { ambiguous = true; G.y = 0u; }     // marked at once: the reference walk decides
                        has = true;
                    }
                }
                if (base + want >= n) exhausted = true;
            }
        }
        if (__ballot_sync(0xffffffffu, has) == 0) break;        // queue drained and every lane finished

        while (true) {
            if (has && tm == 0u && !(G.y & 0xff000000u) && sp > 0) {
                --sp;
                G = sp < WF_SSTACK ? s_stack[sp][threadIdx.x] : stack[sp - WF_SSTACK];
            }
            // ---- node phase: one wide-node visit for every lane with a node pending and no parked triangles ----
            const bool node_work = has && tm == 0u && (G.y & 0xff000000u);
            if (node_work) {
                const uint32_t bit = 31u - (uint32_t)__clz((int)(G.y & 0xff000000u));
                const uint32_t hits_imask = G.y;
                G.y &= ~(1u << bit);
                bool overflow = false;
                if (G.y & 0xff000000u) {
                    if (sp >= WIDE_STACK) overflow = true;          // deeper than the wide stack: reference walk decides
                    else { if (sp < WF_SSTACK) s_stack[sp][threadIdx.x] = G; else stack[sp - WF_SSTACK] = G; sp++; }
                }
                if (overflow) { ambiguous = true; G.y = 0u; sp = 0; pk = 0u; }
                else {
                    const uint32_t sl = (bit - 24u) ^ wr.octinv;
                    const uint32_t rel = __popc(hits_imask & ~(0xffffffffu << sl) & 0xffu);
                    uint32_t child_base, imask, lmask;
                    if (COUNT) { c.nodes++; dbg_nodes++; }
                    Ray rr; rr.o = ro; rr.d = ro;                  // (wide_node_hits only reads the origin)
                    const uint32_t hits = wide_node_hits(WB.nodes, G.x + rel, rr, wr, closest, child_base, tri_base, imask, lmask);
                    G = make_uint2(child_base, ((uint32_t)s_lut[(wr.octinv << 8) | (hits & imask)] << 24) | imask);
                    tm = hits & lmask; tri_lmask = lmask;
                    if (tm != 0u && (pk & 0xffu) == 0u) { pk = tm | (lmask << 8); ptri_base = tri_base; tm = 0u; }
                }
            }
            // ---- triangle phase: one parked triangle per lane, once enough lanes have one ----
            const uint32_t m_tri = __ballot_sync(0xffffffffu, has && (pk & 0xffu) != 0u);
            const uint32_t m_node = __ballot_sync(0xffffffffu, has && tm == 0u && ((G.y & 0xff000000u) || sp > 0));
            const int n_tri = __popc(m_tri), n_node = __popc(m_node);
            if (n_tri >= WF_TRI_BATCH || (n_tri > 0 && n_tri * WF_TRI_ADAPT >= n_node)) {
                if (has && (pk & 0xffu) != 0u) {
                    const int b = __ffs((int)pk) - 1;
                    const float4* t = &WB.tris[3 * (size_t)(ptri_base + (uint32_t)__popc((pk >> 8) & ~(0xffffffffu << b)))];
                    pk &= pk - 1u;
                    if ((pk & 0xffu) == 0u) { pk = tm | (tri_lmask << 8); ptri_base = tri_base; tm = 0u; }     // next set (possibly empty)
                    const float4 A = __ldg(&t[0]), B = __ldg(&t[1]), C = __ldg(&t[2]);
                    if (__float_as_int(B.w) != oc.mesh) {           // another mesh than the cached one: its object-space ray
                        const float4 d4 = P.pool.e1[slot];
                        Ray ray; ray.o = ro; ray.d = V3(d4.x, d4.y, d4.z);
                        oc.fetch(S, __float_as_int(B.w), ray);
                    }
                    float dst, uu, vv;
                    if (COUNT) c.tris++;
                    if (tri_test(V3(A.x, A.y, A.z), V3(B.x, B.y, B.z), V3(C.x, C.y, C.z), oc.oray, WGPT_RAY_TMIN, closest, dst, uu, vv)) {
                        tie = hit_t && dst == closest;                      // equal-distance triangles: visit order decides (resolved
                                                                            // by the shade stage); a strictly closer hit clears it
                        hit_t = true; closest = dst;
                        h_index = __float_as_int(A.w); h_uu = uu; h_vv = vv;
                        h_mat = oc.m.material_id;
                    }
                }
            }
            // a lane whose walk is over idles until the warp refills; its result is written back then, together
            // with the other finished lanes (done one by one as each lane finished, that code ran 2 lanes wide)
            // lanes with work left, as of BEFORE the triangle phase (a lane blocked on tm != 0 also has a parked set, so the
            // two masks cover it): the triangle phase only removes work, so this never undercounts; at worst the warp runs
            // one empty pass before it refills, which is cheaper than a third vote in every pass
            const int active = __popc(m_tri | m_node);
            if (active == 0) break;
            if (!exhausted && active < WF_REFILL_BELOW) break;
        }
    }

    if (COUNT) {
        unsigned v[8] = {c.rays, c.nodes, c.tris, c.spheres, c.quads, c.rng, c.samples, c.aborts};
        #pragma unroll
        for (int k = 0; k < 8; k++) {
            unsigned long long x = v[k];
            for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
            if (lane == 0 && x) atomicAdd(&S.counters[k], x);
        }
    }
}

// End of an iteration, after shade: publish the next ray queue, find the frames that completed, advance the ring window.
// One warp; every lane reads what it needs up front (one round of loads instead of a chain of ~40 by one thread: this
// launch sits between two big ones on the stream, its length is pure bubble).
// (Measured and dropped: running this inside the shade launch, in the last block to leave it -- the arrival counter is one
// more atomic per block on a single L2 address, 33 K per launch: shade +10..15 %, against ~2.5 us for this launch.)
__global__ void k_wf_ctl_end(WfParams P) {
    if (blockIdx.x != 0 || threadIdx.x >= 32) return;
    const uint32_t lane = threadIdx.x;
    WfControl* c = P.ctl;
    const uint32_t ntb = c->n_tie_blocks, n_next = c->n_next, n_frames = c->n_frames, acc_done0 = c->acc_done;
    const uint32_t drain0 = c->drain, n_live_next = c->n_live_next, live_read = c->live_read;
    const unsigned long long total_work = c->total_work;
    uint32_t acc_k = c->acc_k;
    const unsigned long long dealt = lane < WF_WORK_SHARDS ? P.work_shards[16 * lane] : 0ull;
    {   // the marks of this iteration have been consumed by the shade launch: clear them
        if (ntb <= WF_FRONT_BLOCKS) {
            for (uint32_t i = lane; i < ntb; i += 32u) P.tie_bitmap[P.tie_blocks[i] >> 5] = 0u;
        } else {
            const uint32_t words = P.pool.capacity / WF_SHADE_THREADS / 32u + 1u;
            for (uint32_t i = lane; i < words; i += 32u) P.tie_bitmap[i] = 0u;
        }
    }
    // frames whose every item finished (in frame order), 32 candidates per round; they are FOLDED by the accumulate pass,
    // which the host launches behind this step every few iterations (P.do_acc): until then they wait in the staging ring
    while (true) {
        const uint32_t k = acc_k + lane;
        const bool ok = k < n_frames && k - acc_done0 < P.ring && c->done[k % P.ring] == P.items_per_frame;
        const uint32_t m = __ballot_sync(0xffffffffu, ok);
        const uint32_t take = m == 0xffffffffu ? 32u : (uint32_t)__ffs((int)~m) - 1u;      // leading complete frames
        if (lane < take) c->done[k % P.ring] = 0;
        acc_k += take;
        if (take < 32u) break;
    }
    const uint32_t acc_done = P.do_acc ? acc_k : acc_done0;
    const unsigned long long lim = (unsigned long long)(acc_done + P.ring) * P.items_per_frame;
    const unsigned long long work_limit = lim < total_work ? lim : total_work;
    bool all = true;
    if (lane < WF_WORK_SHARDS) {
        c->shard_lim[lane] = wf_shard_count(work_limit, lane);
        all = dealt >= wf_shard_count(total_work, lane);          // all work handed out <=> every shard has dealt its last item
    }
    all = __all_sync(0xffffffffu, all);
    if (lane != 0) return;
    c->n_tie_blocks = 0;
    c->n_extend = n_next; c->n_next = 0; c->extend_cursor = 0;
    if (c->iterations + 1u < 64u) c->hist_n_extend[c->iterations + 1u] = n_next;
    c->acc_k = acc_k;
    if (P.do_acc) { c->ready_first = acc_done0; c->n_ready = acc_k - acc_done0; c->acc_done = acc_k; }
    else c->n_ready = 0;
    c->work_limit = work_limit;
    if (acc_done == n_frames) c->finished = 1;
    // drain bookkeeping: the shade launch that just ran listed its live slots iff it saw drain == 1
    c->n_live = n_live_next; c->n_live_next = 0;
    if (drain0) c->live_read = live_read ^ 1u;
    c->sparse = (drain0 && (unsigned long long)n_live_next * WF_SPARSE_BELOW <= P.pool.capacity) ? 1u : 0u;
    if (all) c->drain = 1;
    c->iterations++;
}

// ---------------------------------------------------------------- shade (+ generate + compact)
__device__ __forceinline__ void start_sample(const DeviceScene& S, uint32_t pixel, uint32_t scnt, int side,
                                             uint32_t& rng, Ray& ray) {
    const PixelSetup ps = pixel_setup(S, pixel);
    const int si = S.K.STRATIFY ? (int)(scnt / (uint32_t)side) : 0;
    const int sj = S.K.STRATIFY ? (int)(scnt % (uint32_t)side) : 0;
    ray = sample_ray(S, ps, si, sj, rng);
}

#ifndef WF_SHADE_THREADS_MIXED
#define WF_SHADE_THREADS_MIXED 512     // shade block of scenes with several shade classes (WfParams::shade_threads)
#endif
#ifndef WF_SHADE_MINB
#define WF_SHADE_MINB 4
#endif
// SPEC: what the launcher knows about the scene, so that code the scene cannot reach is not in the kernel at all
// (the generic kernel is 170 KB of instructions; 9 % of its issue stalls were instruction fetches): -1 = nothing
// (everything decided at run time), else bit 0 = binding 1 holds spheres, bit 1 = IMPORTANCE_SAMPLING is on.
template <bool COUNT, int SPEC, int THREADS, bool TOL>
__global__ void __launch_bounds__(THREADS, WF_SHADE_MINB * 256 / THREADS) k_wf_shade(DeviceScene S, WfParams P) {
    constexpr bool SPH = SPEC < 0 || (SPEC & 1) != 0;
    constexpr int IS = SPEC < 0 ? -1 : ((SPEC >> 1) & 1);
    __shared__ uint32_t s_done[WF_MAX_RING];
    __shared__ uint32_t s_key_cnt[32];                 // slots per shade key
    __shared__ float4 s_x[THREADS];            // (tag, hit index, u, v) of the block's slots
    __shared__ uint32_t s_perm[THREADS];
    __shared__ uint32_t s_q[THREADS];          // slot ids for the next ray queue
    __shared__ uint32_t s_slot[THREADS];       // sparse launches: the slot of list entry (block base + i)
    __shared__ uint32_t s_live[THREADS];       // drain phase: slot ids for the next live list
    __shared__ uint32_t s_qn, s_arrived, s_ln;
    WfControl* ctl = P.ctl;
    if (ctl->finished) return;                             // the host notices the flag one batch late: those launches are empty
    // drain phase (WfControl): list the slots that stay alive; sparse launch: shade the previous list, not every slot
    const bool drain = ctl->drain != 0u, sparse = ctl->sparse != 0u;
    const uint32_t n_live = ctl->n_live;
    const uint32_t* __restrict__ live_in = P.q_live[ctl->live_read & 1u];
    uint32_t* __restrict__ live_out = P.q_live[(ctl->live_read & 1u) ^ 1u];
    // which 256 slots (or list entries) this block takes: the leading WF_FRONT_BLOCKS blocks run the blocks that hold
    // marked slots (wavefront.cuh, tie_bitmap), everyone else its natural position unless a leading block has it
    uint32_t nat;
    {
        const uint32_t ntb = ctl->n_tie_blocks;
        const bool front = !sparse && ntb > 0u && ntb <= WF_FRONT_BLOCKS;
        if (blockIdx.x < WF_FRONT_BLOCKS) {
            if (!front || blockIdx.x >= ntb) return;
            nat = P.tie_blocks[blockIdx.x];
        } else {
            nat = blockIdx.x - WF_FRONT_BLOCKS;
            if (front && ((P.tie_bitmap[nat >> 5] >> (nat & 31u)) & 1u)) return;
        }
    }
    if (sparse && nat * blockDim.x >= n_live) return;
    const uint32_t work_shard = nat % WF_WORK_SHARDS;
    const uint32_t lane = threadIdx.x & 31u;
    for (uint32_t k = threadIdx.x; k < WF_MAX_RING; k += blockDim.x) s_done[k] = 0;
    if (threadIdx.x == 0) { s_qn = 0; s_arrived = 0; s_ln = 0; }

    // ---- material sort inside the block: slot (block base + j) -> thread rank by shade key ----
    // key = (material class, hit kind) for hits, then miss / idle / dead / trace: warps end up
    // shading one material AND one primitive type (finalize_hit differs per kind).  Counting sort
    // with warp-aggregated shared atomics; the order inside a key does not matter.
    const uint32_t block_base = nat * blockDim.x;
    if (threadIdx.x < 32) s_key_cnt[threadIdx.x] = 0;
    {
        uint32_t s0 = block_base + threadIdx.x;
        if (sparse) { s0 = s0 < n_live ? live_in[s0] : 0xffffffffu; s_slot[threadIdx.x] = s0; }
        float4 xv0 = make_float4(__uint_as_float(wf_tag(HIT_NONE, WF_CLS_DEAD, -1)), 0.0f, 0.0f, 0.0f);
        if (s0 < P.pool.capacity) xv0 = P.pool.x[s0];
        s_x[threadIdx.x] = xv0;
        const int cls0 = wf_tag_cls(__float_as_uint(xv0.x)), kind0 = wf_tag_kind(__float_as_uint(xv0.x));
        const uint32_t key = cls0 < WF_CLS_MISS ? (uint32_t)(cls0 * 4 + (kind0 & 3)) : (uint32_t)(12 + cls0);
        const uint32_t peers = __match_any_sync(0xffffffffu, key);
        const int leader = __ffs((int)peers) - 1;
        // (the barrier the sort needs anyway) a block whose slots are all dead has nothing to shade, start or publish:
        // the last iterations of a render launch 32 K such blocks each
        if (!__syncthreads_or(cls0 != WF_CLS_DEAD)) return;
        uint32_t warp_base = 0;
        if ((int)lane == leader) warp_base = atomicAdd(&s_key_cnt[key], (uint32_t)__popc(peers));
        warp_base = __shfl_sync(0xffffffffu, warp_base, leader);
        __syncthreads();
        // exclusive scan over the 32 key counters, redundantly per warp (no extra barrier)
        const uint32_t mine = s_key_cnt[lane];
        uint32_t incl = mine;
        #pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if ((int)lane >= o) incl += t; }
        const uint32_t key_off = __shfl_sync(0xffffffffu, incl - mine, (int)key);
        s_perm[key_off + warp_base + (uint32_t)__popc(peers & lanemask_lt())] = threadIdx.x;
        __syncthreads();
    }
    const uint32_t j = s_perm[threadIdx.x];
    const uint32_t slot = sparse ? s_slot[j] : block_base + j;
    const float4 xv = s_x[j];
    // (hit kind, hit index, material id, shade class) of the slot
    int4 hid = make_int4(wf_tag_kind(__float_as_uint(xv.x)), __float_as_int(xv.y), wf_tag_mat(__float_as_uint(xv.x)), wf_tag_cls(__float_as_uint(xv.x)));
    int cls = hid.w;
    // The slot's records and -- the hit is known from the tag -- the triangle's normals are requested NOW,
    // in one batch of independent loads, so that shading waits for one memory round trip instead of a
    // chain of three.  (Also preloading the material and the pixel sum cost more in registers than the
    // shorter chain gave back: measured.)
    float4 e0v = make_float4(0.0f, 0.0f, 0.0f, 0.0f), e1v = e0v, s0v = e0v, s1v = e0v;
    TriNormals tn; tn.a = tn.b = tn.c = e0v;
    if (cls <= WF_CLS_MISS || cls >= WF_CLS_TIE) {
        e0v = P.pool.e0[slot]; e1v = P.pool.e1[slot]; s0v = P.pool.s0[slot]; s1v = P.pool.s1[slot];
        if (cls < WF_CLS_MISS) {
            if (hid.x == HIT_TRI) tn = load_tri_normals(S, hid.y);
        }
    }
    float4 h4_resolved = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    bool resolved = false;
    if (cls >= WF_CLS_TIE) {
        // ---- tie resolver: closest hit in reference order (hitRay.wgsl:42-110) for the rays extend marked ----
        const float4 o4 = e0v;
        Ray r; r.o = V3(o4.x, o4.y, o4.z); r.d = V3(e1v.x, e1v.y, e1v.z);
        HitLite h; h.t = o4.w; h.uu = 0.0f; h.vv = 0.0f; h.kind = hid.x; h.index = hid.y; h.mat_id = hid.z;
        const bool hit_a = hid.x != HIT_NONE;
        float closest = hit_a ? o4.w : WGPT_MAX_FLOAT;
        ObjRayCache oc; oc.reset();
        // Tie with known closest distance D = h4.y: every triangle at distance D lies in nodes whose slab
        // interval reaches D, and the reference reaches those nodes in the same order whatever else it
        // visits, so the walk only needs the nodes overlapping [D(1 - 1e-4), closest].  (Not when the
        // reference's stack guard could fire: its abort depends on everything visited.)
        const bool restricted = cls == WF_CLS_TIE && P.bvh_depth < S.K.STACK_SIZE && P.bvh_depth < WGPT_MAX_STACK;
        const float box_tmin = restricted ? fmaxf(WGPT_RAY_TMIN, xv.z * 0.9999f) : WGPT_RAY_TMIN;
        const bool hit = hit_bvh_binary(S, r, nullptr, oc, closest, h, box_tmin) | hit_a;
        cls = hit ? (int)__ldg(&S.materials[h.mat_id].material_type) : WF_CLS_MISS;
        hid = make_int4(hit ? h.kind : (int)HIT_NONE, h.index, h.mat_id, cls);
        h4_resolved = make_float4(h.t, h.uu, h.vv, 1.0f);
        resolved = true;
        if (hit) {                                              // the batch above was fetched for the unresolved hit
            if (hid.x == HIT_TRI) tn = load_tri_normals(S, hid.y);
        }
        atomicAdd(&ctl->fallbacks, 1u);
        if (P.prim_type) {
            if ((__float_as_uint(e1v.w) & WF_CTR_FIRST_QUERY_MASK) == 0u) write_probe(S, P, __float_as_uint(s1v.w), hit, h);
        }
    }

    const uint32_t npix = (uint32_t)S.W * (uint32_t)S.H;
    const int side = S.K.STRATIFY ? strat_side(S) : 1;
    const uint32_t samples_per_item = S.K.STRATIFY ? (uint32_t)(side * side) : (uint32_t)S.K.NUM_SAMPLES;

    bool alive = false, need_work = false, store_pix = false, gen = false;
    uint32_t rng = 0, pix_i = 0, ring_slot = 0, frame_k = 0, bounce = 0, scnt = 0, gen_pixel = 0;
    unsigned n_samples = 0, n_draws = 0;
    Counts cnt = {0, 0, 0, 0, 0, 0, 0, 0};
    Ray ray; ray.o = ray.d = V3s(0.0f);
    v3 thr = V3s(1.0f), acc = V3s(0.0f);
    float4 pix = make_float4(0.0f, 0.0f, 0.0f, 0.0f);

    // probes only: hash of the path's hit sequence so far (kind, index of every closest hit; a miss too).  Kept in its
    // per-slot array, not in a register: a value live across the whole kernel cost 0.5 % of shade with no probe running.
    if (P.path_sig && cls <= WF_CLS_MISS) P.path_sig[slot] = (P.path_sig[slot] ^ ((uint32_t)(hid.x + 1) << 28) ^ (uint32_t)hid.y) * 0x9E3779B1u;
    if (cls == WF_CLS_IDLE) {
        need_work = true;
    } else if (cls <= WF_CLS_MISS) {
        bool ended = false;
        const uint32_t ctr = __float_as_uint(e1v.w);
        rng = __float_as_uint(s1v.z); pix_i = __float_as_uint(s1v.w);
        ring_slot = ctr >> 24; bounce = (ctr >> 14) & 0x3ffu; scnt = ctr & 0x3fffu;
        thr = V3(s0v.x, s0v.y, s0v.z); acc = V3(s0v.w, s1v.x, s1v.y);
        if (cls == WF_CLS_MISS) {                             // traceRay.wgsl:12-16
            acc = acc + V3(0.0f, 1.0f, 1.0f) * thr;
            ended = true;
        } else {
            const float4 o4 = e0v;
            ray.o = V3(o4.x, o4.y, o4.z); ray.d = V3(e1v.x, e1v.y, e1v.z);
            HitLite h; h.t = o4.w; h.uu = xv.z; h.vv = xv.w; h.kind = hid.x; h.index = hid.y; h.mat_id = hid.z;
            if (resolved) { h.t = h4_resolved.x; h.uu = h4_resolved.y; h.vv = h4_resolved.z; h.index = hid.y; }
            // the shading of the hit: pinned arithmetic, or (WGPT_MODE_FAST_TOL) the same code on fast arithmetic (shade.cuh)
            const bool cont = shade_bounce<SPH, IS, TOL>(S, h, tn, (int)bounce, ray, rng, n_draws, acc, thr);
            bounce++;
            ended = !cont || (int)bounce >= S.K.MAX_BOUNCES;           // `for (i < MAX_BOUNCES)`
            alive = !ended;
        }
        if (ended) {
            // the sample's ray_color() value is `acc`: shootRay.wgsl:25,42 `pixColor += ray_color(ray)`
            pix = P.pool.pix[slot];
            pix.x = fa(pix.x, acc.x); pix.y = fa(pix.y, acc.y); pix.z = fa(pix.z, acc.z); pix.w += 1.0f;
            scnt++;
            const uint32_t pixel = shard_pixel(S.shard, S.W, pix_i);
            if (scnt < samples_per_item) {
                gen = true; gen_pixel = pixel;                             // next sample, same RNG stream
            } else {
                const float div = S.K.STRATIFY ? pix.w : (float)S.K.NUM_SAMPLES;
                // x / 1 is x (HEAD has NUM_SAMPLES = 1): no need to run three IEEE divisions for it
                const v3 mean = div == 1.0f ? V3(pix.x, pix.y, pix.z) : V3(pix.x, pix.y, pix.z) / div;
                P.stage[(size_t)ring_slot * npix + pixel] = make_float4(mean.x, mean.y, mean.z, 1.0f);
                if (P.prim_rng) {                                              // path signature: draws made + hits seen
                    P.prim_rng[pixel] = rng ^ ((P.path_sig ? P.path_sig[slot] : 0u) * 0x85EBCA6Bu);
                    if (P.path_sig) P.path_sig[slot] = 0u;                 // the slot's next path starts a new hash
                }
                atomicAdd(&s_done[ring_slot], 1u);
                need_work = true;
            }
        }
    }

    // ---- generate: pull the next (frame, pixel) work items, warp-aggregated ----
    int new_cls = WF_CLS_DEAD;
    {
        const uint32_t m = __ballot_sync(0xffffffffu, need_work);
        if (m) {
            const int leader = __ffs((int)m) - 1;
            const uint32_t want = (uint32_t)__popc(m);
            unsigned long long old = 0; uint32_t granted = 0, seg_shard = work_shard; int more_later = 0;
            if ((int)lane == leader) {
                // the block's own shard first, then the others in turn (a pool with fewer blocks than shards, and the end
                // of a window, leave items in foreign shards); a plain load tells whether a foreign shard has anything
                // below its limit, so a starved warp does not pay 16 atomics.  One segment per claim: lanes beyond it
                // stay idle and ask again in the next iteration.
                for (uint32_t k = 0; k < WF_WORK_SHARDS && granted == 0u; k++) {
                    const uint32_t c = (work_shard + k) % WF_WORK_SHARDS;
                    unsigned long long* ctr = &P.work_shards[16u * c];
                    const unsigned long long lim = ctl->shard_lim[c];
                    if (k > 0u && *(volatile unsigned long long*)ctr >= lim) continue;
                    const unsigned long long o = atomicAdd(ctr, (unsigned long long)want);
                    const uint32_t g = o >= lim ? 0u : (uint32_t)((lim - o) < want ? (lim - o) : want);
                    if (g < want) atomicAdd(ctr, (unsigned long long)0 - (unsigned long long)(want - g));
                    if (g) { old = o; granted = g; seg_shard = c; }
                }
                // a lane left without an item waits (idle) if a later window advance can release work or if this claim
                // stopped at its first segment; it is dead only when every shard was found empty at the end of the work
                more_later = (granted != 0u || ctl->work_limit < ctl->total_work) ? 1 : 0;
            }
            old = __shfl_sync(0xffffffffu, old, leader);
            granted = __shfl_sync(0xffffffffu, granted, leader);
            seg_shard = __shfl_sync(0xffffffffu, seg_shard, leader);
            more_later = __shfl_sync(0xffffffffu, more_later, leader);
            if (need_work) {
                const uint32_t r = (uint32_t)__popc(m & lanemask_lt());
                new_cls = more_later ? WF_CLS_IDLE : WF_CLS_DEAD;
                if (r < granted) {
                    // (frame ordinal, pixel) of the shard's item `old + r` without a 64-bit division: quotient
                    // estimate in binary64 (exact to +-1 for g < 2^52), fixed up with integers
                    const unsigned long long g = wf_shard_item(old + r, seg_shard);
                    uint32_t fk0 = (uint32_t)__double2ull_rz(__ull2double_rz(g) * P.inv_items_per_frame);
                    long long rem = (long long)(g - (unsigned long long)fk0 * P.items_per_frame);
                    if (rem < 0) { fk0--; rem += P.items_per_frame; }
                    while (rem >= (long long)P.items_per_frame) { fk0++; rem -= P.items_per_frame; }
                    frame_k = fk0;
                    ring_slot = fk0 % P.ring;
                    pix_i = (uint32_t)rem;
                    const uint32_t pixel = shard_pixel(S.shard, S.W, pix_i);
                    const uint32_t lim_pix = npix < S.shard.pixel_limit ? npix : S.shard.pixel_limit;
                    if (pixel < lim_pix) {
                        const float frame = (float)(P.frame0 + (int)frame_k * P.frame_stride);
                        rng = pixel_seed(pixel, frame);
                        scnt = 0;
                        pix = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                        gen = true; gen_pixel = pixel;
                    } else {
                        atomicAdd(&s_done[ring_slot], 1u);             // tile tail past the image: nothing to render
                        new_cls = WF_CLS_IDLE;                         // try again next iteration
                    }
                }
            }
        }
    }

    // ---- camera ray of a new sample: next sample of the same item, or first sample of a new item ----
    if (gen) {
        start_sample(S, gen_pixel, scnt, side, rng, ray);
        n_draws += 2; n_samples++;
        thr = V3s(1.0f); acc = V3s(0.0f); bounce = 0;
        alive = true; store_pix = true;
    }

    // ---- the new ray's sphere/quad part of hitScene (hitRay.wgsl:6-40) runs here, where all lanes are
    // in lock step; only rays that can still reach a triangle (root box within reach) go to extend ----
    bool trace = false;
    if (alive) {
        HitLite h; h.t = 0.0f; h.kind = HIT_NONE; h.index = -1; h.uu = h.vv = 0.0f; h.mat_id = -1;
        float closest = WGPT_MAX_FLOAT;
        if (COUNT) cnt.rays++;
        const bool hit_a = hit_analytic<SPH>(S, ray, rng, COUNT ? &cnt : nullptr, closest, h);
        trace = S.NUM_AABB > 0 && root_box_hit(S, ray, closest);
        const int ncls = trace ? WF_CLS_TRACE : (hit_a ? (int)__ldg(&S.materials[h.mat_id].material_type) : WF_CLS_MISS);
        const uint32_t ctr = wf_ctr(ring_slot, bounce, scnt);
        P.pool.e0[slot] = make_float4(ray.o.x, ray.o.y, ray.o.z, h.t);
        P.pool.e1[slot] = make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float(ctr));
        P.pool.s0[slot] = make_float4(thr.x, thr.y, thr.z, acc.x);
        P.pool.s1[slot] = make_float4(acc.y, acc.z, __uint_as_float(rng), __uint_as_float(pix_i));
        P.pool.x[slot] = make_float4(__uint_as_float(wf_tag(hit_a ? h.kind : (int)HIT_NONE, ncls, h.mat_id)), __int_as_float(h.index), 0.0f, 0.0f);
        if (store_pix) P.pool.pix[slot] = pix;
        if (!trace && P.prim_type && (ctr & WF_CTR_FIRST_QUERY_MASK) == 0u) write_probe(S, P, pix_i, hit_a, h);
    } else if (cls <= WF_CLS_IDLE && new_cls != cls) {
        reinterpret_cast<uint32_t*>(&P.pool.x[slot])[0] = wf_tag(HIT_NONE, new_cls, -1);    // idle (waiting for the window) or dead (no work left)
    }

    // ---- compact: rays that need the BVH into the next ray queue, one global atomic per block ----
    // No block barrier here: warps finish at very different times (a miss warp is done long before
    // a glossy one), so every warp parks its slot ids in shared memory and leaves; the LAST warp
    // to arrive publishes the block's entries and its finished-item counts.
    {
        const uint32_t m = __ballot_sync(0xffffffffu, trace);
        if (m) {
            uint32_t wbase = 0;
            if (lane == 0) wbase = atomicAdd(&s_qn, (uint32_t)__popc(m));
            wbase = __shfl_sync(0xffffffffu, wbase, 0);
            if (trace) s_q[wbase + (uint32_t)__popc(m & lanemask_lt())] = slot;
        }
        if (drain) {
            const uint32_t ml = __ballot_sync(0xffffffffu, alive);
            if (ml) {
                uint32_t lbase = 0;
                if (lane == 0) lbase = atomicAdd(&s_ln, (uint32_t)__popc(ml));
                lbase = __shfl_sync(0xffffffffu, lbase, 0);
                if (alive) s_live[lbase + (uint32_t)__popc(ml & lanemask_lt())] = slot;
            }
        }
        __syncwarp();
        uint32_t last = 0;
        if (lane == 0) { __threadfence_block(); last = atomicAdd(&s_arrived, 1u) == THREADS / 32 - 1 ? 1u : 0u; }
        last = __shfl_sync(0xffffffffu, last, 0);
        if (last) {
            __threadfence_block();
            const uint32_t total = *(volatile uint32_t*)&s_qn;
            uint32_t base = 0;
            if (lane == 0 && total) base = atomicAdd(&ctl->n_next, total);
            base = __shfl_sync(0xffffffffu, base, 0);
            for (uint32_t i = lane; i < total; i += 32) P.q_extend[base + i] = ((volatile uint32_t*)s_q)[i];
            if (drain) {
                const uint32_t nl = *(volatile uint32_t*)&s_ln;
                uint32_t lb = 0;
                if (lane == 0 && nl) lb = atomicAdd(&ctl->n_live_next, nl);
                lb = __shfl_sync(0xffffffffu, lb, 0);
                for (uint32_t i = lane; i < nl; i += 32) live_out[lb + i] = ((volatile uint32_t*)s_live)[i];
            }
            for (uint32_t k = lane; k < P.ring; k += 32) {
                const uint32_t v = ((volatile uint32_t*)s_done)[k];
                if (v) atomicAdd(&ctl->done[k], v);
            }
        }
    }

    if (COUNT) {
        cnt.samples += n_samples; cnt.rng += n_draws;
        unsigned v[8] = {cnt.rays, cnt.nodes, cnt.tris, cnt.spheres, cnt.quads, cnt.rng, cnt.samples, cnt.aborts};
        #pragma unroll
        for (int k = 0; k < 8; k++) {
            unsigned long long x = v[k];
            for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
            if (lane == 0 && x) atomicAdd(&S.counters[k], x);
        }
    }
}

// ---------------------------------------------------------------- accumulate (K4, ring-aware)
__global__ void __launch_bounds__(256) k_wf_accumulate(WfParams P, float4* __restrict__ fb, uint32_t npix, int reset_first) {
    const uint32_t n_ready = P.ctl->n_ready;
    if (n_ready == 0) return;
    const uint32_t first = P.ctl->ready_first;
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < npix; p += gridDim.x * blockDim.x) {
        float4 s = fb[p];
        bool touched = false;
        for (uint32_t r = 0; r < n_ready; r++) {
            const uint32_t k = first + r;
            const float4 c = P.stage[(size_t)(k % P.ring) * npix + p];
            if (c.w == 0.0f) continue;                       // not a pixel of this shard
            if (reset_first && k == 0) s = make_float4(c.x, c.y, c.z, 1.0f);
            else s = make_float4(fa(s.x, c.x), fa(s.y, c.y), fa(s.z, c.z), 1.0f);
            touched = true;
        }
        if (touched) fb[p] = s;
    }
}

// ---------------------------------------------------------------- host side
void wf_launch_init(const WfParams& P, uint32_t n_frames, cudaStream_t stream) {
    k_wf_init<<<(P.pool.capacity + 255) / 256, 256, 0, stream>>>(P, n_frames);
}

void wf_launch_append(const WfParams& P, uint32_t more_frames, cudaStream_t stream) {
    k_wf_append<<<(P.pool.capacity + 255) / 256, 256, 0, stream>>>(P, more_frames);
}

// persistent grid of the extend stage: exactly as many CTAs as fit on the CURRENT device (SM count x resident CTAs per SM)
unsigned wf_extend_grid(bool counters) {
    int dev = 0, sms = 148, per_sm = 4;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (counters) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_wf_extend<true>, 128, 0);
    else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_wf_extend<false>, 128, 0);
    return (unsigned)(sms * (per_sm > 0 ? per_sm : 1));
}

int wf_launch_iteration(const DeviceScene& S, const WideBvh& WB, const WfParams& P_in, float4* fb, int reset_first,
                        unsigned extend_blocks, bool do_acc, cudaStream_t stream, cudaEvent_t* ev) {
    WfParams P = P_in;
    P.do_acc = do_acc ? 1 : 0;
    P.shade_threads = (!S.counters && (S.NUM_SPHERES > 0 || S.K.IMPORTANCE_SAMPLING)) ? WF_SHADE_THREADS_MIXED : 256u;
    if (ev) cudaEventRecord(ev[0], stream);
    if (S.counters) k_wf_extend<true><<<extend_blocks, 128, 0, stream>>>(S, WB, P);
    else k_wf_extend<false><<<extend_blocks, 128, 0, stream>>>(S, WB, P);
    if (ev) cudaEventRecord(ev[1], stream);
    const int spec = (S.NUM_SPHERES > 0 ? 1 : 0) | (S.K.IMPORTANCE_SAMPLING ? 2 : 0);
    const unsigned st = P.shade_threads;
    const unsigned shade_blocks = (P.pool.capacity + st - 1u) / st + WF_FRONT_BLOCKS;
#define WF_SHADE_LAUNCH(C, SP, TH) do { if (P.tol) k_wf_shade<C, SP, TH, true><<<shade_blocks, TH, 0, stream>>>(S, P); \
                                        else k_wf_shade<C, SP, TH, false><<<shade_blocks, TH, 0, stream>>>(S, P); } while (0)
    if (S.counters) WF_SHADE_LAUNCH(true, -1, 256);
    else switch (spec) {
        case 0: WF_SHADE_LAUNCH(false, 0, 256); break;
        case 1: WF_SHADE_LAUNCH(false, 1, WF_SHADE_THREADS_MIXED); break;
        case 2: WF_SHADE_LAUNCH(false, 2, WF_SHADE_THREADS_MIXED); break;
        default: WF_SHADE_LAUNCH(false, 3, WF_SHADE_THREADS_MIXED); break;
    }
#undef WF_SHADE_LAUNCH
    if (ev) cudaEventRecord(ev[2], stream);
    k_wf_ctl_end<<<1, 32, 0, stream>>>(P);
    if (do_acc) {      // a whole-image launch even when no frame is ready: only every few iterations (fast.cu WF_ACC_EVERY)
        const uint32_t npix = (uint32_t)S.W * (uint32_t)S.H;
        unsigned acc_blocks = (npix + 255u) / 256u;
        if (acc_blocks > 148u * 8u) acc_blocks = 148u * 8u;
        k_wf_accumulate<<<acc_blocks, 256, 0, stream>>>(P, fb, npix, reset_first);
    }
    if (ev) cudaEventRecord(ev[3], stream);
    return do_acc ? 4 : 3;
}

// closest-hit hook over the wide BVH (wgpt_trace_rays in fast mode)
template <bool COUNT>
__global__ void __launch_bounds__(128) k_trace_rays_wide(DeviceScene S, WideBvh WB, int n, const float* __restrict__ rays,
                                                         const uint32_t* __restrict__ seeds, int* __restrict__ out_type,
                                                         int* __restrict__ out_index, float* __restrict__ out_t,
                                                         uint32_t* __restrict__ fallbacks_out) {
    __shared__ unsigned char s_lut[WIDE_LUT_BYTES];
    wide_lut_init(s_lut);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    Counts c = {0, 0, 0, 0, 0, 0, 0, 0};
    unsigned fallbacks = 0;
    if (i < n) {
        Ray r; r.o = V3(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]); r.d = V3(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]);
        uint32_t rng = seeds ? seeds[i] : 0u;
        ObjRayCache oc; HitLite h; h.t = 0.0f; h.kind = HIT_NONE; h.index = -1; h.mat_id = -1; h.uu = h.vv = 0.0f;
        const bool hit = hitScene_wide(S, WB, s_lut, r, rng, COUNT ? &c : nullptr, oc, h, &fallbacks);
        out_type[i] = hit ? (h.kind == HIT_VOLUME ? 0 : h.kind) : -1;
        out_index[i] = hit ? (h.kind == HIT_TRI ? (int)S.triangles[h.index].local_id : h.index) : -1;
        out_t[i] = hit ? h.t : 0.0f;
    }
    if (COUNT) {
        unsigned v[8] = {c.rays, c.nodes, c.tris, c.spheres, c.quads, c.rng, c.samples, c.aborts};
        for (int k = 0; k < 8; k++) {
            unsigned long long x = v[k];
            for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
            if ((threadIdx.x & 31) == 0 && x) atomicAdd(&S.counters[k], x);
        }
    }
    if (fallbacks && fallbacks_out) atomicAdd(fallbacks_out, fallbacks);
}

void launch_trace_rays_wide(const DeviceScene& S, const WideBvh& WB, int n, const float* rays, const uint32_t* seeds,
                            int* out_type, int* out_index, float* out_t, uint32_t* fallbacks, cudaStream_t stream) {
    const unsigned threads = 128, blocks = (n + threads - 1) / threads;
    if (blocks == 0) return;
    if (S.counters) k_trace_rays_wide<true><<<blocks, threads, 0, stream>>>(S, WB, n, rays, seeds, out_type, out_index, out_t, fallbacks);
    else k_trace_rays_wide<false><<<blocks, threads, 0, stream>>>(S, WB, n, rays, seeds, out_type, out_index, out_t, fallbacks);
}

}  // namespace wg

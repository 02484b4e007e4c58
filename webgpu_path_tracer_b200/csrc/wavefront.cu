// wavefront.cu -- K3: the persistent-thread wavefront integrator.
//
// Stages per iteration (one bounce of every live path), all sized from the device-resident
// control block, no host round trip:
//   extend   persistent warps pull batches of 32 ray slots from the extend queue, run the
//            closest-hit query (spheres/quads + 8-wide BVH, trace_wide.cuh) and append the slot to
//            the shade queue of its material class with one warp-aggregated atomic per class
//            (ballot + popc compaction) => material-sorted shading;
//   shade    one thread per queued slot, classes laid out back to back so warps are homogeneous:
//            miss -> background; hit -> emission, BSDF sample, mixture pdf against the light quad
//            (the reference's "shadow" term is the analytic light_pdf, importanceSampling.wgsl:88-125:
//            it casts no occlusion ray), Russian roulette; finished samples are folded into the
//            pixel, finished (pixel, frame) items are written to the staging ring, and the slot
//            REGENERATES from the next work item (generate stage) before being compacted
//            (block-level ballot/prefix) into the next extend queue;
//   accumulate  K4: frames whose every item finished are added to binding 3 in frame order.
// The RNG stream of a (pixel, frame) item lives in its slot, so the draw order equals the
// reference's whatever the schedule; all arithmetic is pinned.
#include "wavefront.cuh"
#include "shade.cuh"
#include "camera.cuh"
#include "trace_wide.cuh"

namespace wg {

__device__ __forceinline__ uint32_t lanemask_lt() { uint32_t m; asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m)); return m; }

__global__ void k_wf_init(WfParams P, uint32_t n_frames) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < P.pool.capacity) {
        P.q_extend[i] = i;
        P.pool.hid[i] = make_int4(WF_KIND_IDLE, -1, -1, 0);
    }
    if (i == 0) {
        WfControl* c = P.ctl;
        c->next_work = 0;
        c->total_work = (unsigned long long)n_frames * P.items_per_frame;
        const unsigned long long win = (unsigned long long)P.ring * P.items_per_frame;
        c->work_limit = win < c->total_work ? win : c->total_work;
        c->n_extend = P.pool.capacity; c->extend_cursor = 0;
        for (int k = 0; k < WF_CLASSES; k++) { c->n_shade[k] = 0; c->shade_off[k] = 0; }
        c->shade_off[WF_CLASSES] = 0; c->shade_cursor = 0;
        for (int k = 0; k < WF_MAX_RING; k++) c->done[k] = 0;
        c->acc_k = 0; c->ready_first = 0; c->n_ready = 0; c->n_frames = n_frames;
        c->finished = n_frames == 0 ? 1u : 0u; c->fallbacks = 0; c->iterations = 0;
    }
}

// ---------------------------------------------------------------- extend
template <bool COUNT>
__global__ void __launch_bounds__(128) k_wf_extend(DeviceScene S, WideBvh WB, WfParams P) {
    WfControl* ctl = P.ctl;
    const uint32_t n = ctl->n_extend;
    const uint32_t lane = threadIdx.x & 31u;
    Counts c = {0, 0, 0, 0, 0, 0, 0, 0};
    unsigned fallbacks = 0;
    while (true) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(&ctl->extend_cursor, 32u);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) break;
        const uint32_t idx = base + lane;
        int cls = -1;
        uint32_t slot = 0;
        if (idx < n) {
            slot = P.q_extend[idx];
            const int4 hid_in = P.pool.hid[slot];
            if (hid_in.x == WF_KIND_IDLE) {
                cls = 5;
            } else {
                const float4 o4 = P.pool.ro[slot], d4 = P.pool.rd[slot];
                Ray ray; ray.o = V3(o4.x, o4.y, o4.z); ray.d = V3(d4.x, d4.y, d4.z);
                const uint4 misc = P.pool.misc[slot];
                uint32_t rng = misc.x;
                ObjRayCache oc;
                HitLite h; h.t = 0.0f; h.kind = HIT_NONE; h.index = -1; h.uu = h.vv = 0.0f; h.mat_id = -1;
                const bool hit = hitScene_wide(S, WB, ray, rng, COUNT ? &c : nullptr, oc, h, &fallbacks);
                if (rng != misc.x) P.pool.misc[slot].x = rng;
                P.pool.hit[slot] = make_float4(h.t, h.uu, h.vv, 0.0f);
                P.pool.hid[slot] = make_int4(hit ? h.kind : (int)HIT_NONE, h.index, h.mat_id, 0);
                cls = hit ? (int)__ldg(&S.materials[h.mat_id].material_type) : 4;
                if (P.prim_type && misc.w == 0u) {           // first query of the item: the primary ray
                    const uint32_t p = shard_pixel(S.shard, S.W, misc.y);
                    P.prim_type[p] = hit ? (h.kind == HIT_VOLUME ? 0 : h.kind) : -1;
                    P.prim_index[p] = hit ? (h.kind == HIT_TRI ? (int)S.triangles[h.index].local_id : h.index) : -1;
                    P.prim_t[p] = hit ? h.t : 0.0f;
                }
            }
        }
        // warp-ballot compaction into the per-class shade queues
        #pragma unroll
        for (int k = 0; k < WF_CLASSES; k++) {
            const uint32_t m = __ballot_sync(0xffffffffu, cls == k);
            if (m == 0) continue;
            uint32_t qb = 0;
            const int leader = __ffs((int)m) - 1;
            if ((int)lane == leader) qb = atomicAdd(&ctl->n_shade[k], (uint32_t)__popc(m));
            qb = __shfl_sync(0xffffffffu, qb, leader);
            if (cls == k) P.q_shade[k][qb + __popc(m & lanemask_lt())] = slot;
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
    for (int o = 16; o > 0; o >>= 1) fallbacks += __shfl_down_sync(0xffffffffu, fallbacks, o);
    if (lane == 0 && fallbacks) atomicAdd(&ctl->fallbacks, fallbacks);
}

// between extend and shade: lay the class queues out back to back, reset the extend queue
__global__ void k_wf_ctl_mid(WfParams P) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    WfControl* c = P.ctl;
    uint32_t off = 0;
    for (int k = 0; k < WF_CLASSES; k++) { c->shade_off[k] = off; off += c->n_shade[k]; }
    c->shade_off[WF_CLASSES] = off;
    c->n_extend = 0; c->extend_cursor = 0;
}

// after shade: reset the shade queues, find frames that completed, advance the ring window
__global__ void k_wf_ctl_end(WfParams P) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    WfControl* c = P.ctl;
    for (int k = 0; k < WF_CLASSES; k++) c->n_shade[k] = 0;
    uint32_t n_ready = 0;
    while (c->acc_k + n_ready < c->n_frames && n_ready < P.ring &&
           c->done[(c->acc_k + n_ready) % P.ring] == P.items_per_frame) {
        c->done[(c->acc_k + n_ready) % P.ring] = 0;
        n_ready++;
    }
    c->ready_first = c->acc_k; c->n_ready = n_ready;
    c->acc_k += n_ready;
    unsigned long long lim = (unsigned long long)(c->acc_k + P.ring) * P.items_per_frame;
    c->work_limit = lim < c->total_work ? lim : c->total_work;
    if (c->acc_k == c->n_frames) c->finished = 1;
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

template <bool COUNT>
__global__ void __launch_bounds__(256) k_wf_shade(DeviceScene S, WfParams P) {
    __shared__ uint32_t s_done[WF_MAX_RING];
    __shared__ uint32_t s_warp[8];
    __shared__ uint32_t s_base;
    WfControl* ctl = P.ctl;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    if (threadIdx.x < WF_MAX_RING) s_done[threadIdx.x] = 0;
    __syncthreads();

    const uint32_t total = ctl->shade_off[WF_CLASSES];
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t npix = (uint32_t)S.W * (uint32_t)S.H;
    const int side = S.K.STRATIFY ? strat_side(S) : 1;
    const uint32_t samples_per_item = S.K.STRATIFY ? (uint32_t)(side * side) : (uint32_t)S.K.NUM_SAMPLES;

    bool alive = false, need_work = false, store_pix = false;
    uint32_t slot = 0, rng = 0, pix_i = 0, frame_k = 0, bounce = 0, scnt = 0;
    unsigned n_samples = 0, n_draws = 0;
    Ray ray; ray.o = ray.d = V3s(0.0f);
    v3 thr = V3s(1.0f), acc = V3s(0.0f);
    float4 pix = make_float4(0.0f, 0.0f, 0.0f, 0.0f);

    if (t < total) {
        int cls = 0;
        while (t >= ctl->shade_off[cls + 1]) cls++;
        slot = P.q_shade[cls][t - ctl->shade_off[cls]];
        bool ended = false;
        if (cls == 5) {
            need_work = true;
        } else {
            const uint4 misc = P.pool.misc[slot];
            rng = misc.x; pix_i = misc.y; frame_k = misc.z; bounce = misc.w >> 16; scnt = misc.w & 0xffffu;
            const float4 t4 = P.pool.thr[slot], a4 = P.pool.acc[slot];
            thr = V3(t4.x, t4.y, t4.z); acc = V3(a4.x, a4.y, a4.z);
            if (cls == 4) {                                   // miss: traceRay.wgsl:12-16
                acc = acc + V3(0.0f, 1.0f, 1.0f) * thr;
                ended = true;
            } else {
                const float4 o4 = P.pool.ro[slot], d4 = P.pool.rd[slot], h4 = P.pool.hit[slot];
                const int4 hid = P.pool.hid[slot];
                ray.o = V3(o4.x, o4.y, o4.z); ray.d = V3(d4.x, d4.y, d4.z);
                HitLite h; h.t = h4.x; h.uu = h4.y; h.vv = h4.z; h.kind = hid.x; h.index = hid.y; h.mat_id = hid.z;
                ObjRayCache oc; oc.reset();
                ShadeState st;
                st.p = st.n = st.unit_w = st.u = st.v = V3s(0.0f);
                st.front_face = false; st.skip_pdf = false; st.doSpecular = 0.0f;
                st.skip_ray.o = st.skip_ray.d = V3s(0.0f);
                finalize_hit(S, ray, h, oc, st);
                const Material m = S.materials[h.mat_id];
                bool early;
                const bool cont = bounce_update(S, m, (int)bounce, ray, st, rng, n_draws, acc, thr, early);
                bounce++;
                ended = !cont || (int)bounce >= S.K.MAX_BOUNCES;       // `for (i < MAX_BOUNCES)`
                alive = !ended;
            }
            if (ended) {
                // the sample's ray_color() value is `acc`: shootRay.wgsl:25,42 `pixColor += ray_color(ray)`
                pix = P.pool.pix[slot];
                pix.x = fa(pix.x, acc.x); pix.y = fa(pix.y, acc.y); pix.z = fa(pix.z, acc.z); pix.w += 1.0f;
                scnt++;
                const uint32_t pixel = shard_pixel(S.shard, S.W, pix_i);
                if (scnt < samples_per_item) {
                    start_sample(S, pixel, scnt, side, rng, ray);          // next sample, same RNG stream
                    n_draws += 2; n_samples++;
                    thr = V3s(1.0f); acc = V3s(0.0f); bounce = 0;
                    alive = true; store_pix = true;
                } else {
                    const float div = S.K.STRATIFY ? pix.w : (float)S.K.NUM_SAMPLES;
                    P.stage[(size_t)(frame_k % P.ring) * npix + pixel] = make_float4(fd(pix.x, div), fd(pix.y, div), fd(pix.z, div), 1.0f);
                    atomicAdd(&s_done[frame_k % P.ring], 1u);
                    need_work = true;
                }
            }
        }
    }

    // ---- generate: pull the next (frame, pixel) work items, warp-aggregated ----
    bool requeue_idle = false;
    {
        const uint32_t m = __ballot_sync(0xffffffffu, need_work);
        if (m) {
            const int leader = __ffs((int)m) - 1;
            const uint32_t want = (uint32_t)__popc(m);
            unsigned long long old = 0; uint32_t granted = 0; int more_later = 0;
            if ((int)lane == leader) {
                old = atomicAdd(&ctl->next_work, (unsigned long long)want);
                const unsigned long long lim = ctl->work_limit;
                granted = old >= lim ? 0u : (uint32_t)((lim - old) < want ? (lim - old) : want);
                if (granted < want) atomicAdd(&ctl->next_work, (unsigned long long)0 - (unsigned long long)(want - granted));
                more_later = lim < ctl->total_work;      // a later window advance can still release work
            }
            old = __shfl_sync(0xffffffffu, old, leader);
            granted = __shfl_sync(0xffffffffu, granted, leader);
            more_later = __shfl_sync(0xffffffffu, more_later, leader);
            if (need_work) {
                const uint32_t r = (uint32_t)__popc(m & lanemask_lt());
                if (r < granted) {
                    const unsigned long long w = old + r;
                    frame_k = (uint32_t)(w / P.items_per_frame);
                    pix_i = (uint32_t)(w - (unsigned long long)frame_k * P.items_per_frame);
                    const uint32_t pixel = shard_pixel(S.shard, S.W, pix_i);
                    uint32_t lim_pix = npix < S.shard.pixel_limit ? npix : S.shard.pixel_limit;
                    if (pixel < lim_pix) {
                        const float frame = (float)(P.frame0 + (int)frame_k * P.frame_stride);
                        rng = pixel_seed(pixel, frame);
                        scnt = 0; bounce = 0;
                        start_sample(S, pixel, 0, side, rng, ray);
                        n_draws += 2; n_samples++;
                        thr = V3s(1.0f); acc = V3s(0.0f);
                        pix = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                        alive = true; store_pix = true;
                    } else {
                        atomicAdd(&s_done[frame_k % P.ring], 1u);      // tile tail past the image: nothing to render
                        requeue_idle = more_later != 0;
                    }
                } else {
                    requeue_idle = more_later != 0;                    // window is full: wait for a frame to retire
                }
            }
        }
    }

    if (alive) {
        P.pool.ro[slot] = make_float4(ray.o.x, ray.o.y, ray.o.z, 0.0f);
        P.pool.rd[slot] = make_float4(ray.d.x, ray.d.y, ray.d.z, 0.0f);
        P.pool.thr[slot] = make_float4(thr.x, thr.y, thr.z, 0.0f);
        P.pool.acc[slot] = make_float4(acc.x, acc.y, acc.z, 0.0f);
        P.pool.misc[slot] = make_uint4(rng, pix_i, frame_k, (bounce << 16) | (scnt & 0xffffu));
        if (store_pix) P.pool.pix[slot] = pix;
        P.pool.hid[slot].x = HIT_NONE;
    } else if (requeue_idle) {
        P.pool.hid[slot].x = WF_KIND_IDLE;
    }

    // ---- compact: survivors (and waiting slots) into the next extend queue, one atomic per block ----
    const bool push = alive || requeue_idle;
    const uint32_t m = __ballot_sync(0xffffffffu, push);
    if (lane == 0) s_warp[warp] = (uint32_t)__popc(m);
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t sum = 0;
        for (int k = 0; k < 8; k++) { const uint32_t v = s_warp[k]; s_warp[k] = sum; sum += v; }
        s_base = sum ? atomicAdd(&ctl->n_extend, sum) : 0u;
    }
    __syncthreads();
    if (push) P.q_extend[s_base + s_warp[warp] + (uint32_t)__popc(m & lanemask_lt())] = slot;
    if (threadIdx.x < WF_MAX_RING && s_done[threadIdx.x]) atomicAdd(&ctl->done[threadIdx.x], s_done[threadIdx.x]);

    if (COUNT) {
        unsigned long long a = n_samples, b = n_draws;
        for (int o = 16; o > 0; o >>= 1) { a += __shfl_down_sync(0xffffffffu, a, o); b += __shfl_down_sync(0xffffffffu, b, o); }
        if (lane == 0) { if (a) atomicAdd(&S.counters[6], a); if (b) atomicAdd(&S.counters[5], b); }
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

void wf_launch_iteration(const DeviceScene& S, const WideBvh& WB, const WfParams& P, float4* fb, int reset_first,
                         cudaStream_t stream, cudaEvent_t* ev) {
    const unsigned extend_blocks = 148u * 8u;                // persistent: 8 CTAs of 4 warps per SM
    if (ev) cudaEventRecord(ev[0], stream);
    if (S.counters) k_wf_extend<true><<<extend_blocks, 128, 0, stream>>>(S, WB, P);
    else k_wf_extend<false><<<extend_blocks, 128, 0, stream>>>(S, WB, P);
    if (ev) cudaEventRecord(ev[1], stream);
    k_wf_ctl_mid<<<1, 32, 0, stream>>>(P);
    const unsigned shade_blocks = (P.pool.capacity + 255u) / 256u;
    if (S.counters) k_wf_shade<true><<<shade_blocks, 256, 0, stream>>>(S, P);
    else k_wf_shade<false><<<shade_blocks, 256, 0, stream>>>(S, P);
    if (ev) cudaEventRecord(ev[2], stream);
    k_wf_ctl_end<<<1, 32, 0, stream>>>(P);
    const uint32_t npix = (uint32_t)S.W * (uint32_t)S.H;
    unsigned acc_blocks = (npix + 255u) / 256u;
    if (acc_blocks > 148u * 8u) acc_blocks = 148u * 8u;
    k_wf_accumulate<<<acc_blocks, 256, 0, stream>>>(P, fb, npix, reset_first);
    if (ev) cudaEventRecord(ev[3], stream);
}

// closest-hit hook over the wide BVH (wgpt_trace_rays in fast mode)
template <bool COUNT>
__global__ void __launch_bounds__(128) k_trace_rays_wide(DeviceScene S, WideBvh WB, int n, const float* __restrict__ rays,
                                                         const uint32_t* __restrict__ seeds, int* __restrict__ out_type,
                                                         int* __restrict__ out_index, float* __restrict__ out_t,
                                                         uint32_t* __restrict__ fallbacks_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    Counts c = {0, 0, 0, 0, 0, 0, 0, 0};
    unsigned fallbacks = 0;
    if (i < n) {
        Ray r; r.o = V3(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]); r.d = V3(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]);
        uint32_t rng = seeds ? seeds[i] : 0u;
        ObjRayCache oc; HitLite h; h.t = 0.0f; h.kind = HIT_NONE; h.index = -1; h.mat_id = -1; h.uu = h.vv = 0.0f;
        const bool hit = hitScene_wide(S, WB, r, rng, COUNT ? &c : nullptr, oc, h, &fallbacks);
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

// wide_bvh.cu -- K2: flatten the reference's binary BVH into the 8-wide quantised layout of
// wide_bvh.cuh, on the device, level by level (one thread per wide node, work queues in HBM).
//
// Input contract (/root/reference/lib/BVH/bvhBuilder.js:37-54):
// pre-order nodes; left child = i+1, right child = i32(right_offset); leaf <=> i32(prim_type)==2
// with prim_count == 1; the subtree of i spans [i, skip_link >= 0 ? skip_link : N) and therefore
// holds (end - i + 1) / 2 triangles.
#include "wide_bvh.cuh"

namespace wg {

#define WIDE_LEAF_MAX 1      // a leaf slot holds exactly one triangle (= one leaf of the binary tree)

struct BuildCtx {
    const BvhNode* bvh; int n_bvh;
    const Triangle* tris; int n_tris;
    uint4* out_nodes; float4* out_tris;
    uint32_t max_nodes;
    uint32_t* counters;      // [0] nodes, [1] tris, [2],[3] queue counts, [4] levels, [5] overflow
    uint2* queue[2];
};

__device__ __forceinline__ int sub_end(const BuildCtx& B, int i) {
    const int s = (int)B.bvh[i].skip_link;
    return s >= 0 ? s : B.n_bvh;
}
__device__ __forceinline__ int sub_tris(const BuildCtx& B, int i) { return (sub_end(B, i) - i + 1) / 2; }
__device__ __forceinline__ bool is_leaf(const BuildCtx& B, int i) { return (int)B.bvh[i].prim_type == 2; }
__device__ __forceinline__ float half_area(const BvhNode& n) {
    const float dx = n.mx[0] - n.mn[0], dy = n.mx[1] - n.mn[1], dz = n.mx[2] - n.mn[2];
    return dx * dy + dy * dz + dz * dx;
}

__global__ void k_wide_init(BuildCtx B) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        B.counters[0] = 1; B.counters[1] = 0; B.counters[2] = 1; B.counters[3] = 0; B.counters[4] = 0; B.counters[5] = 0;
        B.queue[0][0] = make_uint2(0u, 0u);      // (binary root, wide node 0)
    }
}

__global__ void k_wide_prep(BuildCtx B, int parity) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        B.counters[3 - parity] = 0;
        if (B.counters[2 + parity] > 0) B.counters[4] += 1;
    }
}

__global__ void __launch_bounds__(128) k_wide_level(BuildCtx B, int parity) {
    const uint32_t n_in = B.counters[2 + parity];
    for (uint32_t it = blockIdx.x * blockDim.x + threadIdx.x; it < n_in; it += gridDim.x * blockDim.x) {
        const uint2 item = B.queue[parity][it];
        const int root = (int)item.x;
        const uint32_t wide_idx = item.y;

        // ---- 1. choose up to 8 children by expanding the largest box first ----
        int ch[8]; int n = 1; ch[0] = root;
        // phase 0 expands anything bigger than a leaf child; phase 1 then splits leaf children
        // further while slots remain (a box test is cheaper than a triangle test)
        for (int phase = 0; phase < 2; phase++) {
            while (n < 8) {
                int best = -1; float best_a = -1.0f;
                for (int k = 0; k < n; k++) {
                    if (is_leaf(B, ch[k])) continue;
                    if (phase == 0 && ch[k] != root && sub_tris(B, ch[k]) <= WIDE_LEAF_MAX) continue;
                    const float a = half_area(B.bvh[ch[k]]);
                    if (a > best_a) { best_a = a; best = k; }
                }
                if (best < 0) break;
                const int c = ch[best];
                ch[best] = c + 1;
                ch[n++] = (int)B.bvh[c].right_offset;
            }
        }

        // ---- 2. node box, slot assignment (greedy on dot(offset, slot diagonal)) ----
        float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
        for (int k = 0; k < n; k++)
            for (int a = 0; a < 3; a++) { lo[a] = fminf(lo[a], B.bvh[ch[k]].mn[a]); hi[a] = fmaxf(hi[a], B.bvh[ch[k]].mx[a]); }
        int slot_of[8]; bool slot_used[8]; bool assigned[8];
        for (int k = 0; k < 8; k++) { slot_used[k] = false; assigned[k] = false; slot_of[k] = -1; }
        for (int round = 0; round < n; round++) {
            float best = -3.0e38f; int bc = -1, bs = -1;
            for (int k = 0; k < n; k++) {
                if (assigned[k]) continue;
                const BvhNode& c = B.bvh[ch[k]];
                const float ox = (c.mn[0] + c.mx[0]) - (lo[0] + hi[0]);
                const float oy = (c.mn[1] + c.mx[1]) - (lo[1] + hi[1]);
                const float oz = (c.mn[2] + c.mx[2]) - (lo[2] + hi[2]);
                for (int s = 0; s < 8; s++) {
                    if (slot_used[s]) continue;
                    const float cost = ((s & 4) ? ox : -ox) + ((s & 2) ? oy : -oy) + ((s & 1) ? oz : -oz);
                    if (cost > best) { best = cost; bc = k; bs = s; }
                }
            }
            assigned[bc] = true; slot_used[bs] = true; slot_of[bc] = bs;
        }
        int child_in_slot[8];
        for (int s = 0; s < 8; s++) child_in_slot[s] = -1;
        for (int k = 0; k < n; k++) child_in_slot[slot_of[k]] = ch[k];

        // ---- 3. quantisation frame: per-axis power-of-two scale with 255 steps covering the extent ----
        uint32_t ebits[3]; float scale[3];
        for (int a = 0; a < 3; a++) {
            const double ext = (double)hi[a] - (double)lo[a];
            int e = -126;
            if (ext > 0.0) { int k; frexp(ext, &k); e = k - 8; }       // ext in [2^(k-1), 2^k)
            if (e < -126) e = -126;
            while (e < 127 && ldexp(255.0, e) < ext) e++;
            if (e > 100) B.counters[5] = 1;                 // the traversal folds 2^24 into the scale (trace_wide.cuh cvt2)
            ebits[a] = (uint32_t)(e + 127);
            scale[a] = __uint_as_float(ebits[a] << 23);
        }

        // ---- 4. count inner / leaf children, allocate, emit ----
        uint32_t imask = 0, n_inner = 0, n_leaf_tris = 0;
        for (int s = 0; s < 8; s++) {
            const int c = child_in_slot[s];
            if (c < 0) continue;
            const int nt = sub_tris(B, c);
            if (!is_leaf(B, c) && nt > WIDE_LEAF_MAX) { imask |= 1u << s; n_inner++; }
            else n_leaf_tris += (uint32_t)nt;
        }
        uint32_t child_base = 0, tri_base = 0;
        if (n_inner) {
            child_base = atomicAdd(&B.counters[0], n_inner);
            if (child_base + n_inner > B.max_nodes) { B.counters[5] = 1; continue; }
        }
        if (n_leaf_tris) tri_base = atomicAdd(&B.counters[1], n_leaf_tris);
        uint32_t q_base = 0;
        if (n_inner) q_base = atomicAdd(&B.counters[3 - parity], n_inner);

        unsigned char qlo[3][8], qhi[3][8];
        uint32_t inner_i = 0, tri_off = 0, lmask = 0;
        for (int s = 0; s < 8; s++) {
            for (int a = 0; a < 3; a++) { qlo[a][s] = 255; qhi[a][s] = 0; }     // empty: inverted box, never hit
            const int c = child_in_slot[s];
            if (c < 0) continue;
            const BvhNode& cb = B.bvh[c];
            for (int a = 0; a < 3; a++) {
                double l = floor(((double)cb.mn[a] - (double)lo[a]) / (double)scale[a]);
                double h = ceil(((double)cb.mx[a] - (double)lo[a]) / (double)scale[a]);
                l = l < 0.0 ? 0.0 : (l > 255.0 ? 255.0 : l);
                h = h < 0.0 ? 0.0 : (h > 255.0 ? 255.0 : h);
                qlo[a][s] = (unsigned char)l; qhi[a][s] = (unsigned char)h;
            }
            if (imask & (1u << s)) {
                B.queue[1 - parity][q_base + inner_i] = make_uint2((uint32_t)c, child_base + inner_i);
                inner_i++;
            } else {
                // leaf slot: the triangle of this binary leaf; the slot's triangles are packed in slot order
                const int src = (int)B.bvh[c].prim_id;
                const Triangle& t = B.tris[src];
                float4* o = &B.out_tris[3 * (size_t)(tri_base + tri_off)];
                o[0] = make_float4(t.A[0], t.A[1], t.A[2], __int_as_float(src));
                o[1] = make_float4(t.B[0], t.B[1], t.B[2], __int_as_float((int)t.mesh_id));
                o[2] = make_float4(t.C[0], t.C[1], t.C[2], __int_as_float((int)t.local_id));
                lmask |= 1u << s;
                tri_off++;
            }
        }

        auto pack4 = [](const unsigned char* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); };
        uint4* o = &B.out_nodes[5 * (size_t)wide_idx];
        o[0] = make_uint4(__float_as_uint(lo[0]), __float_as_uint(lo[1]), __float_as_uint(lo[2]),
                          ebits[0] | (ebits[1] << 8) | (ebits[2] << 16) | (imask << 24));
        o[1] = make_uint4(child_base, tri_base, lmask, 0u);
        o[2] = make_uint4(pack4(qlo[0]), pack4(qlo[0] + 4), pack4(qlo[1]), pack4(qlo[1] + 4));
        o[3] = make_uint4(pack4(qlo[2]), pack4(qlo[2] + 4), pack4(qhi[0]), pack4(qhi[0] + 4));
        o[4] = make_uint4(pack4(qhi[1]), pack4(qhi[1] + 4), pack4(qhi[2]), pack4(qhi[2] + 4));
    }
}

size_t wide_scratch_bytes(uint32_t max_nodes) { return 64 + 2 * (size_t)(max_nodes + 16) * sizeof(uint2); }

void launch_wide_build(const BvhNode* bvh, int n_bvh, const Triangle* tris, int n_tris, uint4* out_nodes,
                       float4* out_tris, uint32_t max_nodes, uint32_t* d_scratch, cudaStream_t stream) {
    BuildCtx B;
    B.bvh = bvh; B.n_bvh = n_bvh; B.tris = tris; B.n_tris = n_tris;
    B.out_nodes = out_nodes; B.out_tris = out_tris; B.max_nodes = max_nodes;
    B.counters = d_scratch;
    B.queue[0] = reinterpret_cast<uint2*>(d_scratch + 16);
    B.queue[1] = B.queue[0] + (max_nodes + 16);
    k_wide_init<<<1, 32, 0, stream>>>(B);
    // a wide level consumes >= 1 binary level, and stack_size <= 64 bounds the useful depth
    for (int level = 0; level < 48; level++) {
        const int parity = level & 1;
        k_wide_prep<<<1, 32, 0, stream>>>(B, parity);
        k_wide_level<<<148 * 4, 128, 0, stream>>>(B, parity);
    }
}

WideBuildStats wide_build_stats(const uint32_t* d_scratch, cudaStream_t stream) {
    uint32_t h[8] = {0};
    cudaMemcpyAsync(h, d_scratch, sizeof(h), cudaMemcpyDeviceToHost, stream);
    cudaStreamSynchronize(stream);
    WideBuildStats s; s.n_nodes = h[0]; s.n_tris = h[1]; s.n_levels = h[4]; s.overflow = h[5];
    return s;
}

}  // namespace wg

// wide_bvh.cu -- K2: flatten the reference's binary BVH into the 8-wide quantised layout of
// wide_bvh.cuh, on the device, level by level (one thread per wide node, work queues in HBM).
//
// Input contract (/root/reference/lib/BVH/bvhBuilder.js:37-54):
// pre-order nodes; left child = i+1, right child = i32(right_offset); leaf <=> i32(prim_type)==2
// with prim_count == 1; the subtree of i spans [i, skip_link >= 0 ? skip_link : N) and therefore
// holds (end - i + 1) / 2 triangles.
#include "wide_bvh.cuh"
#include <cstdlib>

namespace wg {

#define WIDE_LEAF_MAX 1      // a leaf slot holds exactly one triangle (= one leaf of the binary tree)

struct BuildCtx {
    const BvhNode* bvh; int n_bvh;
    const Triangle* tris; int n_tris;
    uint4* out_nodes; float4* out_tris;
    uint32_t max_nodes;
    uint32_t* counters;      // [0] nodes, [1] tris, [2],[3] queue counts, [4] levels, [5] overflow, [6] largest plane exponent byte of the root
    uint2* queue[2];
    // collapse plan (k_plan_*): which binary nodes become the children of a wide node
    int use_plan;
    int* parent;             // binary parent of every node (-1 for the root)
    uint32_t* arrived;       // bottom-up rendezvous counters
    float* cost;             // 8 floats per binary node: cost[i][k-1] = least summed area of the wide nodes that
                             // represent subtree i when it may occupy at most k slots of its wide parent
    uint2* plan;             // x: 4 bits per k = 2..8, the choice behind cost[i][k-1] (1 = i stays one inner child,
                             //    j = i is opened and its two subtrees share j slots); y: 3 bits per j = 2..8, how many
                             //    of the j slots the left subtree gets
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
        B.counters[0] = 1; B.counters[1] = 0; B.counters[2] = 1; B.counters[3] = 0; B.counters[4] = 0; B.counters[5] = 0; B.counters[6] = 127;
        B.queue[0][0] = make_uint2(0u, 0u);      // (binary root, wide node 0)
    }
}

// ---- SAH-optimal collapse plan -------------------------------------------------------------------
// The expected number of wide-node visits of a random ray is the summed surface area of the wide nodes
// (the triangle tests do not depend on the collapse: every triangle is one leaf slot whatever the shape).
// Largest-area-first expansion approximates that sum; the dynamic programme below minimises it exactly
// for the given binary tree (Ylitie, Karras, Laine 2017, section 3.2, with one-triangle leaves):
//   cost(i, 1) = area(i) + open(i, 8)                      i becomes a wide node of its own
//   cost(i, k) = min(cost(i, k-1), open(i, k))             or its subtrees take k slots of the wide parent
//   open(i, j) = min over a + b = j of cost(left, a) + cost(right, b);   cost(leaf, k) = 0
// evaluated bottom-up (one thread per leaf climbs; the second thread to arrive at a node evaluates it).
__global__ void k_plan_parents(BuildCtx B) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B.n_bvh) return;
    if (i == 0) B.parent[0] = -1;
    if (is_leaf(B, i)) return;
    const int r = (int)B.bvh[i].right_offset;
    if (i + 1 < B.n_bvh) B.parent[i + 1] = i;
    if (r > i && r < B.n_bvh) B.parent[r] = i;
}

__global__ void k_plan_costs(BuildCtx B) {
    const int leaf = blockIdx.x * blockDim.x + threadIdx.x;
    if (leaf >= B.n_bvh || !is_leaf(B, leaf)) return;
    float4* cost4 = reinterpret_cast<float4*>(B.cost);
    cost4[2 * (size_t)leaf] = make_float4(0.f, 0.f, 0.f, 0.f);
    cost4[2 * (size_t)leaf + 1] = make_float4(0.f, 0.f, 0.f, 0.f);
    int cur = leaf;
    for (int guard = 0; guard < 4096; guard++) {
        const int p = cur > 0 ? B.parent[cur] : -1;
        if (p < 0 || p >= B.n_bvh) return;
        __threadfence();
        if (atomicAdd(&B.arrived[p], 1u) == 0u) return;          // the sibling subtree is not done yet: its thread goes on
        __threadfence();
        const int l = p + 1, r = (int)B.bvh[p].right_offset;
        if (r <= l || r >= B.n_bvh) return;                      // not the flatten contract: k_wide_level reports it
        float cl[9], cr[9];
        {
            const float4 a0 = __ldcg(&cost4[2 * (size_t)l]), a1 = __ldcg(&cost4[2 * (size_t)l + 1]);
            const float4 b0 = __ldcg(&cost4[2 * (size_t)r]), b1 = __ldcg(&cost4[2 * (size_t)r + 1]);
            cl[1] = a0.x; cl[2] = a0.y; cl[3] = a0.z; cl[4] = a0.w; cl[5] = a1.x; cl[6] = a1.y; cl[7] = a1.z; cl[8] = a1.w;
            cr[1] = b0.x; cr[2] = b0.y; cr[3] = b0.z; cr[4] = b0.w; cr[5] = b1.x; cr[6] = b1.y; cr[7] = b1.z; cr[8] = b1.w;
        }
        float open[9]; uint32_t splits = 0;
        #pragma unroll
        for (int j = 2; j <= 8; j++) {
            float m = 3.0e38f; uint32_t bs = 1;
            #pragma unroll
            for (int a = 1; a < j; a++) { const float v = cl[a] + cr[j - a]; if (v < m) { m = v; bs = (uint32_t)a; } }
            open[j] = m; splits |= bs << (3 * (j - 2));
        }
        float c[9]; uint32_t choices = 0, ch = 1;
        c[1] = half_area(B.bvh[p]) + open[8];
        #pragma unroll
        for (int k = 2; k <= 8; k++) {
            c[k] = c[k - 1];
            if (open[k] < c[k]) { c[k] = open[k]; ch = (uint32_t)k; }
            choices |= ch << (4 * (k - 2));
        }
        cost4[2 * (size_t)p] = make_float4(c[1], c[2], c[3], c[4]);
        cost4[2 * (size_t)p + 1] = make_float4(c[5], c[6], c[7], c[8]);
        B.plan[p] = make_uint2(choices, splits);
        cur = p;
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
        if (B.use_plan && !is_leaf(B, root)) {
            // the planned cut: the root is opened over all 8 slots, every other node follows its own choice
            int st_node[8], st_k[8]; int sp = 0;
            n = 0;
            st_node[0] = root; st_k[0] = 8; sp = 1;
            bool first = true;
            while (sp > 0 && n < 8) {
                sp--;
                const int node = st_node[sp], k = st_k[sp];
                if (is_leaf(B, node)) { ch[n++] = node; first = false; continue; }
                const uint2 pl = B.plan[node];
                const uint32_t j = first ? 8u : (k >= 2 ? ((pl.x >> (4 * (k - 2))) & 15u) : 1u);
                first = false;
                if (j < 2u || j > (uint32_t)k) { ch[n++] = node; continue; }       // stays one inner child
                uint32_t a = (pl.y >> (3 * (j - 2u))) & 7u;
                if (a < 1u) a = 1u;
                if (a > j - 1u) a = j - 1u;
                if (sp + 2 > 8) { ch[n++] = node; continue; }
                st_node[sp] = (int)B.bvh[node].right_offset; st_k[sp] = (int)(j - a); sp++;
                st_node[sp] = node + 1; st_k[sp] = (int)a; sp++;
            }
        } else
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

        // ---- 3. quantisation frame: per-axis power-of-two cell with 255 cells covering the extent ----
        // The frame starts 1/8 cell below the node box and every child plane is pushed 1/16 cell outwards before
        // floor / ceil: that margin is the error budget of the traversal's binary16 cell size (trace_wide.cuh).
        uint32_t ebits[3]; float scale[3]; float org[3];
        for (int a = 0; a < 3; a++) {
            const double ext = (double)hi[a] - (double)lo[a];
            int e = -126;
            if (ext > 0.0) { int k; frexp(ext, &k); e = k - 8; }       // ext in [2^(k-1), 2^k)
            if (e < -126) e = -126;
            for (;;) {
                const double sc = ldexp(1.0, e);
                org[a] = __double2float_rd((double)lo[a] - 0.125 * sc);
                if (e >= 127 || ((double)hi[a] - (double)org[a]) / sc + 0.0625 <= 255.0) break;
                e++;
            }
            if (e > 100) B.counters[5] = 1;                 // far outside any scene scale the ray scaling can hold
            ebits[a] = (uint32_t)(e + 127);
            scale[a] = __uint_as_float(ebits[a] << 23);
        }
        if (wide_idx == 0u) B.counters[6] = max(ebits[0], max(ebits[1], ebits[2]));

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
                double l = floor(((double)cb.mn[a] - (double)org[a]) / (double)scale[a] - 0.0625);
                double h = ceil(((double)cb.mx[a] - (double)org[a]) / (double)scale[a] + 0.0625);
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
        o[0] = make_uint4(__float_as_uint(org[0]), __float_as_uint(org[1]), __float_as_uint(org[2]),
                          ebits[0] | (ebits[1] << 8) | (ebits[2] << 16) | (imask << 24));
        o[1] = make_uint4(child_base, tri_base, lmask, 0u);
        o[2] = make_uint4(pack4(qlo[0]), pack4(qlo[0] + 4), pack4(qlo[1]), pack4(qlo[1] + 4));
        o[3] = make_uint4(pack4(qlo[2]), pack4(qlo[2] + 4), pack4(qhi[0]), pack4(qhi[0] + 4));
        o[4] = make_uint4(pack4(qhi[1]), pack4(qhi[1] + 4), pack4(qhi[2]), pack4(qhi[2] + 4));
    }
}

// queues + the collapse plan's arrays for up to 2 * max_nodes binary nodes (max_nodes >= n_bvh / 2 + 16)
static size_t plan_offset(uint32_t max_nodes) { return (64 + 2 * (size_t)(max_nodes + 16) * sizeof(uint2) + 255) & ~(size_t)255; }
size_t wide_scratch_bytes(uint32_t max_nodes) { return plan_offset(max_nodes) + 2 * (size_t)max_nodes * (32 + 8 + 4 + 4) + 1024; }

int launch_wide_build(const BvhNode* bvh, int n_bvh, const Triangle* tris, int n_tris, uint4* out_nodes,
                      float4* out_tris, uint32_t max_nodes, uint32_t* d_scratch, cudaStream_t stream) {
    BuildCtx B;
    B.bvh = bvh; B.n_bvh = n_bvh; B.tris = tris; B.n_tris = n_tris;
    B.out_nodes = out_nodes; B.out_tris = out_tris; B.max_nodes = max_nodes;
    B.counters = d_scratch;
    B.queue[0] = reinterpret_cast<uint2*>(d_scratch + 16);
    B.queue[1] = B.queue[0] + (max_nodes + 16);
    // WGPT_COLLAPSE=greedy keeps the largest-area-first expansion (for A/B measurements)
    const char* how = getenv("WGPT_COLLAPSE");
    B.use_plan = !(how && how[0] == 'g') && n_bvh >= 3 && (size_t)n_bvh <= 2 * (size_t)max_nodes;
    {
        char* p = reinterpret_cast<char*>(d_scratch) + plan_offset(max_nodes);
        const size_t cap = 2 * (size_t)max_nodes;
        B.cost = reinterpret_cast<float*>(p); p += cap * 32;
        B.plan = reinterpret_cast<uint2*>(p); p += cap * 8;
        B.parent = reinterpret_cast<int*>(p); p += cap * 4;
        B.arrived = reinterpret_cast<uint32_t*>(p);
    }
    if (B.use_plan) {
        cudaMemsetAsync(B.arrived, 0, (size_t)n_bvh * 4, stream);
        const unsigned blocks = (unsigned)((n_bvh + 255) / 256);
        k_plan_parents<<<blocks, 256, 0, stream>>>(B);
        k_plan_costs<<<blocks, 256, 0, stream>>>(B);
    }
    k_wide_init<<<1, 32, 0, stream>>>(B);
    // a wide level consumes >= 1 binary level, and stack_size <= 64 bounds the useful depth
    for (int level = 0; level < 48; level++) {
        const int parity = level & 1;
        k_wide_prep<<<1, 32, 0, stream>>>(B, parity);
        k_wide_level<<<148 * 4, 128, 0, stream>>>(B, parity);
    }
    return 1 + 2 * 48 + (B.use_plan ? 2 : 0);
}

WideBuildStats wide_build_stats(const uint32_t* d_scratch, cudaStream_t stream) {
    uint32_t h[8] = {0};
    cudaMemcpyAsync(h, d_scratch, sizeof(h), cudaMemcpyDeviceToHost, stream);
    cudaStreamSynchronize(stream);
    WideBuildStats s; s.n_nodes = h[0]; s.n_tris = h[1]; s.n_levels = h[4]; s.overflow = h[5]; s.root_emax = h[6];
    return s;
}

}  // namespace wg

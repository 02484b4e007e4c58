// bvh_build.cu -- K5: bottom-up BVH construction on the device (parallel locally-ordered
// clustering), emitting the reference's pre-order node records for the flatten kernel K2.
// See bvh_build.cuh for why the fast path may use its own hierarchy.
//
// Pipeline (everything stays on the stream; no host round trip):
//   k_prim_bounds   world-space box of every triangle, padded so that it CONTAINS the box the
//                   reference's host code computes (AABB.js:12-51), scene bounds by atomics
//   k_morton        63-bit Morton code of the box centre; cub radix sort (key, triangle)
//   k_make_leaves   leaf node p = p-th triangle of the sorted sequence; cluster list = 0..N-1
//   repeat:  k_nn       every cluster finds, among its 2R neighbours in the sequence, the one whose
//                       union with it has the smallest surface area (tile of boxes in shared memory)
//            k_merge    mutual nearest neighbours merge into a new inner node (lower index creates it)
//            k_scatter  order-preserving compaction of the surviving clusters (block counts + scan)
//            k_advance  publishes the new cluster count
//   k_tail          one block finishes the last <= 1024 clusters without further launches
//   k_emit          pre-order index of every node (sum of increments on the path to the root),
//                   48-byte records: left = i + 1, right_offset, skip_link, leaves with prim_id
#include <cub/device/device_radix_sort.cuh>
#include "bvh_build.cuh"

namespace wg {

#ifndef PLOC_R
#define PLOC_R 16               // neighbours searched on each side
#endif
#define PLOC_TAIL 1024          // cluster count handed to the single-block kernel
#define PLOC_GLOBAL_ITERS 28
#define NONE 0xffffffffu

struct BuildCounters {
    uint32_t nclus[2];          // cluster count, by parity of the iteration
    uint32_t node_counter;      // next free node id
    uint32_t done_iter;         // completed global iterations
    uint32_t root;
    uint32_t bounds[6];         // scene min xyz / max xyz, order-preserving uint encoding
};

struct BuildArrays {
    float4* lo;                 // node box min, w = int bits: left child (inner) or -1 (leaf)
    float4* hi;                 // node box max, w = int bits: right child (inner) or triangle index (leaf)
    uint32_t* cnt;              // leaves below the node
    uint32_t* parent;
    uint32_t* cid[2];           // cluster sequence (node ids), double buffered
    uint32_t* nn;               // nearest neighbour (position in the sequence)
    uint32_t* newid;            // node id after the merge step, NONE = absorbed
    uint32_t* block_sums;
    float4* plo; float4* phi;   // per-triangle world boxes (input order)
    unsigned long long* key[2]; uint32_t* val[2];
    BuildCounters* ctr;
    int n;
};

__device__ __forceinline__ uint32_t enc(float f) { const uint32_t b = __float_as_uint(f); return (b & 0x80000000u) ? ~b : (b | 0x80000000u); }
__device__ __forceinline__ float dec(uint32_t u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u); }
__device__ __forceinline__ float half_area(float3 a, float3 b) { const float dx = b.x - a.x, dy = b.y - a.y, dz = b.z - a.z; return dx * dy + dy * dz + dz * dx; }

__global__ void k_build_init(BuildArrays A) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        BuildCounters* c = A.ctr;
        c->nclus[0] = (uint32_t)A.n; c->nclus[1] = 0; c->node_counter = (uint32_t)A.n; c->done_iter = 0; c->root = 0;
        for (int k = 0; k < 3; k++) { c->bounds[k] = 0xffffffffu; c->bounds[3 + k] = 0u; }
    }
}

// World-space triangle box.  The reference transforms the vertices with gl-matrix in binary64 and
// rounds to f32 (triangle.js:27-39); here the transform runs in f32, a few ulp away, so the box is
// widened by 8 ulp of its largest coordinate, and by the reference's own +-5e-5 pad whenever the
// reference might apply it (AABB.js:35-51 pads extents below 1e-4/2): the result contains the
// reference's leaf box, which is what keeps the wide traversal a superset search.
__global__ void __launch_bounds__(256) k_prim_bounds(BuildArrays A, const Triangle* __restrict__ tris, const Mesh* __restrict__ meshes,
                                                     const ModelTransform* __restrict__ transforms) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    float3 lo = make_float3(3.0e38f, 3.0e38f, 3.0e38f), hi = make_float3(-3.0e38f, -3.0e38f, -3.0e38f);
    if (t < A.n) {
        const Triangle& T = tris[t];
        const float* M = transforms[meshes[(int)T.mesh_id].global_id].modelMatrix;
        const float* V[3] = {T.A, T.B, T.C};
        #pragma unroll
        for (int k = 0; k < 3; k++) {
            const float x = V[k][0], y = V[k][1], z = V[k][2];
            const float wx = M[0] * x + M[4] * y + M[8] * z + M[12];
            const float wy = M[1] * x + M[5] * y + M[9] * z + M[13];
            const float wz = M[2] * x + M[6] * y + M[10] * z + M[14];
            lo.x = fminf(lo.x, wx); lo.y = fminf(lo.y, wy); lo.z = fminf(lo.z, wz);
            hi.x = fmaxf(hi.x, wx); hi.y = fmaxf(hi.y, wy); hi.z = fmaxf(hi.z, wz);
        }
        const float big = fmaxf(fmaxf(fmaxf(fabsf(lo.x), fabsf(hi.x)), fmaxf(fabsf(lo.y), fabsf(hi.y))), fmaxf(fabsf(lo.z), fabsf(hi.z)));
        const float ulps = big * 9.5367431640625e-7f + 1e-30f;           // 8 * 2^-23 * |largest coordinate|
        const float thin = 0.00005f + 2.0f * ulps;
        const float px = (hi.x - lo.x) < thin ? 0.00005f + ulps : ulps;
        const float py = (hi.y - lo.y) < thin ? 0.00005f + ulps : ulps;
        const float pz = (hi.z - lo.z) < thin ? 0.00005f + ulps : ulps;
        lo.x -= px; lo.y -= py; lo.z -= pz; hi.x += px; hi.y += py; hi.z += pz;
        A.plo[t] = make_float4(lo.x, lo.y, lo.z, 0.0f);
        A.phi[t] = make_float4(hi.x, hi.y, hi.z, 0.0f);
    }
    // scene bounds: warp reduce, one atomic per warp and component
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo.x = fminf(lo.x, __shfl_xor_sync(0xffffffffu, lo.x, o)); lo.y = fminf(lo.y, __shfl_xor_sync(0xffffffffu, lo.y, o));
        lo.z = fminf(lo.z, __shfl_xor_sync(0xffffffffu, lo.z, o)); hi.x = fmaxf(hi.x, __shfl_xor_sync(0xffffffffu, hi.x, o));
        hi.y = fmaxf(hi.y, __shfl_xor_sync(0xffffffffu, hi.y, o)); hi.z = fmaxf(hi.z, __shfl_xor_sync(0xffffffffu, hi.z, o));
    }
    if ((threadIdx.x & 31) == 0 && lo.x <= hi.x) {
        atomicMin(&A.ctr->bounds[0], enc(lo.x)); atomicMin(&A.ctr->bounds[1], enc(lo.y)); atomicMin(&A.ctr->bounds[2], enc(lo.z));
        atomicMax(&A.ctr->bounds[3], enc(hi.x)); atomicMax(&A.ctr->bounds[4], enc(hi.y)); atomicMax(&A.ctr->bounds[5], enc(hi.z));
    }
}

__device__ __forceinline__ unsigned long long spread21(unsigned long long v) {
    v &= 0x1fffffull;
    v = (v | (v << 32)) & 0x1f00000000ffffull;
    v = (v | (v << 16)) & 0x1f0000ff0000ffull;
    v = (v | (v << 8)) & 0x100f00f00f00f00full;
    v = (v | (v << 4)) & 0x10c30c30c30c30c3ull;
    v = (v | (v << 2)) & 0x1249249249249249ull;
    return v;
}

__global__ void __launch_bounds__(256) k_morton(BuildArrays A) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= A.n) return;
    const float3 smin = make_float3(dec(A.ctr->bounds[0]), dec(A.ctr->bounds[1]), dec(A.ctr->bounds[2]));
    const float3 smax = make_float3(dec(A.ctr->bounds[3]), dec(A.ctr->bounds[4]), dec(A.ctr->bounds[5]));
    const float4 lo = A.plo[t], hi = A.phi[t];
    const float ex = fmaxf(smax.x - smin.x, 1e-30f), ey = fmaxf(smax.y - smin.y, 1e-30f), ez = fmaxf(smax.z - smin.z, 1e-30f);
    const float cx = (0.5f * (lo.x + hi.x) - smin.x) / ex, cy = (0.5f * (lo.y + hi.y) - smin.y) / ey, cz = (0.5f * (lo.z + hi.z) - smin.z) / ez;
    const unsigned long long qx = (unsigned long long)fminf(fmaxf(cx * 2097152.0f, 0.0f), 2097151.0f);
    const unsigned long long qy = (unsigned long long)fminf(fmaxf(cy * 2097152.0f, 0.0f), 2097151.0f);
    const unsigned long long qz = (unsigned long long)fminf(fmaxf(cz * 2097152.0f, 0.0f), 2097151.0f);
    A.key[0][t] = (spread21(qx) << 2) | (spread21(qy) << 1) | spread21(qz);
    A.val[0][t] = (uint32_t)t;
}

__global__ void __launch_bounds__(256) k_make_leaves(BuildArrays A, const uint32_t* __restrict__ sorted_tri) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= A.n) return;
    const uint32_t t = sorted_tri[p];
    const float4 lo = A.plo[t], hi = A.phi[t];
    A.lo[p] = make_float4(lo.x, lo.y, lo.z, __int_as_float(-1));
    A.hi[p] = make_float4(hi.x, hi.y, hi.z, __int_as_float((int)t));
    A.cnt[p] = 1; A.parent[p] = NONE;
    A.cid[0][p] = (uint32_t)p;
}

// ---- one clustering iteration over global memory ----
__device__ __forceinline__ bool iter_active(const BuildArrays& A, int it, uint32_t& n) {
    n = A.ctr->nclus[it & 1];
    return A.ctr->done_iter == (uint32_t)it && n > PLOC_TAIL;
}

__global__ void __launch_bounds__(256) k_nn(BuildArrays A, int it) {
    uint32_t n;
    if (!iter_active(A, it, n)) return;
    const int base = blockIdx.x * 256;
    if ((uint32_t)base >= n) return;
    __shared__ float3 s_lo[256 + 2 * PLOC_R], s_hi[256 + 2 * PLOC_R];
    const uint32_t* cid = A.cid[it & 1];
    for (int k = threadIdx.x; k < 256 + 2 * PLOC_R; k += 256) {
        const int g = base - PLOC_R + k;
        if (g >= 0 && (uint32_t)g < n) {
            const uint32_t id = cid[g];
            const float4 l = A.lo[id], h = A.hi[id];
            s_lo[k] = make_float3(l.x, l.y, l.z); s_hi[k] = make_float3(h.x, h.y, h.z);
        }
    }
    __syncthreads();
    const int i = base + threadIdx.x;
    if ((uint32_t)i >= n) return;
    const float3 a = s_lo[threadIdx.x + PLOC_R], b = s_hi[threadIdx.x + PLOC_R];
    float best = 3.0e38f; int best_j = -1;
    for (int d = -PLOC_R; d <= PLOC_R; d++) {
        const int j = i + d;
        if (d == 0 || j < 0 || (uint32_t)j >= n) continue;
        const float3 l = s_lo[threadIdx.x + PLOC_R + d], h = s_hi[threadIdx.x + PLOC_R + d];
        const float ar = half_area(make_float3(fminf(a.x, l.x), fminf(a.y, l.y), fminf(a.z, l.z)),
                                   make_float3(fmaxf(b.x, h.x), fmaxf(b.y, h.y), fmaxf(b.z, h.z)));
        if (ar < best) { best = ar; best_j = j; }      // ties keep the lower index: both sides agree
    }
    A.nn[i] = (uint32_t)best_j;
}

__device__ __forceinline__ uint32_t merge_pair(const BuildArrays& A, uint32_t l, uint32_t r) {
    const uint32_t id = atomicAdd(&A.ctr->node_counter, 1u);
    const float4 ll = A.lo[l], lh = A.hi[l], rl = A.lo[r], rh = A.hi[r];
    A.lo[id] = make_float4(fminf(ll.x, rl.x), fminf(ll.y, rl.y), fminf(ll.z, rl.z), __int_as_float((int)l));
    A.hi[id] = make_float4(fmaxf(lh.x, rh.x), fmaxf(lh.y, rh.y), fmaxf(lh.z, rh.z), __int_as_float((int)r));
    A.cnt[id] = A.cnt[l] + A.cnt[r];
    A.parent[id] = NONE; A.parent[l] = id; A.parent[r] = id;
    return id;
}

__global__ void __launch_bounds__(1024) k_merge(BuildArrays A, int it) {
    uint32_t n;
    if (!iter_active(A, it, n)) return;
    const uint32_t i = blockIdx.x * 1024 + threadIdx.x;
    if (blockIdx.x * 1024u >= n) return;
    bool keep = false;
    if (i < n) {
        const uint32_t* cid = A.cid[it & 1];
        const uint32_t j = A.nn[i];
        uint32_t out = cid[i];
        keep = true;
        if (j != NONE && A.nn[j] == i) {
            if (i < j) out = merge_pair(A, cid[i], cid[j]);
            else { out = NONE; keep = false; }
        }
        A.newid[i] = out;
    }
    const int kept = __syncthreads_count(keep);
    if (threadIdx.x == 0) A.block_sums[blockIdx.x] = (uint32_t)kept;
}

// order-preserving compaction: offset of the block = sum of the counts of the blocks before it
__global__ void __launch_bounds__(1024) k_scatter(BuildArrays A, int it) {
    uint32_t n;
    if (!iter_active(A, it, n)) return;
    if (blockIdx.x * 1024u >= n) return;
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_off;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    // block offset
    uint32_t part = 0;
    for (uint32_t b = threadIdx.x; b < blockIdx.x; b += 1024) part += A.block_sums[b];
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if (lane == 0) s_warp[warp] = part;
    __syncthreads();
    if (warp == 0) {
        uint32_t v = s_warp[lane];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) s_off = v;
    }
    __syncthreads();
    const uint32_t i = blockIdx.x * 1024 + threadIdx.x;
    const uint32_t id = i < n ? A.newid[i] : NONE;
    const bool keep = id != NONE;
    const uint32_t m = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) s_warp[warp] = (uint32_t)__popc(m);
    __syncthreads();
    uint32_t before = 0;
    for (uint32_t w = 0; w < warp; w++) before += s_warp[w];
    if (keep) A.cid[1 - (it & 1)][s_off + before + (uint32_t)__popc(m & ((1u << lane) - 1u))] = id;
}

__global__ void k_advance(BuildArrays A, int it) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    uint32_t n;
    if (!iter_active(A, it, n)) return;
    uint32_t total = 0;
    const uint32_t nb = (n + 1023u) / 1024u;
    for (uint32_t b = 0; b < nb; b++) total += A.block_sums[b];
    A.ctr->nclus[1 - (it & 1)] = total;
    A.ctr->done_iter = (uint32_t)it + 1u;
}

// ---- the last clusters: one block, no launches in between ----
__global__ void __launch_bounds__(1024) k_tail(BuildArrays A) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry, s_n;
    int p = (int)(A.ctr->done_iter & 1u);
    uint32_t n = A.ctr->nclus[p];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    while (n > 1) {
        const uint32_t* cid = A.cid[p];
        uint32_t* out = A.cid[1 - p];
        for (uint32_t i = threadIdx.x; i < n; i += 1024) {
            const uint32_t id = cid[i];
            const float4 a = A.lo[id], b = A.hi[id];
            float best = 3.0e38f; int best_j = -1;
            for (int d = -PLOC_R; d <= PLOC_R; d++) {
                const int j = (int)i + d;
                if (d == 0 || j < 0 || (uint32_t)j >= n) continue;
                const uint32_t jd = cid[j];
                const float4 l = A.lo[jd], h = A.hi[jd];
                const float ar = half_area(make_float3(fminf(a.x, l.x), fminf(a.y, l.y), fminf(a.z, l.z)),
                                           make_float3(fmaxf(b.x, h.x), fmaxf(b.y, h.y), fmaxf(b.z, h.z)));
                if (ar < best) { best = ar; best_j = j; }
            }
            A.nn[i] = (uint32_t)best_j;
        }
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < n; i += 1024) {
            const uint32_t j = A.nn[i];
            uint32_t o = cid[i];
            if (j != NONE && A.nn[j] == i) o = i < j ? merge_pair(A, cid[i], cid[j]) : NONE;
            A.newid[i] = o;
        }
        if (threadIdx.x == 0) s_carry = 0;
        __syncthreads();
        for (uint32_t c = 0; c < n; c += 1024) {
            const uint32_t i = c + threadIdx.x;
            const uint32_t id = i < n ? A.newid[i] : NONE;
            const bool keep = id != NONE;
            const uint32_t m = __ballot_sync(0xffffffffu, keep);
            if (lane == 0) s_warp[warp] = (uint32_t)__popc(m);
            __syncthreads();
            uint32_t before = 0, total = 0;
            for (uint32_t w = 0; w < 32; w++) { const uint32_t v = s_warp[w]; if (w < warp) before += v; total += v; }
            const uint32_t carry = s_carry;
            if (keep) out[carry + before + (uint32_t)__popc(m & ((1u << lane) - 1u))] = id;
            __syncthreads();
            if (threadIdx.x == 0) s_carry = carry + total;
            __syncthreads();
        }
        if (threadIdx.x == 0) s_n = s_carry;
        __syncthreads();
        n = s_n;
        p = 1 - p;
        __threadfence_block();
        __syncthreads();
    }
    if (threadIdx.x == 0) A.ctr->root = n == 1 ? A.cid[p][0] : 0u;
}

// ---- pre-order records ----
__global__ void __launch_bounds__(256) k_emit(BuildArrays A, BvhNode* __restrict__ out) {
    const uint32_t total = 2u * (uint32_t)A.n - 1u;
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= total) return;
    const uint32_t root = A.ctr->root;
    // pre-order index = sum over the path to the root: +1 for a left child, +2 * leaves(left sibling) for a right child
    uint32_t idx = 0, c = v;
    while (c != root) {
        const uint32_t p = A.parent[c];
        if (p == NONE) return;                                     // not part of the tree (cannot happen for a finished build)
        const uint32_t l = (uint32_t)__float_as_int(A.lo[p].w);
        idx += (l == c) ? 1u : 2u * A.cnt[l];
        c = p;
    }
    const float4 lo = A.lo[v], hi = A.hi[v];
    const int left = __float_as_int(lo.w);
    const uint32_t skip = idx + 2u * A.cnt[v] - 1u;
    BvhNode nd;
    nd.mn[0] = lo.x; nd.mn[1] = lo.y; nd.mn[2] = lo.z;
    nd.mx[0] = hi.x; nd.mx[1] = hi.y; nd.mx[2] = hi.z;
    nd.skip_link = skip >= total ? -1.0f : (float)skip;
    nd.axis = 0.0f;
    if (left < 0) {
        nd.right_offset = -1.0f; nd.prim_type = 2.0f; nd.prim_id = (float)__float_as_int(hi.w); nd.prim_count = 1.0f;
    } else {
        nd.right_offset = (float)(idx + 2u * A.cnt[left]); nd.prim_type = -1.0f; nd.prim_id = -1.0f; nd.prim_count = -1.0f;
    }
    out[idx] = nd;
}

// ---- host side ----
static size_t al(size_t x) { return (x + 255) & ~(size_t)255; }

static size_t sort_temp_bytes(int n) {
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                    (const uint32_t*)nullptr, (uint32_t*)nullptr, n, 0, 63, (cudaStream_t)0);
    return bytes;
}

size_t bvh_build_scratch_bytes(int n) {
    const size_t N = (size_t)(n > 0 ? n : 1), M = 2 * N;
    return al(sizeof(BuildCounters)) + 2 * al(M * 16) + 2 * al(M * 4) + 4 * al(N * 4) + al((N / 1024 + 2) * 4) +
           2 * al(N * 16) + 2 * al(N * 8) + 2 * al(N * 4) + al(sort_temp_bytes(n)) + 4096;
}

void launch_bvh_build(const Triangle* tris, int n, const Mesh* meshes, const ModelTransform* transforms,
                      BvhNode* out_nodes, void* d_scratch, cudaStream_t stream) {
    if (n <= 0) return;
    const size_t N = (size_t)n, M = 2 * N;
    char* p = (char*)d_scratch;
    auto carve = [&](size_t b) { char* r = p; p += al(b); return r; };
    BuildArrays A;
    A.n = n;
    A.ctr = (BuildCounters*)carve(sizeof(BuildCounters));
    A.lo = (float4*)carve(M * 16); A.hi = (float4*)carve(M * 16);
    A.cnt = (uint32_t*)carve(M * 4); A.parent = (uint32_t*)carve(M * 4);
    A.cid[0] = (uint32_t*)carve(N * 4); A.cid[1] = (uint32_t*)carve(N * 4);
    A.nn = (uint32_t*)carve(N * 4); A.newid = (uint32_t*)carve(N * 4);
    A.block_sums = (uint32_t*)carve((N / 1024 + 2) * 4);
    A.plo = (float4*)carve(N * 16); A.phi = (float4*)carve(N * 16);
    A.key[0] = (unsigned long long*)carve(N * 8); A.key[1] = (unsigned long long*)carve(N * 8);
    A.val[0] = (uint32_t*)carve(N * 4); A.val[1] = (uint32_t*)carve(N * 4);
    size_t temp_bytes = sort_temp_bytes(n);
    void* temp = carve(temp_bytes);

    const unsigned b256 = (unsigned)((N + 255) / 256), b1024 = (unsigned)((N + 1023) / 1024);
    k_build_init<<<1, 32, 0, stream>>>(A);
    k_prim_bounds<<<b256, 256, 0, stream>>>(A, tris, meshes, transforms);
    k_morton<<<b256, 256, 0, stream>>>(A);
    cub::DeviceRadixSort::SortPairs(temp, temp_bytes, A.key[0], A.key[1], A.val[0], A.val[1], n, 0, 63, stream);
    k_make_leaves<<<b256, 256, 0, stream>>>(A, A.val[1]);
    if (n > PLOC_TAIL) {
        for (int it = 0; it < PLOC_GLOBAL_ITERS; it++) {
            k_nn<<<b256, 256, 0, stream>>>(A, it);
            k_merge<<<b1024, 1024, 0, stream>>>(A, it);
            k_scatter<<<b1024, 1024, 0, stream>>>(A, it);
            k_advance<<<1, 32, 0, stream>>>(A, it);
        }
    }
    k_tail<<<1, 1024, 0, stream>>>(A);
    k_emit<<<(unsigned)((M + 255) / 256), 256, 0, stream>>>(A, out_nodes);
}

}  // namespace wg

// wide_bvh.cuh -- the 8-wide, quantised, 16-byte-aligned node layout the fast path traverses,
// derived on the GPU (K2, wide_bvh.cu) from the reference's binary 12-float pre-order nodes
// (/root/reference/lib/BVH/bvhBuilder.js:37-54).
//
// WideNode, 80 bytes = 5 x uint4 (compressed-wide-BVH style):
//   q0: origin.x, origin.y, origin.z (f32), {ex, ey, ez, imask} (u8 x 4)
//       child box lo = origin + qlo * 2^(e-127), hi = origin + qhi * 2^(e-127)  (conservative)
//       imask bit s = slot s holds an inner child
//   q1: child_base (u32: wide index of the first inner child; inner children are contiguous in
//       slot order), tri_base (u32: first packed triangle of this node's leaf slots, contiguous
//       in slot order), lmask (u8: bit s = slot s holds ONE triangle), spare
//       A slot is empty, inner, or exactly one triangle (a leaf of the reference's binary tree
//       has one triangle, bvhNode.js:28-73), so the triangle of leaf slot s is
//       tri_base + popc(lmask & ((1 << s) - 1)) and the per-slot work at traversal time is just
//       the slab test (no per-slot meta byte to decode).  Empty slots carry an inverted box.
//   q2: qlo_x[8], qlo_y[8]   q3: qlo_z[8], qhi_x[8]   q4: qhi_y[8], qhi_z[8]
// Slots are assigned so that slot s lies towards the diagonal (s&4 ? +x : -x, s&2 ? +y : -y,
// s&1 ? +z : -z) from the node centre; a ray with positive-direction code r visits slot s at
// priority (s ^ r): the hit byte is permuted into that order by one table lookup.
//
// PackedTri, 48 bytes = 3 x float4: (A.xyz, src), (B.xyz, mesh), (C.xyz, local_id); src = position
// in binding 5 (for the normals), mesh = i32(Triangle.mesh_id), local_id = i32(Triangle.local_id),
// stored as int bit patterns.  A, B, C are the SAME bits as in binding 5 (object space), so
// hit_triangle's arithmetic (common.wgsl:199-219) is reproduced exactly.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>
#include "scene.cuh"

namespace wg {

struct WideBvh {
    const uint4* nodes;       // 5 uint4 per node; node 0 is the root
    const float4* tris;       // 3 float4 per packed triangle
    uint32_t n_nodes, n_tris;
    // The reference's stack guard (hitRay.wgsl:100-104) can fire: binding 9 is at least STACK_SIZE deep, so what
    // the reference returns depends on its own visit order for ANY ray.  Every ray then takes the reference walk.
    uint32_t force_reference_walk;
    // 371 - (largest plane exponent byte of the root node): the per-ray power-of-two distance scale of the
    // mixed-precision slab test (trace_wide.cuh make_wide_ray_scaled)
    uint32_t scale_k;
};

struct WideBuildStats { uint32_t n_nodes, n_tris, n_levels, overflow, root_emax; };

// K2 host entry (wide_bvh.cu): builds into caller-provided device storage of capacity
// max_nodes (>= max(1, NUM_AABB/2)) nodes and NUM_TRIANGLES packed triangles.
// d_scratch: wide_scratch_bytes(max_nodes) bytes (counters, two work queues, the collapse plan of up to 2 * max_nodes binary nodes).
size_t wide_scratch_bytes(uint32_t max_nodes);
// returns the number of kernels it launched
int launch_wide_build(const BvhNode* bvh, int n_bvh, const Triangle* tris, int n_tris, uint4* out_nodes,
                       float4* out_tris, uint32_t max_nodes, uint32_t* d_scratch, cudaStream_t stream);
// reads the counters back (synchronises the stream)
WideBuildStats wide_build_stats(const uint32_t* d_scratch, cudaStream_t stream);

}  // namespace wg

// bvh_build.cuh -- K5: device-side BVH construction for the fast path.
//
// The reference builds its BVH on the host (lib/BVH/bvhNode.js:28-73: median split of the range
// sorted by bbox-min on the longest axis; 0.44-5.75 s in benchmarks.txt) and uploads it as binding
// 9.  Exact mode and the tie resolver walk THAT tree, because its visit order is part of the
// reference's semantics.  The wide-BVH traversal of the fast path only needs *a* hierarchy over the
// same triangles whose boxes contain the reference's leaf boxes (superset search, trace_wide.cuh),
// so it is free to use a better one: this builder clusters the triangles bottom-up by surface
// area (parallel locally-ordered clustering over a Morton-sorted sequence), which on the atrium
// stand-in cuts the wide-node visits per ray from 19.8 to ~12.
//
// Output: the same 48-byte pre-order node records as binding 9 (bvhBuilder.js:37-54: left child =
// i + 1, right_offset, skip_link, prim_type == 2 leaves with prim_id = index into binding 5,
// prim_count == 1), so the flatten kernel K2 (wide_bvh.cu) consumes it unchanged.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>
#include "scene.cuh"

namespace wg {

// bytes of scratch for n_tris triangles
size_t bvh_build_scratch_bytes(int n_tris);
// Builds into out_nodes (2 * n_tris - 1 records).  All work is enqueued on `stream`; nothing is
// read back.  d_scratch must hold bvh_build_scratch_bytes(n_tris) bytes.
void launch_bvh_build(const Triangle* tris, int n_tris, const Mesh* meshes, const ModelTransform* transforms,
                      BvhNode* out_nodes, void* d_scratch, cudaStream_t stream);
#define BVH_BUILD_LAUNCHES 118   /* our kernels per build (the cub radix sort passes are not counted) */

}  // namespace wg

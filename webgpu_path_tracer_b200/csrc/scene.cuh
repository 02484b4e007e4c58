// scene.cuh -- byte layouts of the reference's bind-group buffers as they sit in HBM, and the
// by-value scene descriptor every kernel receives.
// Layouts: /root/reference/shaders/header.wgsl:41-117.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

namespace wg {

struct Uniforms { float screenW, screenH, frameNum, resetBuffer; float viewMatrix[16]; };            // 80 B
struct Sphere   { float cx, cy, cz, r, global_id, local_id, material_id, pad; };                    // 32 B
struct Quad     { float Q[3], pad0, u[3], local_id, v[3], global_id, normal[3], D, w[3], material_id; }; // 80 B
struct Triangle { float A[3], p0, B[3], p1, C[3], p2, nA[3], p3, nB[3], local_id, nC[3], mesh_id; };     // 96 B
struct Mesh     { int32_t num_triangles, offset, global_id, material_id; };                         // 16 B
struct ModelTransform { float modelMatrix[16], invModelMatrix[16]; };                               // 128 B
struct Material { float color[3], p0, specularColor[3], p1, emissionColor[3], specularStrength,
                  roughness, eta, material_type, p2; };                                             // 64 B
struct BvhNode  { float mn[3], right_offset, mx[3], prim_type, prim_id, prim_count, skip_link, axis; }; // 48 B

static_assert(sizeof(Uniforms) == 80 && sizeof(Sphere) == 32 && sizeof(Quad) == 80, "layout");
static_assert(sizeof(Triangle) == 96 && sizeof(Mesh) == 16 && sizeof(ModelTransform) == 128, "layout");
static_assert(sizeof(Material) == 64 && sizeof(BvhNode) == 48, "layout");

// WGSL constants of header.wgsl:1-13 that are run-time here.
struct Consts {
    int NUM_SAMPLES, MAX_BOUNCES, STRATIFY, IMPORTANCE_SAMPLING, STACK_SIZE;
};

// Which pixels this context renders (tile split of SURVEY.md §8(e)); world == 1 -> all.
struct Shard {
    int tile_rank, tile_world, tile_rows;
    uint32_t n_pixels;          // shard-local indices to run (pixels of this shard)
    uint32_t pixel_limit;       // invocations dispatched: workgroups * 64 (main.wgsl:3-4); pixels >= it are not run
};

struct DeviceScene {
    Uniforms uniforms;
    float aspect;                   // screenW / screenH as one IEEE f32 division (shootRay.wgsl:36), set with the uniforms
    Consts K;
    Shard shard;
    int W, H;
    const Sphere* spheres;          int NUM_SPHERES;
    const Quad* quads;              int NUM_QUADS;
    const Triangle* triangles;      int NUM_TRIANGLES;
    const Mesh* meshes;             int NUM_MESHES;
    const ModelTransform* transforms; int NUM_TRANSFORMS;
    const Material* materials;      int NUM_MATERIALS;
    const BvhNode* bvh;             int NUM_AABB;
    Quad lights;                    // get_lights(), common.wgsl:258-269 (invocation-invariant)
    unsigned long long* counters;   // 8 x u64 or nullptr
};

// shard-local linear index -> global pixel index (>= W*H means "masked")
__host__ __device__ inline uint32_t shard_pixel(const Shard& s, int W, uint32_t i) {
    if (s.tile_world <= 1) return i;
    const uint32_t per_block = (uint32_t)s.tile_rows * (uint32_t)W;
    const uint32_t jb = i / per_block, within = i - jb * per_block;
    return (jb * (uint32_t)s.tile_world + (uint32_t)s.tile_rank) * per_block + within;
}

}  // namespace wg

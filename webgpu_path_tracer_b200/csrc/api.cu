// api.cu -- the C ABI of include/wgpt.h: context, buffer upload, dispatch, present, parity hooks.
// Mirrors the call sequence of /root/reference/renderer.js:68-124 (buffers + bind group),
// :163-191 (per-frame uniform write, compute pass, render pass) and webgpu-utils.js:15-55,125-148.
#include <stdio.h>
#include <string.h>
#include <new>
#include "context.h"
#include "kernels.cuh"

using namespace wg;

static std::string g_create_err;

namespace wg {

int set_err(wgpt_ctx* c, int code, const std::string& msg) {
    if (c) c->err = msg; else g_create_err = msg;
    return code;
}
int cuda_err(wgpt_ctx* c, cudaError_t e, const char* what) {
    if (e == cudaSuccess) return WGPT_OK;
    return set_err(c, WGPT_E_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

static size_t stride_of(int binding) {
    switch (binding) {
        case 0: return 80; case 1: return 32; case 2: return 80; case 3: return 16; case 5: return 96;
        case 6: return 16; case 7: return 128; case 8: return 64; case 9: return 48; default: return 0;
    }
}

// get_lights(), common.wgsl:258-269: first quad whose material has emission.x > 0 (zeros if none)
static void find_lights(wgpt_ctx* c) {
    memset(&c->lights, 0, sizeof(Quad));
    const size_t nq = c->hbuf[2].size() / sizeof(Quad), nm = c->hbuf[8].size() / sizeof(Material);
    const Quad* q = (const Quad*)c->hbuf[2].data();
    const Material* m = (const Material*)c->hbuf[8].data();
    for (size_t i = 0; i < nq; i++) {
        const int mi = (int)q[i].material_id;
        if (mi >= 0 && (size_t)mi < nm && m[mi].emissionColor[0] > 0.0f) { c->lights = q[i]; break; }
    }
    c->lights_valid = true;
}

DeviceScene make_device_scene(wgpt_ctx* c, uint32_t pixel_limit) {
    DeviceScene S;
    memset(&S, 0, sizeof(S));
    S.uniforms = c->uniforms;
    { volatile float w = c->uniforms.screenW, h = c->uniforms.screenH; S.aspect = w / h; }      // one IEEE binary32 division (SSE divss)
    S.K.NUM_SAMPLES = c->cfg.num_samples; S.K.MAX_BOUNCES = c->cfg.max_bounces; S.K.STRATIFY = c->cfg.stratify;
    S.K.IMPORTANCE_SAMPLING = c->cfg.importance_sampling; S.K.STACK_SIZE = c->cfg.stack_size;
    S.shard.tile_rank = c->cfg.tile_rank; S.shard.tile_world = c->cfg.tile_world; S.shard.tile_rows = c->cfg.tile_rows;
    S.shard.n_pixels = c->shard_pixels; S.shard.pixel_limit = pixel_limit;
    S.W = c->cfg.width; S.H = c->cfg.height;
    S.spheres = (const Sphere*)c->dbuf[1];            S.NUM_SPHERES = (int)(c->dbytes[1] / 32);
    S.quads = (const Quad*)c->dbuf[2];                S.NUM_QUADS = (int)(c->dbytes[2] / 80);
    S.triangles = (const Triangle*)c->dbuf[5];        S.NUM_TRIANGLES = (int)(c->dbytes[5] / 96);
    S.meshes = (const Mesh*)c->dbuf[6];               S.NUM_MESHES = (int)(c->dbytes[6] / 16);
    S.transforms = (const ModelTransform*)c->dbuf[7]; S.NUM_TRANSFORMS = (int)(c->dbytes[7] / 128);
    S.materials = (const Material*)c->dbuf[8];        S.NUM_MATERIALS = (int)(c->dbytes[8] / 64);
    S.bvh = (const BvhNode*)c->dbuf[9];               S.NUM_AABB = (int)(c->dbytes[9] / 48);
    if (!c->lights_valid) find_lights(c);
    S.lights = c->lights;
    S.counters = c->counters_on ? c->d_counters : nullptr;
    return S;
}

// index ranges the kernels dereference without bounds checks must be valid before a launch
int validate_scene(wgpt_ctx* c) {
    const size_t nm = c->hbuf[8].size() / 64, nt = c->hbuf[7].size() / 128, nmesh = c->hbuf[6].size() / 16;
    const Sphere* s = (const Sphere*)c->hbuf[1].data();
    for (size_t i = 0; i < c->hbuf[1].size() / 32; i++)
        if (!(s[i].material_id >= 0.0f) || (size_t)s[i].material_id >= nm) return set_err(c, WGPT_E_INVALID, "sphere material_id out of range");
    const Quad* q = (const Quad*)c->hbuf[2].data();
    for (size_t i = 0; i < c->hbuf[2].size() / 80; i++)
        if (!(q[i].material_id >= 0.0f) || (size_t)q[i].material_id >= nm) return set_err(c, WGPT_E_INVALID, "quad material_id out of range");
    const Mesh* m = (const Mesh*)c->hbuf[6].data();
    for (size_t i = 0; i < nmesh; i++)
        if (m[i].global_id < 0 || (size_t)m[i].global_id >= nt || m[i].material_id < 0 || (size_t)m[i].material_id >= nm)
            return set_err(c, WGPT_E_INVALID, "mesh global_id / material_id out of range");
    if (c->tri_bad || c->tri_mesh_max >= (int)nmesh) return set_err(c, WGPT_E_INVALID, "triangle mesh_id out of range");
    if (c->bvh_bad || c->bvh_prim_end_max > (long long)(c->dbytes[5] / 96)) return set_err(c, WGPT_E_INVALID, "bvh node links out of range");
    if (c->dbytes[5] && !nmesh) return set_err(c, WGPT_E_STATE, "triangles bound without meshes");
    if ((c->dbytes[5] == 0) != (c->dbytes[9] == 0)) return set_err(c, WGPT_E_STATE, "triangles (5) and bvh (9) must be bound together");
    return WGPT_OK;
}

int ready(wgpt_ctx* c) {
    if (!c) return WGPT_E_INVALID;
    if (!c->have_uniforms) return set_err(c, WGPT_E_STATE, "uniforms (binding 0) not written");
    if (c->uniforms.screenW != (float)c->cfg.width || c->uniforms.screenH != (float)c->cfg.height)
        return set_err(c, WGPT_E_INVALID, "uniforms.screenDims does not match the context size");
    return validate_scene(c);
}

}  // namespace wg

extern "C" {

void wgpt_default_config(wgpt_config* cfg) {
    memset(cfg, 0, sizeof(*cfg));
    cfg->struct_size = sizeof(wgpt_config);
    cfg->width = 512; cfg->height = 512; cfg->device = 0; cfg->mode = WGPT_MODE_EXACT;
    cfg->num_samples = 1; cfg->max_bounces = 100; cfg->stratify = 0; cfg->importance_sampling = 0; cfg->stack_size = 20;
    cfg->tile_rank = 0; cfg->tile_world = 1; cfg->tile_rows = 8; cfg->strata_rank = 0; cfg->strata_world = 1;
}

const char* wgpt_last_error(const wgpt_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

int wgpt_create(const wgpt_config* cfg, wgpt_ctx** out) {
    if (!cfg || !out || cfg->struct_size != sizeof(wgpt_config)) return set_err(nullptr, WGPT_E_INVALID, "bad config pointer or struct_size");
    if (cfg->width <= 0 || cfg->height <= 0 || (uint64_t)cfg->width * cfg->height > (1u << 24))
        return set_err(nullptr, WGPT_E_INVALID, "width*height must be in 1..2^24 (f32(pixelIndex) must be exact, main.wgsl:5)");
    if (cfg->num_samples < 1 || cfg->max_bounces < 0 || cfg->stack_size < 1 || cfg->stack_size > 64)
        return set_err(nullptr, WGPT_E_INVALID, "num_samples >= 1, max_bounces >= 0, 1 <= stack_size <= 64");
    if (cfg->tile_world < 1 || cfg->tile_rank < 0 || cfg->tile_rank >= cfg->tile_world || cfg->tile_rows < 1 ||
        cfg->strata_world < 1 || cfg->strata_rank < 0 || cfg->strata_rank >= cfg->strata_world)
        return set_err(nullptr, WGPT_E_INVALID, "bad shard description");
    if (cfg->mode != WGPT_MODE_EXACT && cfg->mode != WGPT_MODE_FAST && cfg->mode != WGPT_MODE_FAST_TOL) return set_err(nullptr, WGPT_E_INVALID, "unknown mode");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return set_err(nullptr, WGPT_E_CUDA, std::string("no CUDA device (there is no CPU fallback): ") + cudaGetErrorString(e));
    if (cfg->device < 0 || cfg->device >= ndev) return set_err(nullptr, WGPT_E_INVALID, "device ordinal out of range");
    if ((e = cudaSetDevice(cfg->device)) != cudaSuccess) return cuda_err(nullptr, e, "cudaSetDevice");

    wgpt_ctx* c = new (std::nothrow) wgpt_ctx();
    if (!c) return set_err(nullptr, WGPT_E_NOMEM, "out of host memory");
    c->cfg = *cfg;
    // pixels of this shard
    const uint32_t W = cfg->width, H = cfg->height;
    if (cfg->tile_world == 1) c->shard_pixels = W * H;
    else {
        const uint32_t nblocks = (H + cfg->tile_rows - 1) / cfg->tile_rows;
        const uint32_t mine = nblocks > (uint32_t)cfg->tile_rank ? (nblocks - 1 - cfg->tile_rank) / cfg->tile_world + 1 : 0;
        c->shard_pixels = mine * cfg->tile_rows * W;      // the last block may run past H: masked in the kernel
    }
    const size_t fb_bytes = (size_t)W * H * 16;
    const size_t npix = (size_t)W * H;
    if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaEventCreate(&c->ev0)) != cudaSuccess || (e = cudaEventCreate(&c->ev1)) != cudaSuccess ||
        (e = cudaMalloc(&c->dbuf[3], fb_bytes)) != cudaSuccess ||
        (e = cudaMemsetAsync(c->dbuf[3], 0, fb_bytes, c->stream)) != cudaSuccess ||
        (e = cudaMalloc((void**)&c->d_counters, 8 * sizeof(unsigned long long))) != cudaSuccess ||
        (e = cudaMemsetAsync(c->d_counters, 0, 64, c->stream)) != cudaSuccess ||
        (e = cudaMalloc((void**)&c->d_radiance, npix * 12)) != cudaSuccess ||
        (e = cudaMalloc((void**)&c->d_ptype, npix * 4)) != cudaSuccess ||
        (e = cudaMalloc((void**)&c->d_pindex, npix * 4)) != cudaSuccess ||
        (e = cudaMalloc((void**)&c->d_pt, npix * 4)) != cudaSuccess) {
        cuda_err(nullptr, e, "wgpt_create");
        wgpt_destroy(c);
        return e == cudaErrorMemoryAllocation ? WGPT_E_NOMEM : WGPT_E_CUDA;
    }
    c->dbytes[3] = fb_bytes;
    *out = c;
    return WGPT_OK;
}

void wgpt_destroy(wgpt_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->cfg.device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    mgpu_destroy(c);
    fast_destroy(c);
    for (int i = 0; i < 10; i++) if (c->dbuf[i]) cudaFree(c->dbuf[i]);
    if (c->d_counters) cudaFree(c->d_counters);
    if (c->d_radiance) cudaFree(c->d_radiance);
    if (c->d_ptype) cudaFree(c->d_ptype);
    if (c->d_pindex) cudaFree(c->d_pindex);
    if (c->d_pt) cudaFree(c->d_pt);
    if (c->d_prng) cudaFree(c->d_prng);
    for (void* q : c->d_present) if (q) cudaFree(q);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->tm_ev[0]) cudaEventDestroy(c->tm_ev[0]);
    if (c->tm_ev[1]) cudaEventDestroy(c->tm_ev[1]);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int wgpt_write_buffer(wgpt_ctx* c, int binding, const void* data, size_t bytes) {
    if (!c) return WGPT_E_INVALID;
    const size_t stride = stride_of(binding);
    if (!stride) return set_err(c, WGPT_E_INVALID, "unknown binding");
    if (!data && bytes) return set_err(c, WGPT_E_INVALID, "null data");
    if (bytes % stride) return set_err(c, WGPT_E_INVALID, "size is not a multiple of the element stride");
    cudaSetDevice(c->cfg.device);
    if (binding == 0) {
        if (bytes != 80) return set_err(c, WGPT_E_INVALID, "uniforms must be 80 bytes");
        memcpy(&c->uniforms, data, 80);
        c->have_uniforms = true;
        return WGPT_OK;
    }
    cudaError_t e;
    {   // frames already dispatched were dispatched with the OLD contents: finish them, and start a new wavefront afterwards
        const int rc = fast_close(c);
        if (rc) return rc;
    }
    if (binding == 3) {
        if (bytes != c->dbytes[3]) return set_err(c, WGPT_E_INVALID, "framebuffer must be width*height*16 bytes");
        if ((e = cudaMemcpyAsync(c->dbuf[3], data, bytes, cudaMemcpyHostToDevice, c->stream)) != cudaSuccess) return cuda_err(c, e, "write framebuffer");
        return cuda_err(c, cudaStreamSynchronize(c->stream), "write framebuffer");
    }
    // in-flight kernels may still read the old copy
    if ((e = cudaStreamSynchronize(c->stream)) != cudaSuccess) return cuda_err(c, e, "sync before write");
    if (bytes != c->dbytes[binding]) {
        if (c->dbuf[binding]) cudaFree(c->dbuf[binding]);
        c->dbuf[binding] = nullptr; c->dbytes[binding] = 0;
        if (bytes && (e = cudaMalloc(&c->dbuf[binding], bytes)) != cudaSuccess) return cuda_err(c, e, "cudaMalloc");
        c->dbytes[binding] = bytes;
    }
    if (bytes && (e = cudaMemcpy(c->dbuf[binding], data, bytes, cudaMemcpyHostToDevice)) != cudaSuccess) return cuda_err(c, e, "upload");
    if (binding == 5) {
        const Triangle* t = (const Triangle*)data;
        c->tri_mesh_max = -1; c->tri_bad = false;
        for (size_t i = 0; i < bytes / 96; i++) {
            if (!(t[i].mesh_id >= 0.0f) || t[i].mesh_id > 1e6f) { c->tri_bad = true; break; }
            if ((int)t[i].mesh_id > c->tri_mesh_max) c->tri_mesh_max = (int)t[i].mesh_id;
        }
    }
    if (binding == 9) {
        const BvhNode* b = (const BvhNode*)data;
        const long long n = (long long)(bytes / 48);
        c->bvh_prim_end_max = 0; c->bvh_bad = false; c->bvh_depth = 0;
        std::vector<unsigned short> depth((size_t)n, 0);     // root = 0; the reference's stack never holds more than the leaf depth
        for (long long i = 0; i < n; i++) {
            if (depth[i] > c->bvh_depth) c->bvh_depth = depth[i];
            if ((int)b[i].prim_type == 2) {
                const long long s0 = (long long)b[i].prim_id, cnt = (long long)b[i].prim_count;
                if (s0 < 0 || cnt < 0) { c->bvh_bad = true; break; }
                if (s0 + cnt > c->bvh_prim_end_max) c->bvh_prim_end_max = s0 + cnt;
            } else {
                const long long r = (long long)b[i].right_offset;
                const int ax = (int)b[i].axis;
                if (r <= i || r >= n || i + 1 >= n || ax < 0 || ax > 2) { c->bvh_bad = true; break; }
                depth[i + 1] = depth[r] = (unsigned short)(depth[i] + 1);
            }
        }
    }
    if (binding != 5 && binding != 9) c->hbuf[binding].assign((const unsigned char*)data, (const unsigned char*)data + bytes);
    if (binding == 2 || binding == 8) c->lights_valid = false;
    if (binding == 5 || binding == 9 || binding == 6 || binding == 7) c->wide_dirty = true;
    return WGPT_OK;
}

}  // extern "C"

// mgpu.cu -- multi-GPU inside the product: the scene is replicated, work is split by image tiles
// and frame strata (wgpt_config), and the partial accumulation buffers (binding 3) are combined by ONE ncclReduce(sum)
// that this library issues itself, on the context's own stream, i.e. ordered after the render kernels without any
// host synchronisation.  main.wgsl:16,27 make every (pixel, frame) independent, so no data-path collective exists
// inside a frame.
//
// Two ways to drive it:
//   * one process per GPU (torchrun / MPI): rank 0 calls wgpt_nccl_unique_id, the host layer ships the 128 bytes to
//     the other ranks by whatever plumbing it has (torch.distributed, MPI, a file), every rank calls wgpt_comm_init
//     on its context, and then wgpt_reduce_framebuffer after its renders;
//   * one process, N devices: wgpt_group_* owns one context per device and one host thread per context for the calls
//     that block (uploads, renders), and reduces with ncclGroupStart / ncclGroupEnd over its communicators.
// NCCL is resolved at run time (dlopen of libnccl.so.2; WGPT_NCCL_LIB overrides the name): the library itself links
// only cudart and still loads where NCCL is absent -- the multi-GPU entry points then fail with WGPT_E_STATE.
#include <dlfcn.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <thread>
#include <vector>
#include "context.h"
#include "kernels.cuh"

using namespace wg;

namespace {

// the handful of NCCL declarations this file needs (nccl.h: ncclUniqueId is 128 opaque bytes, ncclFloat32 = 7, ncclSum = 0)
struct NcclId { char internal[128]; };
typedef void* NcclComm;
struct NcclApi {
    void* lib = nullptr;
    int (*GetUniqueId)(NcclId*) = nullptr;
    int (*CommInitRank)(NcclComm*, int, NcclId, int) = nullptr;
    int (*CommInitAll)(NcclComm*, int, const int*) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*Reduce)(const void*, void*, size_t, int, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    std::string why;
};
const int kNcclFloat32 = 7, kNcclSum = 0;

NcclApi* nccl() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return &api;
    tried = true;
    const char* names[] = {getenv("WGPT_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
        if (!n || !*n) continue;
        api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (api.lib) break;
        api.why = dlerror();
    }
    if (!api.lib) return &api;
    bool ok = true;
    auto sym = [&](const char* name) { void* p = dlsym(api.lib, name); if (!p) { ok = false; api.why = std::string("missing symbol ") + name; } return p; };
    api.GetUniqueId = (int (*)(NcclId*))sym("ncclGetUniqueId");
    api.CommInitRank = (int (*)(NcclComm*, int, NcclId, int))sym("ncclCommInitRank");
    api.CommInitAll = (int (*)(NcclComm*, int, const int*))sym("ncclCommInitAll");
    api.CommDestroy = (int (*)(NcclComm))sym("ncclCommDestroy");
    api.Reduce = (int (*)(const void*, void*, size_t, int, int, int, NcclComm, cudaStream_t))sym("ncclReduce");
    api.AllReduce = (int (*)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t))sym("ncclAllReduce");
    api.GroupStart = (int (*)())sym("ncclGroupStart");
    api.GroupEnd = (int (*)())sym("ncclGroupEnd");
    api.GetErrorString = (const char* (*)(int))sym("ncclGetErrorString");
    if (!ok) { dlclose(api.lib); api.lib = nullptr; }
    return &api;
}

int nccl_err(wgpt_ctx* c, int rc, const char* what) {
    if (rc == 0) return WGPT_OK;
    NcclApi* A = nccl();
    return set_err(c, WGPT_E_CUDA, std::string(what) + ": NCCL " + (A->GetErrorString ? A->GetErrorString(rc) : "error"));
}

int need_nccl(wgpt_ctx* c) {
    NcclApi* A = nccl();
    if (A->lib) return WGPT_OK;
    return set_err(c, WGPT_E_STATE, "NCCL not available (dlopen libnccl.so.2): " + A->why);
}

}  // namespace

namespace wg {
void mgpu_destroy(wgpt_ctx* c) {
    if (c->nccl_comm && nccl()->lib) nccl()->CommDestroy((NcclComm)c->nccl_comm);
    c->nccl_comm = nullptr;
}
}

extern "C" {

int wgpt_nccl_unique_id(void* id128) {
    if (!id128) return WGPT_E_INVALID;
    int rc = need_nccl(nullptr);
    if (rc) return rc;
    NcclId id;
    if ((rc = nccl()->GetUniqueId(&id)) != 0) return nccl_err(nullptr, rc, "ncclGetUniqueId");
    memcpy(id128, id.internal, 128);
    return WGPT_OK;
}

int wgpt_comm_init(wgpt_ctx* c, const void* id128, int rank, int world) {
    if (!c || !id128 || world < 1 || rank < 0 || rank >= world) return WGPT_E_INVALID;
    int rc = need_nccl(c);
    if (rc) return rc;
    cudaSetDevice(c->cfg.device);
    mgpu_destroy(c);
    NcclId id;
    memcpy(id.internal, id128, 128);
    NcclComm comm = nullptr;
    if ((rc = nccl()->CommInitRank(&comm, world, id, rank)) != 0) return nccl_err(c, rc, "ncclCommInitRank");
    c->nccl_comm = comm; c->nccl_rank = rank; c->nccl_world = world;
    return WGPT_OK;
}

// fix_alpha: launch the w-lane fix behind the reduce (false inside ncclGroupStart / End, where the reduce itself is only
// issued at GroupEnd: the caller launches the fix afterwards)
static int reduce_framebuffer(wgpt_ctx* c, int root, bool fix_alpha) {
    if (!c) return WGPT_E_INVALID;
    if (!c->nccl_comm) return set_err(c, WGPT_E_STATE, "wgpt_comm_init has not been called on this context");
    if (root < -1 || root >= c->nccl_world) return set_err(c, WGPT_E_INVALID, "root out of range (-1 = all ranks get the sum)");
    cudaSetDevice(c->cfg.device);
    // every frame submitted so far is enqueued on the stream first; the reduce then runs behind them on the SAME stream
    int rc = fast_flush(c);
    if (rc) return rc;
    const size_t n = c->dbytes[3] / 4;
    if (root < 0) rc = nccl()->AllReduce(c->dbuf[3], c->dbuf[3], n, kNcclFloat32, kNcclSum, (NcclComm)c->nccl_comm, c->stream);
    else rc = nccl()->Reduce(c->dbuf[3], c->dbuf[3], n, kNcclFloat32, kNcclSum, root, (NcclComm)c->nccl_comm, c->stream);
    if (rc) return nccl_err(c, rc, "ncclReduce(binding 3)");
    // the w lane now counts the shards that rendered the pixel: back to the reference's 1.0 where the sum landed
    if (fix_alpha && (root < 0 || root == c->nccl_rank)) launch_fix_alpha((float4*)c->dbuf[3], (uint32_t)(c->dbytes[3] / 16), c->stream);
    return cuda_err(c, cudaGetLastError(), "reduce_framebuffer");
}

int wgpt_reduce_framebuffer(wgpt_ctx* c, int root) { return reduce_framebuffer(c, root, true); }

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// wgpt_group: one process, N devices.  Rank r of the group renders tile shard (r % tile_groups) of strata shard
// (r / tile_groups): wgpt_config.tile_* / strata_* are filled in per context.
struct wgpt_group {
    std::vector<wgpt_ctx*> ctx;
    int tile_groups = 1;
    std::string err;
};

namespace {

// runs fn(rank) on one host thread per context (the calls block on their device: uploads, renders with their flag polling)
template <class F>
int for_each_ctx(wgpt_group* g, F fn) {
    std::vector<int> rc(g->ctx.size(), 0);
    std::vector<std::thread> th;
    for (size_t i = 1; i < g->ctx.size(); i++) th.emplace_back([&, i] { rc[i] = fn((int)i); });
    rc[0] = fn(0);
    for (auto& t : th) t.join();
    for (size_t i = 0; i < rc.size(); i++)
        if (rc[i]) { g->err = "rank " + std::to_string(i) + ": " + wgpt_last_error(g->ctx[i]); return rc[i]; }
    return WGPT_OK;
}

}  // namespace

extern "C" {

const char* wgpt_group_last_error(const wgpt_group* g) { return g ? g->err.c_str() : wgpt_last_error(nullptr); }

void wgpt_group_destroy(wgpt_group* g) {
    if (!g) return;
    for (wgpt_ctx* c : g->ctx) wgpt_destroy(c);
    delete g;
}

int wgpt_group_create(const wgpt_config* cfg, int n_devices, const int* devices, int tile_groups, wgpt_group** out) {
    if (!cfg || !out || n_devices < 1 || tile_groups < 1 || n_devices % tile_groups) return WGPT_E_INVALID;
    int rc = n_devices > 1 ? need_nccl(nullptr) : WGPT_OK;
    if (rc) return rc;
    wgpt_group* g = new wgpt_group();
    g->tile_groups = tile_groups;
    std::vector<int> devs(n_devices);
    for (int r = 0; r < n_devices; r++) {
        wgpt_config k = *cfg;
        k.device = devs[r] = devices ? devices[r] : r;
        k.tile_rank = r % tile_groups; k.tile_world = tile_groups;
        k.strata_rank = r / tile_groups; k.strata_world = n_devices / tile_groups;
        wgpt_ctx* c = nullptr;
        if ((rc = wgpt_create(&k, &c)) != WGPT_OK) { wgpt_group_destroy(g); return rc; }
        g->ctx.push_back(c);
    }
    if (n_devices > 1) {
        std::vector<NcclComm> comms(n_devices, nullptr);
        if ((rc = nccl()->CommInitAll(comms.data(), n_devices, devs.data())) != 0) {
            rc = nccl_err(nullptr, rc, "ncclCommInitAll");
            wgpt_group_destroy(g);
            return rc;
        }
        for (int r = 0; r < n_devices; r++) { g->ctx[r]->nccl_comm = comms[r]; g->ctx[r]->nccl_rank = r; g->ctx[r]->nccl_world = n_devices; }
    }
    *out = g;
    return WGPT_OK;
}

int wgpt_group_size(const wgpt_group* g) { return g ? (int)g->ctx.size() : 0; }
wgpt_ctx* wgpt_group_ctx(wgpt_group* g, int rank) { return g && rank >= 0 && rank < (int)g->ctx.size() ? g->ctx[rank] : nullptr; }

// the scene is replicated: every context gets its own copy (binding 3 included)
int wgpt_group_write_buffer(wgpt_group* g, int binding, const void* data, size_t bytes) {
    if (!g) return WGPT_E_INVALID;
    return for_each_ctx(g, [&](int r) { return wgpt_write_buffer(g->ctx[r], binding, data, bytes); });
}

// every context renders ITS frames of [first_frame, first_frame + n_frames) on ITS tiles, concurrently
int wgpt_group_render(wgpt_group* g, int first_frame, int n_frames) {
    if (!g) return WGPT_E_INVALID;
    return for_each_ctx(g, [&](int r) { return wgpt_render(g->ctx[r], first_frame, n_frames); });
}

// sum of the shards' accumulation buffers on `root` (each communicator's reduce is enqueued on its context's stream,
// behind that context's render kernels); returns when the sum is in root's binding 3
int wgpt_group_reduce(wgpt_group* g, int root) {
    if (!g || root < 0 || root >= (int)g->ctx.size()) return WGPT_E_INVALID;
    int rc = WGPT_OK;
    if (g->ctx.size() > 1) {
        nccl()->GroupStart();
        for (size_t r = 0; r < g->ctx.size() && rc == WGPT_OK; r++) rc = reduce_framebuffer(g->ctx[r], root, false);
        const int ge = nccl()->GroupEnd();
        if (rc == WGPT_OK && ge) rc = nccl_err(g->ctx[root], ge, "ncclGroupEnd");
        if (rc) { g->err = wgpt_last_error(g->ctx[root]); return rc; }
        // the reduces were issued at GroupEnd: NOW the w lane can go back to 1.0 behind root's reduce
        wgpt_ctx* rc_ctx = g->ctx[root];
        cudaSetDevice(rc_ctx->cfg.device);
        launch_fix_alpha((float4*)rc_ctx->dbuf[3], (uint32_t)(rc_ctx->dbytes[3] / 16), rc_ctx->stream);
    }
    return for_each_ctx(g, [&](int r) { return wgpt_synchronize(g->ctx[r]); });
}

}  // extern "C"

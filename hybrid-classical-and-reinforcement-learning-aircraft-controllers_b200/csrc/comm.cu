// comm.cu -- the ONE collective of the path, reachable from the C ABI: the all-reduce of the 16-double episode / mission
// statistics over the ranks of a sharded fleet (NCCL over NVLink 5 / NVSwitch, once per report interval,
// latency-bound; no per-step communication).  One process per GPU: every rank owns a context; rank 0 draws an NCCL unique
// id (acb200_comm_unique_id), the caller's own plumbing carries its 128 bytes to the other ranks (torch.distributed, MPI, a
// file...), every rank calls acb200_comm_init, and acb200_episode_stats then enqueues ncclAllReduce on the context's stream.
//
// NCCL is resolved at run time (dlopen), not linked: the library has no NCCL dependency for single-GPU users, and in a
// process that already carries an NCCL (PyTorch bundles one) that copy is the one used.  Order: $ACB200_NCCL_LIB, an
// already-loaded libnccl.so.2 (RTLD_NOLOAD), then the loader's search path.
#include <dlfcn.h>
#include <stdlib.h>
#include <string.h>

#include <string>

#include "ctx.h"

namespace {

// the slice of nccl.h this file needs (stable across NCCL 2.x)
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;                       // 0 = ncclSuccess
enum { kNcclFloat64 = 8, kNcclSum = 0 };        // ncclDataType_t ncclFloat64 / ncclDouble, ncclRedOp_t ncclSum

struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
    std::string origin, error;
};

NcclApi* nccl() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return &api;
    tried = true;
    const char* env = getenv("ACB200_NCCL_LIB");
    if (env && *env) { api.handle = dlopen(env, RTLD_NOW | RTLD_GLOBAL); api.origin = env; }
    if (!api.handle) { api.handle = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD); api.origin = "libnccl.so.2 (already loaded in this process)"; }
    if (!api.handle) { api.handle = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL); api.origin = "libnccl.so.2 (loader search path)"; }
    if (!api.handle) { api.error = std::string("NCCL not found: ") + (dlerror() ? dlerror() : "dlopen failed") + " (set ACB200_NCCL_LIB)"; return &api; }
    api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(api.handle, "ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))dlsym(api.handle, "ncclCommInitRank");
    api.AllReduce = (decltype(api.AllReduce))dlsym(api.handle, "ncclAllReduce");
    api.CommDestroy = (decltype(api.CommDestroy))dlsym(api.handle, "ncclCommDestroy");
    api.GetErrorString = (decltype(api.GetErrorString))dlsym(api.handle, "ncclGetErrorString");
    api.GetVersion = (decltype(api.GetVersion))dlsym(api.handle, "ncclGetVersion");
    if (!api.GetUniqueId || !api.CommInitRank || !api.AllReduce || !api.CommDestroy || !api.GetErrorString) {
        api.error = "NCCL symbols missing in " + api.origin;
        api.handle = nullptr;
    }
    return &api;
}

int cfail(acb200_ctx* ctx, int code, const std::string& msg) {
    if (ctx) ctx->error = msg;
    return code;
}
int nccl_fail(acb200_ctx* ctx, NcclApi* a, ncclResult_t r, const char* where) {
    return cfail(ctx, ACB200_ECUDA, std::string(where) + ": " + (a->GetErrorString ? a->GetErrorString(r) : "NCCL error"));
}

}  // namespace

extern "C" {

int acb200_comm_unique_id(acb200_ctx* ctx, void* id_out) {
    if (!ctx || !id_out) return ACB200_ENULL;
    NcclApi* a = nccl();
    if (!a->handle) return cfail(ctx, ACB200_ENOSUP, a->error);
    static_assert(sizeof(ncclUniqueId) == ACB200_COMM_ID_BYTES, "NCCL unique id is 128 bytes");
    ncclUniqueId id;
    const ncclResult_t r = a->GetUniqueId(&id);
    if (r != 0) return nccl_fail(ctx, a, r, "ncclGetUniqueId");
    memcpy(id_out, &id, sizeof id);
    return ACB200_OK;
}

int acb200_comm_init(acb200_ctx* ctx, int world, int rank, const void* id) {
    if (!ctx || !id) return ACB200_ENULL;
    if (world < 1 || rank < 0 || rank >= world) return cfail(ctx, ACB200_EVALUE, "comm_init: rank outside the world");
    if (ctx->comm) return cfail(ctx, ACB200_EVALUE, "comm_init: this context already has a communicator (acb200_comm_destroy first)");
    NcclApi* a = nccl();
    if (!a->handle) return cfail(ctx, ACB200_ENOSUP, a->error);
    cudaSetDevice(ctx->device);
    ncclUniqueId uid;
    memcpy(&uid, id, sizeof uid);
    ncclComm_t comm = nullptr;
    const ncclResult_t r = a->CommInitRank(&comm, world, uid, rank);        // collective: every rank calls it
    if (r != 0) return nccl_fail(ctx, a, r, "ncclCommInitRank");
    ctx->comm = comm; ctx->comm_world = world; ctx->comm_rank = rank;
    return ACB200_OK;
}

int acb200_comm_destroy(acb200_ctx* ctx) {
    if (!ctx) return ACB200_ENULL;
    if (!ctx->comm) return ACB200_OK;
    NcclApi* a = nccl();
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    const ncclResult_t r = a->CommDestroy((ncclComm_t)ctx->comm);
    ctx->comm = nullptr; ctx->comm_world = 0; ctx->comm_rank = 0;
    return r == 0 ? ACB200_OK : nccl_fail(ctx, a, r, "ncclCommDestroy");
}

int acb200_comm_info(const acb200_ctx* ctx, int* world, int* rank, int* nccl_version) {
    if (!ctx) return ACB200_ENULL;
    if (world) *world = ctx->comm ? ctx->comm_world : 0;
    if (rank) *rank = ctx->comm ? ctx->comm_rank : 0;
    if (nccl_version) {
        *nccl_version = 0;
        NcclApi* a = nccl();
        if (a->handle && a->GetVersion) a->GetVersion(nccl_version);
    }
    return ACB200_OK;
}

int acb200_episode_stats(acb200_ctx* ctx, const double* stats, double* out, int32_t count, int32_t do_allreduce) {
    if (!ctx) return ACB200_ENULL;
    if (!stats || !out) return cfail(ctx, ACB200_ENULL, "episode_stats: stats / out is NULL");
    if (count < 1 || count > 1024) return cfail(ctx, ACB200_EVALUE, "episode_stats: count outside [1, 1024]");
    cudaSetDevice(ctx->device);
    if (!do_allreduce) {
        if (out == stats) return ACB200_OK;
        const cudaError_t e = cudaMemcpyAsync(out, stats, (size_t)count * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
        return e == cudaSuccess ? ACB200_OK : cfail(ctx, ACB200_ECUDA, std::string("episode_stats copy: ") + cudaGetErrorString(e));
    }
    if (!ctx->comm) return cfail(ctx, ACB200_ENOSUP, "episode_stats: do_allreduce without a communicator (acb200_comm_init)");
    NcclApi* a = nccl();
    const ncclResult_t r = a->AllReduce(stats, out, (size_t)count, kNcclFloat64, kNcclSum, (ncclComm_t)ctx->comm, ctx->stream);
    ctx->launches += 1;
    return r == 0 ? ACB200_OK : nccl_fail(ctx, a, r, "ncclAllReduce");
}

}  // extern "C"

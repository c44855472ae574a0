// gk_comm.cu -- the engine's own NCCL communicator (one process per GPU): the collectives of the row-sharded sweeps are
// enqueued by libgk on its own stream between its kernels, so a sharded sweep runs without host round trips
// (K x K statistics / per-block partials all-reduced, block labels and features broadcast by their owner).
//
// libnccl is opened at run time (dlopen "libnccl.so.2": the copy the process already holds -- PyTorch's -- or the system
// one), so libgk.so has no link-time dependency on it and single-GPU users never load it.
#include <dlfcn.h>
#include <nccl.h>
#include <string.h>
#include "gk_internal.cuh"

struct gk_comm {
    void *lib = nullptr;
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};

static int comm_load(gk_ctx *ctx, gk_comm **out) {
    gk_comm *c = new gk_comm();
    c->lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!c->lib) c->lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!c->lib) { delete c; return gk_fail(ctx, GK_EUNSUPPORTED, "libnccl.so.2 not found: %s", dlerror()); }
#define GK_SYM(field, name) *(void **)(&c->field) = dlsym(c->lib, name); if (!c->field) { delete c; return gk_fail(ctx, GK_EUNSUPPORTED, "libnccl: missing symbol %s", name); }
    GK_SYM(GetUniqueId, "ncclGetUniqueId") GK_SYM(CommInitRank, "ncclCommInitRank") GK_SYM(CommDestroy, "ncclCommDestroy")
    GK_SYM(AllReduce, "ncclAllReduce") GK_SYM(Broadcast, "ncclBroadcast") GK_SYM(GetErrorString, "ncclGetErrorString")
#undef GK_SYM
    *out = c;
    return GK_OK;
}

#define GK_NCCL(ctx, c, call)                                                                                       \
    do {                                                                                                            \
        ncclResult_t _r = (call);                                                                                   \
        if (_r != ncclSuccess) return gk_fail((ctx), GK_ECUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, (c)->GetErrorString(_r)); \
    } while (0)

extern "C" int gk_comm_unique_id(gk_ctx *ctx, uint8_t *id128) {
    if (!ctx || !id128) return GK_EINVAL;
    gk_comm *c = nullptr;
    int rc = comm_load(ctx, &c); if (rc) return rc;
    ncclUniqueId id;
    ncclResult_t r = c->GetUniqueId(&id);
    if (r != ncclSuccess) { rc = gk_fail(ctx, GK_ECUDA, "ncclGetUniqueId: %s", c->GetErrorString(r)); delete c; return rc; }
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    memcpy(id128, &id, 128);
    delete c;                    // the library stays loaded
    return GK_OK;
}

extern "C" int gk_comm_init(gk_ctx *ctx, const uint8_t *id128, int rank, int world) {
    if (!ctx || !id128 || world < 1 || rank < 0 || rank >= world) return GK_EINVAL;
    if (ctx->comm) return gk_fail(ctx, GK_EINVAL, "gk_comm_init: this context already has a communicator");
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    gk_comm *c = nullptr;
    int rc = comm_load(ctx, &c); if (rc) return rc;
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    ncclResult_t r = c->CommInitRank(&c->comm, world, id, rank);
    if (r != ncclSuccess) { rc = gk_fail(ctx, GK_ECUDA, "ncclCommInitRank: %s", c->GetErrorString(r)); delete c; return rc; }
    c->rank = rank; c->world = world;
    ctx->comm = c;
    return GK_OK;
}

void gk_comm_release(gk_ctx *ctx) {
    if (!ctx || !ctx->comm) return;
    if (ctx->comm->comm) ctx->comm->CommDestroy(ctx->comm->comm);
    delete ctx->comm;
    ctx->comm = nullptr;
}

extern "C" int gk_comm_destroy(gk_ctx *ctx) {
    if (!ctx) return GK_EINVAL;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    gk_comm_release(ctx);
    return GK_OK;
}

extern "C" int gk_comm_rank(gk_ctx *ctx) { return ctx && ctx->comm ? ctx->comm->rank : 0; }
extern "C" int gk_comm_world(gk_ctx *ctx) { return ctx && ctx->comm ? ctx->comm->world : 1; }

int gk_comm_allreduce_f64(gk_ctx *ctx, double *buf_dev, size_t n) {
    gk_comm *c = ctx->comm;
    if (!c || c->world == 1) return GK_OK;
    GK_NCCL(ctx, c, c->AllReduce(buf_dev, buf_dev, n, ncclDouble, ncclSum, c->comm, ctx->stream));
    return GK_OK;
}
int gk_comm_allreduce_i64(gk_ctx *ctx, long long *buf_dev, size_t n) {
    gk_comm *c = ctx->comm;
    if (!c || c->world == 1) return GK_OK;
    GK_NCCL(ctx, c, c->AllReduce(buf_dev, buf_dev, n, ncclInt64, ncclSum, c->comm, ctx->stream));
    return GK_OK;
}
int gk_comm_broadcast(gk_ctx *ctx, void *buf_dev, size_t bytes, int root) {
    gk_comm *c = ctx->comm;
    if (!c || c->world == 1) return GK_OK;
    GK_NCCL(ctx, c, c->Broadcast(buf_dev, buf_dev, bytes, ncclUint8, root, c->comm, ctx->stream));
    return GK_OK;
}

// rows [r0, r1) of rank `rank`: whole blocks of 128 rows, spread as evenly as possible (== sharding.row_range)
void gk_comm_row_range(int N, int rank, int world, int *r0, int *r1) {
    const long long nblk = (N + 127) / 128;
    const long long b0 = nblk * rank / world, b1 = nblk * (rank + 1) / world;
    *r0 = (int)(b0 * 128);
    *r1 = (int)std::min<long long>(N, b1 * 128);
}

// in-stream SUM over the ranks of a device buffer of doubles: what the sharded Python paths call between two libgk calls
extern "C" int gk_comm_allreduce(gk_ctx *ctx, double *buf_dev, size_t n) {
    if (!ctx || !buf_dev) return GK_EINVAL;
    if (!ctx->comm) return gk_fail(ctx, GK_EINVAL, "gk_comm_allreduce: no communicator (gk_comm_init)");
    return gk_comm_allreduce_f64(ctx, buf_dev, n);
}

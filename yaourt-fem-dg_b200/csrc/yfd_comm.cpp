/*
 * yfd_comm.cpp -- NCCL attachment of a context (multi-GPU).
 * The reference has no distributed layer.  NCCL is loaded at run time from the copy PyTorch
 * ships (libnccl.so.2 is already mapped into any process that imported torch), so the library
 * itself has no link-time dependency on it and single-GPU users never touch it.
 */
#include "yfd_internal.hpp"
#include <dlfcn.h>
#include <cstring>

namespace {

typedef struct { char internal[128]; } nccl_unique_id;
typedef int (*fn_get_unique_id)(nccl_unique_id *);
typedef int (*fn_comm_init_rank)(void **, int, nccl_unique_id, int);
typedef int (*fn_comm_destroy)(void *);
typedef const char *(*fn_get_error_string)(int);
typedef int (*fn_all_reduce)(const void *, void *, size_t, int, int, void *, cudaStream_t);
typedef int (*fn_send)(const void *, size_t, int, int, void *, cudaStream_t);
typedef int (*fn_recv)(void *, size_t, int, int, void *, cudaStream_t);
typedef int (*fn_group)(void);
typedef int (*fn_all_gather)(const void *, void *, size_t, int, void *, cudaStream_t);
constexpr int NCCL_INT8 = 0;        /* ncclInt8 / ncclChar */
constexpr int NCCL_FLOAT64 = 8;     /* ncclFloat64 (nccl.h) */
constexpr int NCCL_SUM = 0;         /* ncclSum */

struct nccl_api {
    void *handle = nullptr;
    fn_get_unique_id get_unique_id = nullptr;
    fn_comm_init_rank comm_init_rank = nullptr;
    fn_comm_destroy comm_destroy = nullptr;
    fn_get_error_string error_string = nullptr;
    fn_all_reduce all_reduce = nullptr;
    fn_send send = nullptr;
    fn_recv recv = nullptr;
    fn_group group_start = nullptr, group_end = nullptr;
    fn_all_gather all_gather = nullptr;
    bool ok = false;
};

nccl_api &api()
{
    static nccl_api a;
    if (a.handle || a.ok) return a;
    const char *names[] = { "libnccl.so.2", "libnccl.so", nullptr };
    for (int i = 0; names[i] && !a.handle; i++) a.handle = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
    if (!a.handle) return a;
    a.get_unique_id = (fn_get_unique_id) dlsym(a.handle, "ncclGetUniqueId");
    a.comm_init_rank = (fn_comm_init_rank) dlsym(a.handle, "ncclCommInitRank");
    a.comm_destroy = (fn_comm_destroy) dlsym(a.handle, "ncclCommDestroy");
    a.error_string = (fn_get_error_string) dlsym(a.handle, "ncclGetErrorString");
    a.all_reduce = (fn_all_reduce) dlsym(a.handle, "ncclAllReduce");
    a.send = (fn_send) dlsym(a.handle, "ncclSend");
    a.recv = (fn_recv) dlsym(a.handle, "ncclRecv");
    a.group_start = (fn_group) dlsym(a.handle, "ncclGroupStart");
    a.group_end = (fn_group) dlsym(a.handle, "ncclGroupEnd");
    a.all_gather = (fn_all_gather) dlsym(a.handle, "ncclAllGather");
    a.ok = a.get_unique_id && a.comm_init_rank && a.comm_destroy && a.all_reduce && a.send && a.recv && a.group_start && a.group_end;
    return a;
}

yfd_status nccl_check(yfd_ctx *ctx, int rc, const char *what)
{
    if (rc == 0) return YFD_OK;
    return yfd::fail(ctx, YFD_ERR_COMM, "%s: %s", what, api().error_string ? api().error_string(rc) : "NCCL error");
}

} // namespace

namespace yfd {

yfd_status comm_allreduce_sum_f64(yfd_ctx *ctx, double *d_buf, size_t count)
{
    if (ctx->nranks <= 1) return YFD_OK;
    if (!ctx->comm) return fail(ctx, YFD_ERR_COMM, "no communicator attached");
    return nccl_check(ctx, api().all_reduce(d_buf, d_buf, count, NCCL_FLOAT64, NCCL_SUM, ctx->comm, ctx->stream), "ncclAllReduce");
}
yfd_status comm_allgather_bytes(yfd_ctx *ctx, const void *d_send, void *d_recv, size_t bytes_per_rank)
{
    if (!ctx->comm || !api().all_gather) return fail(ctx, YFD_ERR_COMM, "ncclAllGather unavailable");
    return nccl_check(ctx, api().all_gather(d_send, d_recv, bytes_per_rank, NCCL_INT8, ctx->comm, ctx->stream), "ncclAllGather");
}
yfd_status comm_group_begin(yfd_ctx *ctx) { return nccl_check(ctx, api().group_start(), "ncclGroupStart"); }
yfd_status comm_group_end(yfd_ctx *ctx) { return nccl_check(ctx, api().group_end(), "ncclGroupEnd"); }
yfd_status comm_send_f64(yfd_ctx *ctx, const double *d_buf, size_t count, int peer)
{
    return nccl_check(ctx, api().send(d_buf, count, NCCL_FLOAT64, peer, ctx->comm, ctx->stream), "ncclSend");
}
yfd_status comm_recv_f64(yfd_ctx *ctx, double *d_buf, size_t count, int peer)
{
    return nccl_check(ctx, api().recv(d_buf, count, NCCL_FLOAT64, peer, ctx->comm, ctx->stream), "ncclRecv");
}

} // namespace yfd

extern "C" {

void yfd_comm_destroy_internal(yfd_ctx *ctx)
{
    if (ctx) yfd::peer_release(ctx);
    if (ctx && ctx->comm && api().ok) api().comm_destroy(ctx->comm);
    if (ctx) ctx->comm = nullptr;
}

yfd_status yfd_comm_unique_id(unsigned char id[YFD_COMM_ID_BYTES])
{
    if (!id) return YFD_ERR_INVALID_ARGUMENT;
    if (!api().ok) return YFD_ERR_COMM;
    nccl_unique_id u;
    if (api().get_unique_id(&u) != 0) return YFD_ERR_COMM;
    static_assert(sizeof(u) == YFD_COMM_ID_BYTES, "ncclUniqueId is 128 bytes");
    std::memcpy(id, &u, sizeof(u));
    return YFD_OK;
}

yfd_status yfd_ctx_attach_comm(yfd_ctx *ctx, int rank, int nranks, const unsigned char id[YFD_COMM_ID_BYTES])
{
    if (!ctx || !id || nranks < 1 || rank < 0 || rank >= nranks) return YFD_ERR_INVALID_ARGUMENT;
    if (!api().ok) return yfd::fail(ctx, YFD_ERR_COMM, "libnccl.so.2 could not be loaded: %s", dlerror());
    yfd::scoped_device dev(ctx->device);
    nccl_unique_id u;
    std::memcpy(&u, id, sizeof(u));
    void *comm = nullptr;
    const int rc = api().comm_init_rank(&comm, nranks, u, rank);
    if (rc != 0) return yfd::fail(ctx, YFD_ERR_COMM, "ncclCommInitRank: %s", api().error_string ? api().error_string(rc) : "error");
    ctx->comm = comm;
    ctx->rank = rank;
    ctx->nranks = nranks;
    return YFD_OK;
}

} /* extern "C" */

/*
 * yfd_peer.cu -- peer-memory transport over NVLink / NVSwitch for the multi-GPU CG.
 *
 * NCCL gives correct halo exchanges and scalar all-reduces, but each call costs ~20-30 us of launch + protocol
 * latency, and a CG iteration needs three of them: at 8 GPUs that is most of an iteration whose kernels take
 * ~110 us.  Here the ranks map each other's memory once (cudaMalloc + CUDA IPC handles, swapped with one
 * ncclAllGather) and the solver's own kernels communicate:
 *   - a reduction's last block sends its partial sums to every rank as 8-byte packets { 32 data bits | epoch tag }, then
 *     spins until the packets of all ranks for this epoch are in and adds them IN RANK ORDER -- every rank
 *     gets the same bits, so `done` flips on all ranks in the same iteration (cg_shared.cuh: xrank_sum);
 *   - there is no per-iteration halo exchange of the direction vector at all: every rank keeps halo copies of r, z
 *     and d and advances them itself with the recurrences the owner uses (same operands, same FMAs, same bits).
 *     What it needs for that is y = A d on its halo cells, which the owner stores straight into this rank's landing
 *     zone from cg_update_xr (yfd_cg.cu) BEFORE that kernel's cross-rank sum of r.r -- the sum's packets, sent after
 *     a system fence, announce the halo data too.  One iteration = 3 kernels and 2 synchronisations (the two dot
 *     products CG cannot do without), against 4 kernels and 3 synchronisations with an explicit exchange.
 * Epoch counters live on the device and advance inside the kernels, so the whole iteration stays inside one
 * CUDA graph with no host or NCCL involvement.  Every rank runs one kernel at a time on its own GPU and
 * writes before it waits, so the spins cannot deadlock; they are bounded anyway and raise comm_error.
 * The landing zone is single-buffered: a neighbour's next store into it follows that neighbour's next d.y sum, which
 * needs this rank's SpMV of the next iteration to have finished, i.e. comes after this rank has consumed the zone.
 */
#include "yfd_internal.hpp"
#include "yfd_cg_shared.cuh"
#include <cstdlib>
#include <cstring>

namespace yfd {

static size_t region_bytes(size_t land_doubles)
{
    return ((sizeof(peer_ctl) + 4095) & ~(size_t) 4095) + land_doubles * sizeof(double) + 256;
}

/* one word per rank through ncclAllGather; returns the OR of the words (doubles as a barrier: every rank has passed this
 * point of its stream when the call returns) */
static yfd_status agree_any(yfd_ctx *ctx, int mine, int *any)
{
    dbuf<int> d_send, d_recv;
    YFD_TRY(d_send.alloc(ctx, 1)); YFD_TRY(d_recv.alloc(ctx, (size_t) ctx->nranks));
    YFD_TRY(d_send.upload(ctx, &mine, 1));
    YFD_TRY(comm_allgather_bytes(ctx, d_send.p, d_recv.p, sizeof(int)));
    std::vector<int> all((size_t) ctx->nranks);
    YFD_TRY(d_recv.download(ctx, all.data(), all.size()));
    *any = 0;
    for (int v : all) *any |= v;
    return YFD_OK;
}

/* collective = every rank of the communicator is inside this call (rebuild, agreed tear-down): the ranks then meet between
 * closing their imported mappings and freeing the exported region, so that no region is freed while a peer still maps it.
 * The non-collective form (context destruction, possibly after a peer died) cannot wait for anybody. */
void peer_release(yfd_ctx *ctx, bool collective)
{
    if (!ctx) return;
    const bool had_region = ctx->peer_region != nullptr;
    if (had_region) cudaStreamSynchronize(ctx->stream);
    for (int q = 0; q < YFD_MAX_RANKS; q++)
        if (ctx->peer_opened[q]) { cudaIpcCloseMemHandle(ctx->peer_opened[q]); ctx->peer_opened[q] = nullptr; }
    if (collective && ctx->comm && ctx->nranks > 1) {
        int any = 0;
        if (agree_any(ctx, 0, &any) != YFD_OK) cudaGetLastError();      /* the free below is all that is left to do anyway */
    }
    if (ctx->peer_region) { cudaFree(ctx->peer_region); ctx->peer_region = nullptr; }
    if (ctx->d_peer_view) { cudaFree(ctx->d_peer_view); ctx->d_peer_view = nullptr; }
    ctx->peer_land_cap = 0;
    ctx->peer_land = nullptr;
    ctx->peer_enabled = false;
}

yfd_status ensure_peer_region(yfd_ctx *ctx, size_t land_doubles)
{
    if (ctx->nranks <= 1 || ctx->nranks > YFD_MAX_RANKS) return YFD_OK;
    const char *mode = getenv("YFD_COMM");
    if (mode && mode[0] == 'n') { ctx->peer_enabled = false; return YFD_OK; }       /* YFD_COMM=nccl */
    /* a rank whose kernels ran into the spin bound (cg_device: comm_error) has epochs and flags nobody can trust any more:
     * the ranks agree on that here, drop the mappings together and stay on NCCL for the rest of the context's life */
    if (ctx->peer_off) return YFD_OK;
    int any_failed = 0;
    YFD_TRY(agree_any(ctx, ctx->peer_failed ? 1 : 0, &any_failed));
    if (any_failed) {
        ctx->peer_failed = ctx->peer_off = true;
        peer_release(ctx, /*collective=*/true);
        return YFD_OK;
    }
    if (ctx->peer_region && ctx->peer_land_cap >= land_doubles) return YFD_OK;
    /* (re)build, collectively: `land_doubles` is the same on every rank */
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    peer_release(ctx, /*collective=*/true);
    const size_t cap = land_doubles + land_doubles / 2 + 1024;
    const size_t bytes = region_bytes(cap);
    bool ok = true;
    void *region = nullptr;
    cudaIpcMemHandle_t mine;
    std::memset(&mine, 0, sizeof(mine));
    if (cudaMalloc(&region, bytes) != cudaSuccess) { cudaGetLastError(); ok = false; region = nullptr; }
    if (ok && cudaMemset(region, 0, bytes) != cudaSuccess) ok = false;
    if (ok && cudaIpcGetMemHandle(&mine, region) != cudaSuccess) { cudaGetLastError(); ok = false; }
    /* handles + an "ok" byte travel through one all-gather, so every rank takes the same decision */
    struct wire { cudaIpcMemHandle_t h; int ok; int pad[3]; };
    wire w;
    std::memset(&w, 0, sizeof(w));
    w.h = mine; w.ok = ok ? 1 : 0;
    dbuf<wire> d_send, d_recv;
    YFD_TRY(d_send.alloc(ctx, 1)); YFD_TRY(d_recv.alloc(ctx, (size_t) ctx->nranks));
    YFD_TRY(d_send.upload(ctx, &w, 1));
    YFD_TRY(comm_allgather_bytes(ctx, d_send.p, d_recv.p, sizeof(wire)));
    std::vector<wire> all((size_t) ctx->nranks);
    YFD_TRY(d_recv.download(ctx, all.data(), all.size()));
    for (auto &a : all) ok = ok && a.ok;
    peer_view pv;
    std::memset(&pv, 0, sizeof(pv));
    pv.me = ctx->rank; pv.nranks = ctx->nranks;
    const size_t land_off = (sizeof(peer_ctl) + 4095) & ~(size_t) 4095;
    int opened_ok = ok ? 1 : 0;
    for (int q = 0; q < ctx->nranks && opened_ok; q++) {
        void *base = region;
        if (q != ctx->rank) {
            if (cudaIpcOpenMemHandle(&base, all[(size_t) q].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                cudaGetLastError(); opened_ok = 0; break;
            }
            ctx->peer_opened[q] = base;
        }
        pv.ctl[q] = reinterpret_cast<peer_ctl *>(base);
        pv.land[q] = reinterpret_cast<double *>((char *) base + land_off);
    }
    /* second agreement round: did every rank manage to map every peer? */
    w.ok = opened_ok;
    YFD_TRY(d_send.upload(ctx, &w, 1));
    YFD_TRY(comm_allgather_bytes(ctx, d_send.p, d_recv.p, sizeof(wire)));
    YFD_TRY(d_recv.download(ctx, all.data(), all.size()));
    bool all_ok = true;
    for (auto &a : all) all_ok = all_ok && a.ok;
    ctx->peer_region = region;
    if (!all_ok) { peer_release(ctx, /*collective=*/true); return YFD_OK; }       /* stay on NCCL */
    YFD_CUDA(ctx, cudaMalloc((void **) &ctx->d_peer_view, sizeof(peer_view)));
    YFD_CUDA(ctx, cudaMemcpy(ctx->d_peer_view, &pv, sizeof(pv), cudaMemcpyHostToDevice));
    ctx->peer_land_cap = cap;
    ctx->peer_land = pv.land[ctx->rank];
    ctx->peer_enabled = true;
    return YFD_OK;
}

} // namespace yfd

/*
 * yfd_peer.cu -- peer-memory transport over NVLink / NVSwitch for the multi-GPU CG.
 *
 * NCCL gives correct halo exchanges and scalar all-reduces, but each call costs ~20-30 us of launch + protocol
 * latency, and a CG iteration needs three of them: at 8 GPUs that is most of an iteration whose kernels take
 * ~110 us.  Here the ranks map each other's memory once (cudaMalloc + CUDA IPC handles, swapped with one
 * ncclAllGather) and the solver's own kernels communicate:
 *   - a reduction's last block stores its partial sums into a slot on every rank, fences, raises a flag, then
 *     spins until the flags of all ranks for this epoch are up and adds the slots IN RANK ORDER -- every rank
 *     gets the same bits, so `done` flips on all ranks in the same iteration (cg_shared.cuh: xrank_sum);
 *   - the direction vector d = [owned | halo] of every rank lives IN its landing zone, and halo_put stores the
 *     dofs of the cells a neighbour needs straight into the halo part of that neighbour's d; its last block
 *     signals the neighbours and waits for theirs.
 * Epoch counters live on the device and advance inside the kernels, so the whole iteration stays inside one
 * CUDA graph with no host or NCCL involvement.  Every rank runs one kernel at a time on its own GPU and
 * writes before it waits, so the spins cannot deadlock; they are bounded anyway and raise comm_error.
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

void peer_release(yfd_ctx *ctx)
{
    if (!ctx) return;
    for (int q = 0; q < YFD_MAX_RANKS; q++)
        if (ctx->peer_opened[q]) { cudaIpcCloseMemHandle(ctx->peer_opened[q]); ctx->peer_opened[q] = nullptr; }
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
    if (ctx->peer_region && ctx->peer_land_cap >= land_doubles) return YFD_OK;
    /* (re)build, collectively: `land_doubles` is the same on every rank */
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    peer_release(ctx);
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
    if (!all_ok) { peer_release(ctx); return YFD_OK; }       /* stay on NCCL */
    YFD_CUDA(ctx, cudaMalloc((void **) &ctx->d_peer_view, sizeof(peer_view)));
    YFD_CUDA(ctx, cudaMemcpy(ctx->d_peer_view, &pv, sizeof(pv), cudaMemcpyHostToDevice));
    ctx->peer_land_cap = cap;
    ctx->peer_land = pv.land[ctx->rank];
    ctx->peer_enabled = true;
    return YFD_OK;
}

/* ---- halo exchange by direct stores ---------------------------------------------------------- */
__global__ void __launch_bounds__(256)
halo_put(const double *d, const int32_t *__restrict__ send_lidx,
         const int32_t *__restrict__ send_rank, const int32_t *__restrict__ send_dst, int64_t nsend, int bs,
         const int32_t *__restrict__ peers, int npeers, cg_scalars *__restrict__ sc)
{
    if (sc->done) return;
    const peer_view *pv = sc->peer;
    const int64_t n = nsend * bs;
    for (int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t) gridDim.x * blockDim.x) {
        const int64_t s = t / bs;
        const int j = (int) (t - s * bs);
        pv->land[send_rank[s]][(int64_t) send_dst[s] * bs + j] = d[(int64_t) send_lidx[s] * bs + j];
    }
    /* last block: all stores of this rank are out -> signal the neighbours, wait for theirs, move the landed dofs */
    __shared__ bool last;
    __shared__ int bad;
    __threadfence_system();
    __syncthreads();
    peer_ctl *me = pv->ctl[pv->me];
    if (threadIdx.x == 0) {
        last = atomicInc(&me->halo_ticket, gridDim.x - 1) == gridDim.x - 1;
        bad = 0;
    }
    __syncthreads();
    if (!last) return;
    __shared__ unsigned long long epoch;
    if (threadIdx.x == 0) { epoch = me->halo_epoch + 1; me->halo_epoch = epoch; }
    __syncthreads();
    __threadfence_system();
    if ((int) threadIdx.x < npeers) {
        const int q = peers[threadIdx.x];
        *((volatile unsigned long long *) &pv->ctl[q]->halo_flag[pv->me]) = epoch;
        long long spins = 0;
        while (*((volatile unsigned long long *) &me->halo_flag[q]) < epoch)
            if (++spins > (1ll << 28)) { bad = 1; break; }
    }
    __syncthreads();
    __threadfence_system();
    if (bad && threadIdx.x == 0) { sc->comm_error = 1; sc->done = 1; }
    /* nothing to move: the neighbours stored straight into the halo part of d (d lives in the landing zone) */
}

yfd_status halo_put_launch(yfd_ctx *ctx, const dist_plan &pl, int bs, double *d, cg_scalars *sc)
{
    if (pl.nranks <= 1 || pl.peers.empty()) return YFD_OK;
    const int64_t n = pl.nsend * bs;
    int64_t grid = ceil_div(n > 0 ? n : 1, 256);
    if (grid > ctx->sm_count) grid = ctx->sm_count;
    YFD_LAUNCH(ctx, halo_put, (unsigned) grid, 256, 0, d, pl.send_lidx.p, pl.send_rank.p, pl.send_dst.p,
               pl.nsend, bs, pl.d_peers.p, (int) pl.peers.size(), sc);
    return YFD_OK;
}

} // namespace yfd

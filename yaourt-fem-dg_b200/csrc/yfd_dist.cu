/*
 * yfd_dist.cu -- device side of the multi-GPU path (no reference counterpart).
 *
 * One process per GPU, the mesh replicated on every rank.  The host planner (yfd_partition.cpp) cuts the
 * internal cell order into contiguous blocks; here its lists become device tables:
 *   g2l        internal cell id -> local index of the x vector  [ owned cells | halo cells ]
 *   send_lidx  owned cells whose dofs a peer needs, grouped by peer
 * and a halo exchange is: one pack kernel, then one grouped ncclSend/ncclRecv per peer straight into the
 * halo part of x (the receiver's halo group "owned by q" has the same order as q's send list).
 */
#include "yfd_internal.hpp"
#include <algorithm>

namespace yfd {

/* nbr[c][e] (internal order and ids) from cell_faces / face_owners (reference ids) */
__global__ void cell_neighbours(const int32_t *__restrict__ cell_faces, const int2 *__restrict__ face_owners,
                                const int32_t *__restrict__ perm, const int32_t *__restrict__ iperm,
                                int64_t ncells, int nf, int32_t *__restrict__ nbr)
{
    const int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ncells * nf) return;
    const int64_t c = t / nf;
    const int e = (int) (t - c * nf);
    const int32_t rc = perm[c];
    const int2 own = face_owners[cell_faces[(int64_t) rc * nf + e]];
    const int32_t nrc = own.x == rc ? own.y : own.x;          /* neighbour_via, core/mesh.hpp:249-272 */
    nbr[t] = nrc >= 0 ? iperm[nrc] : -1;
}

__global__ void fill_g2l(int32_t *__restrict__ g2l, int64_t cell_begin, int64_t nown,
                         const int32_t *__restrict__ halo_gid, int64_t nhalo)
{
    const int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (t < nown) g2l[cell_begin + t] = (int32_t) t;
    else if (t < nown + nhalo) g2l[halo_gid[t - nown]] = (int32_t) t;
}

__global__ void pack_cells(const double *__restrict__ x, const int32_t *__restrict__ lidx, int64_t ncell, int bs,
                           double *__restrict__ out)
{
    const int64_t n = ncell * bs;
    for (int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t) gridDim.x * blockDim.x) {
        const int64_t c = t / bs;
        out[t] = x[(int64_t) lidx[c] * bs + (t - c * bs)];
    }
}

/* the faces between two ranks' blocks: (cell, neighbour) for every cell whose neighbour has another owner; O(sqrt(cells)) of
 * them on a Z-curve order, so only these -- not the neighbour table -- travel to the host planner */
/* yfd_partition.cpp: partition_begin / partition_owner in 64-bit arithmetic (cell ids are int32 here, nranks small) */
__device__ __forceinline__ int64_t begin_of(int64_t ncells, int nranks, int r) { return ncells * r / nranks; }
__device__ __forceinline__ int owner_of(int64_t ncells, int nranks, int64_t c)
{
    int r = (int) (((c + 1) * nranks - 1) / ncells);
    if (r >= nranks) r = nranks - 1;
    while (r > 0 && begin_of(ncells, nranks, r) > c) r--;
    while (r + 1 < nranks && begin_of(ncells, nranks, r + 1) <= c) r++;
    return r;
}
__global__ void cut_faces(const int32_t *__restrict__ nbr, int64_t ncells, int nf, int nranks,
                          int2 *__restrict__ out, unsigned long long cap, unsigned long long *__restrict__ count)
{
    const int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ncells * nf) return;
    const int32_t n = nbr[t];
    if (n < 0) return;
    const int64_t c = t / nf;
    if (owner_of(ncells, nranks, c) == owner_of(ncells, nranks, n)) return;
    const unsigned long long k = atomicAdd(count, 1ull);
    if (k < cap) out[k] = make_int2((int) c, n);
}

yfd_status mesh_neighbours_device(yfd_ctx *ctx, yfd_mesh *m, dbuf<int32_t> &nbr)
{
    const int64_t n = m->ncells * m->nf;
    YFD_TRY(nbr.alloc(ctx, (size_t) n));
    if (n == 0) return YFD_OK;
    YFD_LAUNCH(ctx, cell_neighbours, (unsigned) ceil_div(n, 256), 256, 0, m->cell_faces.p, (const int2 *) m->face_owners.p,
               m->perm.p, m->iperm.p, m->ncells, m->nf, nbr.p);
    return YFD_OK;
}

yfd_status ensure_dist_plan(yfd_ctx *ctx, yfd_mesh *mesh)
{
    if (mesh->plan && mesh->plan->nranks == ctx->nranks && mesh->plan->rank == ctx->rank) return YFD_OK;
    const int64_t nc = mesh->ncells;
    const int nf = mesh->nf;
    dbuf<int32_t> d_nbr;
    YFD_TRY(mesh_neighbours_device(ctx, mesh, d_nbr));
    /* cut faces on the device, planner on the host from those alone (the order of the list does not matter: the planner sorts) */
    std::vector<int64_t> cut_cell, cut_nbr;
    {
        dbuf<unsigned long long> d_count;
        YFD_TRY(d_count.alloc(ctx, 1)); YFD_TRY(d_count.zero(ctx));
        unsigned long long cap = (unsigned long long) (64 * 1024), found = 0;
        dbuf<int2> d_cut;
        for (int pass = 0; pass < 2; pass++) {
            YFD_TRY(d_cut.alloc(ctx, (size_t) cap));
            if (nc * nf > 0)
                YFD_LAUNCH(ctx, cut_faces, (unsigned) ceil_div(nc * nf, 256), 256, 0, d_nbr.p, nc, nf, ctx->nranks, d_cut.p, cap, d_count.p);
            YFD_TRY(d_count.download(ctx, &found, 1));
            if (found <= cap) break;
            cap = found;                                    /* a mesh that is mostly cuts (tiny, or not a Z-curve order) */
            YFD_TRY(d_count.zero(ctx));
        }
        std::vector<int2> cut((size_t) found);
        if (found) YFD_TRY(d_cut.download(ctx, cut.data(), cut.size()));
        cut_cell.resize(cut.size()); cut_nbr.resize(cut.size());
        for (size_t i = 0; i < cut.size(); i++) { cut_cell[i] = cut[i].x; cut_nbr[i] = cut[i].y; }
    }
    yfd_partition *hp = nullptr;
    const yfd_status st = yfd_partition_create_from_cuts(nc, ctx->nranks, ctx->rank, (int64_t) cut_cell.size(), cut_cell.data(),
                                                         cut_nbr.data(), &hp);
    if (st != YFD_OK) return fail(ctx, st, "partition planner failed");
    auto pl = std::make_shared<dist_plan>();
    int npeers = 0;
    pl->nranks = ctx->nranks; pl->rank = ctx->rank; pl->ncells_global = nc;
    yfd_partition_sizes(hp, &pl->cell_begin, &pl->cell_end, &pl->nhalo, &pl->nsend, &npeers);
    pl->nown = pl->cell_end - pl->cell_begin;
    std::vector<int64_t> halo((size_t) pl->nhalo), send((size_t) pl->nsend), rc((size_t) npeers), sc((size_t) npeers);
    pl->peers.resize((size_t) npeers);
    yfd_partition_get(hp, halo.data(), send.data(), pl->peers.data(), rc.data(), sc.data());
    yfd_partition_destroy(hp);
    pl->recv_off.assign((size_t) npeers + 1, 0);
    pl->send_off.assign((size_t) npeers + 1, 0);
    for (int k = 0; k < npeers; k++) { pl->recv_off[k + 1] = pl->recv_off[k] + rc[k]; pl->send_off[k + 1] = pl->send_off[k] + sc[k]; }
    if (pl->nown + pl->nhalo >= 0x7fffffff) return fail(ctx, YFD_ERR_UNSUPPORTED, "local cell count exceeds 32-bit ids");

    /* peer-memory transport: every rank can derive every other rank's plan from the replicated neighbour table,
     * so "where does my block land in q's halo" needs no communication */
    {
        std::vector<int32_t> srank((size_t) pl->nsend), sdst((size_t) pl->nsend);
        pl->max_halo_cells = pl->nhalo;
        pl->max_local_cells = pl->nown + pl->nhalo;
        for (int q = 0; q < ctx->nranks; q++) {
            if (q == ctx->rank) continue;
            yfd_partition *qp = nullptr;
            if (yfd_partition_create_from_cuts(nc, ctx->nranks, q, (int64_t) cut_cell.size(), cut_cell.data(), cut_nbr.data(), &qp) != YFD_OK) return fail(ctx, YFD_ERR_INVALID_STATE, "partition planner failed for rank %d", q);
            int64_t qh = 0; int qn = 0;
            yfd_partition_sizes(qp, nullptr, nullptr, &qh, nullptr, &qn);
            if (qh > pl->max_halo_cells) pl->max_halo_cells = qh;
            const int64_t q_own = partition_begin(nc, ctx->nranks, q + 1) - partition_begin(nc, ctx->nranks, q);
            if (q_own + qh > pl->max_local_cells) pl->max_local_cells = q_own + qh;
            const size_t k = std::lower_bound(pl->peers.begin(), pl->peers.end(), q) - pl->peers.begin();
            if (k < pl->peers.size() && pl->peers[k] == q) {
                std::vector<int32_t> qpeers((size_t) qn);
                std::vector<int64_t> qrc((size_t) qn);
                yfd_partition_get(qp, nullptr, nullptr, qpeers.data(), qrc.data(), nullptr);
                int64_t off = 0;
                for (int j = 0; j < qn && qpeers[(size_t) j] != ctx->rank; j++) off += qrc[(size_t) j];
                for (int64_t i = pl->send_off[k]; i < pl->send_off[k + 1]; i++) {
                    srank[(size_t) i] = q;
                    /* the landing zone of rank q IS its direction vector d = [owned | halo]: slot = its owned count + position in its halo */
                    sdst[(size_t) i] = (int32_t) (q_own + off + (i - pl->send_off[k]));
                }
            }
            yfd_partition_destroy(qp);
        }
        YFD_TRY(pl->send_rank.alloc(ctx, (size_t) pl->nsend)); YFD_TRY(pl->send_rank.upload(ctx, srank.data(), srank.size()));
        YFD_TRY(pl->send_dst.alloc(ctx, (size_t) pl->nsend));   YFD_TRY(pl->send_dst.upload(ctx, sdst.data(), sdst.size()));
        YFD_TRY(pl->d_peers.alloc(ctx, pl->peers.size()));      YFD_TRY(pl->d_peers.upload(ctx, pl->peers.data(), pl->peers.size()));
        YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    std::vector<int32_t> halo32(halo.begin(), halo.end()), send32((size_t) pl->nsend);
    for (int64_t i = 0; i < pl->nsend; i++) send32[(size_t) i] = (int32_t) (send[(size_t) i] - pl->cell_begin);
    dbuf<int32_t> d_halo;
    YFD_TRY(d_halo.alloc(ctx, (size_t) pl->nhalo));
    YFD_TRY(d_halo.upload(ctx, halo32.data(), halo32.size()));
    YFD_TRY(pl->send_lidx.alloc(ctx, (size_t) pl->nsend));
    YFD_TRY(pl->send_lidx.upload(ctx, send32.data(), send32.size()));
    YFD_TRY(pl->g2l.alloc(ctx, (size_t) nc));
    YFD_CUDA(ctx, cudaMemsetAsync(pl->g2l.p, 0xff, (size_t) nc * sizeof(int32_t), ctx->stream));
    const int64_t tot = pl->nown + pl->nhalo;
    if (tot > 0)
        YFD_LAUNCH(ctx, fill_g2l, (unsigned) ceil_div(tot, 256), 256, 0, pl->g2l.p, pl->cell_begin, pl->nown, d_halo.p, pl->nhalo);
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));       /* host staging vectors go out of scope */
    mesh->plan = pl;
    return YFD_OK;
}

yfd_status halo_exchange(yfd_ctx *ctx, const dist_plan &pl, int bs, double *x, double *sendbuf)
{
    if (pl.nranks <= 1 || pl.peers.empty()) return YFD_OK;
    if (pl.nsend > 0) {
        const int64_t n = pl.nsend * bs;
        const int64_t want = ceil_div(n, 256), cap = (int64_t) ctx->sm_count * 8;
        YFD_LAUNCH(ctx, pack_cells, (unsigned) (want < cap ? want : cap), 256, 0, x, pl.send_lidx.p, pl.nsend, bs, sendbuf);
    }
    YFD_TRY(comm_group_begin(ctx));
    yfd_status st = YFD_OK;
    for (size_t k = 0; k < pl.peers.size() && st == YFD_OK; k++) {
        const int64_t ns = pl.send_off[k + 1] - pl.send_off[k], nr = pl.recv_off[k + 1] - pl.recv_off[k];
        if (ns > 0) st = comm_send_f64(ctx, sendbuf + pl.send_off[k] * bs, (size_t) (ns * bs), pl.peers[k]);
        if (st == YFD_OK && nr > 0) st = comm_recv_f64(ctx, x + (pl.nown + pl.recv_off[k]) * bs, (size_t) (nr * bs), pl.peers[k]);
    }
    const yfd_status st2 = comm_group_end(ctx);
    return st != YFD_OK ? st : st2;
}

} // namespace yfd

using namespace yfd;

extern "C" yfd_status yfd_mesh_get_neighbours(yfd_mesh *m, int64_t *neighbours)
{
    if (!m || !neighbours) return YFD_ERR_INVALID_ARGUMENT;
    yfd_ctx *ctx = m->ctx;
    scoped_device dev(ctx->device);
    dbuf<int32_t> nbr;
    YFD_TRY(mesh_neighbours_device(ctx, m, nbr));
    return download_widen(ctx, nbr.p, m->ncells * m->nf, neighbours);
}

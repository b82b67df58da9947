/*
 * yfd_partition.cpp -- host-side partition plan for the multi-GPU path.
 *
 * No reference counterpart: the reference is a single process.  Cells are split into `nranks` contiguous
 * blocks of the INTERNAL (Z-curve) order, so every block is spatially compact and cuts O(sqrt(cells/P)) faces
 * (contiguous blocks of the reference's own numbering would cut 55-99 % of all interior faces).  From the
 * cell-neighbour table every rank derives, with no communication (the mesh is replicated):
 *   - its halo: the non-owned neighbours of owned cells, grouped by owner rank (ascending), ids ascending;
 *   - its send lists: for each peer q, the owned cells that have a neighbour owned by q, ids ascending.
 * By symmetry of the neighbour relation the halo group "owned by q" on rank p equals the send list p <- q
 * on rank q, entry for entry, so a halo exchange is one packed message per peer pair with no index traffic.
 * Pure integer logic, no CUDA: this is what the world-size-2 gloo tests exercise on the CPU.
 */
#include <algorithm>
#include <cstdint>
#include <new>
#include <vector>
#include "yfd.h"

struct yfd_partition {
    int nranks = 1, rank = 0;
    int64_t ncells = 0, cell_begin = 0, cell_end = 0;
    std::vector<int64_t> halo_cells;        /* global internal ids, grouped by owner */
    std::vector<int64_t> send_cells;        /* global internal ids, grouped by destination */
    std::vector<int32_t> peers;             /* ranks exchanged with, ascending */
    std::vector<int64_t> recv_counts, send_counts;   /* cells per peer */
};

namespace yfd {

int64_t partition_begin(int64_t ncells, int nranks, int r)
{
    return (int64_t) (((__int128) ncells * r) / nranks);
}

int partition_owner(int64_t ncells, int nranks, int64_t cell)
{
    /* inverse of partition_begin: the largest r with begin(r) <= cell */
    int r = (int) (((__int128) (cell + 1) * nranks - 1) / ncells);
    if (r >= nranks) r = nranks - 1;
    while (r > 0 && partition_begin(ncells, nranks, r) > cell) r--;
    while (r + 1 < nranks && partition_begin(ncells, nranks, r + 1) <= cell) r++;
    return r;
}

} // namespace yfd

/* the plan of `rank` from the CUT FACES alone: pairs (cell, neighbour) whose two cells have different owners.  Pairs whose
 * cell belongs to another rank are skipped, so one list of all cuts of the mesh serves every rank. */
static yfd_status plan_from_pairs(yfd_partition *p, int64_t ncut, const int64_t *cut_cell, const int64_t *cut_nbr)
{
    const int64_t ncells = p->ncells;
    const int nranks = p->nranks;
    /* (peer, cell) pairs, then sort + unique */
    std::vector<std::pair<int32_t, int64_t>> halo, send;
    for (int64_t i = 0; i < ncut; i++) {
        const int64_t c = cut_cell[i], n = cut_nbr[i];
        if (c < 0 || c >= ncells || n < 0 || n >= ncells) return YFD_ERR_INVALID_ARGUMENT;
        if (c < p->cell_begin || c >= p->cell_end) continue;
        if (n >= p->cell_begin && n < p->cell_end) continue;
        const int32_t q = (int32_t) yfd::partition_owner(ncells, nranks, n);
        halo.emplace_back(q, n);
        send.emplace_back(q, c);
    }
    std::sort(halo.begin(), halo.end()); halo.erase(std::unique(halo.begin(), halo.end()), halo.end());
    std::sort(send.begin(), send.end()); send.erase(std::unique(send.begin(), send.end()), send.end());
    for (auto &h : halo) if (p->peers.empty() || p->peers.back() != h.first) p->peers.push_back(h.first);
    p->recv_counts.assign(p->peers.size(), 0);
    p->send_counts.assign(p->peers.size(), 0);
    for (auto &h : halo) {
        const size_t k = std::lower_bound(p->peers.begin(), p->peers.end(), h.first) - p->peers.begin();
        p->recv_counts[k]++; p->halo_cells.push_back(h.second);
    }
    for (auto &s : send) {
        /* every peer I send to also sends to me (the relation is symmetric), so it is in `peers` */
        const size_t k = std::lower_bound(p->peers.begin(), p->peers.end(), s.first) - p->peers.begin();
        if (k >= p->peers.size() || p->peers[k] != s.first) return YFD_ERR_INVALID_ARGUMENT;
        p->send_counts[k]++; p->send_cells.push_back(s.second);
    }
    return YFD_OK;
}

static yfd_partition *new_partition(int64_t ncells, int nranks, int rank)
{
    yfd_partition *p = new (std::nothrow) yfd_partition();
    if (!p) return nullptr;
    p->nranks = nranks; p->rank = rank; p->ncells = ncells;
    p->cell_begin = yfd::partition_begin(ncells, nranks, rank);
    p->cell_end = yfd::partition_begin(ncells, nranks, rank + 1);
    return p;
}

extern "C" {

yfd_status yfd_partition_create(int64_t ncells, int nf, const int64_t *nbr, int nranks, int rank, yfd_partition **out)
{
    if (!out) return YFD_ERR_INVALID_ARGUMENT;
    *out = nullptr;
    if (ncells < 0 || nf < 1 || nf > 4 || (ncells && !nbr) || nranks < 1 || rank < 0 || rank >= nranks) return YFD_ERR_INVALID_ARGUMENT;
    yfd_partition *p = new_partition(ncells, nranks, rank);
    if (!p) return YFD_ERR_OUT_OF_MEMORY;
    std::vector<int64_t> cc, cn;
    for (int64_t c = p->cell_begin; c < p->cell_end; c++)
        for (int e = 0; e < nf; e++) {
            const int64_t n = nbr[c * nf + e];
            if (n < 0) continue;
            if (n >= ncells) { delete p; return YFD_ERR_INVALID_ARGUMENT; }
            if (n >= p->cell_begin && n < p->cell_end) continue;
            cc.push_back(c); cn.push_back(n);
        }
    const yfd_status st = plan_from_pairs(p, (int64_t) cc.size(), cc.data(), cn.data());
    if (st != YFD_OK) { delete p; return st; }
    *out = p;
    return YFD_OK;
}

yfd_status yfd_partition_create_from_cuts(int64_t ncells, int nranks, int rank, int64_t ncut, const int64_t *cut_cell,
                                          const int64_t *cut_nbr, yfd_partition **out)
{
    if (!out) return YFD_ERR_INVALID_ARGUMENT;
    *out = nullptr;
    if (ncells < 0 || ncut < 0 || (ncut && (!cut_cell || !cut_nbr)) || nranks < 1 || rank < 0 || rank >= nranks) return YFD_ERR_INVALID_ARGUMENT;
    yfd_partition *p = new_partition(ncells, nranks, rank);
    if (!p) return YFD_ERR_OUT_OF_MEMORY;
    const yfd_status st = plan_from_pairs(p, ncut, cut_cell, cut_nbr);
    if (st != YFD_OK) { delete p; return st; }
    *out = p;
    return YFD_OK;
}

void yfd_partition_destroy(yfd_partition *p) { delete p; }

void yfd_partition_sizes(const yfd_partition *p, int64_t *cell_begin, int64_t *cell_end, int64_t *nhalo, int64_t *nsend, int *npeers)
{
    if (cell_begin) *cell_begin = p->cell_begin;
    if (cell_end) *cell_end = p->cell_end;
    if (nhalo) *nhalo = (int64_t) p->halo_cells.size();
    if (nsend) *nsend = (int64_t) p->send_cells.size();
    if (npeers) *npeers = (int) p->peers.size();
}

yfd_status yfd_partition_get(const yfd_partition *p, int64_t *halo_cells, int64_t *send_cells, int32_t *peers,
                             int64_t *recv_counts, int64_t *send_counts)
{
    if (!p) return YFD_ERR_INVALID_ARGUMENT;
    if (halo_cells) std::copy(p->halo_cells.begin(), p->halo_cells.end(), halo_cells);
    if (send_cells) std::copy(p->send_cells.begin(), p->send_cells.end(), send_cells);
    if (peers) std::copy(p->peers.begin(), p->peers.end(), peers);
    if (recv_counts) std::copy(p->recv_counts.begin(), p->recv_counts.end(), recv_counts);
    if (send_counts) std::copy(p->send_counts.begin(), p->send_counts.end(), send_counts);
    return YFD_OK;
}

} /* extern "C" */

/*
 * yfd_internal.hpp -- shared declarations of libyfd_b200.so (not part of the ABI).
 *
 * Everything behind include/yfd.h lives on ONE device per context and is launched on
 * ONE stream per context.  No CPU fallback exists anywhere in this library.
 */
#pragma once
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>
#include <cstdint>
#include <cstdio>
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <memory>
#include <map>
#include "yfd.h"

/* ---- peer-memory (NVLink / NVSwitch) control structures: every rank allocates one region with cudaMalloc,
 * the ranks swap CUDA IPC handles once, and from then on device code stores straight into its peers' regions
 * (yfd_peer.cu).  Layout of a region: [ peer_ctl | landing zone for the neighbours' y = A d on this rank's halo cells ]. ---- */
constexpr int YFD_MAX_RANKS = 16;
struct peer_slot { unsigned long long pk[4]; };       /* two values x two 8-byte packets { 32 data bits | epoch tag }: yfd_cg_shared.cuh */
struct peer_ctl {
    unsigned long long red_epoch[4];            /* local counters, one per reduction kind            */
    unsigned long long pad1[4];
    peer_slot          red[2][4][YFD_MAX_RANKS]; /* [epoch parity][kind][source rank], written by peers */
};
struct peer_view {                              /* lives in device memory, one per context             */
    peer_ctl *ctl[YFD_MAX_RANKS];               /* ctl[q]: rank q's control block as seen from here    */
    double   *land[YFD_MAX_RANKS];              /* land[q]: rank q's landing zone                      */
    int       me, nranks;
};

struct yfd_ctx {
    int           device = 0;
    cudaStream_t  stream = nullptr;
    bool          own_stream = false;
    int           sm_count = 0;
    int           cc_major = 0, cc_minor = 0;
    size_t        smem_optin = 0, smem_per_sm = 0;
    std::string   name;
    std::string   err;
    int64_t       launches = 0;
    /* quadrature tables on the device (yfd_tables.cu), freed by yfd_ctx_destroy */
    double       *tab_gauss_x = nullptr, *tab_gauss_w = nullptr, *tab_dunavant = nullptr;
    /* NCCL (yfd_comm.cpp) */
    void         *comm = nullptr;
    int           rank = 0, nranks = 1;
    /* device-memory cache: blocks released by dbuf are kept and handed out again (all work of a context is ordered
     * on its one stream, so reuse is safe); cudaMalloc / cudaFree synchronise the device and, once peer mappings
     * exist, cost milliseconds each */
    std::multimap<size_t, void *> pool;
    size_t        pool_bytes = 0;
    /* peer-memory transport (yfd_peer.cu); null -> NCCL calls are used instead */
    void         *peer_region = nullptr;       /* this rank's region (cudaMalloc)                     */
    size_t        peer_land_cap = 0;           /* doubles in the landing zone                         */
    double       *peer_land = nullptr;         /* this rank's landing zone                            */
    void         *peer_opened[YFD_MAX_RANKS] = {};   /* cudaIpcOpenMemHandle results, to close later  */
    peer_view    *d_peer_view = nullptr;       /* device copy handed to the kernels                   */
    bool          peer_enabled = false;
    bool          peer_failed = false;          /* a spin bound was hit on this rank ...                  */
    bool          peer_off = false;             /* ... and the ranks have agreed to stay on NCCL          */
};

namespace yfd {

yfd_status fail(yfd_ctx *ctx, yfd_status s, const char *fmt, ...);
/* yfd_ctx.cu: cached device allocations (bytes are rounded up to 512) */
void *pool_alloc(yfd_ctx *ctx, size_t bytes);
void  pool_free(yfd_ctx *ctx, void *p, size_t bytes);
void  pool_trim(yfd_ctx *ctx);

#define YFD_CUDA(ctx, call)                                                                   \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess)                                                                \
            return yfd::fail((ctx), e_ == cudaErrorMemoryAllocation ? YFD_ERR_OUT_OF_MEMORY   \
                                                                    : YFD_ERR_CUDA,           \
                             "%s:%d: %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
    } while (0)

#define YFD_TRY(call)                                  \
    do {                                               \
        yfd_status s_ = (call);                        \
        if (s_ != YFD_OK) return s_;                   \
    } while (0)

/* dynamic shared memory from which a kernel is opted in (cudaFuncAttributeMaxDynamicSharedMemorySize).  The 48 KiB a kernel
 * gets without opting in cover its STATIC shared memory too (reduction scratch: a few hundred bytes), so a request of exactly
 * 48 KiB -- 6 classes of bs = 15 triangle fragments -- fails to launch unless it is opted in as well. */
constexpr size_t YFD_SMEM_OPTIN_FROM = 40 * 1024;

/* kernel launch on the context's stream, counted (bench.py reports gpu_launches) */
#define YFD_LAUNCH(ctx, kernel, grid, block, smem, ...)                                        \
    do {                                                                                      \
        kernel<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__);                      \
        (ctx)->launches++;                                                                    \
        cudaError_t e_ = cudaPeekAtLastError();                                               \
        if (e_ != cudaSuccess) {                                                              \
            cudaGetLastError();                       /* reported once, not again by the next launch */ \
            return yfd::fail((ctx), YFD_ERR_CUDA, "%s:%d: launch %s: %s", __FILE__, __LINE__,  \
                             #kernel, cudaGetErrorString(e_));                                \
        }                                                                                     \
    } while (0)

/* device array with explicit lifetime (freed on the context's device) */
template <typename T>
struct dbuf {
    T     *p = nullptr;
    size_t n = 0;
    yfd_ctx *owner = nullptr;
    size_t bytes_held = 0;
    dbuf() = default;
    dbuf(const dbuf &) = delete;
    dbuf &operator=(const dbuf &) = delete;
    dbuf(dbuf &&o) noexcept : p(o.p), n(o.n), owner(o.owner), bytes_held(o.bytes_held) { o.p = nullptr; o.n = 0; o.owner = nullptr; o.bytes_held = 0; }
    dbuf &operator=(dbuf &&o) noexcept {
        if (this != &o) {
            release();
            p = o.p; n = o.n; owner = o.owner; bytes_held = o.bytes_held;
            o.p = nullptr; o.n = 0; o.owner = nullptr; o.bytes_held = 0;
        }
        return *this;
    }
    ~dbuf() { release(); }
    void release() {
        if (p) { if (owner) pool_free(owner, p, bytes_held); else cudaFree(p); }
        p = nullptr; n = 0; owner = nullptr; bytes_held = 0;
    }
    /* `pad` extra elements are allocated (and zeroed) past n so that aligned vector
     * loads of the last quad never leave the allocation */
    yfd_status alloc(yfd_ctx *ctx, size_t count, size_t pad = 0) {
        release();
        size_t bytes = (count + pad) * sizeof(T);
        if (bytes == 0) bytes = sizeof(T);
        bytes = (bytes + 511) & ~(size_t) 511;
        p = (T *) pool_alloc(ctx, bytes);
        if (!p) return fail(ctx, YFD_ERR_OUT_OF_MEMORY, "cudaMalloc(%zu bytes) failed", bytes);
        owner = ctx; bytes_held = bytes;
        cudaError_t e = cudaSuccess;
        n = count;
        if (pad) {
            e = cudaMemsetAsync(p + count, 0, pad * sizeof(T), ctx->stream);
            if (e != cudaSuccess) return fail(ctx, YFD_ERR_CUDA, "cudaMemsetAsync: %s", cudaGetErrorString(e));
        }
        return YFD_OK;
    }
    yfd_status zero(yfd_ctx *ctx) {
        if (!n) return YFD_OK;
        cudaError_t e = cudaMemsetAsync(p, 0, n * sizeof(T), ctx->stream);
        if (e != cudaSuccess) return fail(ctx, YFD_ERR_CUDA, "cudaMemsetAsync: %s", cudaGetErrorString(e));
        return YFD_OK;
    }
    yfd_status upload(yfd_ctx *ctx, const T *host, size_t count) {
        if (!count) return YFD_OK;
        cudaError_t e = cudaMemcpyAsync(p, host, count * sizeof(T), cudaMemcpyHostToDevice, ctx->stream);
        if (e != cudaSuccess) return fail(ctx, YFD_ERR_CUDA, "H2D copy: %s", cudaGetErrorString(e));
        return YFD_OK;
    }
    yfd_status download(yfd_ctx *ctx, T *host, size_t count) const {
        if (!count) return YFD_OK;
        cudaError_t e = cudaMemcpyAsync(host, p, count * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream);
        if (e != cudaSuccess) return fail(ctx, YFD_ERR_CUDA, "D2H copy: %s", cudaGetErrorString(e));
        e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) return fail(ctx, YFD_ERR_CUDA, "sync after D2H: %s", cudaGetErrorString(e));
        return YFD_OK;
    }
};

/* NVTX range per phase (header-only NVTX 3: a no-op unless a profiler injects itself): nsys / ncu timelines show
 * mesh -> structure -> elements -> classes -> CSR -> CG set-up -> CG iterations, per rank */
struct nvtx_range {
    explicit nvtx_range(const char *name) { nvtxRangePushA(name); }
    ~nvtx_range() { nvtxRangePop(); }
    nvtx_range(const nvtx_range &) = delete;
    nvtx_range &operator=(const nvtx_range &) = delete;
};

struct scoped_device {
    int prev = -1;
    explicit scoped_device(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~scoped_device() { if (prev >= 0) cudaSetDevice(prev); }
};

struct gpu_timer {
    cudaEvent_t a = nullptr, b = nullptr;
    gpu_timer() { cudaEventCreate(&a); cudaEventCreate(&b); }
    ~gpu_timer() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); }
    void start(cudaStream_t s) { cudaEventRecord(a, s); }
    void stop(cudaStream_t s) { cudaEventRecord(b, s); }
    double ms() { float t = 0; cudaEventSynchronize(b); cudaEventElapsedTime(&t, a, b); return (double) t; }
};

inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
/* diagnostic switches: read on every launch, so a test can change them between calls */
inline int env_int(const char *name, int dflt)
{
    const char *e = getenv(name);
    return e && *e ? atoi(e) : dflt;
}
inline int basis_size_2d(int k) { return (k + 2) * (k + 1) / 2; }

/* ---- yfd_sort.cu: scans, radix sort, small device utilities ---- */
/* out[i] = sum_{j<i} in[j]; out has n+1 entries (out[n] = total) */
yfd_status exclusive_scan_i32_to_i64(yfd_ctx *ctx, const int32_t *d_in, int64_t *d_out, int64_t n);
/* stable LSD radix sort of (key, value) pairs on key bits [0, key_bits); result ends in (keys, vals);
 * keys_tmp / vals_tmp are scratch of the same size */
yfd_status radix_sort_pairs(yfd_ctx *ctx, uint64_t *d_keys, uint32_t *d_vals,
                            uint64_t *d_keys_tmp, uint32_t *d_vals_tmp, int64_t n, int key_bits);
yfd_status fill_iota_u32(yfd_ctx *ctx, uint32_t *d, int64_t n);

/* ---- yfd_mesh.cu ---- */
yfd_status upload_narrow(yfd_ctx *ctx, const int64_t *host, int64_t n, int64_t lo, int64_t hi, dbuf<int32_t> &out, int *d_err);
yfd_status download_widen(yfd_ctx *ctx, const int32_t *d_in, int64_t n, int64_t *host);

/* ---- yfd_tables.cu: the reference's tabulated rules (include/yfd_quad_tables.h) on the device ---- */
struct quad_rule { const double *x = nullptr, *w = nullptr; int n = 0; };   /* 1D Gauss: x[n], w[n]          */
struct tri_rule  { const double *cw = nullptr; int n = 0; };                /* Dunavant: (c0,c1,c2,w)[n]     */
yfd_status ensure_tables(yfd_ctx *ctx);
/* gauss_legendre_tab(degree): rule degree/2 -> degree/2+1 points (core/quadratures.hpp:296-311) */
yfd_status gauss_rule_for_degree(yfd_ctx *ctx, int degree, quad_rule *out);
/* triangle_quadrature_dunavant(degree): rule max(degree,1) (core/quadratures.hpp:570-581) */
yfd_status dunavant_rule_for_degree(yfd_ctx *ctx, int degree, tri_rule *out);

/* ---- yfd_partition.cpp ---- */
int64_t partition_begin(int64_t ncells, int nranks, int r);
int     partition_owner(int64_t ncells, int nranks, int64_t cell);

/* ---- yfd_comm.cpp: NCCL on the context's stream (all capturable into CUDA graphs) ---- */
yfd_status comm_allreduce_sum_f64(yfd_ctx *ctx, double *d_buf, size_t count);          /* in place */
yfd_status comm_group_begin(yfd_ctx *ctx);
yfd_status comm_send_f64(yfd_ctx *ctx, const double *d_buf, size_t count, int peer);
yfd_status comm_recv_f64(yfd_ctx *ctx, double *d_buf, size_t count, int peer);
yfd_status comm_group_end(yfd_ctx *ctx);
yfd_status comm_allgather_bytes(yfd_ctx *ctx, const void *d_send, void *d_recv, size_t bytes_per_rank);

/* ---- yfd_peer.cu ---- */
/* collective: make sure every rank's landing zone holds at least `land_doubles` (same value on all ranks);
 * on failure the context silently stays on the NCCL transport (peer_enabled = false) */
yfd_status ensure_peer_region(yfd_ctx *ctx, size_t land_doubles);
void       peer_release(yfd_ctx *ctx, bool collective = false);

/* ---- yfd_dist.cu: this rank's share of a replicated mesh ---- */
struct dist_plan {
    int      nranks = 1, rank = 0;
    int64_t  ncells_global = 0;
    int64_t  cell_begin = 0, cell_end = 0;       /* owned block of the internal order */
    int64_t  nown = 0, nhalo = 0, nsend = 0;
    std::vector<int32_t> peers;
    std::vector<int64_t> recv_off, send_off;     /* [npeers+1], in cells */
    dbuf<int32_t> g2l;                           /* [ncells_global] internal id -> local index (owned: id-cell_begin, halo: nown+k), -1 elsewhere */
    dbuf<int32_t> send_lidx;                     /* [nsend] local indices of the owned cells to pack, grouped by peer */
    /* peer-memory transport: where each packed cell lands on its destination rank */
    int64_t  max_halo_cells = 0;                 /* max over ALL ranks                                        */
    int64_t  max_local_cells = 0;                /* max over ALL ranks of owned + halo (sizes the landing zones identically) */
    dbuf<int32_t> send_rank;                     /* [nsend] destination rank                                 */
    dbuf<int32_t> send_dst;                      /* [nsend] cell slot in the destination's landing zone      */
    dbuf<int32_t> d_peers;                       /* [npeers]                                                 */
};

} // namespace yfd

struct yfd_mesh;
namespace yfd {
yfd_status ensure_dist_plan(yfd_ctx *ctx, yfd_mesh *mesh);
/* fill the halo part of x ([ (nown+nhalo)*bs ]) from the owners; sendbuf holds nsend*bs doubles */
yfd_status halo_exchange(yfd_ctx *ctx, const dist_plan &pl, int bs, double *x, double *sendbuf);
yfd_status mesh_neighbours_device(yfd_ctx *ctx, yfd_mesh *m, dbuf<int32_t> &nbr);
} // namespace yfd

/* ---- objects behind the opaque handles ---- */
struct yfd_mesh {
    yfd_ctx *ctx = nullptr;
    int      cell_type = 0, nv = 3, nf = 3;
    int64_t  npoints = 0, ncells = 0, nfaces = 0;
    int      reorder = 0;
    yfd::dbuf<double>   points;       /* [npoints][2]                               */
    yfd::dbuf<int32_t>  cells;        /* [ncells][nv]    REFERENCE cell order       */
    yfd::dbuf<int32_t>  faces;        /* [nfaces][2]                                */
    yfd::dbuf<uint8_t>  face_bnd;     /* [nfaces]                                   */
    yfd::dbuf<int32_t>  cell_faces;   /* [ncells][nf]    REFERENCE cell order       */
    yfd::dbuf<int32_t>  face_owners;  /* [nfaces][2]     REFERENCE cell ids, -1 none */
    yfd::dbuf<int32_t>  perm;         /* [ncells] internal -> reference             */
    yfd::dbuf<int32_t>  iperm;        /* [ncells] reference -> internal             */
    std::shared_ptr<yfd::dist_plan> plan;   /* built on first distributed use (yfd_dist.cu) */
};

struct yfd_csr {
    yfd_ctx *ctx = nullptr;
    int64_t  nrows = 0, ncols = 0, nnz = 0;
    int64_t  max_chunk_nnz = -1;      /* largest 32-row chunk (yfd_spmv_csr.cu), -1 = not computed yet */
    yfd::dbuf<int64_t> rowptr;        /* [nrows+1]                                  */
    yfd::dbuf<int32_t> colidx;        /* [nnz] (+4 pad)                             */
    yfd::dbuf<double>  values;        /* [nnz] (+4 pad)                             */
};

enum yfd_system_kind { YFD_SYS_DG = 1, YFD_SYS_CFEM = 2 };

namespace yfd {
/* congruence-class operator cache of a DG system (yfd_spmv_class.cu): one block row per class instead of one per cell */
struct dg_classes {
    bool     ready = false;           /* the solvers use the cache                                                    */
    int      reject = 0;              /* why not: 1 = too many classes (> 4096 or tables beyond 8 MB), 2 = too much
                                         padding (many small classes), 3 = a block row deviates from its class by > tol */
    int      ncls = 0;
    int      table_doubles = 0;
    int64_t  npos = 0, ntiles = 0;    /* work list length (classes padded to 32), tiles of 32                           */
    double   max_dev = 0.0, tol = 0.0;/* largest |W_cell - W_class| / max|W_class| found by the verification pass        */
    dbuf<double>  table;              /* [ncls][nf+1][slot stride]: block rows in canonical slot order (self, faces)     */
    dbuf<double>  table_mma;          /* the same blocks as mma.sync fragments (yfd_class_mma.cuh); bs = 6, 10, 15 only        */
    dbuf<int32_t> work_cell;          /* [npos] local cell index, -1 = padding                                           */
    dbuf<int32_t> work_nbr;           /* [nf][npos] local x cell index of the neighbour across each local face           */
    dbuf<int32_t> tile_cls;           /* [ntiles]                                                                        */
    dbuf<int32_t> cell_cls;           /* [ncells]                                                                        */
};

/* cells grouped into classes by a key (yfd_spmv_class.cu: group_cells_by_key) */
struct class_groups {
    int64_t  ncls = 0, npos = 0, ntiles = 0;
    int      reject = 0;
    dbuf<uint32_t> sorted_cell;       /* [ncells] cells class by class, ascending id inside a class                    */
    dbuf<int64_t>  start;             /* [ncls+1] into sorted_cell: start[k] is class k's representative                */
    dbuf<int32_t>  tile_cls, work_cell, cell_cls;
};

/* p-multigrid preconditioner (yfd_pmg.cu): levels k, k-1, ..., 1 over the class cache */
struct pmg_level {
    int    k = 0, bs = 0, nA = 0, nD = 0;
    double lmax = 0.0;                /* 1.1 x the power-iteration estimate of lambda_max(D^-1 A_l)                     */
    dbuf<double> tabA, tabD;          /* per class: the level's block row (leading sub-blocks), its inverse diagonal block */
    dbuf<double> x[2], d, b;          /* work vectors of the level (the finest level's b and second x are the CG's r, z) */
};
struct pmg_hierarchy {
    int    nlevels = 0, nu = 2, coarse_degree = 0;
    double smooth_ratio = 4.0;
    pmg_level lv[4];
};
}

struct yfd_system {
    yfd_ctx *ctx = nullptr;
    int      kind = 0;
    int64_t  n = 0, nnz = 0;
    /* matrix, rhs, Jacobi diagonal in INTERNAL numbering */
    yfd_csr  A;
    yfd::dbuf<double>  rhs;
    yfd::dbuf<double>  pc;            /* 1/diag, empty if not built                 */
    bool     symmetric = true;        /* false: advection-reaction; CG then needs use_normal_eqns and the transposed blocks */
    yfd::dbuf<double>  values_t;      /* DG: the blocks of A^T in the same layout, built on first use (yfd_cg.cu)            */
    /* DG block-Jacobi preconditioner, built on first use (yfd_block_jacobi.cuh): inverse diagonal blocks of the owned
     * (+ halo) cells as a one-block-per-cell matrix in the layout of A */
    yfd::dbuf<double>  pc_block;
    yfd::dbuf<double>  pc_block_halo; /* several ranks: the halo cells' blocks again, in an allocation of their own (16-byte aligned) */
    yfd::dbuf<int64_t> pcb_cellptr;
    yfd::dbuf<int32_t> pcb_cell;
    /* DG block structure: cell c owns rows [c*bs, (c+1)*bs); its blocks are listed in
     * blk_cell[cellptr[c] .. cellptr[c+1]) in ascending internal cell id */
    int      bs = 0, nf = 0;
    int64_t  ncells = 0;              /* cells whose rows this rank holds (all of them on one GPU) */
    int64_t  ncells_global = 0;
    int64_t  n_local = 0;             /* rows held by this rank                     */
    std::shared_ptr<yfd::dist_plan> plan;   /* null on one GPU                      */
    yfd::dbuf<int64_t> cellptr;       /* [ncells+1]                                 */
    yfd::dbuf<int32_t> blk_cell;      /* [nblocks]                                  */
    yfd::dbuf<int8_t>  slots;         /* [ncells][nf+1]: position of the diagonal block and of the block across each local
                                         face inside the cell's (id-sorted) block list, -1 on boundary faces */
    yfd::dg_classes    classes;       /* SIP diffusion on meshes of congruent cells (yfd_spmv_class.cu)      */
    int      operator_mode = 0;       /* YFD_OPERATOR_AUTO / _STREAMED / _CLASSES (yfd_system_set_operator_mode) */
    std::unique_ptr<yfd::pmg_hierarchy> pmg;   /* built on first YFD_PC_PMG solve (yfd_pmg.cu)                   */
    int      pmg_nu = 0, pmg_coarse_degree = 0; /* 0 = defaults (yfd_system_set_pmg_params)                      */
    double   pmg_smooth_ratio = 0.0;
    yfd::dbuf<int32_t> perm, iperm;   /* copies of the mesh's cell permutation      */
    /* CFEM */
    int64_t  npoints = 0;
    yfd::dbuf<int32_t> compress_map;  /* [npoints], -1 on Dirichlet nodes           */
};

namespace yfd {
/* ---- yfd_system.cu ---- */
/* DG: this rank's rows (internal order) -> the FULL vector in reference numbering, on the device (d_out_ref [n]);
 * with several ranks the parts are combined by an all-reduce, so every rank ends up with the whole vector */
yfd_status system_vector_to_ref(yfd_system *sys, const double *d_local, double *d_out_ref);
/* ---- yfd_spmv_class.cu ---- */
yfd_status group_cells_by_key(yfd_ctx *ctx, int64_t nc, uint64_t *d_keys, int key_bits, size_t table_bytes_per_class, size_t table_budget,
                              class_groups &g);
yfd_status build_dg_classes(yfd_ctx *ctx, yfd_system *sys, uint64_t *d_keys, int key_bits, double tol);
yfd_status expand_class_values(yfd_ctx *ctx, const yfd_system *sys, double *d_values_out);
inline bool system_uses_classes(const yfd_system *sys) { return sys->classes.ready && sys->operator_mode != 1 /* YFD_OPERATOR_STREAMED */; }
/* ---- yfd_pmg.cu ---- */
yfd_status ensure_pmg(yfd_ctx *ctx, yfd_system *sys);
/* ---- yfd_triplets.cu ---- */
int key_bits_for(int64_t n);
/* keys = row << col_bits | col, already on the device (clobbered); folds duplicates in input order.  Entries of
 * `discard_row` (-1: none) keep their slot but are not summed (the CFEM assemblers' sentinel row) */
yfd_status csr_from_cell_rows(yfd_ctx *ctx, int64_t n, int64_t ncells, int ld, const uint64_t *d_mkeys, int col_bits, int row_bits,
                              const double *d_mvals, const uint64_t *d_rkeys, const double *d_rvals, yfd_csr *A, double *d_rhs, bool *built);
yfd_status csr_from_device_triplets(yfd_ctx *ctx, int64_t nrows, int64_t ncols, int64_t count,
                                    uint64_t *d_keys, int col_bits, int row_bits, const double *d_vals, yfd_csr *A,
                                    int64_t discard_row = -1);
} // namespace yfd

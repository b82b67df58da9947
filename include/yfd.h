/*
 * include/yfd.h -- C ABI of libyfd_b200.so: the B200 (sm_100a) implementation of
 * Yaourt-FEM-DG's assemble-and-solve hot path.
 *
 * The reference (datafl4sh/yaourt-fem-dg) has no FFI: its "API" for this path is a
 * set of header-only C++ templates.  Each entry point below states the reference
 * interface (file:line, relative to the reference root) whose work it replaces; the
 * C++ shim in include/yaourt_b200/ puts the reference's own names back on top of
 * this ABI, and INTEGRATION.md shows the binding a maintainer would add.
 *
 * Conventions
 *   - plain pointers and sizes only; every array argument is a caller-owned HOST
 *     buffer unless its name starts with d_ (device pointer, for resident pipelines);
 *   - objects live on the GPU behind opaque handles with explicit *_destroy;
 *   - every function returns a yfd_status; yfd_last_error(ctx) gives the message.
 *     There is NO CPU fallback: without a usable CUDA device yfd_ctx_create fails
 *     with YFD_ERR_NO_DEVICE and nothing else can be called;
 *   - externally visible numbering is the reference's, bit for bit: cell id = index
 *     in the sorted cell list, face id = index in the sorted face list, DG dof =
 *     cell*bs + local (methods/dg.hpp:75-84), CFEM dof = rank among non-Dirichlet
 *     nodes (methods/cfem.hpp:140-148).  Internally cells are renumbered along a
 *     space-filling curve (yfd_mesh_create, `reorder`), invisible at this boundary.
 */
#ifndef YFD_H
#define YFD_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define YFD_VERSION 100

typedef enum yfd_status {
    YFD_OK = 0,
    YFD_ERR_INVALID_ARGUMENT = 1,   /* std::invalid_argument at the C++ boundary            */
    YFD_ERR_INVALID_STATE    = 2,   /* "Assembler in invalid state" (methods/dg.hpp:72-73)  */
    YFD_ERR_CELL_NOT_FOUND   = 3,   /* "Mesh cell not found"  (core/mesh.hpp:281)           */
    YFD_ERR_FACE_NOT_FOUND   = 4,   /* "Mesh face not found"  (core/mesh.hpp:293)           */
    YFD_ERR_DEGREE_TOO_HIGH  = 5,   /* "Dunavant quadrature: degree too high" (core/quadratures.hpp:579) */
    YFD_ERR_NO_NEIGHBOURS    = 6,   /* "No neighbour information." (core/mesh.hpp:256)      */
    YFD_ERR_TOO_MANY_OWNERS  = 7,   /* "BUG: a face has max 2 owners" (core/mesh.hpp:199)   */
    YFD_ERR_SHAPE_MISMATCH   = 8,   /* CG: non-square A / wrong vector size -> `return false` (core/solvers.hpp:53-77) */
    YFD_ERR_UNSUPPORTED      = 9,
    YFD_ERR_NO_DEVICE        = 10,  /* no CUDA device / driver: there is no CPU path        */
    YFD_ERR_CUDA             = 11,
    YFD_ERR_COMM             = 12,  /* NCCL failure                                          */
    YFD_ERR_OUT_OF_MEMORY    = 13
} yfd_status;

const char *yfd_status_string(yfd_status s);

/* ---------------------------------------------------------------------------------- */
/* Context: one per GPU rank.  One CUDA stream per context, re-entrant per handle.     */
/* ---------------------------------------------------------------------------------- */
typedef struct yfd_ctx yfd_ctx;

/* `stream` = a cudaStream_t to launch on (e.g. torch's current stream) or NULL to let
 * the context create its own non-blocking stream. */
yfd_status  yfd_ctx_create(int device, void *stream, yfd_ctx **out);
void        yfd_ctx_destroy(yfd_ctx *ctx);
const char *yfd_last_error(const yfd_ctx *ctx);
yfd_status  yfd_ctx_synchronize(yfd_ctx *ctx);
/* Device buffers released by destroyed meshes / systems / solves are cached in the context and reused (cudaMalloc
 * and cudaFree synchronise the device); this returns the cache to the driver. released_bytes may be NULL. */
yfd_status  yfd_ctx_trim(yfd_ctx *ctx, size_t *released_bytes);
/* kernels launched by this context so far (graph replays count their kernel nodes) */
int64_t     yfd_ctx_launch_count(const yfd_ctx *ctx);
/* name and SM count of the device (buf may be NULL) */
yfd_status  yfd_ctx_device_info(const yfd_ctx *ctx, char *name, size_t cap, int *sm_count, int *cc_major, int *cc_minor);

/* FP64 FMA throughput of the device, measured now (16 independent DFMA chains per thread, 8 x 256 threads per SM, best of
 * 5 launches, CUDA events): the denominator for the compute fraction of the element kernels and of the class SpMV */
yfd_status  yfd_ctx_measure_fp64(yfd_ctx *ctx, double *tflops);
/* the same for the FP64 TENSOR cores (mma.sync m8n8k4 f64 = DMMA.8x8x4, 8 independent accumulator chains per warp): the
 * evidence north_star asks for before any DMMA path ("only if ncu shows a win over FP64 FMA") */
yfd_status  yfd_ctx_measure_fp64_tensor(yfd_ctx *ctx, double *tflops);

/* Multi-GPU: attach an NCCL communicator (one rank per context).  `unique_id` is the
 * 128-byte ncclUniqueId made by rank 0 (yfd_comm_unique_id) and shipped by the host
 * program (torch.distributed broadcast in bench.py).  The reference has no
 * distributed layer; this is new. */
#define YFD_COMM_ID_BYTES 128
yfd_status  yfd_comm_unique_id(unsigned char id[YFD_COMM_ID_BYTES]);
yfd_status  yfd_ctx_attach_comm(yfd_ctx *ctx, int rank, int nranks, const unsigned char id[YFD_COMM_ID_BYTES]);
int         yfd_ctx_rank(const yfd_ctx *ctx);
int         yfd_ctx_nranks(const yfd_ctx *ctx);

/* ---------------------------------------------------------------------------------- */
/* Mesher (host): replaces mesher<Mesh>::create_mesh + refine_mesh                     */
/* (core/meshers.hpp:130-179 triangles, :302-361 quadrangles).  Same unit-square base   */
/* meshes, same red refinement, same point / face / cell numbering.                    */
/* ---------------------------------------------------------------------------------- */
#define YFD_CELL_TRIANGLE   0
#define YFD_CELL_QUADRANGLE 1

typedef struct yfd_host_mesh yfd_host_mesh;

yfd_status yfd_mesher_create(int cell_type, int refinement_levels, yfd_host_mesh **out);
/* New (the reference mesher only offers 4^(r+1) cells): nx x ny structured mesh of the unit square in the reference's
 * conventions -- cells (p00, p10, p11, p01) or the two triangles (p00, p10, p11), (p00, p11, p01), sorted cells and faces,
 * boundary ids 0 bottom / 1 right / 2 top / 3 left.  The reference consumes such a mesh unchanged (its public
 * points / faces / cells vectors, core/mesh.hpp:239-243), which is how parity on these meshes is checked. */
yfd_status yfd_mesher_create_structured(int cell_type, int64_t nx, int64_t ny, yfd_host_mesh **out);
void       yfd_host_mesh_destroy(yfd_host_mesh *m);
void       yfd_host_mesh_sizes(const yfd_host_mesh *m, int64_t *npoints, int64_t *ncells,
                               int64_t *nfaces, int *verts_per_cell);
/* borrowed pointers, valid until destroy:
 *   points [npoints][2] f64 ; cells [ncells][nv] i64 (vertex order as in the reference,
 *   NOT sorted inside a cell) ; faces [nfaces][2] i64 with p0 < p1, sorted ;
 *   face_boundary [nfaces] u8 ; face_boundary_id [nfaces] i64 (0 for interior faces,
 *   where the reference leaves it unset). */
const double  *yfd_host_mesh_points(const yfd_host_mesh *m);
const int64_t *yfd_host_mesh_cells(const yfd_host_mesh *m);
const int64_t *yfd_host_mesh_faces(const yfd_host_mesh *m);
const uint8_t *yfd_host_mesh_face_boundary(const yfd_host_mesh *m);
const int64_t *yfd_host_mesh_face_boundary_id(const yfd_host_mesh *m);

/* ---------------------------------------------------------------------------------- */
/* Device mesh: replaces the mesh<T,2,CellT,FaceT> the drivers pass around plus         */
/* mesh::compute_connectivity (core/mesh.hpp:176-204), face_ids / offset(face)          */
/* (core/mesh.hpp:287-296, :329-353).                                                  */
/* ---------------------------------------------------------------------------------- */
typedef struct yfd_mesh yfd_mesh;

#define YFD_REORDER_NONE   0    /* internal cell order = reference order                */
#define YFD_REORDER_MORTON 1    /* internal cell order = Z-curve of cell barycentres    */

/* cells must be sorted as the reference sorts them (lexicographic on the vertex tuple,
 * core/meshers.hpp:178,:360) and faces sorted by (p0,p1) with p0<p1 (core/mesh.hpp:77-81);
 * both are checked on the device: YFD_ERR_INVALID_ARGUMENT if not, YFD_ERR_FACE_NOT_FOUND
 * if a cell edge is missing from `faces`, YFD_ERR_TOO_MANY_OWNERS as in the reference. */
yfd_status yfd_mesh_create(yfd_ctx *ctx, int cell_type,
                           int64_t npoints, const double *points,
                           int64_t ncells, const int64_t *cells,
                           int64_t nfaces, const int64_t *faces, const uint8_t *face_boundary,
                           int reorder, yfd_mesh **out);
void       yfd_mesh_destroy(yfd_mesh *m);
void       yfd_mesh_sizes(const yfd_mesh *m, int64_t *npoints, int64_t *ncells, int64_t *nfaces, int *verts_per_cell);
/* face_owners [nfaces][2] (i64, -1 = NO_OWNER) exactly as compute_connectivity leaves them */
yfd_status yfd_mesh_get_face_owners(yfd_mesh *m, int64_t *owners);
/* cell_faces [ncells][nf]: face_ids(msh, cell) in the reference's local face order
 * (tri: p0p1,p1p2,p0p2 ; quad: p0p1,p1p2,p2p3,p0p3 ; core/mesh.hpp:329-353) */
yfd_status yfd_mesh_get_cell_faces(yfd_mesh *m, int64_t *cell_faces);
/* perm[internal cell] = reference cell id */
yfd_status yfd_mesh_get_permutation(yfd_mesh *m, int64_t *perm);
/* replace the coordinates (shatter_mesh output etc.); same count as at creation */
yfd_status yfd_mesh_set_points(yfd_mesh *m, const double *points);

/* neighbours [ncells][nf] in INTERNAL cell order and ids: the cell across each local face, -1 on the
 * boundary (neighbour_via, core/mesh.hpp:249-272, for every cell at once) */
yfd_status yfd_mesh_get_neighbours(yfd_mesh *m, int64_t *neighbours);

/* ---------------------------------------------------------------------------------- */
/* Partition plan (host, no CUDA): new, SURVEY.md section 8e.  Cells are split into     */
/* `nranks` contiguous blocks of the internal order; from the neighbour table each rank  */
/* derives its halo (grouped by owner rank) and its per-peer send lists.  The assembly   */
/* and solve entry points below build and use this plan themselves when the context has  */
/* a communicator attached; it is exposed so host programs and tests can inspect it.     */
/* ---------------------------------------------------------------------------------- */
typedef struct yfd_partition yfd_partition;
yfd_status yfd_partition_create(int64_t ncells, int nf, const int64_t *neighbours, int nranks, int rank, yfd_partition **out);
/* The same plan from the cut faces alone: pairs (cut_cell[i], cut_nbr[i]) of face-adjacent cells owned by different ranks
 * (both directions listed; pairs of other ranks are ignored, so one list of all cuts serves every rank).  This is what the
 * library uses itself: the device finds the O(sqrt(cells)) cut faces and only those reach the host. */
yfd_status yfd_partition_create_from_cuts(int64_t ncells, int nranks, int rank, int64_t ncut, const int64_t *cut_cell,
                                          const int64_t *cut_nbr, yfd_partition **out);
void       yfd_partition_destroy(yfd_partition *p);
void       yfd_partition_sizes(const yfd_partition *p, int64_t *cell_begin, int64_t *cell_end, int64_t *nhalo,
                               int64_t *nsend, int *npeers);
/* halo_cells [nhalo], send_cells [nsend]: internal cell ids grouped by peer (peers ascending), ids ascending;
 * peers / recv_counts / send_counts [npeers] (counts in cells).  Any output may be NULL. */
yfd_status yfd_partition_get(const yfd_partition *p, int64_t *halo_cells, int64_t *send_cells, int32_t *peers,
                             int64_t *recv_counts, int64_t *send_counts);

/* ---------------------------------------------------------------------------------- */
/* Sparse matrix handle (row-major CSR, fp64 values, int32 columns, int64 row offsets)  */
/* = blaze::CompressedMatrix<double> at this boundary.                                  */
/* ---------------------------------------------------------------------------------- */
typedef struct yfd_csr yfd_csr;

/* Replaces blaze::init_from_triplets + reduce_triplets (core/blaze_sparse_init.hpp:101-155):
 * sort by (row,col), fold equal positions, build CSR.  Deterministic: equal positions are
 * added in the order they appear in the input (a stable sort), explicit zeros are kept as
 * the reference keeps them.  rows/cols/vals have `count` entries. */
yfd_status yfd_csr_from_triplets(yfd_ctx *ctx, int64_t nrows, int64_t ncols, int64_t count,
                                 const int64_t *rows, const int64_t *cols, const double *vals,
                                 yfd_csr **out);
/* Adopt an existing CSR (e.g. a blaze::CompressedMatrix walked row by row). */
yfd_status yfd_csr_from_arrays(yfd_ctx *ctx, int64_t nrows, int64_t ncols, int64_t nnz,
                               const int64_t *rowptr, const int64_t *colidx, const double *values,
                               yfd_csr **out);
void       yfd_csr_destroy(yfd_csr *A);
void       yfd_csr_sizes(const yfd_csr *A, int64_t *nrows, int64_t *ncols, int64_t *nnz);
/* any of the three outputs may be NULL */
yfd_status yfd_csr_get(yfd_csr *A, int64_t *rowptr, int64_t *colidx, double *values);
/* y = A x (host vectors): Blaze's `A*d` (blaze/math/expressions/SMatDVecMultExpr.h:204-207) */
yfd_status yfd_csr_spmv(yfd_ctx *ctx, yfd_csr *A, const double *x, double *y);

/* ---------------------------------------------------------------------------------- */
/* DG-SIP diffusion: replaces the whole cell loop of run_diffusion_solver               */
/* (dg2d_diffusion.cpp:133-201: make_basis, integrate, eval / eval_grads, K, Att, Atn,  */
/* loc_rhs), assembler<Mesh>::assemble x2 (methods/dg.hpp:67-123) and                   */
/* assembler::finalize (methods/dg.hpp:125-144).                                        */
/* ---------------------------------------------------------------------------------- */
#define YFD_FN_ZERO    0   /* data::dirichlet (dg2d_diffusion.cpp:61-66)                  */
#define YFD_FN_SINSIN  1   /* data::rhs = 2 pi^2 sin(pi x) sin(pi y) (dg2d_diffusion.cpp:53-59) */
#define YFD_FN_ONE     2
#define YFD_FN_ADVECTION_RHS 4 /* data::rhs of dg2d_advection.cpp:62-73: beta . grad u + mu u with u = sin(pi x), beta = (1,0), mu = 1 */
#define YFD_FN_SIN_X_SOL 5   /* data::advection_ref_sol = sin(pi x) (dg2d_advection.cpp:75-80); post-processing only */
#define YFD_FN_SINSIN_SOL 3 /* data::diffusion_ref_sol = sin(pi x) sin(pi y) (dg2d_diffusion.cpp:68-76); post-processing only */

typedef struct yfd_dg_config {
    int     degree;             /* cfg.degree, k >= 0; basis size (k+2)(k+1)/2 (core/bases.hpp:63-74) */
    double  eta;                /* cfg.eta; penalty = 3 k^2 eta / diameter(face) (dg2d_diffusion.cpp:127,162) */
    int     rhs_fn;             /* YFD_FN_*                                                */
    int     dirichlet_fn;       /* YFD_FN_*                                                */
    int     build_preconditioner; /* assembler(msh, degree, bpc): Jacobi pc (methods/dg.hpp:88-89,133-142) */
} yfd_dg_config;

typedef struct yfd_system yfd_system;   /* lhs + rhs (+ pc) of one assembled problem, on the device */

typedef struct yfd_assembly_info {
    int64_t system_size;        /* assembler::system_size()                                */
    int64_t nnz;
    int64_t basis_size;
    double  ms_elements;        /* device time: local blocks -> CSR values (CUDA events)   */
    double  ms_structure;       /* device time: sparsity pattern (rowptr / colidx)         */
} yfd_assembly_info;

/* With a communicator attached to ctx (yfd_ctx_attach_comm, nranks > 1) every rank passes the SAME
 * replicated mesh and assembles only the block rows of its own contiguous block of cells; the solve
 * entry points then exchange halos and reduce scalars over NCCL.  system_size stays the global size. */
yfd_status yfd_dg_assemble_sip(yfd_ctx *ctx, yfd_mesh *mesh, const yfd_dg_config *cfg,
                               yfd_system **out, yfd_assembly_info *info);

/* Advection-reaction (dg2d_advection.cpp:137-213: K = int mu phi phi^T + phi (beta . grad phi)^T, face terms with
 * fi_coeff = beta.n [- eta |beta.n| with use_upwinding], inflow boundary term; beta = (1,0), mu = 1 as params:: there).
 * cfg->eta is used unscaled (:133), cfg->rhs_fn should be YFD_FN_ADVECTION_RHS, dirichlet_fn is ignored.  The system is
 * NOT symmetric: solve it as the driver does, yfd_cg_params.use_normal_eqns = 1 (:225).  One GPU. */
yfd_status yfd_dg_assemble_advection(yfd_ctx *ctx, yfd_mesh *mesh, const yfd_dg_config *cfg, int use_upwinding,
                                     yfd_system **out, yfd_assembly_info *info);

/* Per-cell local blocks BEFORE the scatter, for unit-level parity with the reference loop:
 *   K   [ncells][bs][bs]      cell stiffness            (dg2d_diffusion.cpp:141-149)
 *   Att [ncells][nf][bs][bs]  per local face            (:171-173 interior, :186-188 boundary)
 *   Atn [ncells][nf][bs][bs]  zero where no neighbour   (:178-180)
 *   rhs [ncells][bs]                                    (:148, :190-191)
 * all indexed by REFERENCE cell id.  Any output may be NULL. */
yfd_status yfd_dg_local_blocks(yfd_ctx *ctx, yfd_mesh *mesh, const yfd_dg_config *cfg,
                               double *K, double *Att, double *Atn, double *rhs);

/* Post-processing of a DG solution (dg2d_diffusion.cpp:225-268): sol [ncells * bs] in REFERENCE numbering,
 *   *l2_errsq_qp = sum_cells sum_q w (ref(x_q) - u_h(x_q))^2          status.L2_errsq_qp (:261-262)
 *   *l2_errsq_mm = sum_cells (proj - u_c)^T M (proj - u_c)            status.L2_errsq_mm (:258-259,:265-266; the
 *                  reference's solve_LU -> LAPACK is an in-kernel pivoted LU here)
 * ref_fn: YFD_FN_SINSIN_SOL for the driver's manufactured solution.  Either output may be NULL.  The _device variant
 * takes sol as a DEVICE pointer (e.g. the d_x of yfd_system_solve_cg_device). */
yfd_status yfd_dg_postprocess(yfd_ctx *ctx, yfd_mesh *mesh, int degree, int ref_fn, const double *sol,
                              double *l2_errsq_qp, double *l2_errsq_mm);
yfd_status yfd_dg_postprocess_device(yfd_ctx *ctx, yfd_mesh *mesh, int degree, int ref_fn, const double *d_sol,
                                     double *l2_errsq_qp, double *l2_errsq_mm);

/* ---------------------------------------------------------------------------------- */
/* Continuous P1 FEM: replaces cfem::stiffness_matrix / eval_basis_grad                 */
/* (methods/cfem.hpp:53-89), compute_rhs (continuous_fem.cpp:33-50), and                */
/* cfem::assembler ctor / assemble / finalize (methods/cfem.hpp:119-186).               */
/* ---------------------------------------------------------------------------------- */
yfd_status yfd_cfem_assemble_p1(yfd_ctx *ctx, yfd_mesh *mesh, int rhs_fn,
                                yfd_system **out, yfd_assembly_info *info);
/* Continuous P2 on triangles -- an EXTENSION (BASELINE configs[1]); the reference is P1-only (StaticMatrix<T,3,3>,  */
/* methods/cfem.hpp:77-89; get_assembler ignores `degree`, :200-204), so results are "parity unpinned".  Written in   */
/* the reference's conventions: nodes = mesh points followed by edge midpoints (node id = point id | npoints + face    */
/* id, faces in the reference's sorted order), local order (v0,v1,v2,e(p0p1),e(p1p2),e(p0p2)) (core/mesh.hpp:299-312), */
/* Dirichlet nodes = endpoints and midpoints of boundary faces with the compress map of methods/cfem.hpp:134-148 over  */
/* the larger node set, triplets pushed in (cell,i,j) order and folded in that order.  Quadrature: the reference's     */
/* Dunavant rules of degree 2 (stiffness, exact) and 4 (load vector).  expand / get_compress_map work on               */
/* yfd_cfem_node_count() nodes.                                                                                        */
yfd_status yfd_cfem_assemble_p2(yfd_ctx *ctx, yfd_mesh *mesh, int rhs_fn,
                                yfd_system **out, yfd_assembly_info *info);
/* nodes behind the compress map: npoints (P1), npoints + nfaces (P2) */
yfd_status yfd_cfem_node_count(const yfd_system *sys, int64_t *nnodes);
/* sum over cells of int (u_h - u)^2 with the Dunavant rule of degree 6 (fixed summation order); order = 1 | 2,        */
/* expanded [node count], ref_fn = YFD_FN_SINSIN_SOL for the driver's problem.  The reference's continuous_fem.cpp     */
/* reports no error norm; this is the validation hook of the P2 extension.                                             */
yfd_status yfd_cfem_l2_error(yfd_ctx *ctx, yfd_mesh *mesh, int order, int ref_fn, const double *expanded, double *l2_errsq);
/* cfem::assembler::expand (methods/cfem.hpp:190-197): sol[system_size] -> expanded[node count] */
yfd_status yfd_cfem_expand(yfd_system *sys, const double *sol, double *expanded);
/* compress_map[node count]: system row of each node, -1 on Dirichlet nodes */
yfd_status yfd_cfem_get_compress_map(yfd_system *sys, int64_t *compress_map);

/* ---------------------------------------------------------------------------------- */
/* Assembled system accessors (REFERENCE numbering).                                   */
/* ---------------------------------------------------------------------------------- */
void       yfd_system_destroy(yfd_system *sys);
void       yfd_system_sizes(const yfd_system *sys, int64_t *n, int64_t *nnz);
/* distributed systems: rows / non-zeros / halo dofs held by this rank, and the first owned row in internal order */
void       yfd_system_local_sizes(const yfd_system *sys, int64_t *n_local, int64_t *nnz_local, int64_t *n_halo, int64_t *row_begin);
/* assm.lhs as CSR; any output may be NULL */
yfd_status yfd_system_get_csr(yfd_system *sys, int64_t *rowptr, int64_t *colidx, double *values);
yfd_status yfd_system_get_rhs(yfd_system *sys, double *rhs);                 /* assm.rhs         */
yfd_status yfd_system_get_pc_diag(yfd_system *sys, double *diag);            /* diag of assm.pc  */

/* ---------------------------------------------------------------------------------- */
/* Conjugate gradient: replaces conjugated_gradient (core/solvers.hpp:46-138 plain,     */
/* :141-243 with the inverse-preconditioner matrix iM; the DG assembler's iM is         */
/* diagonal, methods/dg.hpp:133-142, and is applied here as a vector scale).            */
/* ---------------------------------------------------------------------------------- */
typedef struct yfd_cg_params {            /* conjugated_gradient_params<T> (core/solvers.hpp:24-43) */
    double  rr_tol;                       /* 1e-8  */
    double  rr_max;                       /* 20    */
    int64_t max_iter;                     /* 100   */
    int     use_initial_guess;            /* false */
    int     use_normal_eqns;              /* false; true: CG on A^T A x = A^T b (core/solvers.hpp:88-91,112-113), x0 = 0, one GPU */
    int     save_iteration_history;       /* false */
    int     check_every;                  /* extension: iterations per device-side batch between
                                             host polls of the stop flag (0 = default 64).  The
                                             stop test itself runs on the device EVERY iteration,
                                             so the iteration count is the reference's. */
} yfd_cg_params;

void yfd_cg_params_default(yfd_cg_params *p);

typedef struct yfd_cg_info {
    int64_t iterations;                   /* `iter` on exit                                 */
    double  nr0;                          /* ||b - A x0||                                   */
    double  nr;                           /* recurrence residual norm on exit               */
    int     converged;                    /* nr/nr0 <= rr_tol (the reference returns true regardless) */
    double  ms_solve;                     /* device time of the whole solve (CUDA events)   */
    double  ms_spmv;                      /* average device time of one y = A d launch (measured on
                                             a separate timing pass of `spmv_timing_reps` launches;
                                             0 if not requested)                             */
    int64_t history_len;                  /* entries written to `history` (iterations+1)    */
    int     transport;                    /* 0 = one GPU, 1 = NCCL calls, 2 = peer-memory stores over NVLink (yfd_peer.cu) */
} yfd_cg_info;

/* values of `use_pc` below */
#define YFD_PC_NONE         0   /* conjugated_gradient(cgp, A, b, x)                 (core/solvers.hpp:46-138)            */
#define YFD_PC_JACOBI       1   /* conjugated_gradient(cgp, A, b, x, assm.pc)        (:141-243; needs build_preconditioner) */
#define YFD_PC_BLOCK_JACOBI 2   /* EXTENSION, DG systems: iM = blockdiag(A_cc)^-1, one bs x bs block */
                                /* per cell, inverted once on the device (in-warp pivoted LU) and applied by the block     */
                                /* SpMV; about half the iterations of the point-Jacobi CG.  Same recurrences (:196-227),   */
                                /* same stopping rule, same solution to the solver tolerance; works on several GPUs.       */

#define YFD_PC_PMG          3   /* EXTENSION, DG-SIP systems on meshes of congruent cells, one GPU: iM = one  */
                                /* symmetric p-multigrid V-cycle over the degrees k, k-1, ..., 1 (yfd_pmg.cu).  The reference's   */
                                /* basis is hierarchical by total degree (core/bases.hpp:96-122), so the coarse spaces are index */
                                /* subsets and the coarse operators sub-blocks of A; smoother: Chebyshev-block-Jacobi; the P1     */
                                /* level is "solved" by a fixed Chebyshev polynomial.  Same recurrences (core/solvers.hpp:196-227) */
                                /* and stopping rule as the reference's preconditioned CG.                                         */
/* p-multigrid parameters (0 = default): nu sweeps before and after each coarse correction (2), Chebyshev degree of the P1
 * solve (3 ncells^(1/4) clamped to [8, 400]), smoothing interval [lambda_max / smooth_ratio, lambda_max] (4) */
yfd_status yfd_system_set_pmg_params(yfd_system *sys, int nu, int coarse_degree, double smooth_ratio);
typedef struct yfd_pmg_info {
    int    nlevels, nu, coarse_degree;
    double smooth_ratio;
    int    degree[4];                 /* finest first                                               */
    double lambda_max[4];             /* the bound used for D^-1 A_l (1.1 x power-iteration estimate) */
} yfd_pmg_info;
yfd_status yfd_system_pmg_info(yfd_ctx *ctx, yfd_system *sys, yfd_pmg_info *info);   /* builds the hierarchy if needed */

/* Solve sys.lhs x = sys.rhs on the device, no host round trip of the matrix.
 * use_pc = YFD_PC_JACOBI -> the 5-argument overload with assm.pc.  x [n] in REFERENCE numbering, in/out
 * (input read only if use_initial_guess).  history may be NULL, else cap >= max_iter+1 is not
 * required: writing stops at history_cap. */
yfd_status yfd_system_solve_cg(yfd_ctx *ctx, yfd_system *sys, int use_pc, const yfd_cg_params *p,
                               double *x, yfd_cg_info *info, double *history, int64_t history_cap);

/* Same solve for resident pipelines: x0 = 0, the solution is written to the DEVICE buffer d_x
 * ([n] doubles, REFERENCE numbering; NULL = discard) and nothing but the scalars in `info`
 * crosses PCIe. */
yfd_status yfd_system_solve_cg_device(yfd_ctx *ctx, yfd_system *sys, int use_pc, const yfd_cg_params *p,
                                      double *d_x, yfd_cg_info *info);

/* Several ranks: yfd_system_solve_cg hands EVERY rank the whole solution (gather + n doubles over PCIe per rank).  A distributed
 * host program usually wants its own rows only: x_owned [n_local] (yfd_system_local_sizes) receives the rows of the cells this
 * rank owns, cell after cell in the library's internal order; yfd_system_owned_cells gives the REFERENCE id of each of them
 * ([n_local / bs] ids; DG systems), so that x_owned[c * bs + i] = x_ref[ref_cell[c] * bs + i].  On one GPU: all rows, in the
 * internal order.  New, no reference counterpart. */
yfd_status yfd_system_solve_cg_owned(yfd_ctx *ctx, yfd_system *sys, int use_pc, const yfd_cg_params *p,
                                     double *x_owned, yfd_cg_info *info);
yfd_status yfd_system_owned_cells(yfd_system *sys, int64_t *ref_cell);

/* The drop-in shape: A, b, x and optional DIAGONAL iM given as host data. */
yfd_status yfd_cg_solve(yfd_ctx *ctx, yfd_csr *A, const double *b, double *x, const double *pc_diag,
                        const yfd_cg_params *p, yfd_cg_info *info, double *history, int64_t history_cap);

/* BiCGStab: replaces yaourt::bicgstab, both overloads (core/solvers.hpp:246-354 without, :356-467 with iM), for   */
/* non-symmetric systems (yfd_dg_assemble_advection).  Same parameter struct and stopping rule as the CG            */
/* (conjugated_gradient_params, :311), same history convention, and the reference's restart r = b - A x; r0 = r     */
/* whenever |r.r0| < 1e-9 (:327-335, an ABSOLUTE threshold: the common path once the residual is small), decided     */
/* and executed on the device.  iM is a diagonal.  YFD_ERR_UNSUPPORTED: use_normal_eqns, distributed systems.        */
yfd_status yfd_system_solve_bicgstab(yfd_ctx *ctx, yfd_system *sys, int use_pc, const yfd_cg_params *p,
                                     double *x, yfd_cg_info *info, double *history, int64_t history_cap);
yfd_status yfd_bicgstab_solve(yfd_ctx *ctx, yfd_csr *A, const double *b, double *x, const double *pc_diag,
                              const yfd_cg_params *p, yfd_cg_info *info, double *history, int64_t history_cap);

/* Operator storage of a DG system.  The reference assembles each cell's blocks from the cell's physical frame
 * (core/bases.hpp:96-155), so on refinement / structured meshes most cells hold the same (1+nf) blocks up to coordinate
 * rounding.  yfd_dg_assemble_sip detects such congruence classes (hash of the quantised shape signature, then an
 * entry-by-entry comparison of every cell's assembled block row with its class representative against a tolerance of
 * 16 eps |x|/h, floor 2e-13 of the block row's max-norm; YFD_CLASS_TOL overrides) and, when the classes fit in shared
 * memory, the solvers multiply by ONE block row per class (yfd_spmv_class.cu) instead of streaming 8 B per non-zero.
 * Otherwise (shattered meshes, advection-reaction) the streamed kernel is used.  AUTO = classes when available. */
#define YFD_OPERATOR_AUTO     0
#define YFD_OPERATOR_STREAMED 1    /* always the per-cell values (yfd_spmv_tma.cu) */
#define YFD_OPERATOR_CLASSES  2    /* YFD_ERR_INVALID_STATE if the system has no usable cache */
typedef struct yfd_operator_info {
    int     uses_classes;         /* what the next solve / apply will do                              */
    int     classes_available;
    int     reject_reason;        /* 0 none; 1 too many classes (> 4096, or tables beyond 8 MB); 2 too many small classes; 3 deviation > tolerance */
    int64_t nclasses;
    double  max_deviation;        /* max over cells of |W_cell - W_class|_max / |W_class|_max          */
    double  tolerance;
    int64_t table_bytes;          /* all class block rows                                              */
    int64_t list_entries;         /* class-sorted cell list incl. padding to 32 per class              */
    int     tensor_pipe;          /* 1: the class kernel is the mma.sync (FP64 tensor pipe) one, dg_spmv_class_mma */
} yfd_operator_info;
yfd_status yfd_system_set_operator_mode(yfd_system *sys, int mode);
yfd_status yfd_system_operator_info(const yfd_system *sys, yfd_operator_info *info);
/* yfd_system_get_csr exports the values the solvers APPLY (class representatives when the cache is in use). */

/* y = A x for the assembled system as conjugated_gradient forms it (`y = A*d`, core/solvers.hpp:115), host vectors
 * in REFERENCE numbering, one GPU.  kernel = YFD_APPLY_SOLVER: the kernel the solvers launch (the TMA-streamed
 * SpMV); YFD_APPLY_PLAIN: an independent one-thread-per-row kernel that reads the matrix straight from global
 * memory and sums each row left to right (SMatDVecMultExpr.h:204-207) -- the device-side cross-check of the former. */
#define YFD_APPLY_SOLVER 0
#define YFD_APPLY_PLAIN  1
yfd_status yfd_system_apply(yfd_ctx *ctx, yfd_system *sys, int kernel, const double *x, double *y);

/* ||b - A x||_2 / ||b||_2 of a candidate solution (the TRUE residual; the solvers stop on the recurrence residual,
 * core/solvers.hpp:99,124), formed with the YFD_APPLY_PLAIN kernel.  x [n] in REFERENCE numbering: x_host, or the
 * device vector d_x when x_host is NULL.  Works on several ranks (every rank passes the full vector, as
 * yfd_system_solve_cg returns it; the sums are combined over the ranks). */
yfd_status yfd_system_true_residual(yfd_ctx *ctx, yfd_system *sys, const double *x_host, const double *d_x, double *rel_residual);

/* The same true residual for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * The call allocates nothing -- no cudaMalloc / cudaFree, no pool, no allocating library routine: its scratch (y = A x, the
 * partial sums) lives in one static device array of YFD_NOALLOC_MAX_ROWS doubles + 16 KiB (4 MiB in all) that every GPU
 * loading this library holds until the process exits.  d_x [n]: DEVICE vector in REFERENCE numbering; *rel_residual: host.
 * One GPU only (YFD_ERR_UNSUPPORTED on a partitioned system); YFD_ERR_UNSUPPORTED with a message naming the limit when the
 * system has more than YFD_NOALLOC_MAX_ROWS rows.  Same arithmetic in the same order as yfd_system_true_residual: the two
 * agree bit for bit.  The array is shared by all calls on a device and keeps the last call's contents (every word read
 * is written first by the same call); calls that OVERLAP on one GPU race on it -- serialise them. */
#define YFD_NOALLOC_MAX_ROWS 524288
yfd_status yfd_system_true_residual_noalloc(yfd_ctx *ctx, yfd_system *sys, const double *d_x, double *rel_residual);

/* Time `reps` launches of y = A d for this system with CUDA events on the context's stream
 * (bench.py's roofline leg).  Returns average ms per launch and the algorithmic CSR bytes of one
 * launch (12 nnz + 8 (n+1) + 16 n) plus the bytes the kernel actually
 * requests (the block kernel does not read colidx). */
yfd_status yfd_system_time_spmv(yfd_ctx *ctx, yfd_system *sys, int reps, double *ms_per_launch,
                                double *csr_bytes, double *kernel_bytes);

/* ---------------------------------------------------------------------------------- */
/* DG Maxwell TM, explicit time domain: replaces maxwell_context's storage + assemble   */
/* (methods/dg_maxwell_2D.hpp:75-152, :220-461), apply_initial_condition (:513-554),    */
/* do_timestep / apply_operator (:558-675) and the driver's report_errors               */
/* (maxwell_driver.cpp:133-191).                                                        */
/*                                                                                      */
/* The reference stores, per cell, a dense 3bs x 3bs on-diagonal block and one dense    */
/* 3bs x 3bs block per face (gOp_ondiag / gOp_offdiag, :90) although only 4 (centred)   */
/* of the 9 bs x bs sub-blocks are non-zero and all of them are scalar multiples of     */
/* M^-1 Sx, M^-1 Sy, M^-1 tmass_f, M^-1 nmass_f.  Here exactly those 2 + 2 nf blocks    */
/* per cell are kept (4.5x fewer bytes per operator application);  */
/* yfd_maxwell_get_operator re-emits the reference's dense layout for parity checks.    */
/* Unknowns at this boundary are numbered as the reference numbers gDofs:               */
/* 3 bs cell + bs field + i, field 0 = Hx, 1 = Hy, 2 = Ez (:541-547).                   */
/* ---------------------------------------------------------------------------------- */
#define YFD_TI_RK4    0     /* time_integrator_type::RUNGE_KUTTA_4 (default, :66)       */
#define YFD_TI_EULER  1     /* time_integrator_type::EXPLICIT_EULER                     */

/* analytic fields: the driver's TM_mn cavity mode (maxwell_driver.cpp:202-216) */
#define YFD_MAXWELL_FIELDS_CAVITY_11  0    /* m = n = 1                                 */
#define YFD_MAXWELL_FIELDS_ZERO       1

typedef struct yfd_maxwell_config {   /* maxwell_config<T> (methods/dg_maxwell_2D.hpp:46-72), solver-relevant fields */
    int     degree;                   /* cfg.degree                                     */
    double  delta_t;                  /* cfg.delta_t                                    */
    double  eta;                      /* cfg.eta (upwind penalisation)                  */
    int     upwind;                   /* cfg.upwind                                     */
    int     time_integrator;          /* YFD_TI_*                                       */
} yfd_maxwell_config;

void yfd_maxwell_config_default(yfd_maxwell_config *cfg);   /* degree 1, delta_t 0.1, eta 1, centred, RK4 (:63-69) */

typedef struct yfd_maxwell yfd_maxwell;

typedef struct yfd_maxwell_info {
    int64_t ncells, basis_size, faces_per_elem;
    int64_t ndofs;                    /* 3 bs ncells                                    */
    int64_t operator_bytes;           /* bytes of the factored operator held in HBM     */
    int64_t reference_operator_bytes; /* (1 + nf) (3bs)^2 8 ncells: gOp_ondiag + gOp_offdiag */
    double  ms_assemble;              /* device time of the element/face kernel incl. the M^-1 solves */
    /* congruence-class cache of the factored operator (see YFD_OPERATOR_*): the 2 + 2 nf blocks of a cell depend on its
     * shape, size and neighbours' frames only, so on refinement meshes one block set per class serves every cell and the
     * operator application streams vectors only (BASELINE configs[3] "matrix-free element+face operator") */
    int     uses_classes;
    int     class_reject_reason;      /* as yfd_operator_info.reject_reason                 */
    int64_t nclasses;
    double  class_max_deviation, class_tolerance;
    int64_t class_table_bytes;
    int     class_tensor_pipe;        /* 1: the class kernel is maxwell_stage_class_mma (mma.sync m8n8k4 f64)   */
} yfd_maxwell_info;
yfd_status yfd_maxwell_set_operator_mode(yfd_maxwell *m, int mode);   /* YFD_OPERATOR_AUTO / _STREAMED / _CLASSES */

/* maxwell_context ctor (storage) + assemble(ctx).  mu_r / eps_r: [ncells] in REFERENCE cell order, or NULL
 * for 1.0 everywhere (the context's default, :149-150).  The mesh may be destroyed afterwards. */
yfd_status yfd_maxwell_create(yfd_ctx *ctx, yfd_mesh *mesh, const yfd_maxwell_config *cfg,
                              const double *mu_r, const double *eps_r, yfd_maxwell **out, yfd_maxwell_info *info);
void       yfd_maxwell_destroy(yfd_maxwell *m);
void       yfd_maxwell_get_info(const yfd_maxwell *m, yfd_maxwell_info *info);
/* change delta_t / the integrator without re-assembling (the operator does not depend on them) */
yfd_status yfd_maxwell_set_time_integration(yfd_maxwell *m, double delta_t, int time_integrator);

/* The operator in the REFERENCE's layout and numbering, for parity tests on small meshes:
 *   ondiag  [ncells][3bs][3bs]       gOp_ondiag   (:90)
 *   offdiag [ncells][nf][3bs][3bs]   gOp_offdiag, zero where the face is on the boundary
 *   neigh   [ncells][nf], has [ncells][nf]   offdiag_neigh_offsets .first / .second (:106-107)
 *   gM      [ncells][bs][bs]         elemental mass matrices (:87)
 * Any output may be NULL. */
yfd_status yfd_maxwell_get_operator(yfd_maxwell *m, double *ondiag, double *offdiag, int64_t *neigh,
                                    int32_t *has, double *gM);

/* apply_initial_condition: per-cell L2 projection of the analytic fields at time t into gDofs */
yfd_status yfd_maxwell_apply_initial_condition(yfd_maxwell *m, int fields, double t);
/* gDofs, host, reference numbering */
yfd_status yfd_maxwell_get_dofs(yfd_maxwell *m, double *dofs);
yfd_status yfd_maxwell_set_dofs(yfd_maxwell *m, const double *dofs);
/* the unknowns of the cells this rank owns, [ncells_owned][3][bs] cell after cell in the internal order, and the REFERENCE id of
 * each of those cells (no gather, ncells_owned * 3 bs doubles over PCIe; one GPU: all cells).  *ncells_owned may be NULL. */
yfd_status yfd_maxwell_get_dofs_owned(yfd_maxwell *m, double *dofs_owned);
yfd_status yfd_maxwell_owned_cells(yfd_maxwell *m, int64_t *ncells_owned, int64_t *ref_cell);
/* v_next = apply_operator(v) (:586-625) on host vectors in reference numbering */
yfd_status yfd_maxwell_apply_operator(yfd_maxwell *m, const double *v, double *v_next);

typedef struct yfd_maxwell_step_info {
    int64_t steps;
    double  ms_total;                 /* device time of all steps (CUDA events)                     */
    double  ms_per_step;
    int64_t operator_applications;    /* 4 per RK4 step, 1 per Euler step                           */
} yfd_maxwell_step_info;

/* nsteps x do_timestep (:558-675); gDofs stays on the device.  info may be NULL. */
yfd_status yfd_maxwell_timesteps(yfd_maxwell *m, int64_t nsteps, yfd_maxwell_step_info *info);
/* report_errors: L2 norms of (numerical - analytic(t)) for Hx, Hy, Ez.  The reference driver calls it with
 * t = cycle * delta_t AFTER do_timestep of that cycle (maxwell_driver.cpp:249,266); callers wanting its
 * numbers pass the same t. */
yfd_status yfd_maxwell_report_errors(yfd_maxwell *m, int fields, double t, double errors[3]);

/* Time `reps` operator applications with CUDA events (bench.py's roofline leg): average ms per application,
 * the bytes the reference's dense layout would stream per application ((1+nf)(3bs)^2 8 per cell + 2 vectors)
 * and the bytes this kernel requests. */
yfd_status yfd_maxwell_time_apply(yfd_maxwell *m, int reps, double *ms_per_apply, double *reference_bytes,
                                  double *kernel_bytes);

#ifdef __cplusplus
}
#endif
#endif /* YFD_H */

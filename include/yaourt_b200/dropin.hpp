/*
 * include/yaourt_b200/dropin.hpp -- the reference's own names on top of the C ABI (include/yfd.h).
 *
 * Include this header AFTER the reference's headers (core/mesh.hpp, core/solvers.hpp, ...) in a translation unit that
 * is otherwise unchanged, and yaourt_b200/dropin_triplets.hpp between core/blaze_sparse_init.hpp and methods/dg.hpp
 * (that call is qualified, so it binds where the header is parsed).  Together they declare NON-template overloads
 *
 *   blaze::init_from_triplets(CompressedMatrix<double>&, vector<triplet<double>>::iterator, ...::iterator)
 *                                                        replaces core/blaze_sparse_init.hpp:125-155 (+ reduce_triplets :101-122)
 *   conjugated_gradient(cgp, A, b, x)                    replaces core/solvers.hpp:46-138
 *   conjugated_gradient(cgp, A, b, x, iM)                replaces core/solvers.hpp:141-243 (iM diagonal, methods/dg.hpp:133-142)
 *   bicgstab(cgp, A, b, x[, iM])                         replaces core/solvers.hpp:246-354 / :356-467
 *
 * with exactly the reference's signatures.  A non-template function beats a template of equal conversion rank,
 * and the reference's call sites (methods/dg.hpp:130,141; methods/cfem.hpp:184; dg2d_diffusion.cpp:217-219;
 * continuous_fem.cpp:80) sit inside templates whose calls are resolved at instantiation, so the reference's
 * drivers pick these up without a source change.  Plus the batched entry the reference lacks:
 *
 *   yaourt::b200::assemble_dg_sip(msh, degree, eta, rhs_fn_id, dirichlet_fn_id[, use_pc, solve_with])
 *                                                        the whole cell loop of run_diffusion_solver
 *                                                        (dg2d_diffusion.cpp:133-201) + assembler::finalize on the GPU
 *   yaourt::b200::assemble_dg_advection(msh, degree, eta, use_upwinding[, solve_with])   dg2d_advection.cpp:137-215
 *   yaourt::b200::assemble_cfem_p1(msh[, rhs_fn_id, solve_with])                          continuous_fem.cpp:64-80
 *
 * Error behaviour follows the reference: std::invalid_argument / std::logic_error with its messages for the
 * mesh / assembler errors, `false` from conjugated_gradient on a shape mismatch, `true` otherwise -- also when
 * not converged (core/solvers.hpp:45,137).  There is no CPU fallback: without a CUDA device the first call throws.
 */
#pragma once
#include "yaourt_b200/core.hpp"
#include "yaourt_b200/dropin_triplets.hpp"

/* ------------------------------------------------------------------------------------------------ */
/* conjugated_gradient (core/solvers.hpp:46-138 and :141-243)                                         */
/* ------------------------------------------------------------------------------------------------ */
namespace yaourt {
namespace b200 {

struct cg_report { size_t iterations = 0; double nr0 = 0, nr = 0; bool converged = false; double ms_solve = 0; };
inline cg_report &last_cg_report()
{
    static cg_report r;
    return r;
}

inline bool cg_impl(const conjugated_gradient_params<double> &cgp, const blaze::CompressedMatrix<double> &A,
                    const blaze::DynamicVector<double> &b, blaze::DynamicVector<double> &x,
                    const blaze::CompressedMatrix<double> *iM, bool use_bicgstab = false)
{
    /* the reference's shape checks and messages (core/solvers.hpp:53-77) */
    if (A.rows() != A.columns()) { if (cgp.verbose) std::cout << "[CG solver] A square matrix is required" << std::endl; return false; }
    const size_t N = A.columns();
    if (b.size() != N) { if (cgp.verbose) std::cout << "[CG solver] Wrong size of RHS vector" << std::endl; return false; }
    if (x.size() != N) { if (cgp.verbose) std::cout << "[CG solver] Wrong size of solution vector" << std::endl; return false; }
    auto &ctx = context::global();
    std::vector<int64_t> rowptr, colidx;
    std::vector<double> values, pc;
    flatten_compressed(A, rowptr, colidx, values);
    if (iM) {
        /* the DG assembler's iM is diag(1/d) stored as a sparse matrix (methods/dg.hpp:133-142); applied as a vector scale */
        if (iM->rows() != N || iM->columns() != N) return false;
        pc.assign(N, 0.0);
        for (size_t r = 0; r < N; r++)
            for (auto it = iM->cbegin(r); it != iM->cend(r); ++it) {
                if (it->index() != r) throw std::invalid_argument("yaourt_b200: conjugated_gradient takes a DIAGONAL inverse preconditioner");
                pc[r] = it->value();
            }
    }
    yfd_csr *dA = nullptr;
    check(ctx.get(), yfd_csr_from_arrays(ctx.get(), (int64_t) N, (int64_t) N, (int64_t) values.size(), rowptr.data(), colidx.data(),
                                         values.data(), &dA));
    yfd_cg_params p;
    yfd_cg_params_default(&p);
    p.rr_tol = cgp.rr_tol; p.rr_max = cgp.rr_max; p.max_iter = (int64_t) cgp.max_iter;
    p.use_initial_guess = cgp.use_initial_guess ? 1 : 0;
    p.use_normal_eqns = cgp.use_normal_eqns ? 1 : 0;
    p.save_iteration_history = cgp.save_iteration_history ? 1 : 0;
    std::vector<double> hist(cgp.save_iteration_history ? cgp.max_iter + 1 : 0);
    yfd_cg_info info;
    std::vector<double> xs(x.data(), x.data() + N);
    const yfd_status st = (use_bicgstab ? yfd_bicgstab_solve : yfd_cg_solve)(ctx.get(), dA, b.data(), xs.data(), iM ? pc.data() : nullptr, &p, &info,
                                                                           hist.empty() ? nullptr : hist.data(), (int64_t) hist.size());
    yfd_csr_destroy(dA);
    check(ctx.get(), st);
    for (size_t i = 0; i < N; i++) x[i] = xs[i];
    if (cgp.save_iteration_history) {                       /* one rr per line, as core/solvers.hpp:109-110,128-132 */
        std::ofstream ofs(cgp.history_filename);
        for (int64_t i = 0; i < info.history_len; i++) ofs << hist[(size_t) i] << std::endl;
    }
    if (cgp.verbose) std::cout << " -> Iteration " << info.iterations << ", rr = " << info.nr / info.nr0 << std::endl;
    auto &rep = last_cg_report();
    rep.iterations = (size_t) info.iterations; rep.nr0 = info.nr0; rep.nr = info.nr; rep.converged = info.converged != 0; rep.ms_solve = info.ms_solve;
    return true;                                            /* also when not converged (core/solvers.hpp:45,137) */
}

/* The batched entry: cell loop + finalize of run_diffusion_solver on the GPU, results in the assembler's shapes. */
template <typename Mesh>
struct dg_sip_system {
    blaze::CompressedMatrix<double> lhs;
    blaze::DynamicVector<double>    rhs;
    blaze::DynamicVector<double>    pc_diag;      /* empty unless use_pc */
    blaze::DynamicVector<double>    sol;          /* filled by solve()   */
};

/* physics of the batched DG entries */
enum dg_physics { DG_SIP_DIFFUSION = 0, DG_ADVECTION_REACTION = 1 };

/* assemble_dg_sip(msh, degree, eta, rhs_fn_id, dirichlet_fn_id) -- the whole cell loop of
 * run_diffusion_solver (dg2d_diffusion.cpp:133-201) + assembler::finalize on the GPU, the data functions chosen by id
 * (YFD_FN_*: per-cell host callbacks cannot feed a GPU).  use_pc builds assm.pc (methods/dg.hpp:133-142); solve_with, if
 * given, also runs conjugated_gradient with the matrix resident on the device. */
template <typename Mesh>
dg_sip_system<Mesh> assemble_dg_sip(const Mesh &msh, size_t degree, double eta, int rhs_fn_id, int dirichlet_fn_id, bool use_pc = false,
                                    const conjugated_gradient_params<double> *solve_with = nullptr, int physics = DG_SIP_DIFFUSION,
                                    bool use_upwinding = false)
{
    auto &ctx = context::global();
    yfd_mesh *dm = upload_mesh(ctx, msh);
    yfd_dg_config cfg;
    cfg.degree = (int) degree; cfg.eta = eta; cfg.rhs_fn = rhs_fn_id; cfg.dirichlet_fn = dirichlet_fn_id; cfg.build_preconditioner = use_pc ? 1 : 0;
    yfd_system *sys = nullptr;
    yfd_assembly_info ai;
    yfd_status st = physics == DG_ADVECTION_REACTION ? yfd_dg_assemble_advection(ctx.get(), dm, &cfg, use_upwinding ? 1 : 0, &sys, &ai)
                                                     : yfd_dg_assemble_sip(ctx.get(), dm, &cfg, &sys, &ai);
    if (st != YFD_OK) { yfd_mesh_destroy(dm); check(ctx.get(), st); }
    dg_sip_system<Mesh> out;
    const size_t n = (size_t) ai.system_size, nnz = (size_t) ai.nnz;
    std::vector<int64_t> rowptr(n + 1), colidx(nnz);
    std::vector<double> values(nnz), v(n);
    st = yfd_system_get_csr(sys, rowptr.data(), colidx.data(), values.data());
    if (st == YFD_OK) { fill_compressed(out.lhs, n, n, rowptr, colidx, values); st = yfd_system_get_rhs(sys, v.data()); }
    if (st == YFD_OK) { out.rhs.resize(n); for (size_t i = 0; i < n; i++) out.rhs[i] = v[i]; }
    if (st == YFD_OK && use_pc) { st = yfd_system_get_pc_diag(sys, v.data()); out.pc_diag.resize(n); for (size_t i = 0; i < n; i++) out.pc_diag[i] = v[i]; }
    if (st == YFD_OK && solve_with) {
        yfd_cg_params p;
        yfd_cg_params_default(&p);
        p.rr_tol = solve_with->rr_tol; p.rr_max = solve_with->rr_max; p.max_iter = (int64_t) solve_with->max_iter;
        p.use_normal_eqns = solve_with->use_normal_eqns ? 1 : 0;          /* the advection driver's solve (dg2d_advection.cpp:225) */
        yfd_cg_info info;
        st = yfd_system_solve_cg(ctx.get(), sys, use_pc ? 1 : 0, &p, v.data(), &info, nullptr, 0);   /* matrix never leaves the GPU */
        out.sol.resize(n);
        for (size_t i = 0; i < n; i++) out.sol[i] = v[i];
        auto &rep = last_cg_report();
        rep.iterations = (size_t) info.iterations; rep.nr0 = info.nr0; rep.nr = info.nr; rep.converged = info.converged != 0; rep.ms_solve = info.ms_solve;
    }
    yfd_system_destroy(sys);
    yfd_mesh_destroy(dm);
    check(ctx.get(), st);
    return out;
}

/* the reference driver's own data functions (dg2d_diffusion.cpp:53-66) */
template <typename Mesh>
dg_sip_system<Mesh> assemble_dg_sip(const Mesh &msh, size_t degree, double eta, bool use_pc,
                                    const conjugated_gradient_params<double> *solve_with = nullptr)
{
    return assemble_dg_sip(msh, degree, eta, YFD_FN_SINSIN, YFD_FN_ZERO, use_pc, solve_with);
}

/* the cell loop of run_advection_reaction_solver (dg2d_advection.cpp:137-213) + finalize; solve_with->use_normal_eqns = true
 * reproduces the driver's solve (:225) */
template <typename Mesh>
dg_sip_system<Mesh> assemble_dg_advection(const Mesh &msh, size_t degree, double eta, bool use_upwinding,
                                          const conjugated_gradient_params<double> *solve_with = nullptr)
{
    return assemble_dg_sip(msh, degree, eta, YFD_FN_ADVECTION_RHS, YFD_FN_ZERO, false, solve_with, DG_ADVECTION_REACTION, use_upwinding);
}

/* continuous P1: the cell loop of continuous_fem.cpp:64-71 + cfem::assembler (methods/cfem.hpp:119-186) + finalize;
 * `expanded` = assm.expand(sol) (:189-197) when solve_with is given */
template <typename Mesh>
struct cfem_system {
    blaze::CompressedMatrix<double> lhs;
    blaze::DynamicVector<double>    rhs, sol, expanded;
    std::vector<int64_t>            compress_map;     /* [points], -1 on Dirichlet nodes (methods/cfem.hpp:140-148) */
};

template <typename Mesh>
cfem_system<Mesh> assemble_cfem_p1(const Mesh &msh, int rhs_fn_id = YFD_FN_SINSIN,
                                   const conjugated_gradient_params<double> *solve_with = nullptr)
{
    auto &ctx = context::global();
    yfd_mesh *dm = upload_mesh(ctx, msh);
    yfd_system *sys = nullptr;
    yfd_assembly_info ai;
    yfd_status st = yfd_cfem_assemble_p1(ctx.get(), dm, rhs_fn_id, &sys, &ai);
    if (st != YFD_OK) { yfd_mesh_destroy(dm); check(ctx.get(), st); }
    cfem_system<Mesh> out;
    const size_t n = (size_t) ai.system_size, nnz = (size_t) ai.nnz, np = msh.points.size();
    std::vector<int64_t> rowptr(n + 1), colidx(nnz);
    std::vector<double> values(nnz), v(n), ex(np);
    out.compress_map.resize(np);
    st = yfd_system_get_csr(sys, rowptr.data(), colidx.data(), values.data());
    if (st == YFD_OK) { fill_compressed(out.lhs, n, n, rowptr, colidx, values); st = yfd_system_get_rhs(sys, v.data()); }
    if (st == YFD_OK) { out.rhs.resize(n); for (size_t i = 0; i < n; i++) out.rhs[i] = v[i]; st = yfd_cfem_get_compress_map(sys, out.compress_map.data()); }
    if (st == YFD_OK && solve_with) {
        yfd_cg_params p;
        yfd_cg_params_default(&p);
        p.rr_tol = solve_with->rr_tol; p.rr_max = solve_with->rr_max; p.max_iter = (int64_t) solve_with->max_iter;
        yfd_cg_info info;
        st = yfd_system_solve_cg(ctx.get(), sys, 0, &p, v.data(), &info, nullptr, 0);
        if (st == YFD_OK) st = yfd_cfem_expand(sys, v.data(), ex.data());
        out.sol.resize(n); out.expanded.resize(np);
        for (size_t i = 0; i < n; i++) out.sol[i] = v[i];
        for (size_t i = 0; i < np; i++) out.expanded[i] = ex[i];
        auto &rep = last_cg_report();
        rep.iterations = (size_t) info.iterations; rep.nr0 = info.nr0; rep.nr = info.nr; rep.converged = info.converged != 0; rep.ms_solve = info.ms_solve;
    }
    yfd_system_destroy(sys);
    yfd_mesh_destroy(dm);
    check(ctx.get(), st);
    return out;
}

} // namespace b200
} // namespace yaourt

inline bool conjugated_gradient(const conjugated_gradient_params<double> &cgp, const blaze::CompressedMatrix<double> &A,
                                const blaze::DynamicVector<double> &b, blaze::DynamicVector<double> &x)
{
    return yaourt::b200::cg_impl(cgp, A, b, x, nullptr);
}

inline bool conjugated_gradient(const conjugated_gradient_params<double> &cgp, const blaze::CompressedMatrix<double> &A,
                                const blaze::DynamicVector<double> &b, blaze::DynamicVector<double> &x,
                                const blaze::CompressedMatrix<double> &iM)
{
    return yaourt::b200::cg_impl(cgp, A, b, x, &iM);
}

/* bicgstab (core/solvers.hpp:246-354 and :356-467): same parameter struct, shape checks and return convention */
inline bool bicgstab(const conjugated_gradient_params<double> &cgp, const blaze::CompressedMatrix<double> &A,
                     const blaze::DynamicVector<double> &b, blaze::DynamicVector<double> &x)
{
    return yaourt::b200::cg_impl(cgp, A, b, x, nullptr, true);
}

inline bool bicgstab(const conjugated_gradient_params<double> &cgp, const blaze::CompressedMatrix<double> &A,
                     const blaze::DynamicVector<double> &b, blaze::DynamicVector<double> &x,
                     const blaze::CompressedMatrix<double> &iM)
{
    return yaourt::b200::cg_impl(cgp, A, b, x, &iM, true);
}

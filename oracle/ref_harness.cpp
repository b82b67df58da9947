/*
 * oracle/ref_harness.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Builds the UNMODIFIED reference (headers and drivers where they lie under
 * /root/reference, pulled in with #include at compile time; nothing is copied
 * into this repository) into oracle/_ref/libyfd_ref.so and exposes a small
 * C ABI so the tests / bench cpu_baseline can run the reference's own
 *   mesh -> per-cell loop -> triplets -> CSR -> CG
 * path and look at every intermediate result.
 *
 * How the reference's own code path is observed without editing it:
 *   - dg2d_diffusion.cpp and continuous_fem.cpp are #included verbatim with
 *     `main` renamed, so run_diffusion_solver<Mesh>() (dg2d_diffusion.cpp:114)
 *     and the CFEM driver body (continuous_fem.cpp:52) run exactly as shipped.
 *   - `init_from_triplets` (core/blaze_sparse_init.hpp:125) and
 *     `conjugated_gradient` (core/solvers.hpp:46,141) are renamed by macro at
 *     their CALL SITES only (methods/dg.hpp:130,141, methods/cfem.hpp:184,
 *     dg2d_diffusion.cpp:217-219, continuous_fem.cpp:80) to thin hooks below
 *     that record inputs/outputs + wall time and then forward to the real
 *     reference function.
 *   - LAPACK dgetrf_/dgetrs_ (needed only by the driver's post-process,
 *     dg2d_diffusion.cpp:265 -> core/blaze_sparse_init.hpp:161) come from the
 *     unblocked partial-pivot LU in oracle/lu_shim.c (system LAPACK is absent).
 *
 * Nothing under yaourt-fem-dg_b200/ may include or link this file.
 */
#include "ref_common.hpp"
#include "core/quadratures.hpp"
#include "core/bases.hpp"

#define init_from_triplets yfd_hook_ift
#include "methods/dg.hpp"
#undef init_from_triplets

#define YFD_CG_NAME yfd_dg_cg
YFD_DECLARE_CG_HOOKS(yfd_dg_cg)

/* -DYFD_REF_ADVECTION builds the same harness around dg2d_advection.cpp instead (libyfd_ref_adv.so): the two drivers
 * define the same global names (params::, data::rhs, dg_config, solver_status), so they cannot share a library. */
#define conjugated_gradient yfd_dg_cg
#ifdef YFD_REF_ADVECTION
#define main yfd_ref_dg2d_advection_main
#include "dg2d_advection.cpp"
#else
#define main yfd_ref_dg2d_diffusion_main
#include "dg2d_diffusion.cpp"
#endif
#undef main
#undef conjugated_gradient

#include "ref_cg_hooks.inl"

#ifdef YFD_REF_ADVECTION
extern "C" int yfd_ref_cfem_driver(int) { return 1; }            /* not in this library */
static int yfd_adv_upwinding = 0;
#else
extern "C" int yfd_ref_cfem_driver(int levels);   /* ref_harness_cfem.cpp */
#endif

/* ======================================================================= */
/*                                C ABI                                    */
/* ======================================================================= */
namespace yfd_ref {

using tri_mesh  = yaourt::simplicial_mesh<double>;
using quad_mesh = yaourt::quad_mesh<double>;

struct mesh_box {
    int       type = 0;     /* 0 = triangles, 1 = quadrangles */
    tri_mesh  tri;
    quad_mesh quad;
};

static mesh_box M;

struct quiet {
    std::streambuf *old;
    std::ostringstream sink;
    char cwd[4096];
    quiet() {
        old = std::cout.rdbuf(sink.rdbuf());
        if (!getcwd(cwd, sizeof(cwd))) cwd[0] = 0;
        if (chdir("/tmp") != 0) {}
    }
    ~quiet() {
        std::cout.rdbuf(old);
        if (cwd[0]) { if (chdir(cwd) != 0) {} }
    }
};

template<typename Mesh>
static void make_mesh(Mesh& msh, int levels, int shatter)
{
    msh = Mesh();
    auto mesher = yaourt::get_mesher(msh);
    mesher.create_mesh(msh, (size_t) levels);     /* core/meshers.hpp:130 / :302 */
    if (shatter)
        shatter_mesh(msh, 0.2);                   /* dg2d_diffusion.cpp:311-312 */
    msh.compute_connectivity();                   /* core/mesh.hpp:176 */
}

template<typename Mesh>
static void copy_mesh(const Mesh& msh, double *pts, int64_t *cells,
                      int64_t *faces, int32_t *fbnd, int64_t *owners)
{
    const size_t nv = msh.cells.empty() ? 0 : msh.cells[0].p.size();
    if (pts)
        for (size_t i = 0; i < msh.points.size(); i++) {
            pts[2*i] = msh.points[i].x(); pts[2*i+1] = msh.points[i].y();
        }
    if (cells)
        for (size_t i = 0; i < msh.cells.size(); i++)
            for (size_t v = 0; v < nv; v++)
                cells[nv*i + v] = (int64_t) msh.cells[i].p[v];
    for (size_t i = 0; i < msh.faces.size(); i++) {
        if (faces) { faces[2*i] = (int64_t) msh.faces[i].p0; faces[2*i+1] = (int64_t) msh.faces[i].p1; }
        if (fbnd)  fbnd[i] = msh.faces[i].is_boundary ? 1 : 0;
        if (owners) {
            owners[2*i]   = msh.face_owners[i][0] == NO_OWNER ? -1 : (int64_t) msh.face_owners[i][0];
            owners[2*i+1] = msh.face_owners[i][1] == NO_OWNER ? -1 : (int64_t) msh.face_owners[i][1];
        }
    }
}

template<typename Mesh>
static void run_dg_on(Mesh& msh, int degree, double eta, int use_pc)
{
    dg_config<double> cfg;                 /* dg2d_diffusion.cpp:81 */
    cfg.eta = eta;
    cfg.degree = degree;
    cfg.use_preconditioner = use_pc != 0;
    G.ift_calls = 0;
    G.t_enter = now_s();
#ifdef YFD_REF_ADVECTION
    cfg.use_upwinding = yfd_adv_upwinding != 0;
    auto st = run_advection_reaction_solver(msh, cfg);     /* dg2d_advection.cpp:120, unmodified */
#else
    auto st = run_diffusion_solver(msh, cfg);     /* dg2d_diffusion.cpp:114, unmodified */
#endif
    G.l2_qp = std::sqrt(st.L2_errsq_qp);
    G.l2_mm = std::sqrt(st.L2_errsq_mm);
    G.mesh_h = st.mesh_h;
}

} // namespace yfd_ref

extern "C" {

/* ---- mesh ---- */
int yfd_ref_mesh_create(int type, int levels, int shatter)
{
    using namespace yfd_ref;
    M.type = type;
    if (type == 0) make_mesh(M.tri, levels, shatter);
    else           make_mesh(M.quad, levels, shatter);
    return 0;
}

/* Inject a mesh made elsewhere (e.g. yfd_mesher_create_structured) into the reference's own mesh type through its public
 * vectors (core/mesh.hpp:239-243), with the reference's own element constructors; compute_connectivity is the reference's. */
extern "C++" {
template<typename Mesh, typename Cell>
static void set_mesh(Mesh& msh, int64_t np, const double *pts, int64_t nc, const int64_t *cells, int64_t nf,
                     const int64_t *faces, const uint8_t *fbnd, const int64_t *fbid)
{
    using point_type = typename Mesh::point_type;
    using face_type = typename Mesh::face_type;
    msh = Mesh();
    for (int64_t i = 0; i < np; i++) msh.points.push_back(point_type(pts[2*i], pts[2*i+1]));
    for (int64_t i = 0; i < nf; i++)
        msh.faces.push_back(fbnd[i] ? face_type((size_t) faces[2*i], (size_t) faces[2*i+1], (size_t) (fbid ? fbid[i] : 0), true)
                                    : face_type((size_t) faces[2*i], (size_t) faces[2*i+1]));
    for (int64_t i = 0; i < nc; i++) {
        Cell c;
        for (size_t v = 0; v < c.p.size(); v++) c.p[v] = (size_t) cells[c.p.size()*i + v];
        msh.cells.push_back(c);
    }
    msh.compute_connectivity();
}
} /* extern "C++" */

int yfd_ref_mesh_set(int type, int64_t np, const double *pts, int64_t nc, const int64_t *cells, int64_t nf,
                     const int64_t *faces, const uint8_t *fbnd, const int64_t *fbid)
{
    using namespace yfd_ref;
    M.type = type;
    try {
        if (type == 0) set_mesh<tri_mesh, yaourt::triangle>(M.tri, np, pts, nc, cells, nf, faces, fbnd, fbid);
        else           set_mesh<quad_mesh, yaourt::quadrangle>(M.quad, np, pts, nc, cells, nf, faces, fbnd, fbid);
    } catch (const std::exception& e) {
        std::fprintf(stderr, "[yfd_ref] exception: %s\n", e.what());
        return 1;
    }
    return 0;
}

void yfd_ref_mesh_sizes(int64_t *npoints, int64_t *ncells, int64_t *nfaces, int64_t *nverts)
{
    using namespace yfd_ref;
    if (M.type == 0) { *npoints = M.tri.points.size(); *ncells = M.tri.cells.size(); *nfaces = M.tri.faces.size(); *nverts = 3; }
    else             { *npoints = M.quad.points.size(); *ncells = M.quad.cells.size(); *nfaces = M.quad.faces.size(); *nverts = 4; }
}

void yfd_ref_mesh_copy(double *pts, int64_t *cells, int64_t *faces, int32_t *fbnd, int64_t *owners)
{
    using namespace yfd_ref;
    if (M.type == 0) copy_mesh(M.tri, pts, cells, faces, fbnd, owners);
    else             copy_mesh(M.quad, pts, cells, faces, fbnd, owners);
}

/* ---- knobs ---- */
void yfd_ref_set_cg(double rr_tol, double rr_max, long max_iter, int skip_solve)
{
    yfd_ref::G.rr_tol = rr_tol; yfd_ref::G.rr_max = rr_max;
    yfd_ref::G.max_iter = max_iter; yfd_ref::G.skip_solve = skip_solve != 0;
}
void yfd_ref_keep_raw_triplets(int on) { yfd_ref::G.keep_raw_triplets = on != 0; }
#ifdef YFD_REF_ADVECTION
void yfd_ref_set_upwinding(int on) { yfd_adv_upwinding = on; }          /* dg_config::use_upwinding, dg2d_advection.cpp:104 */
#endif

/* ---- run the unmodified DG-SIP driver body on the current mesh ---- */
int yfd_ref_run_dg(int degree, double eta, int use_pc)
{
    using namespace yfd_ref;
    quiet q;
    try {
        if (M.type == 0) run_dg_on(M.tri, degree, eta, use_pc);
        else             run_dg_on(M.quad, degree, eta, use_pc);
    } catch (const std::exception& e) {
        std::fprintf(stderr, "[yfd_ref] exception: %s\n", e.what());
        return 1;
    }
    return 0;
}

/* ---- run the unmodified CFEM driver (builds its own mesh at `levels`) ---- */
int yfd_ref_run_cfem(int levels)
{
    using namespace yfd_ref;
    quiet q;
    G.ift_calls = 0;
    G.t_enter = now_s();
    try {
        yfd_ref_cfem_driver(levels);              /* continuous_fem.cpp:52, unmodified */
    } catch (const std::exception& e) {
        std::fprintf(stderr, "[yfd_ref] exception: %s\n", e.what());
        return 1;
    }
    return 0;
}

/* ---- results of the last run ---- */
void yfd_ref_result_sizes(int64_t *n, int64_t *nnz, int64_t *nraw, int64_t *iters, int64_t *nhist, int32_t *has_pc)
{
    auto& G = yfd_ref::G;
    *n = (int64_t) G.lhs.rows; *nnz = (int64_t) G.lhs.values.size();
    *nraw = (int64_t) G.raw_v.size(); *iters = G.iters; *nhist = (int64_t) G.history.size();
    *has_pc = G.has_pc ? 1 : 0;
}
void yfd_ref_result_csr(int64_t *rowptr, int64_t *colidx, double *values)
{
    auto& L = yfd_ref::G.lhs;
    std::memcpy(rowptr, L.rowptr.data(), L.rowptr.size() * sizeof(int64_t));
    std::memcpy(colidx, L.colidx.data(), L.colidx.size() * sizeof(int64_t));
    std::memcpy(values, L.values.data(), L.values.size() * sizeof(double));
}
void yfd_ref_result_pc_diag(double *diag)
{
    auto& P = yfd_ref::G.pc;
    for (size_t r = 0; r < P.rows; r++) diag[r] = P.values[(size_t) P.rowptr[r]];
}
void yfd_ref_result_vectors(double *rhs, double *sol, double *history)
{
    auto& G = yfd_ref::G;
    if (rhs) std::memcpy(rhs, G.rhs.data(), G.rhs.size() * sizeof(double));
    if (sol) std::memcpy(sol, G.sol.data(), G.sol.size() * sizeof(double));
    if (history) std::memcpy(history, G.history.data(), G.history.size() * sizeof(double));
}
void yfd_ref_result_raw_triplets(int64_t *i, int64_t *j, double *v)
{
    auto& G = yfd_ref::G;
    std::memcpy(i, G.raw_i.data(), G.raw_i.size() * sizeof(int64_t));
    std::memcpy(j, G.raw_j.data(), G.raw_j.size() * sizeof(int64_t));
    std::memcpy(v, G.raw_v.data(), G.raw_v.size() * sizeof(double));
}
/* wall-clock of the three hot regions: local blocks + triplet pushes, finalize
 * (sort + fold + CSR fill), CG.  Plus the driver's printed L2 errors. */
void yfd_ref_result_times(double *t_blocks, double *t_csr, double *t_cg, double *l2_qp, double *l2_mm, double *mesh_h)
{
    auto& G = yfd_ref::G;
    *t_blocks = G.t_ift_in - G.t_enter;
    *t_csr = G.t_ift_out - G.t_ift_in;
    *t_cg = G.t_cg_out - G.t_cg_in;
    *l2_qp = G.l2_qp; *l2_mm = G.l2_mm; *mesh_h = G.mesh_h;
}

/* ---- small probes of the L2 numerics layer, for unit-level oracle checks ---- */
/* integrate(msh, cell, degree): core/quadratures.hpp:748 / :804 */
int yfd_ref_cell_quadrature(int64_t cell, int degree, double *xyw, int cap)
{
    using namespace yfd_ref;
    int n = 0;
    auto put = [&](const auto& qps) {
        for (auto& qp : qps) {
            if (n < cap) { xyw[3*n] = qp.point().x(); xyw[3*n+1] = qp.point().y(); xyw[3*n+2] = qp.weight(); }
            n++;
        }
    };
    if (M.type == 0) put(yaourt::quadratures::integrate(M.tri, M.tri.cells.at(cell), (size_t) degree));
    else             put(yaourt::quadratures::integrate(M.quad, M.quad.cells.at(cell), (size_t) degree));
    return n;
}
/* integrate(msh, face, degree): core/quadratures.hpp:723 */
int yfd_ref_face_quadrature(int64_t face, int degree, double *xyw, int cap)
{
    using namespace yfd_ref;
    int n = 0;
    auto put = [&](const auto& qps) {
        for (auto& qp : qps) {
            if (n < cap) { xyw[3*n] = qp.point().x(); xyw[3*n+1] = qp.point().y(); xyw[3*n+2] = qp.weight(); }
            n++;
        }
    };
    if (M.type == 0) put(yaourt::quadratures::integrate(M.tri, M.tri.faces.at(face), (size_t) degree));
    else             put(yaourt::quadratures::integrate(M.quad, M.quad.faces.at(face), (size_t) degree));
    return n;
}
/* make_basis(msh, cell, k).eval / eval_grads at one point: core/bases.hpp:96-155,438 */
int yfd_ref_basis_eval(int64_t cell, int degree, double x, double y, double *phi, double *dphi)
{
    using namespace yfd_ref;
    point<double,2> pt(x, y);
    auto doit = [&](const auto& msh) {
        auto b = yaourt::bases::make_basis(msh, msh.cells.at(cell), (size_t) degree);
        auto p = b.eval(pt);
        auto d = b.eval_grads(pt);
        for (size_t i = 0; i < b.size(); i++) { phi[i] = p[i]; dphi[2*i] = d(i,0); dphi[2*i+1] = d(i,1); }
        return (int) b.size();
    };
    return M.type == 0 ? doit(M.tri) : doit(M.quad);
}
/* normal(msh, cell, face) + diameter(msh, face): core/mesh.hpp:505, :474 */
void yfd_ref_face_geometry(int64_t cell, int local_face, double *nxy, double *diam)
{
    using namespace yfd_ref;
    auto doit = [&](const auto& msh) {
        auto& cl = msh.cells.at(cell);
        auto fcs = faces(msh, cl);
        auto n = normal(msh, cl, fcs[local_face]);
        nxy[0] = n[0]; nxy[1] = n[1];
        *diam = diameter(msh, fcs[local_face]);
    };
    if (M.type == 0) doit(M.tri); else doit(M.quad);
}

/* bicgstab as shipped (core/solvers.hpp:246-354 without iM, :356-467 with iM), called on a matrix given in CSR.
 * No reference driver calls it, so the harness does: A is rebuilt as a blaze::CompressedMatrix entry by entry, iM as the
 * diagonal matrix the drivers build for the CG (dg2d_diffusion.cpp:200-212).  The iteration count and nr/nr0 per
 * iteration come from the reference's own history file (6 significant digits, :320-321,:349-353).  Returns the
 * number of history entries (iterations + 1), -1 on a shape error. */
int64_t yfd_ref_bicgstab(int64_t n, const int64_t *rowptr, const int64_t *colidx, const double *values, const double *b,
                         double *x, const double *pc_diag, double rr_tol, double rr_max, int64_t max_iter,
                         int use_initial_guess, double *history, int64_t history_cap)
{
    using namespace yfd_ref;
    if (n < 0) return -1;
    blaze::CompressedMatrix<double> A((size_t) n, (size_t) n);
    A.reserve((size_t) rowptr[n]);
    for (int64_t i = 0; i < n; i++) {
        for (int64_t q = rowptr[i]; q < rowptr[i + 1]; q++) A.append((size_t) i, (size_t) colidx[q], values[q]);
        A.finalize((size_t) i);
    }
    blaze::DynamicVector<double> bv((size_t) n), xv((size_t) n);
    for (int64_t i = 0; i < n; i++) { bv[i] = b[i]; xv[i] = use_initial_guess ? x[i] : 0.0; }
    conjugated_gradient_params<double> cgp;
    cgp.rr_tol = rr_tol; cgp.rr_max = rr_max; cgp.max_iter = (size_t) max_iter;
    cgp.verbose = false;
    cgp.use_initial_guess = use_initial_guess != 0;
    cgp.save_iteration_history = true;
    char path[256];
    snprintf(path, sizeof(path), "/tmp/yfd_ref_bicgstab_%ld.txt", (long) getpid());
    cgp.history_filename = path;
    {
        quiet q;
        if (pc_diag) {
            blaze::CompressedMatrix<double> iM((size_t) n, (size_t) n);
            iM.reserve((size_t) n);
            for (int64_t i = 0; i < n; i++) { iM.append((size_t) i, (size_t) i, pc_diag[i]); iM.finalize((size_t) i); }
            ::bicgstab(cgp, A, bv, xv, iM);
        } else {
            ::bicgstab(cgp, A, bv, xv);
        }
    }
    for (int64_t i = 0; i < n; i++) x[i] = xv[i];
    int64_t cnt = 0;
    std::ifstream ifs(path);
    double h;
    while (ifs >> h) { if (history && cnt < history_cap) history[cnt] = h; cnt++; }
    ifs.close();
    unlink(path);
    return cnt;
}

int yfd_ref_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

} /* extern "C" */

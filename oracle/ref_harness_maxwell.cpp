/*
 * oracle/ref_harness_maxwell.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * The UNMODIFIED reference DG Maxwell TM solver (methods/dg_maxwell_2D.hpp and
 * maxwell_driver.cpp, #included where they lie under /root/reference, `main`
 * renamed) behind a small C ABI -> oracle/_ref/libyfd_ref_maxwell.so.
 * A separate library from libyfd_ref.so because core/bases.hpp defines the
 * non-inline scalar_basis_size (core/bases.hpp:63): one TU per library only.
 *
 * Two ways in:
 *   yfd_refmx_run_driver   run_maxwell_solver<Mesh>(cfg) exactly as shipped
 *                          (maxwell_driver.cpp:194-272), error log to a file
 *   yfd_refmx_create/...   the same reference functions called one by one
 *                          (maxwell_context ctor, assemble, apply_initial_condition,
 *                          do_timestep, report_errors) so every intermediate
 *                          (gM, gOp_ondiag, gOp_offdiag, offdiag_neigh_offsets,
 *                          gDofs) can be read.  The analytic fields are the
 *                          driver's lambdas (maxwell_driver.cpp:202-216, m=n=1)
 *                          restated below because they are local to
 *                          run_maxwell_solver.
 * LAPACK dgesv_ (blaze::solve, methods/dg_maxwell_2D.hpp:284-285,426,441,539-541)
 * comes from oracle/lu_shim.c.  The shim is pinned against a real LAPACK: the same
 * sources linked against the OpenBLAS 0.3.15 bundled with the opencv wheel
 * (`make -C oracle lapack_pin`) reproduce the golden fixtures written with the shim
 * bit for bit up to bs = 6 and to 1.6e-14 at k = 3, 4 (tests/test_oracle_lapack.py).
 *
 * Nothing under yaourt-fem-dg_b200/ may include or link this file.
 */
#include <functional>
#include <memory>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <sstream>
#include <iostream>
#include <unistd.h>

#define main yfd_ref_maxwell_driver_main
#include "maxwell_driver.cpp"
#undef main

namespace yfd_refmx {

using T = double;
using tri_mesh  = yaourt::simplicial_mesh<T>;
using quad_mesh = yaourt::quad_mesh<T>;

struct box {
    int type = 0;
    std::unique_ptr<ymax::maxwell_context<tri_mesh>>  tri;
    std::unique_ptr<ymax::maxwell_context<quad_mesh>> quad;
    double t_assemble = 0, t_steps = 0;
};
static box B;

static double now_s()
{
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

struct quiet {
    std::streambuf *old;
    std::ostringstream sink;
    quiet() { old = std::cout.rdbuf(sink.rdbuf()); }
    ~quiet() { std::cout.rdbuf(old); }
};

/* maxwell_driver.cpp:202-216 with m = n = 1 */
template<typename Mesh>
struct fields {
    using point_type = typename Mesh::point_type;
    static constexpr size_t m = 1, n = 1;
    static T omega() { return M_PI*std::sqrt(m*m + n*n); }
    static ymax::ICF_type<Mesh> Hx() { return [](const point_type& pt, T t) -> T {
        return -(M_PI*n/omega()) * std::sin(m*M_PI*pt.x()) * std::cos(n*M_PI*pt.y()) * std::sin(omega()*t); }; }
    static ymax::ICF_type<Mesh> Hy() { return [](const point_type& pt, T t) -> T {
        return (M_PI*m/omega()) * std::cos(m*M_PI*pt.x()) * std::sin(n*M_PI*pt.y()) * std::sin(omega()*t); }; }
    static ymax::ICF_type<Mesh> Ez() { return [](const point_type& pt, T t) -> T {
        return std::sin(m*M_PI*pt.x()) * std::sin(n*M_PI*pt.y()) * std::cos(omega()*t); }; }
};

template<typename F>
static auto with_ctx(F&& f)
{
    if (B.type == 0) return f(*B.tri);
    return f(*B.quad);
}

template<typename Ctx>
static void copy_mesh(Ctx& c, double *pts, int64_t *cells, int64_t *faces, int32_t *fbnd)
{
    auto& msh = c.msh;
    const size_t nv = msh.cells.empty() ? 0 : msh.cells[0].p.size();
    if (pts)
        for (size_t i = 0; i < msh.points.size(); i++) { pts[2*i] = msh.points[i].x(); pts[2*i+1] = msh.points[i].y(); }
    if (cells)
        for (size_t i = 0; i < msh.cells.size(); i++)
            for (size_t v = 0; v < nv; v++) cells[nv*i + v] = (int64_t) msh.cells[i].p[v];
    for (size_t i = 0; i < msh.faces.size(); i++) {
        if (faces) { faces[2*i] = (int64_t) msh.faces[i].p0; faces[2*i+1] = (int64_t) msh.faces[i].p1; }
        if (fbnd)  fbnd[i] = msh.faces[i].is_boundary ? 1 : 0;
    }
}

} // namespace yfd_refmx

extern "C" {

int yfd_refmx_create(int type, int levels, int degree, double dt, double eta, int upwind, int euler, int shatter)
{
    using namespace yfd_refmx;
    quiet q;
    ymax::maxwell_config<T> cfg;                       /* methods/dg_maxwell_2D.hpp:46-72 */
    cfg.degree = degree; cfg.mesh_levels = levels; cfg.delta_t = dt; cfg.eta = eta;
    cfg.upwind = upwind != 0; cfg.shatter_mesh = shatter != 0;
    cfg.time_integrator = euler ? ymax::time_integrator_type::EXPLICIT_EULER : ymax::time_integrator_type::RUNGE_KUTTA_4;
    B.type = type;
    B.tri.reset(); B.quad.reset();
    try {
        if (type == 0) B.tri.reset(new ymax::maxwell_context<tri_mesh>(cfg));      /* :116-152 */
        else           B.quad.reset(new ymax::maxwell_context<quad_mesh>(cfg));
    } catch (const std::exception& e) { std::fprintf(stderr, "[yfd_refmx] %s\n", e.what()); return 1; }
    return 0;
}

void yfd_refmx_destroy(void) { yfd_refmx::B.tri.reset(); yfd_refmx::B.quad.reset(); }

void yfd_refmx_sizes(int64_t *npoints, int64_t *ncells, int64_t *nfaces, int64_t *nverts, int64_t *bs)
{
    using namespace yfd_refmx;
    with_ctx([&](auto& c) {
        *npoints = c.msh.points.size(); *ncells = c.msh.cells.size(); *nfaces = c.msh.faces.size();
        *nverts = c.faces_per_elem(); *bs = c.basis_size; return 0; });
}

void yfd_refmx_mesh_copy(double *pts, int64_t *cells, int64_t *faces, int32_t *fbnd)
{
    using namespace yfd_refmx;
    with_ctx([&](auto& c) { copy_mesh(c, pts, cells, faces, fbnd); return 0; });
}

/* mu_r / eps_r are public members of the context (methods/dg_maxwell_2D.hpp:101); set before assemble */
void yfd_refmx_set_material(const double *mu_r, const double *eps_r)
{
    using namespace yfd_refmx;
    with_ctx([&](auto& c) {
        for (size_t i = 0; i < c.msh.cells.size(); i++) { c.mu_r[i] = mu_r[i]; c.eps_r[i] = eps_r[i]; }
        return 0; });
}

int yfd_refmx_assemble(void)
{
    using namespace yfd_refmx;
    quiet q;
    try {
        const double t0 = now_s();
        with_ctx([&](auto& c) { ymax::assemble(c); return 0; });                  /* :220-461 */
        B.t_assemble = now_s() - t0;
    } catch (const std::exception& e) { std::fprintf(stderr, "[yfd_refmx] %s\n", e.what()); return 1; }
    return 0;
}

/* ondiag [ncells*3bs][3bs], offdiag [ncells*nf*3bs][3bs] row-major as stored by the reference (:87);
 * neigh [ncells*nf] = offdiag_neigh_offsets.first, has [ncells*nf] = .second (:106-107) */
void yfd_refmx_get_operator(double *ondiag, double *offdiag, int64_t *neigh, int32_t *has, double *gM)
{
    using namespace yfd_refmx;
    with_ctx([&](auto& c) {
        const size_t nc = c.msh.cells.size(), bs = c.basis_size, nf = c.faces_per_elem(), L = 3*bs;
        if (ondiag)
            for (size_t i = 0; i < nc*L; i++) for (size_t j = 0; j < L; j++) ondiag[i*L + j] = c.gOp_ondiag(i, j);
        if (offdiag)
            for (size_t i = 0; i < nc*nf*L; i++) for (size_t j = 0; j < L; j++) offdiag[i*L + j] = c.gOp_offdiag(i, j);
        for (size_t i = 0; i < nc*nf; i++) {
            if (neigh) neigh[i] = (int64_t) c.offdiag_neigh_offsets[i].first;
            if (has)   has[i] = c.offdiag_neigh_offsets[i].second ? 1 : 0;
        }
        if (gM)
            for (size_t i = 0; i < nc*bs; i++) for (size_t j = 0; j < bs; j++) gM[i*bs + j] = c.gM(i, j);
        return 0; });
}

int yfd_refmx_init(void)
{
    using namespace yfd_refmx;
    try {
        if (B.type == 0) ymax::apply_initial_condition(*B.tri, fields<tri_mesh>::Hx(), fields<tri_mesh>::Hy(), fields<tri_mesh>::Ez());
        else             ymax::apply_initial_condition(*B.quad, fields<quad_mesh>::Hx(), fields<quad_mesh>::Hy(), fields<quad_mesh>::Ez());
    } catch (const std::exception& e) { std::fprintf(stderr, "[yfd_refmx] %s\n", e.what()); return 1; }
    return 0;
}

void yfd_refmx_get_dofs(double *dofs)
{
    using namespace yfd_refmx;
    with_ctx([&](auto& c) { for (size_t i = 0; i < c.gDofs.size(); i++) dofs[i] = c.gDofs[i]; return 0; });
}

void yfd_refmx_set_dofs(const double *dofs)
{
    using namespace yfd_refmx;
    with_ctx([&](auto& c) { for (size_t i = 0; i < c.gDofs.size(); i++) c.gDofs[i] = dofs[i]; return 0; });
}

/* nsteps x do_timestep (:558-675); returns wall seconds */
double yfd_refmx_step(int nsteps)
{
    using namespace yfd_refmx;
    quiet q;
    const double t0 = now_s();
    with_ctx([&](auto& c) { for (int s = 0; s < nsteps; s++) ymax::do_timestep(c); return 0; });
    B.t_steps = now_s() - t0;
    return B.t_steps;
}

/* report_errors (maxwell_driver.cpp:133-191) at `cycle` (analytic fields at cycle*delta_t, as the driver does) */
void yfd_refmx_errors(int64_t cycle, double out[3])
{
    using namespace yfd_refmx;
    if (B.type == 0) {
        auto ei = report_errors(*B.tri, (size_t) cycle, fields<tri_mesh>::Hx(), fields<tri_mesh>::Hy(), fields<tri_mesh>::Ez());
        out[0] = ei.Hx_err; out[1] = ei.Hy_err; out[2] = ei.Ez_err;
    } else {
        auto ei = report_errors(*B.quad, (size_t) cycle, fields<quad_mesh>::Hx(), fields<quad_mesh>::Hy(), fields<quad_mesh>::Ez());
        out[0] = ei.Hx_err; out[1] = ei.Hy_err; out[2] = ei.Ez_err;
    }
}

double yfd_refmx_time_assemble(void) { return yfd_refmx::B.t_assemble; }

/* the driver body exactly as shipped; error log (one line per cycle) written to `error_fn` */
int yfd_refmx_run_driver(int type, int levels, int degree, double dt, int64_t timesteps, int upwind, int euler, const char *error_fn)
{
    using namespace yfd_refmx;
    quiet q;
    ymax::maxwell_config<T> cfg;
    cfg.degree = degree; cfg.mesh_levels = levels; cfg.delta_t = dt; cfg.timesteps = (size_t) timesteps;
    cfg.upwind = upwind != 0; cfg.error_fn = const_cast<char *>(error_fn);
    cfg.time_integrator = euler ? ymax::time_integrator_type::EXPLICIT_EULER : ymax::time_integrator_type::RUNGE_KUTTA_4;
    try {
        if (type == 0) run_maxwell_solver<tri_mesh>(cfg);       /* maxwell_driver.cpp:194 */
        else           run_maxwell_solver<quad_mesh>(cfg);
    } catch (const std::exception& e) { std::fprintf(stderr, "[yfd_refmx] %s\n", e.what()); return 1; }
    return 0;
}

int yfd_refmx_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

} /* extern "C" */

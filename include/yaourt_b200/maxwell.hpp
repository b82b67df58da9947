/*
 * include/yaourt_b200/maxwell.hpp -- the reference's DG Maxwell names on top of the C ABI (include/yfd.h).
 *
 * Include this header AFTER methods/dg_maxwell_2D.hpp in a translation unit that is otherwise unchanged (for instance
 * after maxwell_driver.cpp itself, see oracle/dropin_maxwell_driver.cpp).  It declares NON-template overloads
 *
 *   yaourt::maxwell_2D::assemble(maxwell_context<simplicial_mesh<double>>&)        replaces methods/dg_maxwell_2D.hpp:220-461
 *   yaourt::maxwell_2D::assemble(maxwell_context<quad_mesh<double>>&)
 *   yaourt::maxwell_2D::do_timestep(maxwell_context<...>&)                         replaces methods/dg_maxwell_2D.hpp:558-675
 *
 * with the reference's signatures.  The driver's calls are unqualified and dependent (maxwell_driver.cpp:239,243), so
 * they are found by ADL at instantiation and the non-template overloads win: run_maxwell_solver runs on the GPU without
 * a source change.  The context object keeps its meaning: after assemble() ctx.gM and ctx.offdiag_neigh_offsets hold
 * what the reference computes (apply_initial_condition, which takes host callbacks, keeps working as shipped), and
 * ctx.gOp_ondiag / ctx.gOp_offdiag are filled in the reference's dense layout unless
 * maxwell_options().materialize_dense is cleared (they are (1+nf)(3bs)^2 doubles per cell and the GPU path never reads
 * them).  do_timestep() uploads ctx.gDofs, advances one step on the device and writes ctx.gDofs back, so host code
 * between steps (report_errors, silo export) sees the reference's state.  For runs that should stay resident use
 *
 *   yaourt::b200::do_timesteps(ctx, n)          n steps, one upload + one download
 *   yaourt::b200::report_errors_device(ctx, t)  the driver's three L2 errors computed on the device
 *
 * Errors: std::invalid_argument / std::runtime_error as in core.hpp.  No CPU fallback.
 */
#pragma once
#include <array>
#include <map>
#include "yaourt_b200/core.hpp"

namespace yaourt {
namespace b200 {

struct maxwell_options_t { bool materialize_dense = true; };
inline maxwell_options_t &maxwell_options()
{
    static maxwell_options_t o;
    return o;
}

inline std::map<const void *, yfd_maxwell *> &maxwell_registry()
{
    static struct holder {
        std::map<const void *, yfd_maxwell *> m;
        ~holder() { for (auto &kv : m) yfd_maxwell_destroy(kv.second); }
    } h;
    (void) context::global();          /* constructed first, destroyed last */
    return h.m;
}

template <typename Ctx>
yfd_maxwell *maxwell_handle(Ctx &ctx)
{
    auto &reg = maxwell_registry();
    auto it = reg.find(&ctx);
    if (it == reg.end()) throw std::invalid_argument("yaourt_b200: do_timestep before assemble");
    return it->second;
}

template <typename Ctx>
void maxwell_release(Ctx &ctx)
{
    auto &reg = maxwell_registry();
    auto it = reg.find(&ctx);
    if (it != reg.end()) { yfd_maxwell_destroy(it->second); reg.erase(it); }
}

template <typename Ctx>
void maxwell_assemble(Ctx &ctx)
{
    namespace ym = yaourt::maxwell_2D;
    auto &g = context::global();
    maxwell_release(ctx);
    yfd_mesh *dm = upload_mesh(g, ctx.msh);
    yfd_maxwell_config cfg;
    yfd_maxwell_config_default(&cfg);
    cfg.degree = (int) ctx.cfg.degree; cfg.delta_t = ctx.cfg.delta_t; cfg.eta = ctx.cfg.eta; cfg.upwind = ctx.cfg.upwind ? 1 : 0;
    cfg.time_integrator = ctx.cfg.time_integrator == ym::time_integrator_type::EXPLICIT_EULER ? YFD_TI_EULER : YFD_TI_RK4;
    const size_t nc = ctx.msh.cells.size();
    std::vector<double> mu(nc), eps(nc);
    for (size_t i = 0; i < nc; i++) { mu[i] = ctx.mu_r[i]; eps[i] = ctx.eps_r[i]; }
    yfd_maxwell *h = nullptr;
    yfd_maxwell_info info;
    const yfd_status st = yfd_maxwell_create(g.get(), dm, &cfg, mu.data(), eps.data(), &h, &info);
    yfd_mesh_destroy(dm);
    check(g.get(), st);
    maxwell_registry()[&ctx] = h;
    /* what host code after assemble() may read: gM (apply_initial_condition, :538), offdiag_neigh_offsets, dense blocks */
    const size_t bs = (size_t) info.basis_size, nf = (size_t) info.faces_per_elem, L = 3 * bs;
    const bool dense = maxwell_options().materialize_dense;
    std::vector<double> gM(nc * bs * bs), on(dense ? nc * L * L : 0), off(dense ? nc * nf * L * L : 0);
    std::vector<int64_t> neigh(nc * nf);
    std::vector<int32_t> has(nc * nf);
    check(g.get(), yfd_maxwell_get_operator(h, dense ? on.data() : nullptr, dense ? off.data() : nullptr, neigh.data(), has.data(), gM.data()));
    for (size_t i = 0; i < nc * bs; i++)
        for (size_t j = 0; j < bs; j++) ctx.gM(i, j) = gM[i * bs + j];
    for (size_t i = 0; i < nc * nf; i++) ctx.offdiag_neigh_offsets[i] = std::make_pair((size_t) neigh[i], has[i] != 0);
    if (dense) {
        for (size_t i = 0; i < nc * L; i++)
            for (size_t j = 0; j < L; j++) ctx.gOp_ondiag(i, j) = on[i * L + j];
        for (size_t i = 0; i < nc * nf * L; i++)
            for (size_t j = 0; j < L; j++) ctx.gOp_offdiag(i, j) = off[i * L + j];
    }
}

/* n x do_timestep with one upload of ctx.gDofs and one download */
template <typename Ctx>
void do_timesteps(Ctx &ctx, size_t n)
{
    namespace ym = yaourt::maxwell_2D;
    auto &g = context::global();
    yfd_maxwell *h = maxwell_handle(ctx);
    check(g.get(), yfd_maxwell_set_time_integration(h, ctx.cfg.delta_t,
          ctx.cfg.time_integrator == ym::time_integrator_type::EXPLICIT_EULER ? YFD_TI_EULER : YFD_TI_RK4));
    check(g.get(), yfd_maxwell_set_dofs(h, ctx.gDofs.data()));
    check(g.get(), yfd_maxwell_timesteps(h, (int64_t) n, nullptr));
    check(g.get(), yfd_maxwell_get_dofs(h, ctx.gDofs.data()));
}

/* the driver's report_errors (maxwell_driver.cpp:133-191) for its TM_11 cavity fields, on the device state */
template <typename Ctx>
std::array<double, 3> report_errors_device(Ctx &ctx, double t)
{
    std::array<double, 3> e{};
    check(context::global().get(), yfd_maxwell_report_errors(maxwell_handle(ctx), YFD_MAXWELL_FIELDS_CAVITY_11, t, e.data()));
    return e;
}

} // namespace b200

namespace maxwell_2D {

inline void assemble(maxwell_context<yaourt::simplicial_mesh<double>> &ctx) { yaourt::b200::maxwell_assemble(ctx); }
inline void assemble(maxwell_context<yaourt::quad_mesh<double>> &ctx) { yaourt::b200::maxwell_assemble(ctx); }
inline void do_timestep(maxwell_context<yaourt::simplicial_mesh<double>> &ctx) { yaourt::b200::do_timesteps(ctx, 1); }
inline void do_timestep(maxwell_context<yaourt::quad_mesh<double>> &ctx) { yaourt::b200::do_timesteps(ctx, 1); }

} // namespace maxwell_2D
} // namespace yaourt

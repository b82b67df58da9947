/*
 * oracle/dropin_driver.cpp -- TEST INFRASTRUCTURE: the reference's dg2d_diffusion.cpp, UNMODIFIED (#included where it
 * lies under /root/reference), recompiled against include/yaourt_b200/dropin.hpp.  Its assembler::finalize()
 * (-> blaze::init_from_triplets) and its conjugated_gradient calls now run on the GPU through libyfd_b200.so; the
 * driver prints its usual L2 errors, which tests/test_gpu_dropin.py compares with the reference's own run.
 * Built by oracle/Makefile into oracle/_ref/dropin_dg2d_diffusion (needs /root/reference at build time only).
 *
 *   dropin_dg2d_diffusion [-r levels] [-k degree] [-m tri|quad] [-p] [-S]        (the reference's own flags)
 *   YFD_DROPIN_BATCHED=1 adds a line comparing the batched yaourt::b200::assemble_dg_sip with the driver's system.
 */
#include <functional>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <type_traits>
#include "core/blaze_sparse_init.hpp"
#include "yaourt_b200/dropin_triplets.hpp"   /* before methods/dg.hpp is parsed: that call is qualified */
#define main yfd_reference_main
#ifdef YFD_DROPIN_ADVECTION          /* -> oracle/_ref/dropin_dg2d_advection: dg2d_advection.cpp, CG on the normal equations */
#include "dg2d_advection.cpp"
#else
#include "dg2d_diffusion.cpp"
#endif
#undef main
#include "yaourt_b200/dropin.hpp"       /* after the reference headers: the overloads win at instantiation */

/* YFD_DROPIN_BATCHED="<tri|quad> <levels> <degree> <pc> [fn|cfem]": the batched entries on a mesh made by the reference's own
 * mesher; prints sizes, iteration count and ||sol||_2 so the test can compare with the reference.  "fn" = the overload with
 * explicit data-function ids; "cfem" = assemble_cfem_p1; the advection build runs assemble_dg_advection
 * (pc = use_upwinding there) */
template <typename Mesh>
static void batched(int levels, int degree, bool pc, const char *variant)
{
    Mesh msh;
    auto mesher = yaourt::get_mesher(msh);
    mesher.create_mesh(msh, (size_t) levels);
    conjugated_gradient_params<double> cgp;
    cgp.rr_max = 10000; cgp.rr_tol = 1e-12;
    cgp.max_iter = 2 * yaourt::bases::scalar_basis_size((size_t) degree, 2) * msh.cells.size();
    std::cout.precision(17);
#ifdef YFD_DROPIN_ADVECTION
    cgp.use_normal_eqns = true; cgp.rr_tol = 1e-8;             /* dg2d_advection.cpp:222-225 */
    auto sys = yaourt::b200::assemble_dg_advection(msh, (size_t) degree, 1.0, pc, &cgp);
#else
    if (std::strcmp(variant, "cfem") == 0) {
        if constexpr (std::is_same<Mesh, yaourt::simplicial_mesh<double>>::value) {
            cgp.rr_max = 1000;
            auto cs = yaourt::b200::assemble_cfem_p1(msh, YFD_FN_SINSIN, &cgp);
            size_t ndir = 0;
            for (auto c : cs.compress_map) ndir += c < 0;
            std::cout << "[yaourt_b200] batched cfem: n " << cs.rhs.size() << " nnz " << cs.lhs.nonZeros() << " iterations "
                      << yaourt::b200::last_cg_report().iterations << " sol_norm " << blaze::norm(cs.sol) << " rhs_norm " << blaze::norm(cs.rhs)
                      << " expanded_norm " << blaze::norm(cs.expanded) << " dirichlet " << ndir << std::endl;
        }
        return;
    }
    auto sys = std::strcmp(variant, "fn") == 0 ? yaourt::b200::assemble_dg_sip(msh, (size_t) degree, 1.0, YFD_FN_SINSIN, YFD_FN_ZERO, pc, &cgp)
                                               : yaourt::b200::assemble_dg_sip(msh, (size_t) degree, 1.0, pc, &cgp);
#endif
    std::cout << "[yaourt_b200] batched: n " << sys.rhs.size() << " nnz " << sys.lhs.nonZeros() << " iterations "
              << yaourt::b200::last_cg_report().iterations << " sol_norm " << blaze::norm(sys.sol) << " rhs_norm " << blaze::norm(sys.rhs)
              << std::endl;
}

int main(int argc, char **argv)
{
    if (const char *b = std::getenv("YFD_DROPIN_BATCHED")) {
        char type[16] = "tri", variant[16] = "";
        int levels = 3, degree = 1, pc = 0;
        std::sscanf(b, "%15s %d %d %d %15s", type, &levels, &degree, &pc, variant);
        if (std::strcmp(type, "quad") == 0) batched<yaourt::quad_mesh<double>>(levels, degree, pc != 0, variant);
        else batched<yaourt::simplicial_mesh<double>>(levels, degree, pc != 0, variant);
        return 0;
    }
    const int rc = yfd_reference_main(argc, argv);
    const auto &rep = yaourt::b200::last_cg_report();
    std::cout << "[yaourt_b200] CG on GPU: iterations " << rep.iterations << ", rr " << (rep.nr0 > 0 ? rep.nr / rep.nr0 : 0.0)
              << ", converged " << rep.converged << ", device ms " << rep.ms_solve << std::endl;
    return rc;
}

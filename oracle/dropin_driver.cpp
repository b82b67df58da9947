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

#ifndef YFD_DROPIN_ADVECTION
/* YFD_DROPIN_BATCHED="<tri|quad> <levels> <degree> <pc>": the batched entry on a mesh made by the reference's own mesher;
 * prints sizes, iteration count and ||sol||_2 so the test can compare with the reference */
template <typename Mesh>
static void batched(int levels, int degree, bool pc)
{
    Mesh msh;
    auto mesher = yaourt::get_mesher(msh);
    mesher.create_mesh(msh, (size_t) levels);
    conjugated_gradient_params<double> cgp;
    cgp.rr_max = 10000; cgp.rr_tol = 1e-12;
    cgp.max_iter = 2 * yaourt::bases::scalar_basis_size((size_t) degree, 2) * msh.cells.size();
    auto sys = yaourt::b200::assemble_dg_sip(msh, (size_t) degree, 1.0, pc, &cgp);
    std::cout.precision(17);
    std::cout << "[yaourt_b200] batched: n " << sys.rhs.size() << " nnz " << sys.lhs.nonZeros() << " iterations "
              << yaourt::b200::last_cg_report().iterations << " sol_norm " << blaze::norm(sys.sol) << " rhs_norm " << blaze::norm(sys.rhs)
              << std::endl;
}

#endif

int main(int argc, char **argv)
{
#ifndef YFD_DROPIN_ADVECTION
    if (const char *b = std::getenv("YFD_DROPIN_BATCHED")) {
        char type[16] = "tri";
        int levels = 3, degree = 1, pc = 0;
        std::sscanf(b, "%15s %d %d %d", type, &levels, &degree, &pc);
        if (std::strcmp(type, "quad") == 0) batched<yaourt::quad_mesh<double>>(levels, degree, pc != 0);
        else batched<yaourt::simplicial_mesh<double>>(levels, degree, pc != 0);
        return 0;
    }
#endif
    const int rc = yfd_reference_main(argc, argv);
    const auto &rep = yaourt::b200::last_cg_report();
    std::cout << "[yaourt_b200] CG on GPU: iterations " << rep.iterations << ", rr " << (rep.nr0 > 0 ? rep.nr / rep.nr0 : 0.0)
              << ", converged " << rep.converged << ", device ms " << rep.ms_solve << std::endl;
    return rc;
}

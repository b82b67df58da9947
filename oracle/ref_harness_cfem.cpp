/*
 * oracle/ref_harness_cfem.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 * Second TU of oracle/_ref/libyfd_ref.so: the UNMODIFIED continuous_fem.cpp
 * driver (#included where it lies, `main` renamed) with the same two hooks as
 * ref_harness.cpp.  See that file for the description.
 */
#include "ref_common.hpp"

#define init_from_triplets yfd_hook_ift
#include "methods/cfem.hpp"
#undef init_from_triplets

#define YFD_CG_NAME yfd_cfem_cg
YFD_DECLARE_CG_HOOKS(yfd_cfem_cg)

/* The CFEM driver hard-codes 2 refinements (continuous_fem.cpp:60); only that
 * literal argument is redirected, the mesher itself is the reference's. */
static size_t yfd_cfem_levels = 2;

#define conjugated_gradient yfd_cfem_cg
#define main yfd_ref_continuous_fem_main
#define create_mesh(m, n) create_mesh(m, yfd_cfem_levels)
#include "continuous_fem.cpp"
#undef create_mesh
#undef main
#undef conjugated_gradient

#include "ref_cg_hooks.inl"

extern "C" int yfd_ref_cfem_driver(int levels)
{
    yfd_cfem_levels = (size_t) levels;
    char a0[] = "continuous_fem";
    char *argv[] = { a0, nullptr };
    return yfd_ref_continuous_fem_main(1, argv);
}

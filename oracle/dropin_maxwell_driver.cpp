/*
 * oracle/dropin_maxwell_driver.cpp -- TEST INFRASTRUCTURE: the reference's maxwell_driver.cpp, UNMODIFIED (#included
 * where it lies under /root/reference), recompiled against include/yaourt_b200/maxwell.hpp.  Its assemble(ctx) and
 * do_timestep(ctx) calls (maxwell_driver.cpp:239,243) now run on the GPU through libyfd_b200.so; mesh creation,
 * apply_initial_condition (host callbacks) and report_errors stay the reference's own code.  The error log it writes
 * (-E file) is compared by tests/test_gpu_dropin.py with the log of the reference's own run.
 * Built by oracle/Makefile into oracle/_ref/dropin_maxwell (needs /root/reference at build time only).
 *
 *   dropin_maxwell [-m tri|quad] [-r levels] [-k degree] [-d dt] [-t steps] [-u] [-E errlog] ...   (the reference's own flags)
 */
#include <functional>
#include <cstdio>
#define main yfd_reference_main
#include "maxwell_driver.cpp"
#undef main
#include "yaourt_b200/maxwell.hpp"      /* after the reference headers: the overloads win at instantiation */

int main(int argc, char **argv)
{
    const int rc = yfd_reference_main(argc, argv);
    std::cout << "[yaourt_b200] maxwell contexts on the GPU: " << yaourt::b200::maxwell_registry().size() << std::endl;
    return rc;
}

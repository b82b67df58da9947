/*
 * include/yaourt_b200/dropin_triplets.hpp -- blaze::init_from_triplets on the GPU, the reference's signature.
 *
 * The reference calls it QUALIFIED (`blaze::init_from_triplets(lhs, triplets.begin(), triplets.end())`,
 * methods/dg.hpp:130,141; methods/cfem.hpp:184), so the overload must be visible where those headers are parsed:
 * include this file after core/blaze_sparse_init.hpp and BEFORE methods/dg.hpp / methods/cfem.hpp.
 */
#pragma once
#include "yaourt_b200/core.hpp"

/* ------------------------------------------------------------------------------------------------ */
/* blaze::init_from_triplets (core/blaze_sparse_init.hpp:125-155): sort + fold + CSR build on the GPU */
/* ------------------------------------------------------------------------------------------------ */
namespace blaze {

inline void init_from_triplets(CompressedMatrix<double> &M, std::vector<triplet<double>>::iterator first,
                               std::vector<triplet<double>>::iterator last)
{
    auto &ctx = yaourt::b200::context::global();
    const size_t count = (size_t) (last - first);
    std::vector<int64_t> ri(count), rj(count);
    std::vector<double> rv(count);
    size_t k = 0;
    for (auto it = first; it != last; ++it, ++k) { ri[k] = (int64_t) it->row(); rj[k] = (int64_t) it->col(); rv[k] = it->value(); }
    yfd_csr *A = nullptr;
    yaourt::b200::check(ctx.get(), yfd_csr_from_triplets(ctx.get(), (int64_t) M.rows(), (int64_t) M.columns(), (int64_t) count,
                                                         ri.data(), rj.data(), rv.data(), &A));
    int64_t nr = 0, ncol = 0, nnz = 0;
    yfd_csr_sizes(A, &nr, &ncol, &nnz);
    std::vector<int64_t> rowptr((size_t) nr + 1), colidx((size_t) nnz);
    std::vector<double> values((size_t) nnz);
    const yfd_status st = yfd_csr_get(A, rowptr.data(), colidx.data(), values.data());
    yfd_csr_destroy(A);
    yaourt::b200::check(ctx.get(), st);
    yaourt::b200::fill_compressed(M, (size_t) nr, (size_t) ncol, rowptr, colidx, values);
}

} // namespace blaze


/*
 * oracle/ref_cg_hooks.inl -- TEST INFRASTRUCTURE.  Included once per TU AFTER the
 * reference driver; defines the non-template CG overloads declared by
 * YFD_DECLARE_CG_HOOKS (ref_common.hpp).  YFD_CG_NAME<double>(...) below is the
 * reference's own conjugated_gradient template (core/solvers.hpp:46,141).
 */
static conjugated_gradient_params<double>
yfd_override(const conjugated_gradient_params<double>& in)
{
    auto& G = yfd_ref::G;
    conjugated_gradient_params<double> p = in;
    p.verbose = false;
    p.rr_tol = G.rr_tol;
    p.rr_max = G.rr_max;
    if (G.max_iter >= 0) p.max_iter = (size_t) G.max_iter;
    p.save_iteration_history = true;
    p.history_filename = yfd_ref::history_path();
    return p;
}

bool YFD_CG_NAME(const conjugated_gradient_params<double>& cgp,
                 const blaze::CompressedMatrix<double>& A,
                 const blaze::DynamicVector<double>& b,
                 blaze::DynamicVector<double>& x)
{
    auto& G = yfd_ref::G;
    G.lhs.take(A);
    G.rhs.assign(b.data(), b.data() + b.size());
    G.has_pc = false;
    bool ok = true;
    if (G.skip_solve) {
        x = blaze::DynamicVector<double>(b.size(), 0.0);
        G.t_cg_in = G.t_cg_out = yfd_ref::now_s();
        G.iters = 0; G.history.clear();
    } else {
        auto p = yfd_override(cgp);
        G.t_cg_in = yfd_ref::now_s();
        ok = YFD_CG_NAME<double>(p, A, b, x);      /* the reference's own template */
        G.t_cg_out = yfd_ref::now_s();
        yfd_ref::read_history();
    }
    G.sol.assign(x.data(), x.data() + x.size());
    return ok;
}

bool YFD_CG_NAME(const conjugated_gradient_params<double>& cgp,
                 const blaze::CompressedMatrix<double>& A,
                 const blaze::DynamicVector<double>& b,
                 blaze::DynamicVector<double>& x,
                 const blaze::CompressedMatrix<double>& iM)
{
    auto& G = yfd_ref::G;
    G.lhs.take(A);
    G.pc.take(iM);
    G.has_pc = true;
    G.rhs.assign(b.data(), b.data() + b.size());
    bool ok = true;
    if (G.skip_solve) {
        x = blaze::DynamicVector<double>(b.size(), 0.0);
        G.t_cg_in = G.t_cg_out = yfd_ref::now_s();
        G.iters = 0; G.history.clear();
    } else {
        auto p = yfd_override(cgp);
        G.t_cg_in = yfd_ref::now_s();
        ok = YFD_CG_NAME<double>(p, A, b, x, iM);  /* the reference's own template */
        G.t_cg_out = yfd_ref::now_s();
        yfd_ref::read_history();
    }
    G.sol.assign(x.data(), x.data() + x.size());
    return ok;
}


"""GPU parity of the advection-reaction path (yfd_dg_assemble_advection + CG on the normal equations) against the
UNMODIFIED dg2d_advection.cpp (oracle/_ref/libyfd_ref_adv.so) through the C ABI.

Bars: pattern bit-exact, values / rhs 1e-12 (as for SIP).  The driver solves the non-symmetric system by CG on
A^T A x = A^T b (dg2d_advection.cpp:225), which squares the condition number: iterates of two correct implementations
drift apart after a few dozen iterations, so the solve is pinned by (a) the early residual history, (b) the true
residual / a direct solve of the same matrix once converged, (c) the driver's L2 error."""
import numpy as np
import pytest

from conftest import host_mesh_from_ref, block_max_rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def aref():
    import refbind
    if not refbind.available(advection=True):
        pytest.skip("oracle/_ref/libyfd_ref_adv.so not built (needs /root/reference at build time)")
    return refbind.Ref(advection=True)


CASES = [(0, 3, 1, 0, 0), (0, 2, 2, 1, 0), (1, 2, 3, 0, 0), (1, 3, 2, 1, 1), (0, 1, 4, 1, 1), (0, 2, 0, 0, 0)]


@pytest.mark.parametrize("mtype,r,k,up,shatter", CASES)
def test_system_matches_reference(yfd, aref, ctx, mtype, r, k, up, shatter):
    m = aref.mesh(mtype, r, shatter=bool(shatter))
    aref.set_cg(rr_tol=1e-8, rr_max=1e4, max_iter=30)
    out = aref.run_dg(k, eta=1.0, upwinding=bool(up))
    mesh = yfd.Mesh(ctx, host_mesh_from_ref(yfd, m))
    sysm = ctx.dg_assemble_advection(mesh, k, eta=1.0, upwinding=bool(up))
    rowptr, colidx, vals = sysm.csr()
    assert np.array_equal(rowptr, out["rowptr"]) and np.array_equal(colidx, out["colidx"])
    bs = (k + 1) * (k + 2) // 2
    assert block_max_rel_err(vals, out["values"], rowptr, bs) <= 1e-12
    assert np.abs(sysm.rhs() - out["rhs"]).max() <= 1e-12 * np.abs(out["rhs"]).max()
    # CG on the normal equations: the reference's own residual history, iteration by iteration (6 printed digits)
    res = sysm.solve_cg(yfd.CgParams(rr_tol=1e-8, rr_max=1e4, max_iter=30, use_normal_eqns=1), history=True)
    nh = min(len(res["history"]), len(out["history"]), 8)
    assert nh >= 2
    assert np.allclose(res["history"][:nh], out["history"][:nh], rtol=1e-4)


@pytest.mark.parametrize("mtype,r,k,up", [(0, 3, 1, 1), (1, 3, 2, 0)])
def test_normal_equations_solve(yfd, aref, ctx, mtype, r, k, up):
    import scipy.sparse as sp
    import scipy.sparse.linalg as spl
    m = aref.mesh(mtype, r)
    aref.set_cg(rr_tol=1e-8, rr_max=1e4, max_iter=-1)
    out = aref.run_dg(k, eta=1.0, upwinding=bool(up))          # the driver as shipped: rr_tol 1e-8, max_iter 2N
    mesh = yfd.Mesh(ctx, host_mesh_from_ref(yfd, m))
    sysm = ctx.dg_assemble_advection(mesh, k, eta=1.0, upwinding=bool(up))
    res = sysm.solve_cg(yfd.CgParams(rr_tol=1e-8, rr_max=1e4, max_iter=2 * sysm.n, use_normal_eqns=1))
    A = sp.csr_matrix((out["values"], out["colidx"], out["rowptr"]))
    direct = spl.spsolve(A.tocsc(), out["rhs"])
    # both stop on the same criterion (normal-equation residual or 2N iterations): same distance class from the direct solve
    e_gpu = np.linalg.norm(res["x"] - direct) / np.linalg.norm(direct)
    e_ref = np.linalg.norm(out["sol"] - direct) / np.linalg.norm(direct)
    assert e_gpu <= max(3 * e_ref, 1e-6)
    qp, mm = ctx.dg_postprocess(mesh, k, res["x"], ref_fn=yfd.FN_SIN_X_SOL)
    assert abs(np.sqrt(qp) - out["l2_qp"]) <= 0.05 * out["l2_qp"]
    # the generic CSR entry (drop-in conjugated_gradient with cgp.use_normal_eqns) on the reference's own matrix
    csr = yfd.Csr.from_arrays(ctx, sysm.n, sysm.n, out["rowptr"], out["colidx"], out["values"])
    res2 = csr.solve_cg(out["rhs"], yfd.CgParams(rr_tol=1e-8, rr_max=1e4, max_iter=2 * sysm.n, use_normal_eqns=1))
    assert np.linalg.norm(res2["x"] - direct) / np.linalg.norm(direct) <= max(3 * e_ref, 1e-6)


def test_transposed_blocks(yfd, ctx):
    """A^T through the block layout: <A^T A d, d> history equals that of the explicit CSR transpose path."""
    mesh = yfd.Mesh(ctx, yfd.HostMesh.create(1, 3))
    sysm = ctx.dg_assemble_advection(mesh, 3, eta=1.0, upwinding=True)
    rowptr, colidx, vals = sysm.csr()
    p = yfd.CgParams(rr_tol=1e-30, rr_max=1e30, max_iter=15, use_normal_eqns=1)
    h1 = sysm.solve_cg(p, history=True)["history"]
    csr = yfd.Csr.from_arrays(ctx, sysm.n, sysm.n, rowptr, colidx, vals)
    h2 = csr.solve_cg(sysm.rhs(), p, history=True)["history"]
    assert np.allclose(h1, h2, rtol=1e-9)
    # from an initial guess (core/solvers.hpp:79-91: r0 = A^T (b - A x0)): x0 = 0 reproduces the cold start bit for bit, and
    # x0 = the cold solution starts from a residual already reduced by the cold solve's factor
    q = dict(rr_tol=1e-8, rr_max=1e4, max_iter=2 * sysm.n, use_normal_eqns=1)
    cold = sysm.solve_cg(yfd.CgParams(**q))
    warm0 = sysm.solve_cg(yfd.CgParams(**q), x0=np.zeros(sysm.n))
    assert warm0["iterations"] == cold["iterations"] and np.array_equal(warm0["x"], cold["x"])
    again = sysm.solve_cg(yfd.CgParams(rr_tol=1e-3, rr_max=1e4, max_iter=5, use_normal_eqns=1), x0=cold["x"])
    assert again["nr0"] <= 1e-6 * cold["nr0"]


def test_normal_equations_with_diagonal_iM(yfd, ctx):
    """conjugated_gradient(cgp, A, b, x, iM) with cgp.use_normal_eqns (core/solvers.hpp:185-227: r0 = A^T (b - A x),
    y = A^T A d, z = iM r) against a numpy restatement of exactly those lines; bicgstab with iM ignores the flag (:356-467)."""
    import scipy.sparse as sp
    mesh = yfd.Mesh(ctx, yfd.HostMesh.create(0, 2))
    sysm = ctx.dg_assemble_advection(mesh, 2, eta=1.0, upwinding=True)
    rowptr, colidx, vals = sysm.csr()
    n = sysm.n
    A = sp.csr_matrix((vals, colidx, rowptr), shape=(n, n))
    b = sysm.rhs()
    iM = 1.0 / (A.T @ A).diagonal()
    csr = yfd.Csr.from_arrays(ctx, n, n, rowptr, colidx, vals)
    res = csr.solve_cg(b, yfd.CgParams(rr_tol=1e-30, rr_max=1e30, max_iter=12, use_normal_eqns=1), pc_diag=iM, history=True)
    x = np.zeros(n); r = A.T @ (b - A @ x); d = iM * r; nr0 = np.linalg.norm(r); hist = [1.0]
    for _ in range(12):
        y = A.T @ (A @ d)
        rho = r @ (iM * r)
        alpha = rho / (d @ y)
        x = x + alpha * d; r = r - alpha * y
        d = iM * r + (r @ (iM * r)) / rho * d
        hist.append(np.linalg.norm(r) / nr0)
    assert np.allclose(res["history"], hist, rtol=1e-7)
    assert np.linalg.norm(res["x"] - x) <= 1e-7 * np.linalg.norm(x)
    a = csr.solve_bicgstab(b, yfd.CgParams(rr_tol=1e-8, rr_max=1e6, max_iter=200, use_normal_eqns=1), pc_diag=1.0 / A.diagonal())
    c = csr.solve_bicgstab(b, yfd.CgParams(rr_tol=1e-8, rr_max=1e6, max_iter=200), pc_diag=1.0 / A.diagonal())
    assert a["iterations"] == c["iterations"] and np.array_equal(a["x"], c["x"])
    with pytest.raises(yfd.YfdError):
        csr.solve_bicgstab(b, yfd.CgParams(max_iter=5, use_normal_eqns=1))

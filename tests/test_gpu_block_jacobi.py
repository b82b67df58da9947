"""GPU: the block-Jacobi preconditioned CG (YFD_PC_BLOCK_JACOBI) -- an EXTENSION (the reference has the
point-Jacobi iM only), so there is no reference run of it.  Checked against what does not depend on the preconditioner
(the reference's own solution of the same system), and against a numpy restatement of the reference's preconditioned
recurrences (core/solvers.hpp:196-227) with iM = blockdiag(A_cc)^-1 built by numpy."""
import numpy as np
import pytest
import scipy.sparse as sp

from conftest import host_mesh_from_ref

pytestmark = pytest.mark.gpu


def pcg_numpy(A, b, apply_m, rr_tol, max_iter):
    x = np.zeros_like(b); r = b.copy(); z = apply_m(r); d = z.copy()
    nr0 = nr = np.linalg.norm(r); it = 0; hist = [1.0]
    while nr / nr0 > rr_tol and it < max_iter:
        y = A @ d
        rho = r @ z
        alpha = rho / (d @ y)
        x += alpha * d; r -= alpha * y
        z = apply_m(r)
        d = z + (r @ z) / rho * d
        nr = np.linalg.norm(r); it += 1; hist.append(nr / nr0)
    return x, it, np.array(hist)


@pytest.mark.parametrize("t,r,k", [(0, 3, 1), (1, 3, 2), (0, 2, 3), (1, 3, 3), (0, 1, 4), (1, 1, 6)])
def test_block_jacobi_cg(yfd, ref, ctx, t, r, k):
    ref.set_cg(rr_tol=1e-12, rr_max=1e4)
    m = ref.mesh(t, r)
    out = ref.run_dg(k, eta=1.0, use_pc=True)
    mesh = yfd.Mesh(ctx, host_mesh_from_ref(yfd, m))
    sysm = ctx.dg_assemble_sip(mesh, k, eta=1.0, build_pc=True)
    n, bs = sysm.n, (k + 1) * (k + 2) // 2
    prm = yfd.CgParams(rr_tol=1e-12, rr_max=1e4, max_iter=4 * n)
    res = sysm.solve_cg(prm, use_pc=yfd.PC_BLOCK_JACOBI, history=True)
    assert res["converged"]
    # the solution does not depend on the preconditioner: the reference's own (point-Jacobi) solve is the yardstick
    assert np.linalg.norm(res["x"] - out["sol"]) <= 1e-10 * np.linalg.norm(out["sol"])
    # iterates: numpy restatement of the same recurrences with numpy's block inverses
    A = sp.csr_matrix((out["values"], out["colidx"], out["rowptr"]), shape=(n, n))
    inv = np.stack([np.linalg.inv(A[c * bs:(c + 1) * bs, c * bs:(c + 1) * bs].toarray()) for c in range(n // bs)])
    apply_m = lambda v: np.einsum("cij,cj->ci", inv, v.reshape(-1, bs)).reshape(-1)
    xs, its, hist = pcg_numpy(A, out["rhs"], apply_m, 1e-12, 4 * n)
    assert abs(res["iterations"] - its) <= max(2, 0.03 * its), (res["iterations"], its)
    m8 = min(8, len(hist), len(res["history"]))
    assert np.allclose(res["history"][:m8], hist[:m8], rtol=1e-6)
    # and it pays from k = 2 on: fewer iterations than the reference's point-Jacobi (k = 1: the scaled monomials of a
    # triangle are nearly orthogonal already, the two are on par; block size 1: they coincide)
    jac = sysm.solve_cg(prm, use_pc=yfd.PC_JACOBI)
    if k >= 2:
        assert res["iterations"] < jac["iterations"]
    else:
        assert res["iterations"] <= 1.1 * jac["iterations"]


def test_block_jacobi_large_and_device_resident(yfd, ctx):
    """sizes no CPU run fits in a test: the true residual of the returned solution and the iteration saving"""
    mesh = yfd.Mesh(ctx, yfd.HostMesh.structured(yfd.QUAD, 96))
    sysm = ctx.dg_assemble_sip(mesh, 3, eta=1.0, build_pc=True)
    prm = yfd.CgParams(rr_tol=1e-10, rr_max=1e4, max_iter=4 * sysm.n)
    blk = sysm.solve_cg(prm, use_pc=yfd.PC_BLOCK_JACOBI)
    none = sysm.solve_cg(prm)
    assert blk["converged"] and none["converged"]
    assert blk["iterations"] < 0.7 * none["iterations"]
    assert np.linalg.norm(blk["x"] - none["x"]) <= 1e-7 * np.linalg.norm(none["x"])
    rowptr, colidx, vals = sysm.csr()
    A = sp.csr_matrix((vals, colidx, rowptr), shape=(sysm.n, sysm.n))
    b = sysm.rhs()
    assert np.linalg.norm(b - A @ blk["x"]) <= 2e-10 * np.linalg.norm(b)
    dev = sysm.solve_cg_device(prm, use_pc=yfd.PC_BLOCK_JACOBI)
    assert dev["iterations"] == blk["iterations"]           # deterministic reductions
    # warm start at the solution: at most a couple of iterations
    again = sysm.solve_cg(yfd.CgParams(rr_tol=1e-10, rr_max=1e4, max_iter=50), use_pc=yfd.PC_BLOCK_JACOBI, x0=blk["x"])
    assert again["iterations"] <= 50


def test_block_jacobi_errors(yfd, ctx):
    tri = yfd.Mesh(ctx, yfd.HostMesh.create(yfd.TRI, 2))
    cf = ctx.cfem_assemble_p1(tri)
    with pytest.raises(yfd.YfdError):
        cf.solve_cg(use_pc=yfd.PC_BLOCK_JACOBI)
    dg = ctx.dg_assemble_sip(tri, 1, eta=1.0, build_pc=False)
    with pytest.raises(yfd.YfdError):
        dg.solve_cg(use_pc=7)
    with pytest.raises(yfd.YfdError):
        dg.solve_bicgstab(use_pc=yfd.PC_BLOCK_JACOBI)
    with pytest.raises(yfd.YfdError):
        dg.solve_cg(use_pc=yfd.PC_JACOBI)                    # assembled without the diagonal
    res = dg.solve_cg(yfd.CgParams(rr_tol=1e-10, rr_max=1e4, max_iter=4 * dg.n), use_pc=yfd.PC_BLOCK_JACOBI)
    assert res["converged"]                                  # no build_pc needed: the blocks come from the matrix
    # degree 0: the reference's penalty vanishes, its matrix has a zero diagonal (its own solve returns NaN); reported
    k0 = ctx.dg_assemble_sip(tri, 0, eta=1.0, build_pc=False)
    with pytest.raises(yfd.YfdError, match="singular diagonal block"):
        k0.solve_cg(use_pc=yfd.PC_BLOCK_JACOBI)

"""GPU: yfd_bicgstab_solve / yfd_system_solve_bicgstab against the reference's own bicgstab (core/solvers.hpp:246-467)
through the golden fixture tests/golden/bicgstab.npz (tools/make_golden_bicgstab.py), with the same acceptance rule as the
CPU restatement (tests/test_oracle_bicgstab.py explains why the iterates are pinned by the early history only)."""
import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spla

from conftest import golden
from test_oracle_bicgstab import SYSTEMS, check_against_reference, load

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", SYSTEMS)
@pytest.mark.parametrize("tag", ["plain", "jacobi"])
def test_device_bicgstab_vs_reference(yfd, ctx, name, tag):
    g, A, n = load(name)
    G = golden("bicgstab.npz")
    pc = 1.0 / A.diagonal() if tag == "jacobi" else None
    dA = yfd.Csr.from_arrays(ctx, n, n, g["rowptr"], g["colidx"], g["values"])
    res = dA.solve_bicgstab(g["rhs"], yfd.CgParams(rr_tol=1e-10, rr_max=1e8, max_iter=4 * n), pc_diag=pc, history=True)
    check_against_reference(res, G, name, tag, A, g["rhs"])


def test_device_bicgstab_follows_oracle_early(yfd, oracle, ctx):
    """first 8 residuals against the C restatement at full double precision (the golden file has 6 digits)"""
    g, A, n = load("adv_tri_r1_k2_up_shatter")
    dA = yfd.Csr.from_arrays(ctx, n, n, g["rowptr"], g["colidx"], g["values"])
    for pc in (None, 1.0 / A.diagonal()):
        o = oracle.bicgstab(g["rowptr"], g["colidx"], g["values"], g["rhs"], pc_diag=pc, rr_tol=1e-10, rr_max=1e8, max_iter=8)
        d = dA.solve_bicgstab(g["rhs"], yfd.CgParams(rr_tol=1e-10, rr_max=1e8, max_iter=8), pc_diag=pc, history=True)
        assert d["iterations"] == o["iterations"] == 8
        assert np.allclose(d["history"], o["history"], rtol=1e-8, atol=0)
        assert np.linalg.norm(d["x"] - o["x"]) <= 1e-8 * np.linalg.norm(o["x"])


def test_system_bicgstab_on_assembled_advection(yfd, ctx):
    """the matrix never leaves the device: upwind advection-reaction, Jacobi iM, solution against a direct solve and
    against the driver's manufactured solution"""
    mesh = yfd.Mesh(ctx, yfd.HostMesh.create(yfd.QUAD, 4))
    sysm = ctx.dg_assemble_advection(mesh, 2, eta=1.0, upwinding=True, build_pc=True)
    rowptr, colidx, vals = sysm.csr()
    A = sp.csr_matrix((vals, colidx, rowptr), shape=(sysm.n, sysm.n))
    xd = spla.spsolve(A.tocsc(), sysm.rhs())
    res = sysm.solve_bicgstab(yfd.CgParams(rr_tol=1e-10, rr_max=1e8, max_iter=4 * sysm.n), use_pc=True, history=True)
    assert res["converged"] and res["history"][-1] <= 1e-10
    assert np.linalg.norm(res["x"] - xd) <= 1e-7 * np.linalg.norm(xd)
    qp, _ = ctx.dg_postprocess(mesh, 2, res["x"], ref_fn=yfd.FN_SIN_X_SOL)
    assert np.sqrt(qp) < 5e-4
    # far fewer operator applications than CG on the normal equations needs for the same system
    cgn = sysm.solve_cg(yfd.CgParams(rr_tol=1e-10, rr_max=1e8, max_iter=40 * sysm.n, use_normal_eqns=1))
    assert res["iterations"] < cgn["iterations"]


def test_bicgstab_limits_and_errors(yfd, ctx):
    g, A, n = load("dg_quad_r1_k3")
    dA = yfd.Csr.from_arrays(ctx, n, n, g["rowptr"], g["colidx"], g["values"])
    res = dA.solve_bicgstab(g["rhs"], yfd.CgParams(rr_tol=1e-8, rr_max=1e8, max_iter=0), history=True)
    assert res["iterations"] == 0 and not res["x"].any() and list(res["history"]) == [1.0]
    xd = spla.spsolve(A.tocsc(), g["rhs"])
    res = dA.solve_bicgstab(g["rhs"], yfd.CgParams(rr_tol=1e-8, rr_max=1e8, max_iter=3), x0=xd)
    assert res["iterations"] <= 3 and np.linalg.norm(res["x"] - xd) <= 1e-6 * np.linalg.norm(xd)
    g2, A2, n2 = load("adv_tri_r2_k1")
    d2 = yfd.Csr.from_arrays(ctx, n2, n2, g2["rowptr"], g2["colidx"], g2["values"])
    res = d2.solve_bicgstab(g2["rhs"], yfd.CgParams(rr_tol=1e-12, rr_max=5.0, max_iter=4 * n2), history=True)
    assert res["iterations"] < 4 * n2 and res["history"][-1] >= 5.0          # rr_max stops a diverging run
    with pytest.raises(yfd.YfdError):
        dA.solve_bicgstab(g["rhs"], yfd.CgParams(use_normal_eqns=1))
    with pytest.raises(yfd.YfdError):
        dA.solve_bicgstab(np.zeros(3))

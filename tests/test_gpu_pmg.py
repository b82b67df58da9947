"""GPU: the p-multigrid preconditioned CG (YFD_PC_PMG, yfd_pmg.cu) -- an EXTENSION (the reference has the
point-Jacobi iM only), so there is no reference run of it.  Checked against what does not depend on the preconditioner
(the unmodified reference's own solution of the same system, 1e-10) and against the numpy restatement of the same V-cycle
(tests/pmg_numpy.py) inside the reference's preconditioned recurrences (core/solvers.hpp:196-227), run with the Chebyshev
bounds the device reports: same iteration count, first residuals to 1e-6."""
import numpy as np
import pytest
import scipy.sparse as sp

from conftest import host_mesh_from_ref
from pmg_numpy import PmgNumpy, pcg

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("t,r,k", [(1, 3, 3), (0, 3, 2), (1, 3, 1), (0, 2, 4), (1, 4, 2), (0, 3, 3)])
def test_pmg_cg_matches_reference_solution_and_numpy_restatement(yfd, ref, ctx, t, r, k):
    ref.set_cg(rr_tol=1e-12, rr_max=1e4)
    m = ref.mesh(t, r)
    out = ref.run_dg(k, eta=1.0, use_pc=True)
    mesh = yfd.Mesh(ctx, host_mesh_from_ref(yfd, m))
    sysm = ctx.dg_assemble_sip(mesh, k, eta=1.0, build_pc=True)
    assert sysm.operator_info()["uses_classes"]
    sysm.set_pmg_params(nu=2, coarse_degree=12)
    pi = sysm.pmg_info()
    assert pi["nlevels"] == k and pi["degree"] == list(range(k, 0, -1)) and pi["coarse_degree"] == 12
    assert all(1.0 < lm < 1.1 * (len(m["cells"][0]) + 1.5) for lm in pi["lambda_max"]), pi        # <= blocks per row
    n = sysm.n
    prm = yfd.CgParams(rr_tol=1e-12, rr_max=1e4, max_iter=4 * n)
    res = sysm.solve_cg(prm, use_pc=yfd.PC_PMG, history=True)
    assert res["converged"]
    assert np.linalg.norm(res["x"] - out["sol"]) <= 1e-10 * np.linalg.norm(out["sol"])
    assert sysm.true_residual(res["x"]) <= 1e-10
    rowptr, colidx, vals = sysm.csr()                               # the values the solver applies
    A = sp.csr_matrix((vals, colidx, rowptr), shape=(n, n))
    M = PmgNumpy(A, len(m["cells"]), k, pi["nu"], pi["coarse_degree"], pi["smooth_ratio"], pi["lambda_max"])
    xs, its, hist = pcg(A, sysm.rhs(), M, 1e-12, 4 * n)
    assert abs(res["iterations"] - its) <= max(2, 0.1 * its), (res["iterations"], its)
    m8 = min(8, len(hist), len(res["history"]))
    assert np.allclose(res["history"][:m8], hist[:m8], rtol=1e-6), (res["history"][:m8], hist[:m8])
    jac = sysm.solve_cg(prm, use_pc=yfd.PC_BLOCK_JACOBI)
    assert res["iterations"] < (0.5 if k > 1 else 1.0) * jac["iterations"], (res["iterations"], jac["iterations"])


def test_pmg_large_default_parameters(yfd, ctx):
    """147,456 quadrangles k=3 (1.5 M dofs): default parameters, against the plain CG of the same system."""
    mesh = yfd.Mesh(ctx, yfd.HostMesh.structured(yfd.QUAD, 384))
    sysm = ctx.dg_assemble_sip(mesh, 3)
    prm = yfd.CgParams(rr_tol=1e-10, rr_max=1e4, max_iter=4 * sysm.n)
    a = sysm.solve_cg(prm, use_pc=yfd.PC_PMG)
    b = sysm.solve_cg(prm)
    pi = sysm.pmg_info()
    assert a["converged"] and b["converged"]
    assert a["iterations"] * (4 * pi["nu"] + 2) < 0.2 * b["iterations"]          # fine-level operator applications
    assert np.linalg.norm(a["x"] - b["x"]) <= 1e-8 * np.linalg.norm(b["x"])
    assert sysm.true_residual(a["x"]) <= 5e-9
    dev = sysm.solve_cg_device(prm, use_pc=yfd.PC_PMG)
    assert dev["iterations"] == a["iterations"]                                  # deterministic
    c = sysm.solve_cg(yfd.CgParams(rr_tol=1e-10, rr_max=1e4, max_iter=4 * sysm.n, check_every=1), use_pc=yfd.PC_PMG)
    assert c["iterations"] == a["iterations"] and np.array_equal(c["x"], a["x"])


def test_pmg_errors(yfd, ref, ctx):
    m = ref.mesh(0, 3, shatter=True)
    sysm = ctx.dg_assemble_sip(yfd.Mesh(ctx, host_mesh_from_ref(yfd, m)), 2)
    with pytest.raises(yfd.YfdError):                               # no congruence classes on a shattered mesh
        sysm.solve_cg(yfd.CgParams(max_iter=5), use_pc=yfd.PC_PMG)
    sysm = ctx.dg_assemble_sip(yfd.Mesh(ctx, yfd.HostMesh.create(yfd.QUAD, 2)), 0)
    with pytest.raises(yfd.YfdError):                               # degree 0: no level below
        sysm.solve_cg(yfd.CgParams(max_iter=5), use_pc=yfd.PC_PMG)
    sysm = ctx.dg_assemble_sip(yfd.Mesh(ctx, yfd.HostMesh.create(yfd.QUAD, 2)), 2)
    with pytest.raises(yfd.YfdError):
        sysm.set_pmg_params(nu=-1)
    sysm.set_operator_mode(yfd.OPERATOR_STREAMED)
    with pytest.raises(yfd.YfdError):
        sysm.solve_cg(yfd.CgParams(max_iter=5), use_pc=yfd.PC_PMG)

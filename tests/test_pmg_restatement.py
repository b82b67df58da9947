"""CPU: the numpy restatement of the p-multigrid preconditioner (tests/pmg_numpy.py) on systems assembled by the C
oracle: the V-cycle is a symmetric positive definite linear operator (what the preconditioned CG of
core/solvers.hpp:141-243 requires), the preconditioned solve returns the solution of the system, and the outer
iteration count is far below the plain CG's."""
import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spla

from pmg_numpy import PmgNumpy, lambda_max_estimate, pcg, bs_of


def system(oracle, cell, r, k):
    m = oracle.mesh(cell, r)
    out = oracle.dg_sip(cell, m["points"], m["cells"], m["faces"], k, 1.0)
    n = len(out["rowptr"]) - 1
    return sp.csr_matrix((out["values"], out["colidx"], out["rowptr"]), shape=(n, n)), out["rhs"], len(m["cells"])


@pytest.mark.parametrize("cell,r,k", [(1, 1, 3), (0, 1, 2)])
def test_vcycle_is_symmetric_positive_definite(oracle, cell, r, k):
    A, b, nc = system(oracle, cell, r, k)
    M = PmgNumpy(A, nc, k, 2, 6, 4.0, lambda_max_estimate(A, nc, k))
    n = A.shape[0]
    Minv = np.column_stack([M(e) for e in np.eye(n)])
    assert np.abs(Minv - Minv.T).max() <= 1e-12 * np.abs(Minv).max()
    assert np.linalg.eigvalsh(0.5 * (Minv + Minv.T)).min() > 0


# (triangles k = 1 with the driver's default eta = 1 are left out: that matrix is INDEFINITE -- smallest eigenvalue
# -0.31 at -r 2: the penalty 3 k^2 eta / h_F is below the coercivity threshold -- so no SPD preconditioner theory applies;
# the reference's plain CG happens to converge on it regardless)
@pytest.mark.parametrize("cell,r,k", [(1, 3, 3), (0, 3, 2), (1, 3, 1), (1, 2, 4)])
def test_pmg_pcg_solves_the_system(oracle, cell, r, k):
    A, b, nc = system(oracle, cell, r, k)
    M = PmgNumpy(A, nc, k, 2, 20, 4.0, lambda_max_estimate(A, nc, k))
    x, its, hist = pcg(A, b, M, 1e-12, 500)
    want = spla.spsolve(A.tocsc(), b)
    assert np.linalg.norm(x - want) <= 1e-9 * np.linalg.norm(want)
    _, plain, _ = pcg(A, b, lambda v: v, 1e-12, 20000)
    assert its < 0.25 * plain, (its, plain)

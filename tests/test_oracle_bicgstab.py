"""CPU: the C restatement of bicgstab (oracle/yfd_oracle.c :: yfo_bicgstab) against the reference's own bicgstab
(core/solvers.hpp:246-467), run by tools/make_golden_bicgstab.py on systems the unmodified drivers assembled.

BiCGStab on these advection systems lets the residual grow by 3-4 orders before it falls (history of quad_r1_k3_up:
1, 1.7, 7.8, 3.6, 64, 28, 3199, ...), so rounding differences between two correct implementations are amplified by
the same factor: the iterates are pinned by the FIRST entries of the reference's history (6 printed digits), the
outcome by its convergence class, iteration count (loosely) and the distance of the solution from a direct solve."""
import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spla

from conftest import golden

SYSTEMS = ["adv_quad_r1_k3_up", "adv_quad_r2_k1_shatter", "adv_tri_r1_k2_up_shatter", "adv_tri_r2_k1", "dg_quad_r1_k3", "dg_tri_r1_k2_pc"]


def load(name):
    g = golden(name + ".npz")
    n = len(g["rhs"])
    A = sp.csr_matrix((g["values"], g["colidx"], g["rowptr"]), shape=(n, n))
    return g, A, n


def check_against_reference(res, G, name, tag, A, rhs, lead=5):
    """shared by the CPU (oracle) and GPU (device) tests"""
    h_ref, it_ref, x_ref = G["%s_%s_history" % (name, tag)], int(G["%s_%s_iters" % (name, tag)]), G["%s_%s_x" % (name, tag)]
    h = res["history"]
    assert len(h) == res["iterations"] + 1
    m = min(lead, len(h_ref), len(h))
    assert np.allclose(h[:m], h_ref[:m], rtol=2e-5, atol=0), (name, tag, h[:m], h_ref[:m])
    # outcome class: where the reference got the residual down, so must we (within 3 orders: the reference's restart
    # rule makes the tail stagnate, and where exactly a run crosses rr_tol is rounding-dependent); where it failed, we
    # must not claim convergence either
    if h_ref[-1] < 1e-3:
        assert abs(np.log10(h[-1]) - np.log10(h_ref[-1])) <= 3.0, (name, tag, h[-1], h_ref[-1])
        xd = spla.spsolve(A.tocsc(), rhs)
        e, e_ref = np.linalg.norm(res["x"] - xd), np.linalg.norm(x_ref - xd)
        assert e <= max(300 * e_ref, 1e-8 * np.linalg.norm(xd)), (name, tag, e, e_ref)
    else:
        assert h[-1] > 1e-6 and res["iterations"] == it_ref, (name, tag, h[-1], res["iterations"], it_ref)
    if h_ref[-1] <= 1e-10 and h[-1] <= 1e-10:
        assert 0.4 * it_ref <= res["iterations"] <= 2.5 * it_ref, (name, tag, res["iterations"], it_ref)


@pytest.mark.parametrize("name", SYSTEMS)
@pytest.mark.parametrize("tag", ["plain", "jacobi"])
def test_oracle_bicgstab_vs_reference(oracle, name, tag):
    g, A, n = load(name)
    G = golden("bicgstab.npz")
    pc = 1.0 / A.diagonal() if tag == "jacobi" else None
    res = oracle.bicgstab(g["rowptr"], g["colidx"], g["values"], g["rhs"], pc_diag=pc, rr_tol=1e-10, rr_max=1e8, max_iter=4 * n)
    check_against_reference(res, G, name, tag, A, g["rhs"])


def test_oracle_bicgstab_initial_guess_and_limits(oracle):
    g, A, n = load("dg_quad_r1_k3")
    xd = spla.spsolve(A.tocsc(), g["rhs"])
    # starting at the solution: nr0 is rounding noise, nr/nr0 = 1 > rr_tol, so the loop runs like the reference's does
    res = oracle.bicgstab(g["rowptr"], g["colidx"], g["values"], g["rhs"], rr_tol=1e-8, rr_max=1e8, max_iter=3, x0=xd)
    assert res["iterations"] <= 3
    res = oracle.bicgstab(g["rowptr"], g["colidx"], g["values"], g["rhs"], rr_tol=1e-8, rr_max=1e8, max_iter=0)
    assert res["iterations"] == 0 and not res["x"].any() and res["history"][0] == 1.0
    # rr_max stops a diverging run
    g2, A2, n2 = load("adv_tri_r2_k1")
    res = oracle.bicgstab(g2["rowptr"], g2["colidx"], g2["values"], g2["rhs"], rr_tol=1e-12, rr_max=5.0, max_iter=4 * n2)
    assert res["iterations"] < 4 * n2 and res["history"][-1] >= 5.0

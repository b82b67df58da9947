"""CPU: the plain-C Maxwell restatement (oracle/yfd_oracle_maxwell.inc) against the golden vectors written by
tools/make_golden_maxwell.py from the UNMODIFIED reference, and live against oracle/_ref where present.

Tolerances: everything downstream of M^-1 goes through an LU whose elimination order differs between the reference
(LAPACK dgesv via blaze::solve, closed forms for N <= 6) and the restatement, so values agree to cond(M) * eps:
1e-11 relative to the operator's max-norm covers k <= 4 (cond(M) <= 1.5e5)."""
import glob
import os

import numpy as np
import pytest

from conftest import GOLDEN

OP_TOL, SOL_TOL = 1e-11, 1e-10
FILES = sorted(f for f in glob.glob(os.path.join(GOLDEN, "maxwell_*.npz")) if "driver_logs" not in f)


def _case(path):
    g = np.load(path)
    return g, int(g["degree"]), bool(g["upwind"]), bool(g["euler"]), float(g["dt"]), int(g["steps"])


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f)[8:-4] for f in FILES])
def test_oracle_matches_reference_fixture(oracle, path):
    g, k, up, euler, dt, steps = _case(path)
    mu = g["mu_r"] if "mu_r" in g else None
    ep = g["eps_r"] if "eps_r" in g else None
    o = oracle.maxwell_assemble(g["points"], g["cells"], g["faces"], k, upwind=up, eta=float(g["eta"]), mu_r=mu, eps_r=ep)
    assert np.array_equal(o["neigh"], g["neigh"]) and np.array_equal(o["has"], g["has"])          # offdiag_neigh_offsets: bit-exact
    assert np.abs(o["gM"] - g["gM"]).max() <= 1e-14 * np.abs(g["gM"]).max()
    sc = np.abs(g["ondiag"]).max()
    assert np.abs(o["ondiag"] - g["ondiag"]).max() <= OP_TOL * sc
    assert np.abs(o["offdiag"] - g["offdiag"]).max() <= OP_TOL * sc
    d0 = oracle.maxwell_init(g["points"], g["cells"], k, o["gM"])
    assert np.linalg.norm(d0 - g["dofs0"]) <= SOL_TOL * np.linalg.norm(g["dofs0"])
    dN = oracle.maxwell_step(o, k, g["dofs0"], dt, steps, euler=euler)
    assert np.linalg.norm(dN - g["dofsN"]) <= SOL_TOL * np.linalg.norm(g["dofsN"])
    if "errors" in g:
        e = oracle.maxwell_errors(g["points"], g["cells"], k, g["dofsN"], (steps - 1) * dt)
        assert np.abs(e - g["errors"][-1]).max() <= 1e-12


def test_oracle_reproduces_driver_error_log(oracle):
    """run_maxwell_solver's error log (maxwell_driver.cpp:249-270) on the reference's own mesh at -r 2."""
    logs = np.load(os.path.join(GOLDEN, "maxwell_driver_logs.npz"))
    for name in ("tri_r2_k2", "quad_r2_k1_up"):
        t, r, k, up, dt, steps = logs[name + "_cfg"]
        m = oracle.mesh(int(t), int(r))
        o = oracle.maxwell_assemble(m["points"], m["cells"], m["faces"], int(k), upwind=bool(up))
        d = oracle.maxwell_init(m["points"], m["cells"], int(k), o["gM"])
        for cyc in range(int(steps)):
            d = oracle.maxwell_step(o, int(k), d, dt, 1)
            e = oracle.maxwell_errors(m["points"], m["cells"], int(k), d, cyc * dt)
            row = logs[name][cyc]
            assert row[0] == cyc
            assert np.allclose(e, row[2:5], rtol=2e-5, atol=0)          # the log holds 6 significant digits


def test_oracle_matches_live_reference():
    import oraclebind
    import refbind
    if not refbind.maxwell_available():
        pytest.skip("oracle/_ref/libyfd_ref_maxwell.so not built")
    O, R = oraclebind.Oracle(), refbind.RefMaxwell()
    for (t, r, k, up, sh) in [(0, 2, 3, 1, 1), (1, 2, 2, 0, 1), (0, 1, 4, 0, 0)]:
        R.create(t, r, k, dt=0.004, upwind=up, shatter=sh)
        m = R.mesh(); R.assemble(); op = R.operator(); R.init()
        o = O.maxwell_assemble(m["points"], m["cells"], m["faces"], k, upwind=up)
        sc = np.abs(op["ondiag"]).max()
        assert np.abs(o["ondiag"] - op["ondiag"]).max() <= OP_TOL * sc and np.abs(o["offdiag"] - op["offdiag"]).max() <= OP_TOL * sc
        d0 = R.dofs()
        R.step(5)
        d5 = O.maxwell_step(o, k, d0, 0.004, 5)
        assert np.linalg.norm(d5 - R.dofs()) <= SOL_TOL * np.linalg.norm(d5)

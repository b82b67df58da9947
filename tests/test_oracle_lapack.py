"""CPU: pin oracle/lu_shim.c against a REAL LAPACK.

The reference links a system LAPACK (CMakeLists.txt:32) this image does not have, so oracle/_ref gets dgesv_/dgetrf_/
dgetrs_ from the unblocked LU in oracle/lu_shim.c -- and everything downstream of an elemental M^-1 (the whole Maxwell
operator; the diffusion driver's "L2 error (mm)") was "parity unpinned against a real LAPACK" (SURVEY.md 8c).  The opencv
wheel in this image bundles an LP64 OpenBLAS 0.3.15 exporting those routines; `make -C oracle lapack_pin` links the same
unmodified reference sources against it.  Here the golden fixtures (written with the shim) are re-derived with that build:
agreement to cond(M) * eps closes the gap.  Skipped where the wheel (or /root/reference at build time) is absent."""
import glob
import os
import sys

import numpy as np
import pytest

from conftest import GOLDEN, ROOT

sys.path.insert(0, os.path.join(ROOT, "oracle"))
os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")
import refbind  # noqa: E402

FIXTURES = sorted(p for p in glob.glob(os.path.join(GOLDEN, "maxwell_*.npz")) if "driver_logs" not in p and "material" not in p)


@pytest.mark.skipif(not refbind.maxwell_available(lapack=True), reason="oracle/_ref/libyfd_ref_maxwell_lapack.so not built")
@pytest.mark.parametrize("path", FIXTURES, ids=[os.path.basename(p)[8:-4] for p in FIXTURES])
def test_maxwell_reference_with_real_lapack_reproduces_the_shim_fixtures(path):
    g = np.load(path)
    R = refbind.RefMaxwell(lapack=True)
    R.create(int(g["mtype"]), int(g["levels"]), int(g["degree"]), dt=float(g["dt"]), eta=float(g["eta"]), upwind=bool(g["upwind"]),
             euler=bool(g["euler"]), shatter=bool(g["shatter"]))
    assert np.array_equal(R.mesh()["points"], g["points"])
    R.assemble()
    op = R.operator()
    assert np.array_equal(op["neigh"], g["neigh"]) and np.array_equal(op["has"], g["has"])
    assert np.array_equal(op["gM"], g["gM"])                       # no LAPACK involved: bit-exact
    # M^-1-dependent blocks: two correct pivoted LUs agree to cond(M) * eps (cond(M) <= 1.4e5 at k = 4)
    scale = max(np.abs(g["ondiag"]).max(), np.abs(g["offdiag"]).max())
    assert np.abs(op["ondiag"] - g["ondiag"]).max() <= 1e-12 * scale              # measured: 0 up to bs = 6, 1.6e-14 at k = 4
    assert np.abs(op["offdiag"] - g["offdiag"]).max() <= 1e-12 * scale
    R.init()
    assert np.linalg.norm(R.dofs() - g["dofs0"]) <= 1e-11 * np.linalg.norm(g["dofs0"])
    R.step(1)
    assert np.linalg.norm(R.dofs() - g["dofs1"]) <= 1e-11 * np.linalg.norm(g["dofs1"])
    e0 = R.errors(0)
    R.step(int(g["steps"]) - 1)
    assert np.linalg.norm(R.dofs() - g["dofsN"]) <= 1e-11 * np.linalg.norm(g["dofsN"])
    eN = R.errors(int(g["steps"]) - 1)
    assert np.abs(np.array([e0, eN]) - g["errors"]).max() <= 1e-11


@pytest.mark.skipif(not (refbind.available(lapack=True) and refbind.available()), reason="oracle/_ref/libyfd_ref_lapack.so not built")
@pytest.mark.parametrize("t,r,k", [(0, 2, 1), (1, 2, 3), (0, 1, 4)])
def test_diffusion_postprocess_with_real_lapack_matches_the_shim(t, r, k):
    """dg2d_diffusion.cpp:225-268: L2_errsq_mm goes through blaze::solve_LU -> dgetrf_/dgetrs_; nothing else does"""
    outs = []
    for lapack in (False, True):
        R = refbind.Ref(lapack=lapack)
        R.mesh(t, r)
        R.set_cg(rr_tol=1e-12, rr_max=1e4, max_iter=-1)
        outs.append(R.run_dg(k, eta=1.0, use_pc=False))
    a, b = outs
    assert np.array_equal(a["values"], b["values"]) and np.array_equal(a["rhs"], b["rhs"]) and np.array_equal(a["sol"], b["sol"])
    assert a["l2_qp"] == b["l2_qp"]
    assert abs(a["l2_mm"] - b["l2_mm"]) <= 1e-9 * abs(b["l2_mm"])

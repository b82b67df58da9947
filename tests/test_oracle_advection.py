"""CPU: the C restatement of the advection-reaction loop and of CG on the normal equations (oracle/yfd_oracle.c,
yfo_set_physics / yfo_cg_normal) against golden vectors written from the UNMODIFIED dg2d_advection.cpp."""
import glob
import os

import numpy as np
import pytest

from conftest import GOLDEN

FILES = sorted(glob.glob(os.path.join(GOLDEN, "adv_*.npz")))


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f)[4:-4] for f in FILES])
def test_oracle_advection_matches_reference_fixture(oracle, path):
    g = np.load(path)
    k, up = int(g["degree"]), bool(g["upwinding"])
    a = oracle.advection(int(g["mtype"]), g["points"], g["cells"], g["faces"], k, eta=float(g["eta"]), upwinding=up)
    assert np.array_equal(a["rowptr"], g["rowptr"]) and np.array_equal(a["colidx"], g["colidx"])        # bit-exact pattern
    assert np.abs(a["values"] - g["values"]).max() <= 1e-13 * np.abs(g["values"]).max()
    assert np.abs(a["rhs"] - g["rhs"]).max() <= 1e-13 * np.abs(g["rhs"]).max()
    c = oracle.cg_normal(g["rowptr"], g["colidx"], g["values"], g["rhs"], rr_tol=1e-8, rr_max=1e4, max_iter=30)
    n = min(len(c["history"]), len(g["history30"]), 8)
    assert np.allclose(c["history"][:n], g["history30"][:n], rtol=1e-4)        # the history file holds 6 digits
    # the SIP physics is untouched by the switch
    s = oracle.dg_sip(int(g["mtype"]), g["points"], g["cells"], g["faces"], k)
    assert not np.array_equal(s["values"], a["values"])

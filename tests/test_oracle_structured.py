"""CPU: structured nx x ny meshes (sizes between the reference mesher's refinement levels) against fixtures the UNMODIFIED
reference produced on exactly those meshes (tools/make_golden_structured.py: the mesh is injected through the reference's
public mesh vectors; connectivity, DG-SIP system and CG solution are its own).  Pins (1) yfd_mesher_create_structured -- host
C++, runs without a GPU -- bit for bit, and (2) the C oracle on such meshes."""
import glob
import os

import numpy as np
import pytest

from conftest import GOLDEN, block_max_rel_err

FIXTURES = sorted(glob.glob(os.path.join(GOLDEN, "struct_*.npz")))


@pytest.mark.parametrize("path", FIXTURES, ids=[os.path.basename(p)[7:-4] for p in FIXTURES])
def test_structured_mesher_reproduces_the_fixture_mesh(yfd, path):
    g = np.load(path)
    hm = yfd.HostMesh.structured(int(g["mtype"]), int(g["nx"]), int(g["ny"]))
    assert np.array_equal(hm.points, g["points"]) and np.array_equal(hm.cells, g["cells"])
    assert np.array_equal(hm.faces, g["faces"]) and np.array_equal(hm.face_boundary, g["face_boundary"])
    assert np.array_equal(hm.face_boundary_id, g["face_boundary_id"])
    # the reference's conventions: cells and faces sorted (core/meshers.hpp:178,:360), faces with sorted endpoints
    assert (np.diff(hm.faces, axis=1) > 0).all()
    key = hm.faces[:, 0] * (len(hm.points) + 1) + hm.faces[:, 1]
    assert (np.diff(key) > 0).all()
    nx, ny = int(g["nx"]), int(g["ny"])
    assert hm.face_boundary.sum() == 2 * (nx + ny)


@pytest.mark.parametrize("path", FIXTURES, ids=[os.path.basename(p)[7:-4] for p in FIXTURES])
def test_oracle_on_structured_mesh_vs_reference(oracle, path):
    g = np.load(path)
    t, k, eta = int(g["mtype"]), int(g["degree"]), float(g["eta"])
    cf, ow = oracle.connectivity(g["cells"], g["faces"])
    assert np.array_equal(ow, g["face_owners"])
    d = oracle.dg_sip(t, g["points"], g["cells"], g["faces"], k, eta)
    assert np.array_equal(d["rowptr"], g["rowptr"]) and np.array_equal(d["colidx"], g["colidx"])
    bs = (k + 1) * (k + 2) // 2
    assert block_max_rel_err(d["values"], g["values"], g["rowptr"], bs) <= 1e-12
    assert np.abs(d["rhs"] - g["rhs"]).max() <= 1e-12 * np.abs(g["rhs"]).max()
    pc = d["pc_diag"] if int(g["use_pc"]) else None
    if pc is not None:
        assert np.abs(pc - g["pc_diag"]).max() <= 1e-12 * np.abs(g["pc_diag"]).max()
    res = oracle.cg(d["rowptr"], d["colidx"], d["values"], d["rhs"], pc_diag=pc, rr_tol=1e-12, rr_max=1e4, max_iter=2 * d["n"])
    assert np.linalg.norm(res["x"] - g["sol"]) <= 1e-10 * np.linalg.norm(g["sol"])
    assert abs(res["iterations"] - int(g["iters"])) <= max(3, 0.05 * int(g["iters"]))

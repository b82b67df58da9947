"""CPU: the C-ABI library loads and exports every symbol include/yfd.h declares; host-side logic (the mesher,
parameter defaults, error mapping) against the reference's golden vectors.  No compute calls: no GPU here."""
import hashlib
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, golden, golden_digests


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_library_exports_every_declared_symbol(yfd):
    L = yfd.lib()
    names = yfd.declared_symbols()
    assert len(names) >= 40
    for s in names:
        assert hasattr(L, s), "libyfd_b200.so does not export " + s
    # and nothing torch-typed in the signatures: plain pointers and sizes only
    hdr = open(os.path.join(ROOT, "include", "yfd.h")).read()
    code = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    assert "torch" not in code and "at::" not in code and "std::" not in code and 'extern "C"' in code


def test_every_entry_point_cites_the_reference(yfd):
    hdr = open(os.path.join(ROOT, "include", "yfd.h")).read()
    cites = re.findall(r"(?:core|methods)/[a-z_0-9]+\.hpp:\d+|dg2d_diffusion\.cpp:\d+|continuous_fem\.cpp:\d+", hdr)
    assert len(cites) >= 30


def test_library_is_cuda_only_and_does_not_link_the_oracle(yfd):
    out = subprocess.run(["ldd", yfd.lib_path()], stdout=subprocess.PIPE, text=True).stdout
    assert "libcudart" in out
    assert "oracle" not in out and "yfd_ref" not in out
    syms = subprocess.run(["nm", "-D", yfd.lib_path()], stdout=subprocess.PIPE, text=True).stdout
    assert "yfo_" not in syms and "yfd_ref_" not in syms


def test_no_device_fails_loudly(yfd):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(yfd.YfdError) as e:
        yfd.Context(0)
    assert e.value.status == 10            # YFD_ERR_NO_DEVICE: there is no CPU path


def test_status_strings_and_cg_defaults(yfd):
    L = yfd.lib()
    assert L.yfd_status_string(0) == b"ok" or L.yfd_status_string(0)
    seen = {L.yfd_status_string(i) for i in range(14)}
    assert len(seen) == 14
    p = yfd.CgParams()                     # conjugated_gradient_params() (core/solvers.hpp:35-42)
    assert (p.rr_tol, p.rr_max, p.max_iter, p.use_initial_guess, p.use_normal_eqns, p.save_iteration_history) == (1e-8, 20.0, 100, 0, 0, 0)
    with pytest.raises(TypeError):
        yfd.CgParams(bogus=1)


@pytest.mark.parametrize("name,t", [("tri", 0), ("quad", 1)])
def test_host_mesher_matches_reference_numbering(yfd, name, t):
    """yfd_mesher_create (host C++) vs mesher<Mesh>::create_mesh outputs of the reference (tests/golden)."""
    dig = golden_digests()
    for r in range(0, 7):
        m = yfd.HostMesh.create(t, r)
        assert dig["%s_r%d_sizes" % (name, r)] == "%d %d %d" % (len(m.points), len(m.cells), len(m.faces))
        assert sha(m.points) == dig["%s_r%d_points" % (name, r)]
        assert sha(m.cells) == dig["%s_r%d_cells" % (name, r)]
        assert sha(m.faces) == dig["%s_r%d_faces" % (name, r)]
        assert sha(m.face_boundary) == dig["%s_r%d_face_boundary" % (name, r)]
    g = golden("mesh_%s_r2.npz" % name)
    m = yfd.HostMesh.create(t, 2)
    assert np.array_equal(m.points, g["points"]) and np.array_equal(m.cells, g["cells"]) and np.array_equal(m.faces, g["faces"])


def test_host_mesher_vs_oracle_larger(yfd, oracle):
    for t in (0, 1):
        m, o = yfd.HostMesh.create(t, 7), oracle.mesh(t, 7)
        assert np.array_equal(m.points, o["points"]) and np.array_equal(m.cells, o["cells"])
        assert np.array_equal(m.faces, o["faces"]) and np.array_equal(m.face_boundary, o["face_boundary"])
    # size formulae of SURVEY.md section 8
    m = yfd.HostMesh.create(0, 6)
    assert len(m.cells) == 4 ** 7 and len(m.points) == 65 ** 2 + 64 ** 2 and len(m.faces) == (3 * 4 ** 7 + 4 * 64) // 2


def test_mesher_rejects_bad_arguments(yfd):
    with pytest.raises(yfd.YfdError):
        yfd.HostMesh.create(7, 1)
    with pytest.raises(yfd.YfdError):
        yfd.HostMesh.create(0, -1)


def test_structured_mesher_follows_the_reference_conventions():
    """yfd_mesher_create_structured (host only): sorted cells / faces, boundary flags, and -- where oracle/_ref is present --
    the reference's own compute_connectivity accepts the mesh and reproduces the same owners as the C restatement."""
    import yaourt_fem_dg_b200 as yfd
    import oraclebind
    for t, nx, ny in ((0, 5, 7), (1, 6, 4), (0, 1, 1), (1, 1, 3)):
        hm = yfd.HostMesh.structured(t, nx, ny)
        nv = 3 if t == 0 else 4
        assert hm.cells.shape == ((2 if t == 0 else 1) * nx * ny, nv)
        assert len(hm.points) == (nx + 1) * (ny + 1)
        assert (np.lexsort(hm.cells.T[::-1]) == np.arange(len(hm.cells))).all()
        assert (np.lexsort(hm.faces.T[::-1]) == np.arange(len(hm.faces))).all() and (hm.faces[:, 0] < hm.faces[:, 1]).all()
        assert hm.face_boundary.sum() == 2 * (nx + ny)
        cf, owners = oraclebind.Oracle().connectivity(hm.cells, hm.faces)
        assert ((owners[:, 1] < 0) == (hm.face_boundary == 1)).all()          # boundary faces are exactly the one-owner faces
        # positive orientation of every cell (counter-clockwise like the reference's base cells)
        p = hm.points[hm.cells]
        area2 = (p[:, 1, 0] - p[:, 0, 0]) * (p[:, 2, 1] - p[:, 0, 1]) - (p[:, 2, 0] - p[:, 0, 0]) * (p[:, 1, 1] - p[:, 0, 1])
        assert (area2 > 0).all()
        import refbind
        if refbind.available():
            m = refbind.Ref().set_mesh(t, hm.points, hm.cells, hm.faces, hm.face_boundary, hm.face_boundary_id)
            assert np.array_equal(m["face_owners"], owners) and np.array_equal(m["cells"], hm.cells)
    with pytest.raises(yfd.YfdError):
        yfd.HostMesh.structured(0, 0, 3)

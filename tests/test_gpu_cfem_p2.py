"""GPU: continuous P2 (yfd_cfem_assemble_p2) against the CPU statement in oracle/yfd_oracle_cfem_p2.inc.

PARITY UNPINNED against the reference, which is P1-only (methods/cfem.hpp:200-204): the comparison partner is the
oracle's restatement of the extension, itself pinned by closed forms and its convergence order
(tests/test_oracle_cfem_p2.py).  Index work is compared bit-exactly, values to 1e-12, solutions to 1e-10."""
import numpy as np
import pytest

from conftest import host_mesh_from_ref

pytestmark = pytest.mark.gpu


def oracle_mesh(oracle, yfd, ctx, r):
    m = oracle.mesh(0, r)
    return m, yfd.Mesh(ctx, host_mesh_from_ref(yfd, m))


@pytest.mark.parametrize("levels,rhs_fn", [(0, 1), (1, 1), (3, 1), (5, 1), (4, 2)])
def test_p2_system_matches_oracle(yfd, oracle, ctx, levels, rhs_fn):
    m, mesh = oracle_mesh(oracle, yfd, ctx, levels)
    o = oracle.cfem_p2(m["points"], m["cells"], m["faces"], m["face_boundary"], rhs_fn)
    sysm = ctx.cfem_assemble_p2(mesh, rhs_fn)
    assert sysm.n == o["n"] and sysm.nnz == o["nnz"]
    assert np.array_equal(sysm.compress_map(), o["compress_map"])
    rowptr, colidx, vals = sysm.csr()
    assert np.array_equal(rowptr, o["rowptr"]) and np.array_equal(colidx, o["colidx"])
    assert np.abs(vals - o["values"]).max() <= 1e-12 * np.abs(o["values"]).max()
    assert np.abs(sysm.rhs() - o["rhs"]).max() <= 1e-12 * np.abs(o["rhs"]).max()


def test_p2_solution_and_error(yfd, oracle, ctx):
    m, mesh = oracle_mesh(oracle, yfd, ctx, 4)
    o = oracle.cfem_p2(m["points"], m["cells"], m["faces"], m["face_boundary"])
    sysm = ctx.cfem_assemble_p2(mesh)
    res = sysm.solve_cg(yfd.CgParams(rr_tol=1e-13, rr_max=1e3, max_iter=4 * sysm.n))
    ref = oracle.cg(o["rowptr"], o["colidx"], o["values"], o["rhs"], rr_tol=1e-13, rr_max=1e3, max_iter=4 * o["n"])
    assert np.linalg.norm(res["x"] - ref["x"]) <= 1e-10 * np.linalg.norm(ref["x"])
    assert abs(res["iterations"] - ref["iterations"]) <= max(3, 0.05 * ref["iterations"])
    ex = sysm.expand(res["x"])
    cm = o["compress_map"]
    assert len(ex) == len(cm) and not ex[cm < 0].any() and np.array_equal(ex[cm >= 0], res["x"])
    err = ctx.cfem_l2_error(mesh, 2, ex)
    err_o = np.sqrt(oracle.cfem_p2_errsq(m["points"], m["cells"], m["faces"], ex, 3))
    assert abs(err - err_o) <= 1e-10 * err_o
    with pytest.raises(yfd.YfdError):
        ctx.cfem_l2_error(mesh, 3, ex)


def test_p2_third_order_on_structured_meshes(yfd, ctx):
    """O(h^3) in L2 on the arbitrary-size structured mesher, P1 beside it at O(h^2)."""
    errs = {1: [], 2: []}
    for nx in (12, 24, 48):
        mesh = yfd.Mesh(ctx, yfd.HostMesh.structured(yfd.TRI, nx))
        for order in (1, 2):
            sysm = ctx.cfem_assemble_p2(mesh) if order == 2 else ctx.cfem_assemble_p1(mesh)
            res = sysm.solve_cg(yfd.CgParams(rr_tol=1e-13, rr_max=1e3, max_iter=4 * sysm.n))
            errs[order].append(ctx.cfem_l2_error(mesh, order, sysm.expand(res["x"])))
            if order == 2:
                assert sysm.n == (2 * nx - 1) ** 2
    o1 = np.log2(np.array(errs[1][:-1]) / np.array(errs[1][1:]))
    o2 = np.log2(np.array(errs[2][:-1]) / np.array(errs[2][1:]))
    assert (o1 > 1.8).all() and (o1 < 2.2).all(), (errs[1], o1)
    assert (o2 > 2.85).all() and (o2 < 3.2).all(), (errs[2], o2)
    assert errs[2][-1] < errs[1][-1] / 50


def test_p2_rejects_quads_and_empty(yfd, ctx):
    quad = yfd.Mesh(ctx, yfd.HostMesh.structured(yfd.QUAD, 4))
    with pytest.raises(yfd.YfdError):
        ctx.cfem_assemble_p2(quad)

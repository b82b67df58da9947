"""GPU: the nx x ny structured mesher (yfd_mesher_create_structured, new -- the reference mesher only offers 4^(r+1) cells).
Parity is pinned by handing the SAME mesh to the unmodified reference through its public mesh vectors
(oracle/ref_harness.cpp: yfd_ref_mesh_set) and running its own driver loop on it."""
import numpy as np
import pytest

from conftest import block_max_rel_err

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mtype,nx,ny,k,pc", [(0, 5, 7, 2, False), (1, 6, 4, 3, False), (0, 8, 8, 1, True), (1, 9, 9, 2, True), (0, 6, 5, 3, False)])
def test_dg_sip_on_structured_mesh_matches_reference(yfd, ref, ctx, mtype, nx, ny, k, pc):
    hm = yfd.HostMesh.structured(mtype, nx, ny)
    m = ref.set_mesh(mtype, hm.points, hm.cells, hm.faces, hm.face_boundary, hm.face_boundary_id)
    ref.set_cg(rr_tol=1e-12, rr_max=1e4, max_iter=-1)
    out = ref.run_dg(k, eta=2.0, use_pc=pc)
    for reorder in (0, 1):
        mesh = yfd.Mesh(ctx, hm, reorder=reorder)
        assert np.array_equal(mesh.face_owners(), m["face_owners"])
        sysm = ctx.dg_assemble_sip(mesh, k, eta=2.0, build_pc=pc)
        rowptr, colidx, vals = sysm.csr()
        assert np.array_equal(rowptr, out["rowptr"]) and np.array_equal(colidx, out["colidx"])
        bs = (k + 1) * (k + 2) // 2
        assert block_max_rel_err(vals, out["values"], rowptr, bs) <= 1e-12
        assert np.abs(sysm.rhs() - out["rhs"]).max() <= 1e-12 * np.abs(out["rhs"]).max()
        res = sysm.solve_cg(yfd.CgParams(rr_tol=1e-12, rr_max=1e4, max_iter=2 * sysm.n), use_pc=pc)
        assert np.linalg.norm(res["x"] - out["sol"]) <= 1e-10 * np.linalg.norm(out["sol"])
        assert abs(res["iterations"] - out["iters"]) <= max(3, 0.05 * out["iters"])


def test_structured_sizes_and_orders(yfd, ctx):
    """configs[0] as written (64 x 64 triangles, k = 1) and the L2 order on non-power-of-two grids."""
    hm = yfd.HostMesh.structured(0, 64)
    assert len(hm.cells) == 2 * 64 * 64 and len(hm.points) == 65 * 65 and len(hm.faces) == 3 * 64 * 64 + 2 * 64
    errs = []
    for nx in (12, 24, 48):
        mesh = yfd.Mesh(ctx, yfd.HostMesh.structured(1, nx))
        sysm = ctx.dg_assemble_sip(mesh, 2, eta=1.0)
        res = sysm.solve_cg(yfd.CgParams(rr_tol=1e-12, rr_max=1e4, max_iter=2 * sysm.n))
        errs.append(np.sqrt(ctx.dg_postprocess(mesh, 2, res["x"])[0]))
    orders = np.log2(np.array(errs[:-1]) / np.array(errs[1:]))
    assert (orders > 2.6).all(), orders

"""GPU parity of the DG-SIP path against the UNMODIFIED reference (oracle/_ref), through the C ABI.

Bars (BASELINE.json north_star): sparsity pattern and dof maps bit-exact, assembled values within
1e-12 relative (to the block max-norm: zeros are stored), solution within 1e-10 relative L2.
"""
import numpy as np
import pytest

from conftest import host_mesh_from_ref, block_max_rel_err

pytestmark = pytest.mark.gpu

CASES = [(0, 3, 1), (0, 2, 2), (0, 3, 3), (0, 1, 4), (1, 3, 1), (1, 2, 2), (1, 2, 3), (0, 2, 0), (1, 1, 5), (0, 1, 6)]


def bs_of(k):
    return (k + 2) * (k + 1) // 2


@pytest.mark.parametrize("mtype,r", [(0, 0), (0, 3), (1, 0), (1, 3)])
@pytest.mark.parametrize("reorder", [0, 1])
def test_connectivity_matches_reference(yfd, ref, ctx, mtype, r, reorder):
    m = ref.mesh(mtype, r)
    mesh = yfd.Mesh(ctx, host_mesh_from_ref(yfd, m), reorder=reorder)
    assert np.array_equal(mesh.face_owners(), m["face_owners"])      # compute_connectivity, bit-exact
    perm = mesh.permutation()
    assert np.array_equal(np.sort(perm), np.arange(mesh.ncells))
    if reorder == 0:
        assert np.array_equal(perm, np.arange(mesh.ncells))
    # face_ids(): every cell edge must be the face with those two point ids
    cf = mesh.cell_faces()
    nv = m["cells"].shape[1]
    for e in range(nv):
        a = m["cells"][:, 0 if e == nv - 1 else e]
        b = m["cells"][:, nv - 1 if e == nv - 1 else e + 1]
        lo, hi = np.minimum(a, b), np.maximum(a, b)
        assert np.array_equal(m["faces"][cf[:, e], 0], lo) and np.array_equal(m["faces"][cf[:, e], 1], hi)


@pytest.mark.parametrize("mtype,r,k", CASES)
def test_local_blocks_match_reference_triplet_stream(yfd, ref, ctx, mtype, r, k):
    """K, Att, Atn per cell against the triplets the reference pushes, in its push order
    (dg2d_diffusion.cpp:195-200: per face Att [, Atn], then K)."""
    m = ref.mesh(mtype, r)
    ref.set_cg(skip_solve=True)
    out = ref.run_dg(k, eta=1.0, raw_triplets=True)
    ri, rj, rv = out["raw"]
    bs = bs_of(k)
    mesh = yfd.Mesh(ctx, host_mesh_from_ref(yfd, m))
    K, Att, Atn, rhs = ctx.dg_local_blocks(mesh, k, eta=1.0)
    owners = m["face_owners"]
    cf = mesh.cell_faces()
    pos = 0
    nf = m["cells"].shape[1]
    worst = 0.0
    for c in range(len(m["cells"])):
        for e in range(nf):
            blk = rv[pos:pos + bs * bs].reshape(bs, bs); pos += bs * bs
            worst = max(worst, np.abs(Att[c, e] - blk).max() / max(np.abs(blk).max(), 1e-300))
            o = owners[cf[c, e]]
            nb = o[1] if o[0] == c else o[0]
            if nb >= 0:
                assert ri[pos] == c * bs and rj[pos] == nb * bs
                blk = rv[pos:pos + bs * bs].reshape(bs, bs); pos += bs * bs
                worst = max(worst, np.abs(Atn[c, e] - blk).max() / max(np.abs(blk).max(), 1e-300))
            else:
                assert not Atn[c, e].any()
        blk = rv[pos:pos + bs * bs].reshape(bs, bs); pos += bs * bs
        worst = max(worst, np.abs(K[c] - blk).max() / max(np.abs(blk).max(), 1e-300))
    assert pos == len(rv)
    assert worst <= 1e-12, worst
    scale = np.abs(out["rhs"]).max()
    assert np.abs(rhs.reshape(-1) - out["rhs"]).max() <= 1e-12 * scale


@pytest.mark.parametrize("mtype,r,k", CASES)
@pytest.mark.parametrize("reorder", [0, 1])
def test_csr_pattern_bit_exact_values_1e12(yfd, ref, ctx, mtype, r, k, reorder):
    m = ref.mesh(mtype, r)
    ref.set_cg(skip_solve=True)
    out = ref.run_dg(k, eta=1.0, use_pc=True)
    mesh = yfd.Mesh(ctx, host_mesh_from_ref(yfd, m), reorder=reorder)
    sysm = ctx.dg_assemble_sip(mesh, k, eta=1.0, build_pc=True)
    assert sysm.n == out["n"] and sysm.nnz == out["nnz"]
    rowptr, colidx, vals = sysm.csr()
    assert np.array_equal(rowptr, out["rowptr"])
    assert np.array_equal(colidx, out["colidx"])
    assert block_max_rel_err(vals, out["values"], rowptr, bs_of(k)) <= 1e-12
    scale = np.abs(out["rhs"]).max()
    assert np.abs(sysm.rhs() - out["rhs"]).max() <= 1e-12 * scale
    pc, rpc = sysm.pc_diag(), out["pc_diag"]
    fin = np.isfinite(rpc)                     # k = 0: eta = 3 k^2 eta = 0, zero diagonal, 1/0 on both sides
    assert np.array_equal(np.isfinite(pc), fin)
    if fin.any():
        assert np.abs(pc[fin] - rpc[fin]).max() <= 1e-12 * np.abs(rpc[fin]).max()


def test_shattered_mesh(yfd, ref, ctx):
    """-S: perturbed interior points (core/mesh.hpp:532-580) as dumped by the reference."""
    for mtype, k in ((0, 2), (1, 2)):
        m = ref.mesh(mtype, 3, shatter=True)
        ref.set_cg(skip_solve=True)
        out = ref.run_dg(k, eta=1.0)
        mesh = yfd.Mesh(ctx, host_mesh_from_ref(yfd, m))
        sysm = ctx.dg_assemble_sip(mesh, k, eta=1.0)
        rowptr, colidx, vals = sysm.csr()
        assert np.array_equal(rowptr, out["rowptr"]) and np.array_equal(colidx, out["colidx"])
        assert block_max_rel_err(vals, out["values"], rowptr, bs_of(k)) <= 1e-12
        assert np.abs(sysm.rhs() - out["rhs"]).max() <= 1e-12 * np.abs(out["rhs"]).max()


@pytest.mark.parametrize("mtype,r,k,use_pc", [(0, 4, 1, False), (0, 4, 1, True), (1, 3, 2, False), (0, 3, 3, True), (1, 3, 3, False)])
def test_cg_solution_matches_reference(yfd, ref, ctx, mtype, r, k, use_pc):
    """Both sides at rr_tol = 1e-12 (1e-8 vs 1e-10 stops differ by up to 4e-10)."""
    m = ref.mesh(mtype, r)
    ref.set_cg(rr_tol=1e-12, rr_max=1e4, max_iter=-1)
    out = ref.run_dg(k, eta=1.0, use_pc=use_pc)
    mesh = yfd.Mesh(ctx, host_mesh_from_ref(yfd, m))
    sysm = ctx.dg_assemble_sip(mesh, k, eta=1.0, build_pc=use_pc)
    p = yfd.CgParams(rr_tol=1e-12, rr_max=1e4, max_iter=2 * sysm.n)
    res = sysm.solve_cg(p, use_pc=use_pc, history=True)
    rel = np.linalg.norm(res["x"] - out["sol"]) / np.linalg.norm(out["sol"])
    assert rel <= 1e-10, rel
    assert res["converged"]
    # same algorithm, different summation order: iteration counts agree to a few percent
    assert abs(res["iterations"] - out["iters"]) <= max(3, 0.05 * out["iters"])
    assert len(res["history"]) == res["iterations"] + 1 and res["history"][0] == 1.0
    # early history must track the reference's closely
    h = min(20, len(res["history"]), len(out["history"]))
    # (the reference writes its history file with 6 significant digits, core/solvers.hpp:109-110)
    assert np.allclose(res["history"][:h], out["history"][:h], rtol=2e-5)


def test_cg_default_params_and_stop_rules(yfd, ref, ctx):
    """rr_tol / max_iter / rr_max semantics of core/solvers.hpp:99; reference returns true even
    when not converged."""
    m = ref.mesh(0, 3)
    mesh = yfd.Mesh(ctx, host_mesh_from_ref(yfd, m))
    sysm = ctx.dg_assemble_sip(mesh, 1)
    p = yfd.CgParams()
    assert (p.rr_tol, p.rr_max, p.max_iter) == (1e-8, 20.0, 100)
    res = sysm.solve_cg(yfd.CgParams(max_iter=7), history=True)
    assert res["iterations"] == 7 and not res["converged"] and len(res["history"]) == 8
    ref.set_cg(rr_tol=1e-8, rr_max=1e4, max_iter=7)
    out = ref.run_dg(1)
    assert out["iters"] == 7
    assert np.linalg.norm(res["x"] - out["sol"]) <= 1e-10 * np.linalg.norm(out["sol"])
    assert np.allclose(res["history"], out["history"], rtol=2e-5)
    # check_every must not change the iteration count
    a = sysm.solve_cg(yfd.CgParams(rr_tol=1e-9, rr_max=1e4, max_iter=10000, check_every=1))
    b = sysm.solve_cg(yfd.CgParams(rr_tol=1e-9, rr_max=1e4, max_iter=10000, check_every=200))
    assert a["iterations"] == b["iterations"] and np.array_equal(a["x"], b["x"])
    # warm start (use_initial_guess, core/solvers.hpp:79-91): r0 = b - A x0, and the stop test is relative to
    # THAT residual, so starting from a converged x the initial residual norm is already tiny
    c = sysm.solve_cg(yfd.CgParams(rr_tol=1e-6, rr_max=1e4, max_iter=3), x0=a["x"])
    assert c["nr0"] <= 1e-8 * np.linalg.norm(sysm.rhs()) and c["iterations"] <= 3
    d = sysm.solve_cg(yfd.CgParams(rr_tol=1e-9, rr_max=1e4, max_iter=10000), x0=np.zeros(sysm.n))
    assert d["iterations"] == a["iterations"] and np.array_equal(d["x"], a["x"])


def test_full_size_properties_tri_k1_r6(yfd, ctx):
    """BASELINE config 1 (tri k=1 -r 6) at full size through size-independent properties:
    symmetry of A, A*1 structure, residual of the returned solution, L2 error order."""
    hm = yfd.HostMesh.create(yfd.TRI, 6)
    mesh = yfd.Mesh(ctx, hm)
    sysm = ctx.dg_assemble_sip(mesh, 1)
    assert sysm.n == 49152 and sysm.nnz == 587520
    rowptr, colidx, vals = sysm.csr()
    import scipy.sparse as sp
    A = sp.csr_matrix((vals, colidx, rowptr), shape=(sysm.n, sysm.n))
    assert abs(A - A.T).max() <= 1e-12 * abs(A).max()
    res = sysm.solve_cg(yfd.CgParams(rr_tol=1e-10, rr_max=1e4, max_iter=2 * sysm.n))
    assert res["converged"]
    b = sysm.rhs()
    assert np.linalg.norm(b - A @ res["x"]) <= 5e-10 * np.linalg.norm(b)


def test_owned_rows_entry_points_on_one_gpu(yfd, ctx):
    """yfd_system_solve_cg_owned / yfd_system_owned_cells and the Maxwell pair (new, SURVEY 8e: a rank reads back its own rows
    instead of the gathered vector): on one GPU "owned" is every cell, in the library's internal order -- the same numbers as the
    full entry points return in reference numbering, bit for bit.  The several-rank form is checked by tools/multi_gpu_check.py."""
    mesh = yfd.Mesh(ctx, yfd.HostMesh.create(yfd.QUAD, 3))
    sysm = ctx.dg_assemble_sip(mesh, 2, eta=1.0, build_pc=True)
    p = yfd.CgParams(rr_tol=1e-10, rr_max=1e4, max_iter=2 * sysm.n)
    full = sysm.solve_cg(p, use_pc=True)
    x_own = np.empty(sysm.local_sizes()["n_local"])
    res = sysm.solve_cg_owned_into(p, x_own, use_pc=True)
    ids = sysm.owned_cells()
    assert res["iterations"] == full["iterations"] and sorted(ids.tolist()) == list(range(mesh.ncells))
    bs = sysm.info.basis_size
    assert np.array_equal(x_own.reshape(-1, bs), full["x"].reshape(-1, bs)[ids])
    mx = yfd.Maxwell(ctx, mesh, 2, delta_t=0.002)
    mx.apply_initial_condition()
    mx.timesteps(3)
    ids = mx.owned_cells()
    assert sorted(ids.tolist()) == list(range(mesh.ncells))
    assert np.array_equal(mx.dofs_owned().reshape(-1, 3 * mx.bs), mx.dofs().reshape(-1, 3 * mx.bs)[ids])
    mx.close(); sysm.close(); mesh.close()

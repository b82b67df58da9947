"""GPU: the congruence-class operator cache (yfd_spmv_class.cu) -- one block row per class of congruent cells instead of
one per cell.  The reference has no such thing (it assembles and stores every cell's blocks, dg2d_diffusion.cpp:133-201),
so parity is the usual one: the values the solver APPLIES (yfd_system_get_csr exports the class representatives when the
cache is in use) against the unmodified reference's CSR within 1e-12 of the block-row max-norm, y = A x within 1e-12 of
|A||x|, CG solution within 1e-10 -- plus the class kernel against the streamed kernel on the same system.
"""
import os

import numpy as np
import pytest

from conftest import host_mesh_from_ref, block_max_rel_err

pytestmark = pytest.mark.gpu


def bs_of(k):
    return (k + 2) * (k + 1) // 2


def row_scaled_err(y, y_ref, A, x):
    scale = abs(A) @ np.abs(x)
    scale[scale == 0] = 1.0
    return float((np.abs(y - y_ref) / scale).max())


class env:
    def __init__(self, **kw):
        self.kw = {k: str(v) for k, v in kw.items()}

    def __enter__(self):
        self.old = {k: os.environ.get(k) for k in self.kw}
        os.environ.update(self.kw)

    def __exit__(self, *a):
        for k, v in self.old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


CASES = [(0, 4, 1), (0, 4, 2), (0, 3, 4), (1, 4, 3), (1, 3, 2), (0, 3, 3), (1, 3, 1), (0, 3, 0)]


@pytest.mark.parametrize("mtype,r,k", CASES)
def test_class_operator_matches_reference_and_streamed(yfd, ref, ctx, mtype, r, k):
    import scipy.sparse as sp
    m = ref.mesh(mtype, r)
    ref.set_cg(rr_tol=1e-12, rr_max=1e4, max_iter=-1)
    if k == 0:
        ref.set_cg(skip_solve=True)          # degree 0: eta = 3 k^2 eta = 0, singular (the reference returns NaN)
    out = ref.run_dg(k, eta=1.0, use_pc=k > 0)
    A = sp.csr_matrix((out["values"], out["colidx"], out["rowptr"]), shape=(out["n"], out["n"]))
    mesh = yfd.Mesh(ctx, host_mesh_from_ref(yfd, m))
    sysm = ctx.dg_assemble_sip(mesh, k, eta=1.0, build_pc=k > 0)
    oi = sysm.operator_info()
    assert oi["classes_available"] and oi["uses_classes"], oi
    assert 1 <= oi["nclasses"] <= 64 and oi["max_deviation"] <= oi["tolerance"] <= 1e-12, oi
    # the values the solver applies, in the reference's CSR: pattern bit-exact, values 1e-12
    rowptr, colidx, vals = sysm.csr()
    assert np.array_equal(rowptr, out["rowptr"]) and np.array_equal(colidx, out["colidx"])
    assert block_max_rel_err(vals, out["values"], rowptr, bs_of(k)) <= 1e-12
    x = np.random.default_rng(k + 10 * r).standard_normal(sysm.n)
    y_cls = sysm.apply(x)
    y_plain = sysm.apply(x, plain=True)        # class table expanded, one thread per row
    sysm.set_operator_mode(yfd.OPERATOR_STREAMED)
    assert not sysm.operator_info()["uses_classes"]
    y_str = sysm.apply(x)
    _, _, vals_str = sysm.csr()
    assert block_max_rel_err(vals_str, out["values"], rowptr, bs_of(k)) <= 1e-12
    sysm.set_operator_mode(yfd.OPERATOR_CLASSES)
    y_ref = A @ x
    for y in (y_cls, y_plain, y_str):
        assert row_scaled_err(y, y_ref, A, x) <= 1e-12
    if k == 0:
        return
    p = yfd.CgParams(rr_tol=1e-12, rr_max=1e4, max_iter=2 * sysm.n)
    for use_pc in (yfd.PC_NONE, yfd.PC_JACOBI, yfd.PC_BLOCK_JACOBI):
        res = sysm.solve_cg(p, use_pc=use_pc, history=True)
        assert res["converged"]
        assert np.linalg.norm(res["x"] - out["sol"]) <= 1e-10 * np.linalg.norm(out["sol"])
        assert sysm.true_residual(res["x"]) <= 1e-10       # recurrence residual 1e-12 + eps x iterations x |A||x| / |b|
        if use_pc == yfd.PC_JACOBI:                    # the reference ran with its Jacobi iM (-p)
            assert abs(res["iterations"] - out["iters"]) <= max(3, 0.05 * out["iters"])
            h = min(20, len(res["history"]), len(out["history"]))
            assert np.allclose(res["history"][:h], out["history"][:h], rtol=2e-5)


@pytest.mark.parametrize("cell,nx,ny,k", [(1, 37, 23, 3), (0, 19, 31, 2), (1, 40, 40, 1), (0, 16, 16, 4)])
def test_class_operator_structured_meshes(yfd, ref, ctx, cell, nx, ny, k):
    """nx x ny structured meshes (anisotropic cells, sizes that are not powers of two) fed to the unmodified reference."""
    import scipy.sparse as sp
    hm = yfd.HostMesh.structured(cell, nx, ny)
    ref.set_mesh(cell, hm.points, hm.cells, hm.faces, hm.face_boundary, hm.face_boundary_id)
    ref.set_cg(rr_tol=1e-12, rr_max=1e4, max_iter=-1)
    out = ref.run_dg(k, eta=1.0)
    mesh = yfd.Mesh(ctx, hm)
    sysm = ctx.dg_assemble_sip(mesh, k, eta=1.0)
    oi = sysm.operator_info()
    assert oi["uses_classes"] and oi["nclasses"] <= 40, oi
    rowptr, colidx, vals = sysm.csr()
    assert np.array_equal(rowptr, out["rowptr"]) and np.array_equal(colidx, out["colidx"])
    assert block_max_rel_err(vals, out["values"], rowptr, bs_of(k)) <= 1e-12
    res = sysm.solve_cg(yfd.CgParams(rr_tol=1e-12, rr_max=1e4, max_iter=2 * sysm.n))
    assert res["converged"] and np.linalg.norm(res["x"] - out["sol"]) <= 1e-10 * np.linalg.norm(out["sol"])
    assert abs(res["iterations"] - out["iters"]) <= max(3, 0.05 * out["iters"])


def test_shattered_mesh_keeps_the_streamed_kernel(yfd, ref, ctx):
    """-S: every cell is its own class; the cache must decline and the solve go through the streamed kernel."""
    m = ref.mesh(0, 4, shatter=True)
    ref.set_cg(rr_tol=1e-12, rr_max=1e4, max_iter=-1)
    out = ref.run_dg(2, eta=1.0)
    mesh = yfd.Mesh(ctx, host_mesh_from_ref(yfd, m))
    sysm = ctx.dg_assemble_sip(mesh, 2, eta=1.0)
    oi = sysm.operator_info()
    assert not oi["classes_available"] and not oi["uses_classes"] and oi["reject_reason"] in (1, 2), oi
    with pytest.raises(yfd.YfdError):
        sysm.set_operator_mode(yfd.OPERATOR_CLASSES)
    res = sysm.solve_cg(yfd.CgParams(rr_tol=1e-12, rr_max=1e4, max_iter=2 * sysm.n))
    # slivers: cond(A) x 1e-12 between two correct solves
    assert np.linalg.norm(res["x"] - out["sol"]) <= 1e-9 * np.linalg.norm(out["sol"])


def test_verification_pass_rejects_and_switch_disables(yfd, ctx):
    hm = yfd.HostMesh.create(yfd.QUAD, 4)
    mesh = yfd.Mesh(ctx, hm)
    with env(YFD_CLASS_TOL="1e-30"):           # nothing but bit-identical block rows would pass
        sysm = ctx.dg_assemble_sip(mesh, 2)
        oi = sysm.operator_info()
        assert (not oi["classes_available"] and oi["reject_reason"] == 3) or oi["max_deviation"] == 0.0, oi
    with env(YFD_DG_CLASSES="0"):
        sysm = ctx.dg_assemble_sip(mesh, 2)
        oi = sysm.operator_info()
        assert not oi["classes_available"] and oi["nclasses"] == 0
    # a mesh whose cells differ by 1e-9 of their size: same hash bucket, but not the same operator -> must be refused
    pts = hm.points.copy()
    rng = np.random.default_rng(0)
    interior = (pts[:, 0] > 0) & (pts[:, 0] < 1) & (pts[:, 1] > 0) & (pts[:, 1] < 1)
    pts[interior] += 1e-9 / 32 * rng.standard_normal((interior.sum(), 2))
    mesh2 = yfd.Mesh(ctx, yfd.HostMesh(hm.cell_type, pts, hm.cells, hm.faces, hm.face_boundary))
    sysm = ctx.dg_assemble_sip(mesh2, 2)
    oi = sysm.operator_info()
    assert not oi["uses_classes"] and oi["reject_reason"] == 3 and oi["max_deviation"] > oi["tolerance"], oi


def test_advection_system_has_no_classes(yfd, ctx):
    mesh = yfd.Mesh(ctx, yfd.HostMesh.create(yfd.TRI, 3))
    sysm = ctx.dg_assemble_advection(mesh, 1)
    assert not sysm.operator_info()["classes_available"]


def test_class_operator_large_mesh(yfd, ctx):
    """302,500 quadrangles k=3: class kernel against the streamed kernel and the plain kernel; the deviation of the
    class representatives from the per-cell blocks is the coordinate-rounding noise eps |x| / h."""
    hm = yfd.HostMesh.structured(yfd.QUAD, 550)
    mesh = yfd.Mesh(ctx, hm)
    sysm = ctx.dg_assemble_sip(mesh, 3)
    oi = sysm.operator_info()
    assert oi["uses_classes"] and oi["nclasses"] <= 40 and oi["max_deviation"] <= oi["tolerance"] <= 2e-11, oi
    rng = np.random.default_rng(4)
    x = rng.standard_normal(sysm.n)
    y_cls = sysm.apply(x)
    y_plain = sysm.apply(x, plain=True)
    sysm.set_operator_mode(yfd.OPERATOR_STREAMED)
    y_str = sysm.apply(x)
    scale = np.abs(y_str).max()
    assert np.abs(y_cls - y_plain).max() <= 1e-12 * scale            # same numbers, two kernels
    assert np.abs(y_cls - y_str).max() <= 50 * oi["tolerance"] * scale  # class representatives vs per-cell blocks
    sysm.set_operator_mode(yfd.OPERATOR_AUTO)
    p = yfd.CgParams(rr_tol=1e-10, rr_max=1e4, max_iter=2 * sysm.n)
    a = sysm.solve_cg(p)
    sysm.set_operator_mode(yfd.OPERATOR_STREAMED)
    b = sysm.solve_cg(p)
    assert a["converged"] and b["converged"] and abs(a["iterations"] - b["iterations"]) <= 0.02 * b["iterations"]
    assert np.linalg.norm(a["x"] - b["x"]) <= 1e-8 * np.linalg.norm(b["x"])
    # the recurrence residual (1e-10) drifts from the true one by ~eps x iterations x |A||x| / |b|: 4e-9 at this size, in BOTH modes
    tb = sysm.true_residual(b["x"])
    sysm.set_operator_mode(yfd.OPERATOR_AUTO)
    ta = sysm.true_residual(a["x"])
    assert ta <= 2e-8 and tb <= 2e-8 and abs(ta - tb) <= 0.2 * tb


@pytest.mark.parametrize("mtype,r,k", [(0, 4, 2), (0, 4, 3), (0, 3, 4), (1, 4, 2), (1, 4, 3), (1, 3, 4), (0, 6, 3)])
def test_tensor_pipe_class_kernel_equals_the_cuda_core_one(yfd, ctx, mtype, r, k):
    """dg_spmv_class_mma (mma.sync m8n8k4 f64, yfd_class_mma.cuh) against dg_spmv_class on the same class cache: the same
    products summed slot by slot, column chunks in a different order -- y agrees to rounding (1e-14 of |A||x| per row), the
    fused d.y too, and a CG solve takes the same number of iterations (+-1) to the same solution.  Both the shared-memory
    (few classes) and the read-through-L1 variant (> 108 KB of fragments: tri k = 3, 4) are covered; ragged tiles come from the
    boundary classes."""
    hm = yfd.HostMesh.create(yfd.TRI if mtype == 0 else yfd.QUAD, r)
    mesh = yfd.Mesh(ctx, hm)
    sysm = ctx.dg_assemble_sip(mesh, k, eta=1.0)
    oi = sysm.operator_info()
    assert oi["uses_classes"], oi
    x = np.random.default_rng(7).standard_normal(sysm.n)
    res = {}
    p = yfd.CgParams(rr_tol=1e-11, rr_max=1e4, max_iter=2 * sysm.n)
    for mode in (0, 2):
        with env(YFD_CLS_MMA=mode):
            res[mode] = (sysm.apply(x), sysm.solve_cg(p))
    y_plain = sysm.apply(x, plain=True)
    scale = np.abs(y_plain).max()
    assert np.abs(res[2][0] - res[0][0]).max() <= 1e-14 * scale
    assert np.abs(res[2][0] - y_plain).max() <= 1e-14 * scale
    a, b = res[0][1], res[2][1]
    assert a["converged"] and b["converged"]
    assert abs(a["iterations"] - b["iterations"]) <= max(1, a["iterations"] // 200), (a["iterations"], b["iterations"])
    assert np.linalg.norm(a["x"] - b["x"]) <= 1e-9 * np.linalg.norm(a["x"])
    sysm.close(); mesh.close()


def test_class_kernels_launch_on_every_small_table_size(yfd, ctx):
    """Tiny meshes give every class count from 1 up, i.e. class tables of every size around the 48 KiB a kernel gets without
    opting in to more dynamic shared memory -- which its static shared memory shares: 6 classes of tri k=4 fragments are exactly
    48 KiB and did not launch (found by tools/multi_gpu_check.py at 2 ranks).  Every solve must run and converge."""
    seen = set()
    for k in (4, 3):
        for nx in range(1, 5):
            for ny in range(nx, 6):
                for cell in (yfd.TRI, yfd.QUAD):
                    mesh = yfd.Mesh(ctx, yfd.HostMesh.structured(cell, nx, ny))
                    sysm = ctx.dg_assemble_sip(mesh, k, eta=1.0, build_pc=True)
                    oi = sysm.operator_info()
                    for mode in (0, 2):
                        with env(YFD_CLS_MMA=mode):
                            res = sysm.solve_cg(yfd.CgParams(rr_tol=1e-10, rr_max=1e6, max_iter=20 * sysm.n), use_pc=True)
                            assert res["converged"], (cell, nx, ny, k, mode, oi)
                    seen.add((cell, k, oi["nclasses"]))
                    sysm.close(); mesh.close()
    # (the exact 48 KiB case needs a rank-local half of a 5 x 7 triangle mesh: tools/multi_gpu_check.py at 2 ranks runs it)
    assert len({n for c, kk, n in seen}) >= 4, sorted(seen)

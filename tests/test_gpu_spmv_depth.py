"""GPU parity of the streamed SpMV kernels WHERE THEY RUN IN THE HEADLINE: rings that refill and flip their
mbarrier phase many times per warp (yfd_spmv_tma.cu `issue(k + NS)` / parity `((k+1)/NS)&1`, yfd_spmv_csr.cu
likewise).  On the small parity meshes a warp of the production launch owns <= 2 chunks, so that path never ran
in a parity test; here the launch is squeezed (YFD_SPMV_CTAS / YFD_SPMV_WARPS / YFD_SPMV_STAGES) so that every
warp walks >= 8 chunks, and large meshes (>= 260k cells) are checked against scipy and against the independent
one-thread-per-row kernel (yfd_system_apply with YFD_APPLY_PLAIN).

Oracle: the UNMODIFIED reference (oracle/_ref) -- its CSR arrays, its rhs, its CG solution.
Bars: y = A x within 1e-12 of |A||x| row-wise, solution within 1e-10 relative L2 (BASELINE north_star).
"""
import os

import numpy as np
import pytest

from conftest import host_mesh_from_ref, block_max_rel_err

pytestmark = pytest.mark.gpu


def bs_of(k):
    return (k + 2) * (k + 1) // 2


class squeezed:
    """Few CTAs of one warp with a ring of `stages`: many chunks per warp on a small mesh."""

    def __init__(self, stages, ctas=2, warps=1):
        self.env = dict(YFD_SPMV_STAGES=str(stages), YFD_SPMV_CTAS=str(ctas), YFD_SPMV_WARPS=str(warps))

    def __enter__(self):
        self.old = {k: os.environ.get(k) for k in self.env}
        os.environ.update(self.env)

    def __exit__(self, *a):
        for k, v in self.old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def row_scaled_err(y, y_ref, A, x):
    """max_i |y_i - yref_i| / (|A| |x|)_i : the rounding bound of a row dot product is eps * nnz_row * (|A||x|)_i"""
    import scipy.sparse as sp
    scale = abs(A) @ np.abs(x)
    scale[scale == 0] = 1.0
    return float((np.abs(y - y_ref) / scale).max())


DEEP = [(0, 4, 1, False), (1, 3, 3, False), (0, 3, 4, False), (0, 3, 2, True), (1, 3, 2, True)]


@pytest.mark.parametrize("stages", [2, 3, 4])
@pytest.mark.parametrize("mtype,r,k,shatter", DEEP)
def test_dg_spmv_deep_ring_matches_reference(yfd, ref, ctx, mtype, r, k, shatter, stages):
    import scipy.sparse as sp
    m = ref.mesh(mtype, r, shatter=shatter)
    ref.set_cg(rr_tol=1e-12, rr_max=1e4, max_iter=-1)
    out = ref.run_dg(k, eta=1.0)
    A = sp.csr_matrix((out["values"], out["colidx"], out["rowptr"]), shape=(out["n"], out["n"]))
    mesh = yfd.Mesh(ctx, host_mesh_from_ref(yfd, m))
    sysm = ctx.dg_assemble_sip(mesh, k, eta=1.0)
    sysm.set_operator_mode(yfd.OPERATOR_STREAMED)                    # this file pins the STREAMED kernels (classes: test_gpu_classes.py)
    bs, ncells = bs_of(k), len(m["cells"])
    cells_per_chunk = max(1, 32 // bs)
    chunks_per_warp = -(-ncells // cells_per_chunk) / 2.0
    assert chunks_per_warp >= 8 * stages / 2, chunks_per_warp       # every stage refilled >= 4 times
    rng = np.random.default_rng(11 + stages)
    x = rng.standard_normal(sysm.n)
    y_prod = sysm.apply(x)                                           # the production launch (<= 2 chunks per warp)
    with squeezed(stages):
        y_deep = sysm.apply(x)
        res = sysm.solve_cg(yfd.CgParams(rr_tol=1e-12, rr_max=1e4, max_iter=2 * sysm.n), history=True)
    y_plain = sysm.apply(x, plain=True)
    y_ref = A @ x
    assert row_scaled_err(y_deep, y_ref, A, x) <= 1e-12
    assert row_scaled_err(y_plain, y_ref, A, x) <= 1e-12
    # same row sums in the same order whatever the launch shape: identical bits
    assert np.array_equal(y_deep, y_prod)
    rel = np.linalg.norm(res["x"] - out["sol"]) / np.linalg.norm(out["sol"])
    # shattered meshes are badly conditioned (slivers): two correct CG runs to rr 1e-12 differ by cond(A) x 1e-12
    # (and CG may hit max_iter = 2N before rr 1e-12 on them, as the reference's own run does)
    assert (res["converged"] or shatter) and rel <= (1e-9 if shatter else 1e-10), (rel, res["converged"], res["iterations"], out["iters"])
    assert abs(res["iterations"] - out["iters"]) <= max(3, 0.05 * out["iters"])
    h = min(20, len(res["history"]), len(out["history"]))
    assert np.allclose(res["history"][:h], out["history"][:h], rtol=2e-5)
    # recurrence residual 1e-12; the true one sits eps x iterations x |A||x| / |b| above it (slivers: more)
    assert sysm.true_residual(res["x"]) <= (1e-9 if shatter else 1e-10)


@pytest.mark.parametrize("stages", [2, 4])
def test_dg_deep_ring_jacobi_and_block_jacobi(yfd, ref, ctx, stages):
    """The preconditioned recurrences (core/solvers.hpp:141-243) and the block-Jacobi extension over deep rings."""
    m = ref.mesh(1, 3)
    ref.set_cg(rr_tol=1e-12, rr_max=1e4, max_iter=-1)
    out = ref.run_dg(3, eta=1.0, use_pc=True)
    mesh = yfd.Mesh(ctx, host_mesh_from_ref(yfd, m))
    sysm = ctx.dg_assemble_sip(mesh, 3, eta=1.0, build_pc=True)
    sysm.set_operator_mode(yfd.OPERATOR_STREAMED)
    p = yfd.CgParams(rr_tol=1e-12, rr_max=1e4, max_iter=2 * sysm.n)
    with squeezed(stages):
        a = sysm.solve_cg(p, use_pc=yfd.PC_JACOBI)
        b = sysm.solve_cg(p, use_pc=yfd.PC_BLOCK_JACOBI)
    for res in (a, b):
        assert res["converged"]
        assert np.linalg.norm(res["x"] - out["sol"]) <= 1e-10 * np.linalg.norm(out["sol"])
    assert abs(a["iterations"] - out["iters"]) <= max(3, 0.05 * out["iters"])


@pytest.mark.parametrize("stages", [2, 3, 4])
def test_csr_spmv_deep_ring_cfem(yfd, ref, ctx, stages):
    """csr_spmv_tma on the reference's own CFEM system (continuous_fem.cpp), 4 CTAs x 1 warp."""
    import scipy.sparse as sp
    ref.set_cg(rr_tol=1e-12, rr_max=1e3)
    out = ref.run_cfem(5)
    m = ref.mesh(0, 5)
    mesh = yfd.Mesh(ctx, host_mesh_from_ref(yfd, m))
    sysm = ctx.cfem_assemble_p1(mesh)
    assert sysm.n == out["n"]
    A = sp.csr_matrix((out["values"], out["colidx"], out["rowptr"]), shape=(out["n"], out["n"]))
    assert -(-sysm.n // 32) / 4.0 >= 8
    x = np.random.default_rng(3).standard_normal(sysm.n)
    y_prod = sysm.apply(x)
    with squeezed(stages, ctas=4):
        y_deep = sysm.apply(x)
        res = sysm.solve_cg(yfd.CgParams(rr_tol=1e-12, rr_max=1e3, max_iter=4 * sysm.n))
    y_ref = A @ x
    assert row_scaled_err(y_deep, y_ref, A, x) <= 1e-12
    assert row_scaled_err(sysm.apply(x, plain=True), y_ref, A, x) <= 1e-12
    assert np.array_equal(y_deep, y_prod)
    assert np.linalg.norm(res["x"] - out["sol"]) <= 1e-10 * np.linalg.norm(out["sol"])
    assert sysm.true_residual(res["x"]) <= 1e-10       # recurrence residual 1e-12; the true one sits eps x iterations x |A||x| / |b| above it


def test_csr_spmv_deep_ring_generic_matrix(yfd, ctx):
    """Any CSR through the drop-in shape: ragged rows incl. empty ones, integer values (sums exact in any order)."""
    import scipy.sparse as sp
    rng = np.random.default_rng(5)
    n, cnt = 20011, 400_000
    rows = rng.integers(0, n, cnt); cols = rng.integers(0, n, cnt)
    rows[rows % 97 == 0] = 1                                         # a long row and many empty ones
    vals = rng.integers(-8, 9, cnt).astype(np.float64)
    S = sp.coo_matrix((vals, (rows, cols)), shape=(n, n)).tocsr()
    S.sort_indices()
    A = yfd.Csr.from_arrays(ctx, n, n, S.indptr, S.indices, S.data)
    x = rng.integers(-4, 5, n).astype(np.float64)
    want = S @ x
    for stages in (2, 3, 4):
        with squeezed(stages, ctas=4):
            assert np.array_equal(A.spmv(x), want)
    assert np.array_equal(A.spmv(x), want)


LARGE = [("quad", 1, dict(nx=550), 3), ("tri", 0, dict(r=8), 2)]


@pytest.mark.parametrize("name,cell,how,k", LARGE)
def test_large_mesh_spmv_vs_scipy_and_plain_kernel(yfd, ctx, name, cell, how, k):
    """>= 260k cells: the production launch walks ~50-90 chunks per warp.  y = A x for a random x: the solver's
    kernel against scipy on the exported CSR (<= 1e-13 relative to |A||x|) and against the independent kernel."""
    import scipy.sparse as sp
    hm = yfd.HostMesh.structured(cell, how["nx"]) if "nx" in how else yfd.HostMesh.create(cell, how["r"])
    mesh = yfd.Mesh(ctx, hm)
    sysm = ctx.dg_assemble_sip(mesh, k, eta=1.0)
    sysm.set_operator_mode(yfd.OPERATOR_STREAMED)
    assert len(hm.cells) >= 260_000
    rowptr, colidx, vals = sysm.csr()
    A = sp.csr_matrix((vals, colidx.astype(np.int32), rowptr), shape=(sysm.n, sysm.n))
    del colidx, vals
    rng = np.random.default_rng(17)
    x = rng.standard_normal(sysm.n)
    y = sysm.apply(x)
    y_plain = sysm.apply(x, plain=True)
    y_ref = A @ x
    assert row_scaled_err(y, y_ref, A, x) <= 1e-13
    assert row_scaled_err(y_plain, y_ref, A, x) <= 1e-13
    # symmetry of the assembled operator through the kernel itself: x.(A z) = z.(A x)
    z = rng.standard_normal(sysm.n)
    assert abs(x @ sysm.apply(z) - z @ y) <= 1e-11 * np.linalg.norm(x) * np.linalg.norm(y)
    # a short solve: the true residual (independent kernel) follows the recurrence residual
    res = sysm.solve_cg(yfd.CgParams(rr_tol=1e-6, rr_max=1e4, max_iter=2 * sysm.n))
    assert res["converged"]
    tr = sysm.true_residual(res["x"])
    assert tr <= 2e-6 and abs(tr - res["nr"] / res["nr0"]) <= 1e-2 * tr


def test_config0_full_size_against_live_reference(yfd, ref, ctx):
    """BASELINE configs[0] at full size (tri k=1 -r 6: 16,384 cells, 49,152 dofs, nnz 587,520) against the
    unmodified reference run here: pattern bit-exact, values 1e-12, solution 1e-10, L2 error of the driver."""
    m = ref.mesh(0, 6)
    ref.set_cg(rr_tol=1e-12, rr_max=1e4, max_iter=-1)
    out = ref.run_dg(1, eta=1.0)
    mesh = yfd.Mesh(ctx, host_mesh_from_ref(yfd, m))
    sysm = ctx.dg_assemble_sip(mesh, 1, eta=1.0)
    sysm.set_operator_mode(yfd.OPERATOR_STREAMED)
    assert sysm.n == 49152 and sysm.nnz == 587520 == out["nnz"]
    rowptr, colidx, vals = sysm.csr()
    assert np.array_equal(rowptr, out["rowptr"]) and np.array_equal(colidx, out["colidx"])
    assert block_max_rel_err(vals, out["values"], rowptr, 3) <= 1e-12
    assert np.abs(sysm.rhs() - out["rhs"]).max() <= 1e-12 * np.abs(out["rhs"]).max()
    res = sysm.solve_cg(yfd.CgParams(rr_tol=1e-12, rr_max=1e4, max_iter=2 * sysm.n))
    assert res["converged"]
    assert np.linalg.norm(res["x"] - out["sol"]) <= 1e-10 * np.linalg.norm(out["sol"])
    assert abs(res["iterations"] - out["iters"]) <= max(3, 0.05 * out["iters"])
    assert sysm.true_residual(res["x"]) <= 1e-10       # recurrence residual 1e-12; the true one sits eps x iterations x |A||x| / |b| above it

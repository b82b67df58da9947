"""GPU parity: triplet->CSR (init_from_triplets), generic CSR SpMV / CG, and the P1 CFEM path."""
import numpy as np
import pytest

from conftest import host_mesh_from_ref

pytestmark = pytest.mark.gpu


def test_init_from_triplets_matches_reference(yfd, ref, ctx):
    """Feed the reference's own raw triplet stream (DG tri k=1 r=3) through the sort-based build."""
    ref.mesh(0, 3)
    ref.set_cg(skip_solve=True)
    out = ref.run_dg(1, raw_triplets=True)
    ri, rj, rv = out["raw"]
    A = yfd.Csr.from_triplets(ctx, out["n"], out["n"], ri, rj, rv)
    rowptr, colidx, vals = A.get()
    assert np.array_equal(rowptr, out["rowptr"]) and np.array_equal(colidx, out["colidx"])
    scale = np.abs(out["values"]).max()
    assert np.abs(vals - out["values"]).max() <= 1e-12 * scale


def test_triplets_edge_cases(yfd, ctx):
    # empty
    A = yfd.Csr.from_triplets(ctx, 5, 5, [], [], [])
    rp, ci, v = A.get()
    assert np.array_equal(rp, np.zeros(6, np.int64)) and len(ci) == 0
    # ragged: empty rows, duplicates, explicit zeros kept, unsorted input
    rows = [4, 0, 4, 2, 0, 4, 2]
    cols = [1, 3, 1, 2, 0, 0, 2]
    vals = [1.0, 2.0, 3.0, 0.0, 5.0, 6.0, 0.0]
    A = yfd.Csr.from_triplets(ctx, 6, 4, rows, cols, vals)
    rp, ci, v = A.get()
    assert rp.tolist() == [0, 2, 2, 3, 3, 5, 5]
    assert ci.tolist() == [0, 3, 2, 0, 1]
    assert v.tolist() == [5.0, 2.0, 0.0, 6.0, 4.0]
    with pytest.raises(yfd.YfdError):
        yfd.Csr.from_triplets(ctx, 3, 3, [3], [0], [1.0])


def test_triplets_random_large_vs_scipy(yfd, ctx):
    import scipy.sparse as sp
    rng = np.random.default_rng(7)
    n, cnt = 20011, 1_500_000
    rows = rng.integers(0, n, cnt); cols = rng.integers(0, n, cnt)
    vals = rng.integers(-8, 9, cnt).astype(np.float64)      # integer values: sums are exact in any order
    A = yfd.Csr.from_triplets(ctx, n, n, rows, cols, vals)
    rp, ci, v = A.get()
    S = sp.coo_matrix((vals, (rows, cols)), shape=(n, n)).tocsr()
    S.sort_indices()
    # scipy drops nothing on sum_duplicates (explicit zeros from cancellation stay)
    assert np.array_equal(rp, S.indptr) and np.array_equal(ci, S.indices) and np.array_equal(v, S.data)
    x = rng.standard_normal(n)
    y = A.spmv(x)
    assert np.allclose(y, S @ x, rtol=1e-12, atol=1e-9)


def test_generic_cg_on_reference_matrix(yfd, ref, ctx):
    """conjugated_gradient(cgp, A, b, x[, iM]) with A handed over as CSR arrays (the drop-in shape)."""
    ref.mesh(1, 3)
    ref.set_cg(rr_tol=1e-12, rr_max=1e4)
    for use_pc in (False, True):
        out = ref.run_dg(2, use_pc=use_pc)
        A = yfd.Csr.from_arrays(ctx, out["n"], out["n"], out["rowptr"], out["colidx"], out["values"])
        res = A.solve_cg(out["rhs"], yfd.CgParams(rr_tol=1e-12, rr_max=1e4, max_iter=2 * out["n"]),
                         pc_diag=out["pc_diag"] if use_pc else None)
        assert np.linalg.norm(res["x"] - out["sol"]) <= 1e-10 * np.linalg.norm(out["sol"])
        assert abs(res["iterations"] - out["iters"]) <= max(3, 0.05 * out["iters"])
    with pytest.raises(yfd.YfdError):
        A.solve_cg(np.zeros(3))


@pytest.mark.parametrize("levels", [2, 4])
def test_cfem_p1_matches_reference(yfd, ref, ctx, levels):
    ref.set_cg(rr_tol=1e-12, rr_max=1e3)
    out = ref.run_cfem(levels)
    m = ref.mesh(0, levels)
    mesh = yfd.Mesh(ctx, host_mesh_from_ref(yfd, m))
    sysm = ctx.cfem_assemble_p1(mesh)
    assert sysm.n == out["n"] and sysm.nnz == out["nnz"]
    rowptr, colidx, vals = sysm.csr()
    assert np.array_equal(rowptr, out["rowptr"]) and np.array_equal(colidx, out["colidx"])
    assert np.abs(vals - out["values"]).max() <= 1e-12 * np.abs(out["values"]).max()
    assert np.abs(sysm.rhs() - out["rhs"]).max() <= 1e-12 * np.abs(out["rhs"]).max()
    res = sysm.solve_cg(yfd.CgParams(rr_tol=1e-12, rr_max=1e3, max_iter=2 * sysm.n))
    assert np.linalg.norm(res["x"] - out["sol"]) <= 1e-10 * np.linalg.norm(out["sol"])
    sysm.npoints = mesh.npoints
    cmap = sysm.compress_map()
    bnd = np.zeros(mesh.npoints, bool)
    bnd[m["faces"][m["face_boundary"] == 1].reshape(-1)] = True
    assert np.array_equal(cmap >= 0, ~bnd)
    assert np.array_equal(cmap[~bnd], np.arange(sysm.n))
    ex = sysm.expand(res["x"])
    assert np.array_equal(ex[~bnd], res["x"]) and not ex[bnd].any()


@pytest.mark.parametrize("order,kind,size", [(1, "r", 5), (2, "r", 4), (2, "s", 37), (1, "s", 200)])
def test_cfem_rows_gathered_from_cells_equal_the_triplet_sort(yfd, ctx, order, kind, size):
    """The continuous-FEM builder gathers each row from the cells around its node (csr_from_cell_rows) instead of sorting all
    triplets; duplicates are added in cell order either way, so pattern, values and right-hand side are the SAME BITS as the
    generic path (YFD_CFEM_SORT=1 = blaze::init_from_triplets restated, core/blaze_sparse_init.hpp:101-155)."""
    import os
    hm = yfd.HostMesh.create(yfd.TRI, size) if kind == "r" else yfd.HostMesh.structured(yfd.TRI, size)
    mesh = yfd.Mesh(ctx, hm)
    res = {}
    for mode in ("0", "1"):
        os.environ["YFD_CFEM_SORT"] = mode
        try:
            sysm = ctx.cfem_assemble_p2(mesh) if order == 2 else ctx.cfem_assemble_p1(mesh)
        finally:
            os.environ.pop("YFD_CFEM_SORT", None)
        res[mode] = (sysm.n, sysm.nnz) + sysm.csr() + (sysm.rhs(),)
        sysm.close()
    a, b = res["0"], res["1"]
    assert a[0] == b[0] and a[1] == b[1] and a[1] > 0
    for u, v in zip(a[2:], b[2:]):
        assert np.array_equal(u, v)
    mesh.close()

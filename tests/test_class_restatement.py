"""CPU: numpy restatement of the congruence-class operator cache (yaourt-fem-dg_b200/csrc/yfd_spmv_class.cu,
yfd_class_keys.cuh) on systems assembled by the C oracle: cells are grouped by the quantised shape signature, every
cell's block row is compared with its class representative's, and y = A x formed from ONE block row per class equals the
CSR product.  Pins the two facts the device path rests on: the reference's own meshes have a handful of classes, and the
deviation inside a class is coordinate-rounding noise (<= 16 eps |x| / h, floor 2e-13)."""
import numpy as np
import pytest
import scipy.sparse as sp


def bs_of(k):
    return (k + 2) * (k + 1) // 2


def frames(points, cells):
    P = points[cells]                                   # [nc][nv][2]
    bar = P.mean(axis=1)
    nv = cells.shape[1]
    h = np.zeros(len(cells))
    for a in range(nv):
        for b in range(a + 1, nv):
            h = np.maximum(h, np.linalg.norm(P[:, a] - P[:, b], axis=1))      # diameter, core/mesh.hpp:474-487
    return P, bar, h


def neighbours(cells, faces):
    """[nc][nf] neighbour across each local face (reference's local face order, core/mesh.hpp:299-325), -1 on the boundary"""
    nc, nv = cells.shape
    key = {}
    for c in range(nc):
        for e in range(nv):
            a = cells[c, 0 if e == nv - 1 else e]
            b = cells[c, nv - 1 if e == nv - 1 else e + 1]
            key.setdefault((min(a, b), max(a, b)), []).append((c, e))
    nbr = -np.ones((nc, nv), np.int64)
    for owners in key.values():
        if len(owners) == 2:
            (c0, e0), (c1, e1) = owners
            nbr[c0, e0] = c1
            nbr[c1, e1] = c0
    return nbr


def class_keys(points, cells, nbr):
    P, bar, h = frames(points, cells)
    q = 65536.0 / h
    parts = [np.rint((P[:, v, d] - bar[:, d]) * q) for v in range(cells.shape[1]) for d in range(2)]
    for e in range(cells.shape[1]):
        nb = nbr[:, e]
        has = nb >= 0
        safe = np.where(has, nb, 0)
        for d in range(2):
            parts.append(np.where(has, np.rint((bar[safe, d] - bar[:, d]) * q), 1e15))
        parts.append(np.where(has, np.rint(h[safe] * q), 1e15))
    sig = np.stack(parts, axis=1)
    _, inv = np.unique(sig, axis=0, return_inverse=True)
    return inv.ravel(), h


def block_rows(A, nc, bs, nbr):
    """per cell the block row in canonical slot order (self, faces): [nc][1+nf][bs][bs], zero blocks on the boundary"""
    B = A.tobsr(blocksize=(bs, bs))
    B.sort_indices()
    nf = nbr.shape[1]
    out = np.zeros((nc, 1 + nf, bs, bs))
    for c in range(nc):
        cols = B.indices[B.indptr[c]:B.indptr[c + 1]]
        data = B.data[B.indptr[c]:B.indptr[c + 1]]
        where = {int(cc): i for i, cc in enumerate(cols)}
        out[c, 0] = data[where[c]]
        for e in range(nf):
            if nbr[c, e] >= 0:
                out[c, 1 + e] = data[where[int(nbr[c, e])]]
    return out


CASES = [(0, 4, 1), (0, 3, 2), (1, 4, 3), (1, 3, 2), (0, 3, 4)]


@pytest.mark.parametrize("cell,r,k", CASES)
def test_reference_meshes_have_few_classes_and_rounding_level_deviation(oracle, cell, r, k):
    m = oracle.mesh(cell, r)
    out = oracle.dg_sip(cell, m["points"], m["cells"], m["faces"], k, 1.0)
    nc, bs = len(m["cells"]), bs_of(k)
    n = nc * bs
    A = sp.csr_matrix((out["values"], out["colidx"], out["rowptr"]), shape=(n, n))
    nbr = neighbours(m["cells"], m["faces"])
    cls, h = class_keys(m["points"], m["cells"], nbr)
    ncls = cls.max() + 1
    assert ncls <= 64, ncls
    rows = block_rows(A, nc, bs, nbr)
    rep = np.array([np.flatnonzero(cls == kk)[0] for kk in range(ncls)])         # lowest id = representative
    scale = np.abs(rows[rep]).reshape(ncls, -1).max(axis=1)
    dev = (np.abs(rows - rows[rep][cls]).reshape(nc, -1).max(axis=1) / scale[cls]).max()
    tol = max(2e-13, 16 * 2.220446049250313e-16 * np.abs(m["points"]).max() / h.min())
    assert dev <= tol, (dev, tol)
    # y = A x from one block row per class
    x = np.random.default_rng(0).standard_normal(n).reshape(nc, bs)
    T = rows[rep]
    y = np.einsum("cij,cj->ci", T[cls, 0], x)
    for e in range(nbr.shape[1]):
        nb = np.where(nbr[:, e] >= 0, nbr[:, e], np.arange(nc))
        y += np.einsum("cij,cj->ci", T[cls, 1 + e], x[nb])
    y_ref = A @ x.ravel()
    bound = abs(A) @ np.abs(x.ravel())
    assert (np.abs(y.ravel() - y_ref) / bound).max() <= 1e-12


def test_perturbed_mesh_hashes_alike_but_is_refused(oracle):
    m = oracle.mesh(1, 3)
    pts = m["points"].copy()
    rng = np.random.default_rng(0)
    interior = (pts[:, 0] > 0) & (pts[:, 0] < 1) & (pts[:, 1] > 0) & (pts[:, 1] < 1)
    pts[interior] += 1e-9 / 16 * rng.standard_normal((interior.sum(), 2))
    out = oracle.dg_sip(1, pts, m["cells"], m["faces"], 2, 1.0)
    nc, bs = len(m["cells"]), 6
    A = sp.csr_matrix((out["values"], out["colidx"], out["rowptr"]), shape=(nc * bs, nc * bs))
    nbr = neighbours(m["cells"], m["faces"])
    cls, h = class_keys(pts, m["cells"], nbr)
    assert cls.max() + 1 <= 64                                     # the 2^-16 hash does not see 1e-9
    rows = block_rows(A, nc, bs, nbr)
    rep = np.array([np.flatnonzero(cls == kk)[0] for kk in range(cls.max() + 1)])
    scale = np.abs(rows[rep]).reshape(len(rep), -1).max(axis=1)
    dev = (np.abs(rows - rows[rep][cls]).reshape(nc, -1).max(axis=1) / scale[cls]).max()
    assert dev > 100 * max(2e-13, 16 * 2.220446049250313e-16 / h.min())          # ... the verification pass does

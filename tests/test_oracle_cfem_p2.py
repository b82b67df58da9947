"""CPU: the continuous P2 statement (oracle/yfd_oracle_cfem_p2.inc).

PARITY UNPINNED -- the reference is P1-only (methods/cfem.hpp:31-89, :200-204), so there is no reference output to
compare with.  What can be checked is checked here: the element matrix against closed forms, the numbering against
the P1 rule it extends, and the discretisation against its theoretical order on the driver's manufactured solution."""
import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spla

TRI = np.array([[0.1, 0.2], [0.9, 0.15], [0.35, 0.8]])


def p2_nodes(v):
    return np.vstack([v, 0.5 * (v[0] + v[1]), 0.5 * (v[1] + v[2]), 0.5 * (v[0] + v[2])])


def test_element_matrix_closed_forms(oracle):
    K, b = oracle.cfem_p2_element(TRI, rhs_fn=2)
    d1, d2 = TRI[1] - TRI[0], TRI[2] - TRI[0]
    area = 0.5 * abs(d1[0] * d2[1] - d1[1] * d2[0])
    assert np.abs(K - K.T).max() <= 1e-15 * np.abs(K).max()
    assert np.abs(K.sum(axis=1)).max() <= 1e-14 * np.abs(K).max()          # constants are in the kernel
    # the load vector of f = 1: vertex functions integrate to 0, edge functions to area / 3
    assert np.allclose(b, [0, 0, 0, area / 3, area / 3, area / 3], rtol=0, atol=1e-14)
    # energy of quadratic polynomials, which P2 interpolates exactly: p^T K p = int |grad p|^2
    X = p2_nodes(TRI)
    x, y = X[:, 0], X[:, 1]
    (x0, y0), (x1, y1), (x2, y2) = TRI
    ixx = area / 6 * (x0 * x0 + x1 * x1 + x2 * x2 + x0 * x1 + x0 * x2 + x1 * x2)
    iyy = area / 6 * (y0 * y0 + y1 * y1 + y2 * y2 + y0 * y1 + y0 * y2 + y1 * y2)
    for p, energy in ((x, area), (x * x, 4 * ixx), (x * y, ixx + iyy), (x + 2 * y, 5 * area)):
        assert abs(p @ K @ p - energy) <= 1e-13 * max(1.0, energy)
    # vertex order must not matter beyond the induced permutation of the local nodes
    perm_v = [1, 2, 0]
    Kp, _ = oracle.cfem_p2_element(TRI[perm_v], rhs_fn=2)
    Xp = p2_nodes(TRI[perm_v])
    idx = [int(np.argmin(np.abs(X - q).sum(axis=1))) for q in Xp]
    assert np.abs(Kp - K[np.ix_(idx, idx)]).max() <= 1e-13 * np.abs(K).max()


def test_numbering_extends_p1(oracle):
    m = oracle.mesh(0, 3)
    p1 = oracle.cfem_p1(m["points"], m["cells"], m["faces"], m["face_boundary"])
    p2 = oracle.cfem_p2(m["points"], m["cells"], m["faces"], m["face_boundary"])
    npts, nf = len(m["points"]), len(m["faces"])
    cm = p2["compress_map"]
    # vertex dofs come first and are numbered exactly like the P1 dofs; edge dofs follow in face order
    assert np.array_equal(cm[:npts], p1["compress_map"])
    interior = m["face_boundary"] == 0
    assert np.array_equal(cm[npts:][interior], p1["n"] + np.arange(interior.sum()))
    assert (cm[npts:][~interior] == -1).all()
    assert p2["n"] == p1["n"] + interior.sum()
    A = sp.csr_matrix((p2["values"], p2["colidx"], p2["rowptr"]), shape=(p2["n"], p2["n"]))
    assert abs(A - A.T).max() <= 1e-14 * abs(A).max()
    # row lengths: a vertex couples to its ring's nodes, an interior edge to the 9 nodes of its two cells
    rl = np.diff(p2["rowptr"])
    assert rl[p1["n"]:].max() == 9 and rl.min() >= 3


def solve(oracle, r, rhs_fn=1):
    m = oracle.mesh(0, r)
    s = oracle.cfem_p2(m["points"], m["cells"], m["faces"], m["face_boundary"], rhs_fn)
    A = sp.csr_matrix((s["values"], s["colidx"], s["rowptr"]), shape=(s["n"], s["n"]))
    x = spla.spsolve(A.tocsc(), s["rhs"])
    ex = np.zeros(len(s["compress_map"]))
    free = s["compress_map"] >= 0
    ex[free] = x[s["compress_map"][free]]
    return m, s, ex


def test_third_order_convergence(oracle):
    errs = []
    for r in (2, 3, 4, 5):
        m, s, ex = solve(oracle, r)
        errs.append(np.sqrt(oracle.cfem_p2_errsq(m["points"], m["cells"], m["faces"], ex, 3)))
    orders = np.log2(np.array(errs[:-1]) / np.array(errs[1:]))
    assert errs[-1] < 5e-6
    assert (orders[1:] > 2.85).all() and (orders[1:] < 3.2).all(), orders


def test_nodal_values_close_to_solution(oracle):
    m, s, ex = solve(oracle, 4)
    npts = len(m["points"])
    mid = 0.5 * (m["points"][m["faces"][:, 0]] + m["points"][m["faces"][:, 1]])
    X = np.vstack([m["points"], mid])
    u = np.sin(np.pi * X[:, 0]) * np.sin(np.pi * X[:, 1])
    assert np.abs(ex - u).max() < 1e-4
    assert len(ex) == npts + len(m["faces"])


def test_broken_mesh_is_reported(oracle):
    m = oracle.mesh(0, 1)
    with pytest.raises(ValueError):
        oracle.cfem_p2(m["points"], m["cells"], m["faces"][:-1], m["face_boundary"][:-1])


def test_element_matrix_and_load_against_exact_symbolic_integrals(oracle):
    """An independent derivation: the P2 Lagrange functions written in x, y for the physical triangle, their gradients
    and the integrals over the triangle done EXACTLY by sympy (rational vertex coordinates), against the oracle's
    quadrature-based element matrix and (for a polynomial right-hand side) load vector."""
    sympy = pytest.importorskip("sympy")
    x, y, s, t = sympy.symbols("x y s t")
    V = [(sympy.Rational(1, 10), sympy.Rational(1, 5)), (sympy.Rational(9, 10), sympy.Rational(3, 20)), (sympy.Rational(7, 20), sympy.Rational(4, 5))]
    (x0, y0), (x1, y1), (x2, y2) = V
    det = (x1 - x0) * (y2 - y0) - (x2 - x0) * (y1 - y0)
    l1 = ((x - x0) * (y2 - y0) - (x2 - x0) * (y - y0)) / det
    l2 = ((x1 - x0) * (y - y0) - (x - x0) * (y1 - y0)) / det
    l0 = 1 - l1 - l2
    phi = [l0 * (2 * l0 - 1), l1 * (2 * l1 - 1), l2 * (2 * l2 - 1), 4 * l0 * l1, 4 * l1 * l2, 4 * l0 * l2]

    def integrate(f):        # map the reference triangle (s, t) -> physical, |det| constant
        fx = f.subs({x: x0 + (x1 - x0) * s + (x2 - x0) * t, y: y0 + (y1 - y0) * s + (y2 - y0) * t}, simultaneous=True)
        return sympy.integrate(sympy.integrate(fx, (t, 0, 1 - s)), (s, 0, 1)) * abs(det)

    Kex = np.array([[float(integrate(sympy.diff(a, x) * sympy.diff(b, x) + sympy.diff(a, y) * sympy.diff(b, y))) for b in phi] for a in phi])
    verts = np.array([[float(a), float(b)] for a, b in V])
    K, b1 = oracle.cfem_p2_element(verts, rhs_fn=2)
    assert np.abs(K - Kex).max() <= 1e-13 * np.abs(Kex).max()
    bex = np.array([float(integrate(p)) for p in phi])                   # f = 1
    assert np.abs(b1 - bex).max() <= 1e-14

"""ctypes binding of oracle/libyfd_oracle.so (oracle/yfd_oracle.c) -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this module.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_i64p = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")


def lib_path():
    return os.path.join(_HERE, "libyfd_oracle.so")


def available():
    return os.path.exists(lib_path())


def _c(a, dt):
    return np.ascontiguousarray(a, dt)


class Oracle:
    def __init__(self):
        self.lib = L = C.CDLL(lib_path())
        i64, dbl, cint, vp = C.c_int64, C.c_double, C.c_int, C.c_void_p
        L.yfo_basis_size.argtypes = [cint]
        L.yfo_face_quadrature.argtypes = [_f64p, _f64p, cint, _f64p, cint]
        L.yfo_cell_quadrature.argtypes = [cint, _f64p, cint, _f64p, cint]
        L.yfo_basis_eval.argtypes = [cint, _f64p, cint, dbl, dbl, _f64p, _f64p]
        L.yfo_face_geometry.argtypes = [cint, _f64p, _f64p, _f64p, _f64p, C.POINTER(dbl)]
        L.yfo_connectivity.argtypes = [cint, i64, _i64p, i64, _i64p, _i64p, _i64p]
        L.yfo_dg_sip_blocks.argtypes = [cint, _f64p, i64, _i64p, i64, _i64p, cint, dbl, cint, cint, _f64p, _f64p, _f64p, _f64p]
        L.yfo_dg_sip_triplets.argtypes = [cint, _f64p, i64, _i64p, i64, _i64p, cint, dbl, cint, cint, _i64p, _i64p, _f64p, _f64p, _f64p]
        L.yfo_dg_sip_triplets.restype = i64
        L.yfo_csr_from_triplets.argtypes = [i64, i64, i64, _i64p, _i64p, _f64p, _i64p, _i64p, _f64p]
        L.yfo_csr_from_triplets.restype = i64
        L.yfo_cg.argtypes = [i64, _i64p, _i64p, _f64p, _f64p, _f64p, vp, dbl, dbl, i64, cint, vp, i64, C.POINTER(dbl), C.POINTER(dbl)]
        L.yfo_cg.restype = i64
        L.yfo_cfem_p1.argtypes = [i64, _f64p, i64, _i64p, i64, _i64p, _u8p, cint, _i64p, C.POINTER(i64), _i64p, _i64p, _f64p, _f64p]
        L.yfo_cfem_p1.restype = i64
        L.yfo_cfem_p2.argtypes = L.yfo_cfem_p1.argtypes
        L.yfo_cfem_p2.restype = i64
        L.yfo_cfem_p2_element.argtypes = [_f64p, cint, _f64p, _f64p]
        L.yfo_cfem_p2_element.restype = None
        L.yfo_cfem_p2_errsq.argtypes = [i64, _f64p, i64, _i64p, i64, _i64p, _f64p, cint]
        L.yfo_cfem_p2_errsq.restype = C.c_double
        L.yfo_mesher.argtypes = [cint, cint, C.POINTER(i64), C.POINTER(i64), C.POINTER(i64), vp, vp, vp, vp]
        _i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
        L.yfo_set_physics.argtypes = [cint, cint]
        L.yfo_cg_normal.argtypes = [i64, _i64p, _i64p, _f64p, _f64p, _f64p, dbl, dbl, i64, vp, i64]
        L.yfo_cg_normal.restype = i64
        L.yfo_maxwell_assemble.argtypes = [cint, _f64p, i64, _i64p, i64, _i64p, _i64p, _i64p, cint, cint, dbl, vp, vp,
                                           _f64p, _f64p, _i64p, _i32p, _f64p]
        L.yfo_maxwell_init.argtypes = [cint, _f64p, i64, _i64p, cint, _f64p, _f64p]
        L.yfo_maxwell_step.argtypes = [cint, i64, cint, _f64p, _f64p, _i64p, _i32p, dbl, cint, i64, _f64p]
        L.yfo_maxwell_errors.argtypes = [cint, _f64p, i64, _i64p, cint, _f64p, dbl, _f64p]

    # ---- mesher (core/meshers.hpp) ----
    def mesh(self, cell_type, levels):
        a, b, c = C.c_int64(), C.c_int64(), C.c_int64()
        self.lib.yfo_mesher(cell_type, levels, C.byref(a), C.byref(b), C.byref(c), None, None, None, None)
        nv = 3 if cell_type == 0 else 4
        pts = np.empty((a.value, 2)); cells = np.empty((b.value, nv), np.int64)
        faces = np.empty((c.value, 2), np.int64); fb = np.empty(c.value, np.uint8)
        vp = lambda x: x.ctypes.data_as(C.c_void_p)
        self.lib.yfo_mesher(cell_type, levels, C.byref(a), C.byref(b), C.byref(c), vp(pts), vp(cells), vp(faces), vp(fb))
        return dict(type=cell_type, points=pts, cells=cells, faces=faces, face_boundary=fb)

    # ---- probes ----
    def cell_quadrature(self, verts, degree):
        verts = _c(verts, np.float64)
        buf = np.empty((128, 3))
        n = self.lib.yfo_cell_quadrature(len(verts), verts, degree, buf, 128)
        if n < 0:
            raise ValueError("degree too high")
        return buf[:n].copy()

    def face_quadrature(self, p0, p1, degree):
        buf = np.empty((32, 3))
        n = self.lib.yfo_face_quadrature(_c(p0, np.float64), _c(p1, np.float64), degree, buf, 32)
        if n < 0:
            raise ValueError("degree too high")
        return buf[:n].copy()

    def basis_eval(self, verts, k, x, y):
        verts = _c(verts, np.float64)
        phi = np.empty(64); dphi = np.empty((64, 2))
        n = self.lib.yfo_basis_eval(len(verts), verts, k, x, y, phi, dphi)
        return phi[:n].copy(), dphi[:n].copy()

    def face_geometry(self, verts, p0, p1):
        verts = _c(verts, np.float64)
        n = np.empty(2); d = C.c_double()
        self.lib.yfo_face_geometry(len(verts), verts, _c(p0, np.float64), _c(p1, np.float64), n, C.byref(d))
        return n, d.value

    def connectivity(self, cells, faces):
        cells = _c(cells, np.int64); faces = _c(faces, np.int64)
        nv = cells.shape[1]
        cf = np.empty_like(cells); ow = np.empty((len(faces), 2), np.int64)
        st = self.lib.yfo_connectivity(nv, len(cells), cells, len(faces), faces, cf, ow)
        if st:
            raise RuntimeError("connectivity status %d" % st)
        return cf, ow

    # ---- DG-SIP (dg2d_diffusion.cpp:133-201 + methods/dg.hpp) ----
    def dg_blocks(self, points, cells, faces, k, eta=1.0, rhs_fn=1, dir_fn=0):
        points = _c(points, np.float64); cells = _c(cells, np.int64); faces = _c(faces, np.int64)
        nv = cells.shape[1]; nc = len(cells); bs = (k + 2) * (k + 1) // 2
        K = np.empty((nc, bs, bs)); Att = np.empty((nc, nv, bs, bs)); Atn = np.empty((nc, nv, bs, bs)); rhs = np.empty((nc, bs))
        st = self.lib.yfo_dg_sip_blocks(nv, points, nc, cells, len(faces), faces, k, eta, rhs_fn, dir_fn, K, Att, Atn, rhs)
        if st:
            raise RuntimeError("dg blocks status %d" % st)
        return K, Att, Atn, rhs

    def dg_triplets(self, points, cells, faces, k, eta=1.0, rhs_fn=1, dir_fn=0):
        points = _c(points, np.float64); cells = _c(cells, np.int64); faces = _c(faces, np.int64)
        nv = cells.shape[1]; nc = len(cells); bs = (k + 2) * (k + 1) // 2
        cap = bs * bs * nc * (2 * nv + 1)
        ri = np.empty(cap, np.int64); rj = np.empty(cap, np.int64); rv = np.empty(cap)
        rhs = np.empty(nc * bs); pct = np.empty(nc * bs)
        cnt = self.lib.yfo_dg_sip_triplets(nv, points, nc, cells, len(faces), faces, k, eta, rhs_fn, dir_fn, ri, rj, rv, rhs, pct)
        if cnt < 0:
            raise RuntimeError("dg triplets status %d" % -cnt)
        return ri[:cnt].copy(), rj[:cnt].copy(), rv[:cnt].copy(), rhs, pct

    def dg_sip(self, cell_type, points, cells, faces, k, eta=1.0, rhs_fn=1, dir_fn=0):
        """The assembled system as the reference's assembler leaves it: lhs CSR, rhs, pc diagonal."""
        ri, rj, rv, rhs, pct = self.dg_triplets(points, cells, faces, k, eta, rhs_fn, dir_fn)
        n = len(rhs)
        rowptr, colidx, values = self.csr_from_triplets(n, n, ri, rj, rv)
        with np.errstate(divide="ignore"):
            pc = 1.0 / pct                                   # methods/dg.hpp:133-142
        return dict(n=n, nnz=len(values), rowptr=rowptr, colidx=colidx, values=values, rhs=rhs, pc_diag=pc, raw=(ri, rj, rv))

    # ---- triplets -> CSR (core/blaze_sparse_init.hpp:101-155) ----
    def csr_from_triplets(self, nrows, ncols, ri, rj, rv):
        ri = _c(ri, np.int64); rj = _c(rj, np.int64); rv = _c(rv, np.float64)
        cnt = len(rv)
        rowptr = np.empty(nrows + 1, np.int64); colidx = np.empty(max(cnt, 1), np.int64); values = np.empty(max(cnt, 1))
        nnz = self.lib.yfo_csr_from_triplets(nrows, ncols, cnt, ri, rj, rv, rowptr, colidx, values)
        if nnz < 0:
            raise ValueError("triplet out of range")
        return rowptr, colidx[:nnz].copy(), values[:nnz].copy()

    # ---- CG (core/solvers.hpp:46-243) ----
    def cg(self, rowptr, colidx, values, b, pc_diag=None, rr_tol=1e-8, rr_max=20.0, max_iter=100, x0=None, history=True):
        rowptr = _c(rowptr, np.int64); colidx = _c(colidx, np.int64); values = _c(values, np.float64); b = _c(b, np.float64)
        n = len(b)
        x = np.zeros(n) if x0 is None else _c(x0, np.float64).copy()
        pc = None if pc_diag is None else _c(pc_diag, np.float64)
        cap = int(max_iter) + 2 if history else 0
        hist = np.empty(max(cap, 1))
        nr0, nr = C.c_double(), C.c_double()
        it = self.lib.yfo_cg(n, rowptr, colidx, values, b, x, None if pc is None else pc.ctypes.data_as(C.c_void_p),
                             rr_tol, rr_max, int(max_iter), int(x0 is not None),
                             hist.ctypes.data_as(C.c_void_p) if history else None, cap, C.byref(nr0), C.byref(nr))
        if it < 0:
            raise ValueError("shape mismatch")
        return dict(x=x, iterations=it, nr0=nr0.value, nr=nr.value, history=hist[:it + 1].copy() if history else None)

    def bicgstab(self, rowptr, colidx, values, b, pc_diag=None, rr_tol=1e-8, rr_max=20.0, max_iter=100, x0=None, history=True):
        """bicgstab, both overloads (core/solvers.hpp:246-467)."""
        rowptr = _c(rowptr, np.int64); colidx = _c(colidx, np.int64); values = _c(values, np.float64); b = _c(b, np.float64)
        n = len(b)
        x = np.zeros(n) if x0 is None else _c(x0, np.float64).copy()
        pc = None if pc_diag is None else _c(pc_diag, np.float64)
        cap = int(max_iter) + 2 if history else 0
        hist = np.empty(max(cap, 1))
        nr0, nr = C.c_double(), C.c_double()
        f = self.lib.yfo_bicgstab
        f.restype = C.c_int64
        f.argtypes = [C.c_int64, _i64p, _i64p, _f64p, _f64p, _f64p, C.c_void_p, C.c_double, C.c_double, C.c_int64, C.c_int,
                      C.c_void_p, C.c_int64, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        it = f(n, rowptr, colidx, values, b, x, None if pc is None else pc.ctypes.data_as(C.c_void_p), rr_tol, rr_max,
               int(max_iter), int(x0 is not None), hist.ctypes.data_as(C.c_void_p) if history else None, cap, C.byref(nr0), C.byref(nr))
        return dict(x=x, iterations=it, nr0=nr0.value, nr=nr.value, history=hist[:it + 1].copy() if history else None)

    # ---- CFEM P1 (methods/cfem.hpp, continuous_fem.cpp) ----
    def cfem_p1(self, points, cells, faces, face_boundary, rhs_fn=1):
        points = _c(points, np.float64); cells = _c(cells, np.int64); faces = _c(faces, np.int64)
        fb = _c(face_boundary, np.uint8)
        npts, nc = len(points), len(cells)
        cmap = np.empty(npts, np.int64); n = C.c_int64()
        ri = np.empty(9 * nc, np.int64); rj = np.empty(9 * nc, np.int64); rv = np.empty(9 * nc); rhs = np.empty(npts)
        cnt = self.lib.yfo_cfem_p1(npts, points, nc, cells, len(faces), faces, fb, rhs_fn, cmap, C.byref(n), ri, rj, rv, rhs)
        n = n.value
        rowptr, colidx, values = self.csr_from_triplets(n, n, ri[:cnt], rj[:cnt], rv[:cnt])
        return dict(n=n, nnz=len(values), rowptr=rowptr, colidx=colidx, values=values, rhs=rhs[:n].copy(), compress_map=cmap,
                    raw=(ri[:cnt].copy(), rj[:cnt].copy(), rv[:cnt].copy()))

    # ---- the drivers' post-process (dg2d_diffusion.cpp:225-268) ----
    def dg_postprocess(self, points, cells, k, sol, ref_fn=3):
        """(L2_errsq_qp, L2_errsq_mm) of a DG solution vector in the reference's numbering."""
        points = _c(points, np.float64); cells = _c(cells, np.int64); sol = _c(sol, np.float64)
        f = self.lib.yfo_dg_postprocess
        f.restype = C.c_int64
        f.argtypes = [C.c_int, _f64p, C.c_int64, _i64p, C.c_int, _f64p, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        qp, mm = C.c_double(), C.c_double()
        bad = f(cells.shape[1], points, len(cells), cells, k, sol, ref_fn, C.byref(qp), C.byref(mm))
        if bad:
            raise ValueError("singular elemental mass matrix in cell %d" % (bad - 1))
        return qp.value, mm.value

    # ---- continuous P2 (extension: the reference is P1-only, parity unpinned) ----
    def cfem_p2_element(self, verts, rhs_fn=1):
        v = _c(verts, np.float64).reshape(6)
        K = np.empty((6, 6)); b = np.empty(6)
        self.lib.yfo_cfem_p2_element(v, rhs_fn, K, b)
        return K, b

    def cfem_p2(self, points, cells, faces, face_boundary, rhs_fn=1):
        points = _c(points, np.float64); cells = _c(cells, np.int64); faces = _c(faces, np.int64)
        fb = _c(face_boundary, np.uint8)
        nn, nc = len(points) + len(faces), len(cells)
        cmap = np.empty(nn, np.int64); n = C.c_int64()
        ri = np.empty(36 * nc, np.int64); rj = np.empty(36 * nc, np.int64); rv = np.empty(36 * nc); rhs = np.empty(nn)
        cnt = self.lib.yfo_cfem_p2(len(points), points, nc, cells, len(faces), faces, fb, rhs_fn, cmap, C.byref(n), ri, rj, rv, rhs)
        if cnt < 0:
            raise ValueError("a cell face is missing from the face list")
        n = n.value
        rowptr, colidx, values = self.csr_from_triplets(n, n, ri[:cnt], rj[:cnt], rv[:cnt])
        return dict(n=n, nnz=len(values), rowptr=rowptr, colidx=colidx, values=values, rhs=rhs[:n].copy(), compress_map=cmap)

    def cfem_p2_errsq(self, points, cells, faces, expanded, ref_fn=3):
        points = _c(points, np.float64); cells = _c(cells, np.int64); faces = _c(faces, np.int64)
        return self.lib.yfo_cfem_p2_errsq(len(points), points, len(cells), cells, len(faces), faces, _c(expanded, np.float64), ref_fn)

    # ---- DG Maxwell TM (methods/dg_maxwell_2D.hpp, maxwell_driver.cpp) ----
    def maxwell_assemble(self, points, cells, faces, k, upwind=False, eta=1.0, mu_r=None, eps_r=None):
        points, cells, faces = _c(points, np.float64), _c(cells, np.int64), _c(faces, np.int64)
        nc, nv = cells.shape
        cf, owners = self.connectivity(cells, faces)
        bs = self.lib.yfo_basis_size(k)
        L = 3 * bs
        on = np.empty((nc, L, L)); off = np.empty((nc, nv, L, L))
        nb = np.empty((nc, nv), np.int64); has = np.empty((nc, nv), np.int32); gM = np.empty((nc, bs, bs))
        mu = None if mu_r is None else _c(mu_r, np.float64)
        ep = None if eps_r is None else _c(eps_r, np.float64)
        vp = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
        rc = self.lib.yfo_maxwell_assemble(nv, points, nc, cells, len(faces), faces, _c(cf, np.int64), _c(owners, np.int64),
                                           k, int(upwind), eta, vp(mu), vp(ep), on, off, nb, has, gM)
        if rc != 0:
            raise RuntimeError("yfo_maxwell_assemble rc=%d" % rc)
        return dict(ondiag=on, offdiag=off, neigh=nb, has=has, gM=gM, bs=bs)

    def maxwell_init(self, points, cells, k, gM):
        points, cells = _c(points, np.float64), _c(cells, np.int64)
        nc, nv = cells.shape
        bs = self.lib.yfo_basis_size(k)
        dofs = np.empty(nc * 3 * bs)
        rc = self.lib.yfo_maxwell_init(nv, points, nc, cells, k, _c(gM, np.float64), dofs)
        if rc != 0:
            raise RuntimeError("yfo_maxwell_init rc=%d" % rc)
        return dofs

    def maxwell_step(self, op, k, dofs, dt, nsteps=1, euler=False):
        nc, nv = op["neigh"].shape
        d = np.array(dofs, np.float64, copy=True)
        self.lib.yfo_maxwell_step(nv, nc, k, op["ondiag"], op["offdiag"], op["neigh"], op["has"], dt, int(euler), nsteps, d)
        return d

    def maxwell_errors(self, points, cells, k, dofs, t):
        points, cells = _c(points, np.float64), _c(cells, np.int64)
        nc, nv = cells.shape
        out = np.empty(3)
        self.lib.yfo_maxwell_errors(nv, points, nc, cells, k, _c(dofs, np.float64), t, out)
        return out

    # ---- advection-reaction (dg2d_advection.cpp) ----
    def advection(self, cell_type, points, cells, faces, k, eta=1.0, upwinding=False):
        """The advection driver's cell loop + finalize: same shapes as dg_sip()."""
        self.lib.yfo_set_physics(1, int(upwinding))
        try:
            return self.dg_sip(cell_type, points, cells, faces, k, eta=eta, rhs_fn=4, dir_fn=0)
        finally:
            self.lib.yfo_set_physics(0, 0)

    def cg_normal(self, rowptr, colidx, values, b, rr_tol=1e-8, rr_max=1e4, max_iter=100):
        n = len(b)
        x = np.zeros(n); hist = np.empty(max_iter + 2)
        it = self.lib.yfo_cg_normal(n, _c(rowptr, np.int64), _c(colidx, np.int64), _c(values, np.float64), _c(b, np.float64), x,
                                    rr_tol, rr_max, max_iter, hist.ctypes.data_as(C.c_void_p), len(hist))
        return dict(x=x, iterations=it, history=hist[:it + 1])

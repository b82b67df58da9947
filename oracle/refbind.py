"""ctypes binding of oracle/_ref/libyfd_ref.so -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

The library is the UNMODIFIED reference (built by oracle/Makefile from the sources
where they lie under /root/reference; see oracle/ref_harness.cpp).  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
TRI, QUAD = 0, 1

_i64p = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")


def lib_path(omp=False, advection=False, lapack=False):
    """lapack=True: the same harness linked against a real LAPACK (the OpenBLAS bundled with the opencv wheel) instead of
    oracle/lu_shim.c; built by `make -C oracle lapack_pin` where that wheel exists, used only to pin the shim."""
    if lapack:
        return os.path.join(_HERE, "_ref", "libyfd_ref_lapack.so")
    if advection:
        return os.path.join(_HERE, "_ref", "libyfd_ref_adv.so")
    return os.path.join(_HERE, "_ref", "libyfd_ref_omp.so" if omp else "libyfd_ref.so")


def available(omp=False, advection=False, lapack=False):
    return os.path.exists(lib_path(omp, advection, lapack))


class Ref:
    """One loaded copy of the reference library (it keeps one mesh + one result)."""

    def __init__(self, omp=False, advection=False, lapack=False):
        """advection=True loads the same harness built around dg2d_advection.cpp (run_dg then runs
        run_advection_reaction_solver, dg2d_advection.cpp:120; CG on the normal equations as that driver sets it)."""
        self.lib = L = C.CDLL(lib_path(omp, advection, lapack))
        self.advection = advection
        if advection:
            L.yfd_ref_set_upwinding.argtypes = [C.c_int]
        L.yfd_ref_mesh_create.argtypes = [C.c_int, C.c_int, C.c_int]
        L.yfd_ref_mesh_sizes.argtypes = [C.POINTER(C.c_int64)] * 4
        L.yfd_ref_mesh_copy.argtypes = [_f64p, _i64p, _i64p, _i32p, _i64p]
        L.yfd_ref_mesh_set.argtypes = [C.c_int, C.c_int64, _f64p, C.c_int64, _i64p, C.c_int64, _i64p,
                                       np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS"), C.c_void_p]
        L.yfd_ref_set_cg.argtypes = [C.c_double, C.c_double, C.c_long, C.c_int]
        L.yfd_ref_keep_raw_triplets.argtypes = [C.c_int]
        L.yfd_ref_run_dg.argtypes = [C.c_int, C.c_double, C.c_int]
        L.yfd_ref_run_cfem.argtypes = [C.c_int]
        L.yfd_ref_result_sizes.argtypes = [C.POINTER(C.c_int64)] * 5 + [C.POINTER(C.c_int32)]
        L.yfd_ref_result_csr.argtypes = [_i64p, _i64p, _f64p]
        L.yfd_ref_result_pc_diag.argtypes = [_f64p]
        L.yfd_ref_result_vectors.argtypes = [_f64p, _f64p, C.c_void_p]
        L.yfd_ref_result_raw_triplets.argtypes = [_i64p, _i64p, _f64p]
        L.yfd_ref_result_times.argtypes = [C.POINTER(C.c_double)] * 6
        L.yfd_ref_cell_quadrature.argtypes = [C.c_int64, C.c_int, _f64p, C.c_int]
        L.yfd_ref_face_quadrature.argtypes = [C.c_int64, C.c_int, _f64p, C.c_int]
        L.yfd_ref_basis_eval.argtypes = [C.c_int64, C.c_int, C.c_double, C.c_double, _f64p, _f64p]
        L.yfd_ref_face_geometry.argtypes = [C.c_int64, C.c_int, _f64p, C.POINTER(C.c_double)]
        L.yfd_ref_num_threads.restype = C.c_int
        L.yfd_ref_bicgstab.restype = C.c_int64
        L.yfd_ref_bicgstab.argtypes = [C.c_int64, _i64p, _i64p, _f64p, _f64p, _f64p, C.c_void_p, C.c_double, C.c_double, C.c_int64,
                                       C.c_int, _f64p, C.c_int64]

    def bicgstab(self, rowptr, colidx, values, b, pc_diag=None, rr_tol=1e-8, rr_max=20.0, max_iter=100, x0=None):
        """The reference's own bicgstab (core/solvers.hpp:246-467) on a CSR matrix; history = its history file (6 digits)."""
        rowptr = np.ascontiguousarray(rowptr, np.int64); colidx = np.ascontiguousarray(colidx, np.int64)
        values = np.ascontiguousarray(values, np.float64); b = np.ascontiguousarray(b, np.float64)
        n = len(b)
        x = np.zeros(n) if x0 is None else np.ascontiguousarray(x0, np.float64).copy()
        pc = None if pc_diag is None else np.ascontiguousarray(pc_diag, np.float64)
        hist = np.empty(int(max_iter) + 2)
        cnt = self.lib.yfd_ref_bicgstab(n, rowptr, colidx, values, b, x, None if pc is None else pc.ctypes.data_as(C.c_void_p),
                                        rr_tol, rr_max, int(max_iter), int(x0 is not None), hist, len(hist))
        if cnt < 0:
            raise ValueError("shape mismatch")
        return dict(x=x, iterations=int(cnt) - 1, history=hist[:cnt].copy())

    # ---- mesh -----------------------------------------------------------
    def mesh(self, mtype, levels, shatter=False):
        self.lib.yfd_ref_mesh_create(mtype, levels, int(shatter))
        return self.mesh_arrays(mtype)

    def mesh_arrays(self, mtype):
        s = [C.c_int64() for _ in range(4)]
        self.lib.yfd_ref_mesh_sizes(*[C.byref(v) for v in s])
        npts, ncells, nfaces, nv = [v.value for v in s]
        pts = np.empty((npts, 2), np.float64)
        cells = np.empty((ncells, nv), np.int64)
        faces = np.empty((nfaces, 2), np.int64)
        fbnd = np.empty(nfaces, np.int32)
        owners = np.empty((nfaces, 2), np.int64)
        self.lib.yfd_ref_mesh_copy(pts, cells, faces, fbnd, owners)
        return dict(type=mtype, points=pts, cells=cells, faces=faces, face_boundary=fbnd, face_owners=owners)

    def set_mesh(self, mtype, points, cells, faces, face_boundary, face_boundary_id=None):
        """Hand the reference a mesh made elsewhere (its public points / faces / cells vectors + its own compute_connectivity)."""
        points = np.ascontiguousarray(points, np.float64); cells = np.ascontiguousarray(cells, np.int64)
        faces = np.ascontiguousarray(faces, np.int64); fb = np.ascontiguousarray(face_boundary, np.uint8)
        fid = None if face_boundary_id is None else np.ascontiguousarray(face_boundary_id, np.int64)
        rc = self.lib.yfd_ref_mesh_set(mtype, len(points), points, len(cells), cells, len(faces), faces, fb,
                                       None if fid is None else fid.ctypes.data_as(C.c_void_p))
        if rc != 0:
            raise RuntimeError("reference rejected the mesh")
        return self.mesh_arrays(mtype)

    # ---- drivers --------------------------------------------------------
    def set_cg(self, rr_tol=1e-8, rr_max=1e4, max_iter=-1, skip_solve=False):
        self.lib.yfd_ref_set_cg(rr_tol, rr_max, max_iter, int(skip_solve))

    def run_dg(self, degree, eta=1.0, use_pc=False, raw_triplets=False, upwinding=False):
        """run_diffusion_solver (dg2d_diffusion.cpp:114) on the current mesh."""
        self.lib.yfd_ref_keep_raw_triplets(int(raw_triplets))
        if self.advection:
            self.lib.yfd_ref_set_upwinding(int(upwinding))
        rc = self.lib.yfd_ref_run_dg(degree, eta, int(use_pc))
        if rc != 0:
            raise RuntimeError("reference DG driver failed")
        return self._results(raw_triplets)

    def run_cfem(self, levels, raw_triplets=False):
        """continuous_fem.cpp main() with its refinement count redirected."""
        self.lib.yfd_ref_keep_raw_triplets(int(raw_triplets))
        rc = self.lib.yfd_ref_run_cfem(levels)
        if rc != 0:
            raise RuntimeError("reference CFEM driver failed")
        return self._results(raw_triplets)

    def _results(self, raw):
        s = [C.c_int64() for _ in range(5)]
        hp = C.c_int32()
        self.lib.yfd_ref_result_sizes(*[C.byref(v) for v in s], C.byref(hp))
        n, nnz, nraw, iters, nhist = [v.value for v in s]
        rowptr = np.empty(n + 1, np.int64)
        colidx = np.empty(nnz, np.int64)
        values = np.empty(nnz, np.float64)
        self.lib.yfd_ref_result_csr(rowptr, colidx, values)
        rhs = np.empty(n, np.float64)
        sol = np.empty(n, np.float64)
        hist = np.empty(max(nhist, 1), np.float64)
        self.lib.yfd_ref_result_vectors(rhs, sol, hist.ctypes.data_as(C.c_void_p))
        out = dict(n=n, nnz=nnz, rowptr=rowptr, colidx=colidx, values=values, rhs=rhs, sol=sol,
                   iters=iters, history=hist[:nhist])
        if hp.value:
            d = np.empty(n, np.float64)
            self.lib.yfd_ref_result_pc_diag(d)
            out["pc_diag"] = d
        if raw and nraw:
            ri = np.empty(nraw, np.int64)
            rj = np.empty(nraw, np.int64)
            rv = np.empty(nraw, np.float64)
            self.lib.yfd_ref_result_raw_triplets(ri, rj, rv)
            out["raw"] = (ri, rj, rv)
        t = [C.c_double() for _ in range(6)]
        self.lib.yfd_ref_result_times(*[C.byref(v) for v in t])
        out.update(t_blocks=t[0].value, t_csr=t[1].value, t_cg=t[2].value,
                   l2_qp=t[3].value, l2_mm=t[4].value, mesh_h=t[5].value)
        return out

    # ---- probes ---------------------------------------------------------
    def cell_quadrature(self, cell, degree):
        buf = np.empty((128, 3), np.float64)
        n = self.lib.yfd_ref_cell_quadrature(cell, degree, buf, 128)
        return buf[:n].copy()

    def face_quadrature(self, face, degree):
        buf = np.empty((32, 3), np.float64)
        n = self.lib.yfd_ref_face_quadrature(face, degree, buf, 32)
        return buf[:n].copy()

    def basis_eval(self, cell, degree, x, y):
        phi = np.empty(64, np.float64)
        dphi = np.empty((64, 2), np.float64)
        n = self.lib.yfd_ref_basis_eval(cell, degree, x, y, phi, dphi)
        return phi[:n].copy(), dphi[:n].copy()

    def face_geometry(self, cell, local_face):
        n = np.empty(2, np.float64)
        d = C.c_double()
        self.lib.yfd_ref_face_geometry(cell, local_face, n, C.byref(d))
        return n, d.value

    def num_threads(self):
        return self.lib.yfd_ref_num_threads()


# --------------------------------------------------------------------------------------------
# DG Maxwell TM (methods/dg_maxwell_2D.hpp + maxwell_driver.cpp), oracle/ref_harness_maxwell.cpp
# --------------------------------------------------------------------------------------------
def maxwell_lib_path(omp=False, lapack=False):
    if lapack:
        return os.path.join(_HERE, "_ref", "libyfd_ref_maxwell_lapack.so")
    return os.path.join(_HERE, "_ref", "libyfd_ref_maxwell_omp.so" if omp else "libyfd_ref_maxwell.so")


def maxwell_available(omp=False, lapack=False):
    return os.path.exists(maxwell_lib_path(omp, lapack))


class RefMaxwell:
    """The unmodified reference Maxwell solver; one context at a time."""

    def __init__(self, omp=False, lapack=False):
        self.lib = L = C.CDLL(maxwell_lib_path(omp, lapack))
        L.yfd_refmx_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_int, C.c_int, C.c_int]
        L.yfd_refmx_sizes.argtypes = [C.POINTER(C.c_int64)] * 5
        L.yfd_refmx_mesh_copy.argtypes = [_f64p, _i64p, _i64p, _i32p]
        L.yfd_refmx_set_material.argtypes = [_f64p, _f64p]
        L.yfd_refmx_get_operator.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.yfd_refmx_get_dofs.argtypes = [_f64p]
        L.yfd_refmx_set_dofs.argtypes = [_f64p]
        L.yfd_refmx_step.argtypes = [C.c_int]
        L.yfd_refmx_step.restype = C.c_double
        L.yfd_refmx_errors.argtypes = [C.c_int64, _f64p]
        L.yfd_refmx_time_assemble.restype = C.c_double
        L.yfd_refmx_run_driver.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_int64, C.c_int, C.c_int, C.c_char_p]
        L.yfd_refmx_num_threads.restype = C.c_int

    def create(self, mtype, levels, degree, dt=0.1, eta=1.0, upwind=False, euler=False, shatter=False):
        if self.lib.yfd_refmx_create(mtype, levels, degree, dt, eta, int(upwind), int(euler), int(shatter)) != 0:
            raise RuntimeError("reference maxwell_context failed")
        s = [C.c_int64() for _ in range(5)]
        self.lib.yfd_refmx_sizes(*[C.byref(v) for v in s])
        self.npoints, self.ncells, self.nfaces, self.nf, self.bs = [v.value for v in s]
        return self

    def mesh(self):
        pts = np.empty((self.npoints, 2), np.float64)
        cells = np.empty((self.ncells, self.nf), np.int64)
        faces = np.empty((self.nfaces, 2), np.int64)
        fbnd = np.empty(self.nfaces, np.int32)
        self.lib.yfd_refmx_mesh_copy(pts, cells, faces, fbnd)
        return dict(points=pts, cells=cells, faces=faces, face_boundary=fbnd)

    def set_material(self, mu_r, eps_r):
        self.lib.yfd_refmx_set_material(np.ascontiguousarray(mu_r, np.float64), np.ascontiguousarray(eps_r, np.float64))

    def assemble(self):
        if self.lib.yfd_refmx_assemble() != 0:
            raise RuntimeError("reference maxwell assemble failed")
        return self.lib.yfd_refmx_time_assemble()

    def operator(self, blocks=True):
        """gOp_ondiag [nc][3bs][3bs], gOp_offdiag [nc][nf][3bs][3bs], neighbour offsets / flags [nc][nf], gM [nc][bs][bs]"""
        nc, nf, bs = self.ncells, self.nf, self.bs
        L = 3 * bs
        on = np.empty((nc, L, L), np.float64) if blocks else None
        off = np.empty((nc, nf, L, L), np.float64) if blocks else None
        nb = np.empty((nc, nf), np.int64)
        has = np.empty((nc, nf), np.int32)
        gM = np.empty((nc, bs, bs), np.float64)
        vp = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
        self.lib.yfd_refmx_get_operator(vp(on), vp(off), vp(nb), vp(has), vp(gM))
        return dict(ondiag=on, offdiag=off, neigh=nb, has=has, gM=gM)

    def init(self):
        if self.lib.yfd_refmx_init() != 0:
            raise RuntimeError("reference apply_initial_condition failed")

    def dofs(self):
        d = np.empty(3 * self.bs * self.ncells, np.float64)
        self.lib.yfd_refmx_get_dofs(d)
        return d

    def set_dofs(self, d):
        self.lib.yfd_refmx_set_dofs(np.ascontiguousarray(d, np.float64))

    def step(self, nsteps=1):
        return self.lib.yfd_refmx_step(nsteps)

    def errors(self, cycle):
        e = np.empty(3, np.float64)
        self.lib.yfd_refmx_errors(cycle, e)
        return e

    def run_driver(self, mtype, levels, degree, dt, timesteps, upwind=False, euler=False):
        """run_maxwell_solver exactly as shipped; returns the error log rows (cycle, time, Hx, Hy, Ez errors)."""
        path = "/tmp/yfd_refmx_err_%d.txt" % os.getpid()
        rc = self.lib.yfd_refmx_run_driver(mtype, levels, degree, dt, timesteps, int(upwind), int(euler), path.encode())
        if rc != 0:
            raise RuntimeError("reference maxwell driver failed")
        rows = [list(map(float, ln.split())) for ln in open(path) if ln.strip() and not ln.startswith("#")]
        os.remove(path)
        return np.array(rows)

    def num_threads(self):
        return self.lib.yfd_refmx_num_threads()

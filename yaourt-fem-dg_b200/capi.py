"""ctypes binding of libyfd_b200.so (include/yfd.h), and small host objects over its handles.

The reference's interface for this path is C++ templates; the objects here
mirror them one to one:
    HostMesh.create(type, r)         mesher<Mesh>::create_mesh            core/meshers.hpp:130,302
    Mesh(ctx, ...)                   mesh + compute_connectivity          core/mesh.hpp:158-204
    Context.dg_assemble_sip(...)     run_diffusion_solver's cell loop     dg2d_diffusion.cpp:133-201
                                     + assembler::finalize                methods/dg.hpp:125-144
    Context.cfem_assemble_p1(...)    continuous_fem.cpp:64-71 + cfem::assembler
    Csr.from_triplets(...)           blaze::init_from_triplets            core/blaze_sparse_init.hpp:125
    System.solve_cg / Csr.solve_cg   conjugated_gradient                  core/solvers.hpp:46,141
"""
import ctypes as C
import os
import re
import weakref

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)

TRI, QUAD = 0, 1
PC_NONE, PC_JACOBI, PC_BLOCK_JACOBI, PC_PMG = 0, 1, 2, 3     # use_pc of System.solve_cg (include/yfd.h YFD_PC_*)
TI_RK4, TI_EULER = 0, 1
FIELDS_CAVITY_11, FIELDS_ZERO = 0, 1
FN_ZERO, FN_SINSIN, FN_ONE, FN_SINSIN_SOL, FN_ADVECTION_RHS, FN_SIN_X_SOL = 0, 1, 2, 3, 4, 5
REORDER_NONE, REORDER_MORTON = 0, 1


class YfdError(RuntimeError):
    def __init__(self, status, message):
        super().__init__("yfd status %d: %s" % (status, message))
        self.status = status


def lib_path():
    return os.path.join(_HERE, "lib", "libyfd_b200.so")


def header_path():
    return os.path.join(_ROOT, "include", "yfd.h")


def declared_symbols():
    """Every function include/yfd.h declares (used by the CPU test that the library exports them)."""
    src = open(header_path()).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(yfd_[a-z0-9_]+)\s*\(", src)))


_i64p = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")


class DgConfig(C.Structure):
    _fields_ = [("degree", C.c_int), ("eta", C.c_double), ("rhs_fn", C.c_int), ("dirichlet_fn", C.c_int),
                ("build_preconditioner", C.c_int)]


class AssemblyInfo(C.Structure):
    _fields_ = [("system_size", C.c_int64), ("nnz", C.c_int64), ("basis_size", C.c_int64),
                ("ms_elements", C.c_double), ("ms_structure", C.c_double)]


class CgParams(C.Structure):
    """conjugated_gradient_params<T> (core/solvers.hpp:24-43) + check_every."""
    _fields_ = [("rr_tol", C.c_double), ("rr_max", C.c_double), ("max_iter", C.c_int64),
                ("use_initial_guess", C.c_int), ("use_normal_eqns", C.c_int),
                ("save_iteration_history", C.c_int), ("check_every", C.c_int)]

    def __init__(self, **kw):
        super().__init__()
        lib().yfd_cg_params_default(C.byref(self))
        for k, v in kw.items():
            if k not in dict(self._fields_):
                raise TypeError("unknown CG parameter " + k)
            setattr(self, k, v)


class MaxwellConfig(C.Structure):
    """maxwell_config<T> (methods/dg_maxwell_2D.hpp:46-72), solver-relevant fields"""
    _fields_ = [("degree", C.c_int), ("delta_t", C.c_double), ("eta", C.c_double), ("upwind", C.c_int),
                ("time_integrator", C.c_int)]


class MaxwellInfo(C.Structure):
    _fields_ = [("ncells", C.c_int64), ("basis_size", C.c_int64), ("faces_per_elem", C.c_int64), ("ndofs", C.c_int64),
                ("operator_bytes", C.c_int64), ("reference_operator_bytes", C.c_int64), ("ms_assemble", C.c_double),
                ("uses_classes", C.c_int), ("class_reject_reason", C.c_int), ("nclasses", C.c_int64),
                ("class_max_deviation", C.c_double), ("class_tolerance", C.c_double), ("class_table_bytes", C.c_int64), ("class_tensor_pipe", C.c_int)]


class MaxwellStepInfo(C.Structure):
    _fields_ = [("steps", C.c_int64), ("ms_total", C.c_double), ("ms_per_step", C.c_double),
                ("operator_applications", C.c_int64)]


class OperatorInfo(C.Structure):
    _fields_ = [("uses_classes", C.c_int), ("classes_available", C.c_int), ("reject_reason", C.c_int), ("nclasses", C.c_int64),
                ("max_deviation", C.c_double), ("tolerance", C.c_double), ("table_bytes", C.c_int64), ("list_entries", C.c_int64), ("tensor_pipe", C.c_int)]


OPERATOR_AUTO, OPERATOR_STREAMED, OPERATOR_CLASSES = 0, 1, 2
NOALLOC_MAX_ROWS = 524288          # YFD_NOALLOC_MAX_ROWS (include/yfd.h)


class PmgInfo(C.Structure):
    _fields_ = [("nlevels", C.c_int), ("nu", C.c_int), ("coarse_degree", C.c_int), ("smooth_ratio", C.c_double),
                ("degree", C.c_int * 4), ("lambda_max", C.c_double * 4)]


class CgInfo(C.Structure):
    _fields_ = [("iterations", C.c_int64), ("nr0", C.c_double), ("nr", C.c_double), ("converged", C.c_int),
                ("ms_solve", C.c_double), ("ms_spmv", C.c_double), ("history_len", C.c_int64), ("transport", C.c_int)]


_lib = None


def lib():
    """Load the CUDA library; fail loudly if it has not been built (no fallback exists)."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise YfdError(-1, "%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(the hot path is CUDA only, there is no CPU fallback)" % path)
    L = C.CDLL(path)
    vp, i64, dbl, cint = C.c_void_p, C.c_int64, C.c_double, C.c_int
    P = C.POINTER
    sig = {
        "yfd_status_string": (C.c_char_p, [cint]),
        "yfd_ctx_create": (cint, [cint, vp, P(vp)]),
        "yfd_ctx_destroy": (None, [vp]),
        "yfd_last_error": (C.c_char_p, [vp]),
        "yfd_ctx_synchronize": (cint, [vp]),
        "yfd_ctx_launch_count": (i64, [vp]),
        "yfd_ctx_trim": (cint, [vp, P(C.c_size_t)]),
        "yfd_ctx_device_info": (cint, [vp, C.c_char_p, C.c_size_t, P(cint), P(cint), P(cint)]),
        "yfd_ctx_measure_fp64": (cint, [vp, P(dbl)]),
        "yfd_ctx_measure_fp64_tensor": (cint, [vp, P(dbl)]),
        "yfd_comm_unique_id": (cint, [C.c_char_p]),
        "yfd_ctx_attach_comm": (cint, [vp, cint, cint, C.c_char_p]),
        "yfd_ctx_rank": (cint, [vp]),
        "yfd_ctx_nranks": (cint, [vp]),
        "yfd_mesher_create": (cint, [cint, cint, P(vp)]),
        "yfd_mesher_create_structured": (cint, [cint, i64, i64, P(vp)]),
        "yfd_host_mesh_destroy": (None, [vp]),
        "yfd_host_mesh_sizes": (None, [vp, P(i64), P(i64), P(i64), P(cint)]),
        "yfd_host_mesh_points": (P(dbl), [vp]),
        "yfd_host_mesh_cells": (P(i64), [vp]),
        "yfd_host_mesh_faces": (P(i64), [vp]),
        "yfd_host_mesh_face_boundary": (P(C.c_uint8), [vp]),
        "yfd_host_mesh_face_boundary_id": (P(i64), [vp]),
        "yfd_mesh_create": (cint, [vp, cint, i64, _f64p, i64, _i64p, i64, _i64p, _u8p, cint, P(vp)]),
        "yfd_mesh_destroy": (None, [vp]),
        "yfd_mesh_sizes": (None, [vp, P(i64), P(i64), P(i64), P(cint)]),
        "yfd_mesh_get_face_owners": (cint, [vp, _i64p]),
        "yfd_mesh_get_cell_faces": (cint, [vp, _i64p]),
        "yfd_mesh_get_permutation": (cint, [vp, _i64p]),
        "yfd_mesh_set_points": (cint, [vp, _f64p]),
        "yfd_mesh_get_neighbours": (cint, [vp, _i64p]),
        "yfd_partition_create": (cint, [i64, cint, _i64p, cint, cint, P(vp)]),
        "yfd_partition_create_from_cuts": (cint, [i64, cint, cint, i64, _i64p, _i64p, P(vp)]),
        "yfd_partition_destroy": (None, [vp]),
        "yfd_partition_sizes": (None, [vp, P(i64), P(i64), P(i64), P(i64), P(cint)]),
        "yfd_partition_get": (cint, [vp, vp, vp, vp, vp, vp]),
        "yfd_system_local_sizes": (None, [vp, P(i64), P(i64), P(i64), P(i64)]),
        "yfd_csr_from_triplets": (cint, [vp, i64, i64, i64, _i64p, _i64p, _f64p, P(vp)]),
        "yfd_csr_from_arrays": (cint, [vp, i64, i64, i64, _i64p, _i64p, _f64p, P(vp)]),
        "yfd_csr_destroy": (None, [vp]),
        "yfd_csr_sizes": (None, [vp, P(i64), P(i64), P(i64)]),
        "yfd_csr_get": (cint, [vp, vp, vp, vp]),
        "yfd_csr_spmv": (cint, [vp, vp, _f64p, _f64p]),
        "yfd_dg_assemble_sip": (cint, [vp, vp, P(DgConfig), P(vp), P(AssemblyInfo)]),
        "yfd_dg_local_blocks": (cint, [vp, vp, P(DgConfig), vp, vp, vp, vp]),
        "yfd_dg_assemble_advection": (cint, [vp, vp, P(DgConfig), cint, P(vp), P(AssemblyInfo)]),
        "yfd_cfem_assemble_p1": (cint, [vp, vp, cint, P(vp), P(AssemblyInfo)]),
        "yfd_cfem_assemble_p2": (cint, [vp, vp, cint, P(vp), P(AssemblyInfo)]),
        "yfd_cfem_node_count": (cint, [vp, P(i64)]),
        "yfd_cfem_l2_error": (cint, [vp, vp, cint, cint, _f64p, P(C.c_double)]),
        "yfd_cfem_expand": (cint, [vp, _f64p, _f64p]),
        "yfd_cfem_get_compress_map": (cint, [vp, _i64p]),
        "yfd_system_destroy": (None, [vp]),
        "yfd_system_sizes": (None, [vp, P(i64), P(i64)]),
        "yfd_system_get_csr": (cint, [vp, vp, vp, vp]),
        "yfd_system_get_rhs": (cint, [vp, _f64p]),
        "yfd_system_get_pc_diag": (cint, [vp, _f64p]),
        "yfd_cg_params_default": (None, [P(CgParams)]),
        "yfd_system_solve_cg": (cint, [vp, vp, cint, P(CgParams), _f64p, P(CgInfo), vp, i64]),
        "yfd_system_solve_cg_device": (cint, [vp, vp, cint, P(CgParams), vp, P(CgInfo)]),
        "yfd_cg_solve": (cint, [vp, vp, _f64p, _f64p, vp, P(CgParams), P(CgInfo), vp, i64]),
        "yfd_bicgstab_solve": (cint, [vp, vp, _f64p, _f64p, vp, P(CgParams), P(CgInfo), vp, i64]),
        "yfd_system_solve_bicgstab": (cint, [vp, vp, cint, P(CgParams), _f64p, P(CgInfo), vp, i64]),
        "yfd_system_set_pmg_params": (cint, [vp, cint, cint, dbl]),
        "yfd_system_pmg_info": (cint, [vp, vp, P(PmgInfo)]),
        "yfd_system_set_operator_mode": (cint, [vp, cint]),
        "yfd_system_operator_info": (cint, [vp, P(OperatorInfo)]),
        "yfd_system_apply": (cint, [vp, vp, cint, _f64p, _f64p]),
        "yfd_system_solve_cg_owned": (cint, [vp, vp, cint, vp, vp, vp]),
        "yfd_system_owned_cells": (cint, [vp, _i64p]),
        "yfd_maxwell_get_dofs_owned": (cint, [vp, vp]),
        "yfd_maxwell_owned_cells": (cint, [vp, P(i64), vp]),
        "yfd_system_true_residual": (cint, [vp, vp, vp, vp, P(dbl)]),
        "yfd_system_true_residual_noalloc": (cint, [vp, vp, vp, P(dbl)]),
        "yfd_system_time_spmv": (cint, [vp, vp, cint, P(dbl), P(dbl), P(dbl)]),
        "yfd_dg_postprocess": (cint, [vp, vp, cint, cint, _f64p, P(dbl), P(dbl)]),
        "yfd_dg_postprocess_device": (cint, [vp, vp, cint, cint, vp, P(dbl), P(dbl)]),
        "yfd_maxwell_config_default": (None, [P(MaxwellConfig)]),
        "yfd_maxwell_create": (cint, [vp, vp, P(MaxwellConfig), vp, vp, P(vp), P(MaxwellInfo)]),
        "yfd_maxwell_destroy": (None, [vp]),
        "yfd_maxwell_get_info": (None, [vp, P(MaxwellInfo)]),
        "yfd_maxwell_set_time_integration": (cint, [vp, dbl, cint]),
        "yfd_maxwell_set_operator_mode": (cint, [vp, cint]),
        "yfd_maxwell_get_operator": (cint, [vp, vp, vp, vp, vp, vp]),
        "yfd_maxwell_apply_initial_condition": (cint, [vp, cint, dbl]),
        "yfd_maxwell_get_dofs": (cint, [vp, _f64p]),
        "yfd_maxwell_set_dofs": (cint, [vp, _f64p]),
        "yfd_maxwell_apply_operator": (cint, [vp, _f64p, _f64p]),
        "yfd_maxwell_timesteps": (cint, [vp, i64, P(MaxwellStepInfo)]),
        "yfd_maxwell_report_errors": (cint, [vp, cint, dbl, _f64p]),
        "yfd_maxwell_time_apply": (cint, [vp, cint, P(dbl), P(dbl), P(dbl)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Context:
    """One GPU rank (yfd_ctx).  `stream` = raw cudaStream_t (int) to launch on, e.g.
    torch.cuda.current_stream().cuda_stream, or None for a private stream."""

    def __init__(self, device=0, stream=None):
        self._h = C.c_void_p()
        st = lib().yfd_ctx_create(device, C.c_void_p(stream) if stream else None, C.byref(self._h))
        if st != 0:
            self._h = None
            raise YfdError(st, lib().yfd_status_string(st).decode())
        self.device = device
        self._children = weakref.WeakSet()     # meshes / systems / matrices: closed before the context goes

    def check(self, st):
        if st != 0:
            raise YfdError(st, lib().yfd_last_error(self._h).decode() or lib().yfd_status_string(st).decode())

    def close(self):
        if self._h:
            for ch in list(self._children):
                ch.close()
            lib().yfd_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def trim(self):
        n = C.c_size_t()
        self.check(lib().yfd_ctx_trim(self._h, C.byref(n)))
        return n.value

    def synchronize(self):
        self.check(lib().yfd_ctx_synchronize(self._h))

    @property
    def launches(self):
        return lib().yfd_ctx_launch_count(self._h)

    def device_info(self):
        name = C.create_string_buffer(256)
        sm, ma, mi = C.c_int(), C.c_int(), C.c_int()
        self.check(lib().yfd_ctx_device_info(self._h, name, 256, C.byref(sm), C.byref(ma), C.byref(mi)))
        return dict(name=name.value.decode(), sm_count=sm.value, cc=(ma.value, mi.value))

    def measure_fp64(self):
        """FP64 FMA peak of this device in TFLOP/s (DFMA microbenchmark, yfd_ctx_measure_fp64)."""
        t = C.c_double()
        self.check(lib().yfd_ctx_measure_fp64(self._h, C.byref(t)))
        return t.value

    def measure_fp64_tensor(self):
        """FP64 tensor-core (DMMA m8n8k4) peak of this device in TFLOP/s (yfd_ctx_measure_fp64_tensor)."""
        t = C.c_double()
        self.check(lib().yfd_ctx_measure_fp64_tensor(self._h, C.byref(t)))
        return t.value

    # ---- multi-GPU ------------------------------------------------------------------
    @staticmethod
    def comm_unique_id():
        buf = C.create_string_buffer(128)
        st = lib().yfd_comm_unique_id(buf)
        if st != 0:
            raise YfdError(st, "ncclGetUniqueId failed")
        return buf.raw

    def attach_comm(self, rank, nranks, unique_id):
        self.check(lib().yfd_ctx_attach_comm(self._h, rank, nranks, unique_id))

    # ---- assembly -------------------------------------------------------------------
    def dg_assemble_sip(self, mesh, degree, eta=1.0, rhs_fn=FN_SINSIN, dirichlet_fn=FN_ZERO, build_pc=False):
        cfg = DgConfig(degree, eta, rhs_fn, dirichlet_fn, int(build_pc))
        h = C.c_void_p()
        info = AssemblyInfo()
        self.check(lib().yfd_dg_assemble_sip(self._h, mesh._h, C.byref(cfg), C.byref(h), C.byref(info)))
        return System(self, h, info)

    def dg_assemble_advection(self, mesh, degree, eta=1.0, upwinding=False, build_pc=False):
        """run_advection_reaction_solver's cell loop + finalize (dg2d_advection.cpp:137-215); solve with
        CgParams(use_normal_eqns=1) as that driver does."""
        cfg = DgConfig(degree, eta, FN_ADVECTION_RHS, FN_ZERO, int(build_pc))
        h, info = C.c_void_p(), AssemblyInfo()
        self.check(lib().yfd_dg_assemble_advection(self._h, mesh._h, C.byref(cfg), int(upwinding), C.byref(h), C.byref(info)))
        return System(self, h, info)

    def dg_local_blocks(self, mesh, degree, eta=1.0, rhs_fn=FN_SINSIN, dirichlet_fn=FN_ZERO):
        cfg = DgConfig(degree, eta, rhs_fn, dirichlet_fn, 0)
        bs = (degree + 2) * (degree + 1) // 2
        nc, nf = mesh.ncells, mesh.nv
        K = np.empty((nc, bs, bs)); Att = np.empty((nc, nf, bs, bs)); Atn = np.empty((nc, nf, bs, bs)); rhs = np.empty((nc, bs))
        self.check(lib().yfd_dg_local_blocks(self._h, mesh._h, C.byref(cfg), _ptr(K), _ptr(Att), _ptr(Atn), _ptr(rhs)))
        return K, Att, Atn, rhs

    def dg_postprocess(self, mesh, degree, sol, ref_fn=FN_SINSIN_SOL, device_ptr=None):
        """(L2_errsq_qp, L2_errsq_mm) of run_diffusion_solver's post-process (dg2d_diffusion.cpp:225-268)."""
        qp, mm = C.c_double(), C.c_double()
        if device_ptr is not None:
            self.check(lib().yfd_dg_postprocess_device(self._h, mesh._h, degree, ref_fn, C.c_void_p(device_ptr), C.byref(qp), C.byref(mm)))
        else:
            sol = np.ascontiguousarray(sol, np.float64)
            self.check(lib().yfd_dg_postprocess(self._h, mesh._h, degree, ref_fn, sol, C.byref(qp), C.byref(mm)))
        return qp.value, mm.value

    def cfem_assemble_p1(self, mesh, rhs_fn=FN_SINSIN):
        h = C.c_void_p()
        info = AssemblyInfo()
        self.check(lib().yfd_cfem_assemble_p1(self._h, mesh._h, rhs_fn, C.byref(h), C.byref(info)))
        return System(self, h, info)

    def cfem_assemble_p2(self, mesh, rhs_fn=FN_SINSIN):
        """Continuous P2 (extension; the reference is P1-only, methods/cfem.hpp:200-204)."""
        h = C.c_void_p()
        info = AssemblyInfo()
        self.check(lib().yfd_cfem_assemble_p2(self._h, mesh._h, rhs_fn, C.byref(h), C.byref(info)))
        return System(self, h, info)

    def cfem_l2_error(self, mesh, order, expanded, ref_fn=FN_SINSIN_SOL):
        """sqrt of the summed cell integrals of (u_h - u)^2."""
        e = C.c_double()
        self.check(lib().yfd_cfem_l2_error(self._h, mesh._h, order, ref_fn, np.ascontiguousarray(expanded, np.float64), C.byref(e)))
        return float(np.sqrt(e.value))


class HostMesh:
    """Host arrays of a mesh in the reference's numbering."""

    def __init__(self, cell_type, points, cells, faces, face_boundary, face_boundary_id=None):
        self.cell_type = cell_type
        self.points = np.ascontiguousarray(points, np.float64)
        self.cells = np.ascontiguousarray(cells, np.int64)
        self.faces = np.ascontiguousarray(faces, np.int64)
        self.face_boundary = np.ascontiguousarray(face_boundary, np.uint8)
        self.face_boundary_id = face_boundary_id

    @classmethod
    def create(cls, cell_type, levels):
        """mesher<Mesh>::create_mesh(msh, levels) (core/meshers.hpp:130-179, :302-361)."""
        L = lib()
        h = C.c_void_p()
        st = L.yfd_mesher_create(cell_type, levels, C.byref(h))
        return cls._adopt(cell_type, h, st)

    @classmethod
    def structured(cls, cell_type, nx, ny=None):
        """nx x ny structured mesh of the unit square in the reference's conventions (yfd_mesher_create_structured;
        new: the reference mesher only offers 4^(r+1) cells)."""
        L = lib()
        h = C.c_void_p()
        st = L.yfd_mesher_create_structured(cell_type, nx, nx if ny is None else ny, C.byref(h))
        return cls._adopt(cell_type, h, st)

    @classmethod
    def _adopt(cls, cell_type, h, st):
        L = lib()
        if st != 0:
            raise YfdError(st, L.yfd_status_string(st).decode())
        try:
            a, b, c, nv = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int()
            L.yfd_host_mesh_sizes(h, C.byref(a), C.byref(b), C.byref(c), C.byref(nv))
            npts, nc, nf, nv = a.value, b.value, c.value, nv.value
            as_arr = np.ctypeslib.as_array
            m = cls(cell_type,
                    as_arr(L.yfd_host_mesh_points(h), (npts, 2)).copy(),
                    as_arr(L.yfd_host_mesh_cells(h), (nc, nv)).copy(),
                    as_arr(L.yfd_host_mesh_faces(h), (nf, 2)).copy(),
                    as_arr(L.yfd_host_mesh_face_boundary(h), (nf,)).copy(),
                    as_arr(L.yfd_host_mesh_face_boundary_id(h), (nf,)).copy())
        finally:
            L.yfd_host_mesh_destroy(h)
        return m


class Mesh:
    """Device mesh + connectivity (yfd_mesh)."""

    def __init__(self, ctx, host, reorder=REORDER_MORTON):
        self.ctx = ctx
        self.host = host
        self._h = C.c_void_p()
        self.nv = host.cells.shape[1]
        self.npoints, self.ncells, self.nfaces = len(host.points), len(host.cells), len(host.faces)
        ctx.check(lib().yfd_mesh_create(ctx._h, host.cell_type, self.npoints, host.points, self.ncells, host.cells,
                                        self.nfaces, host.faces, host.face_boundary, reorder, C.byref(self._h)))
        ctx._children.add(self)

    def close(self):
        if self._h:
            lib().yfd_mesh_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def face_owners(self):
        out = np.empty((self.nfaces, 2), np.int64)
        self.ctx.check(lib().yfd_mesh_get_face_owners(self._h, out))
        return out

    def cell_faces(self):
        out = np.empty((self.ncells, self.nv), np.int64)
        self.ctx.check(lib().yfd_mesh_get_cell_faces(self._h, out))
        return out

    def permutation(self):
        out = np.empty(self.ncells, np.int64)
        self.ctx.check(lib().yfd_mesh_get_permutation(self._h, out))
        return out

    def neighbours(self):
        """[ncells][nf] internal ids of the cell across each local face (-1: boundary), internal cell order."""
        out = np.empty((self.ncells, self.nv), np.int64)
        self.ctx.check(lib().yfd_mesh_get_neighbours(self._h, out))
        return out

    def set_points(self, points):
        points = np.ascontiguousarray(points, np.float64)
        assert points.shape == (self.npoints, 2)
        self.ctx.check(lib().yfd_mesh_set_points(self._h, points))


def _read_partition(h):
    try:
        b, e, nh, ns, npr = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64(), C.c_int()
        lib().yfd_partition_sizes(h, C.byref(b), C.byref(e), C.byref(nh), C.byref(ns), C.byref(npr))
        halo = np.empty(nh.value, np.int64); send = np.empty(ns.value, np.int64)
        peers = np.empty(npr.value, np.int32); rc = np.empty(npr.value, np.int64); sc = np.empty(npr.value, np.int64)
        lib().yfd_partition_get(h, _ptr(halo), _ptr(send), _ptr(peers), _ptr(rc), _ptr(sc))
    finally:
        lib().yfd_partition_destroy(h)
    return dict(cell_begin=b.value, cell_end=e.value, halo_cells=halo, send_cells=send, peers=peers, recv_counts=rc, send_counts=sc)


def partition_plan(neighbours, nranks, rank):
    """yfd_partition_create (host, no CUDA): contiguous blocks of the internal cell order + halo / send lists."""
    nbr = np.ascontiguousarray(neighbours, np.int64)
    h = C.c_void_p()
    st = lib().yfd_partition_create(nbr.shape[0], nbr.shape[1], nbr, nranks, rank, C.byref(h))
    if st != 0:
        raise YfdError(st, lib().yfd_status_string(st).decode())
    return _read_partition(h)


def partition_plan_from_cuts(ncells, nranks, rank, cut_cell, cut_nbr):
    """yfd_partition_create_from_cuts: the same plan from the (cell, neighbour) pairs that straddle two ranks' blocks."""
    cc = np.ascontiguousarray(cut_cell, np.int64); cn = np.ascontiguousarray(cut_nbr, np.int64)
    assert cc.shape == cn.shape and cc.ndim == 1
    h = C.c_void_p()
    st = lib().yfd_partition_create_from_cuts(ncells, nranks, rank, cc.size, cc, cn, C.byref(h))
    if st != 0:
        raise YfdError(st, lib().yfd_status_string(st).decode())
    return _read_partition(h)


def _solve_common(ctx, call, n, params, x0, history):
    params = params or CgParams()
    x = np.zeros(n, np.float64) if x0 is None else np.ascontiguousarray(x0, np.float64).copy()
    info = CgInfo()
    hist = None
    cap = 0
    if history:
        params.save_iteration_history = 1
        cap = int(min(params.max_iter + 1, 50_000_000))
        hist = np.zeros(cap, np.float64)
    ctx.check(call(params, x, info, _ptr(hist), cap))
    out = dict(x=x, iterations=info.iterations, nr0=info.nr0, nr=info.nr, converged=bool(info.converged),
               ms_solve=info.ms_solve, transport=("one GPU", "NCCL", "peer memory")[info.transport])
    if history:
        out["history"] = hist[:info.history_len].copy()
    return out


class Csr:
    """blaze::CompressedMatrix<double> on the device (yfd_csr)."""

    def __init__(self, ctx, handle):
        self.ctx, self._h = ctx, handle
        ctx._children.add(self)
        a, b, c = C.c_int64(), C.c_int64(), C.c_int64()
        lib().yfd_csr_sizes(handle, C.byref(a), C.byref(b), C.byref(c))
        self.nrows, self.ncols, self.nnz = a.value, b.value, c.value

    @classmethod
    def from_triplets(cls, ctx, nrows, ncols, rows, cols, vals):
        """blaze::init_from_triplets (core/blaze_sparse_init.hpp:125-155)."""
        rows = np.ascontiguousarray(rows, np.int64); cols = np.ascontiguousarray(cols, np.int64)
        vals = np.ascontiguousarray(vals, np.float64)
        h = C.c_void_p()
        ctx.check(lib().yfd_csr_from_triplets(ctx._h, nrows, ncols, len(vals), rows, cols, vals, C.byref(h)))
        return cls(ctx, h)

    @classmethod
    def from_arrays(cls, ctx, nrows, ncols, rowptr, colidx, values):
        rowptr = np.ascontiguousarray(rowptr, np.int64); colidx = np.ascontiguousarray(colidx, np.int64)
        values = np.ascontiguousarray(values, np.float64)
        h = C.c_void_p()
        ctx.check(lib().yfd_csr_from_arrays(ctx._h, nrows, ncols, len(values), rowptr, colidx, values, C.byref(h)))
        return cls(ctx, h)

    def close(self):
        if self._h:
            lib().yfd_csr_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def get(self):
        rowptr = np.empty(self.nrows + 1, np.int64); colidx = np.empty(self.nnz, np.int64); values = np.empty(self.nnz, np.float64)
        self.ctx.check(lib().yfd_csr_get(self._h, _ptr(rowptr), _ptr(colidx), _ptr(values)))
        return rowptr, colidx, values

    def spmv(self, x):
        x = np.ascontiguousarray(x, np.float64)
        y = np.empty(self.nrows, np.float64)
        self.ctx.check(lib().yfd_csr_spmv(self.ctx._h, self._h, x, y))
        return y

    def solve_cg(self, b, params=None, pc_diag=None, x0=None, history=False):
        """conjugated_gradient(cgp, A, b, x[, iM]) (core/solvers.hpp:46,141); iM diagonal."""
        b = np.ascontiguousarray(b, np.float64)
        if b.shape != (self.nrows,):
            raise YfdError(8, "[CG solver] Wrong size of RHS vector")
        pc = None if pc_diag is None else np.ascontiguousarray(pc_diag, np.float64)
        call = lambda p, x, info, h, cap: lib().yfd_cg_solve(self.ctx._h, self._h, b, x, _ptr(pc), C.byref(p), C.byref(info), h, cap)
        return _solve_common(self.ctx, call, self.nrows, params, x0, history)

    def solve_bicgstab(self, b, params=None, pc_diag=None, x0=None, history=False):
        """bicgstab(cgp, A, b, x[, iM]) (core/solvers.hpp:246,356); iM diagonal."""
        b = np.ascontiguousarray(b, np.float64)
        if b.shape != (self.nrows,):
            raise YfdError(8, "[CG solver] Wrong size of RHS vector")
        pc = None if pc_diag is None else np.ascontiguousarray(pc_diag, np.float64)
        call = lambda p, x, info, h, cap: lib().yfd_bicgstab_solve(self.ctx._h, self._h, b, x, _ptr(pc), C.byref(p), C.byref(info), h, cap)
        if x0 is not None:
            params = params or CgParams()
            params.use_initial_guess = 1
        return _solve_common(self.ctx, call, self.nrows, params, x0, history)


class System:
    """An assembled problem on the device: assm.lhs / assm.rhs / assm.pc (yfd_system)."""

    def __init__(self, ctx, handle, info):
        self.ctx, self._h, self.info = ctx, handle, info
        ctx._children.add(self)
        self.n, self.nnz = info.system_size, info.nnz

    def close(self):
        if self._h:
            lib().yfd_system_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def csr(self, values=True):
        rowptr = np.empty(self.n + 1, np.int64); colidx = np.empty(self.nnz, np.int64)
        vals = np.empty(self.nnz, np.float64) if values else None
        self.ctx.check(lib().yfd_system_get_csr(self._h, _ptr(rowptr), _ptr(colidx), _ptr(vals)))
        return rowptr, colidx, vals

    def rhs(self):
        out = np.empty(self.n, np.float64)
        self.ctx.check(lib().yfd_system_get_rhs(self._h, out))
        return out

    def pc_diag(self):
        out = np.empty(self.n, np.float64)
        self.ctx.check(lib().yfd_system_get_pc_diag(self._h, out))
        return out

    def solve_cg(self, params=None, use_pc=False, x0=None, history=False):
        call = lambda p, x, info, h, cap: lib().yfd_system_solve_cg(self.ctx._h, self._h, int(use_pc), C.byref(p), x, C.byref(info), h, cap)
        if x0 is not None:
            params = params or CgParams()
            params.use_initial_guess = 1
        return _solve_common(self.ctx, call, self.n, params, x0, history)

    def solve_bicgstab(self, params=None, use_pc=False, x0=None, history=False):
        """bicgstab on the assembled system, matrix resident on the device (core/solvers.hpp:246-467)."""
        call = lambda p, x, info, h, cap: lib().yfd_system_solve_bicgstab(self.ctx._h, self._h, int(use_pc), C.byref(p), x, C.byref(info), h, cap)
        if x0 is not None:
            params = params or CgParams()
            params.use_initial_guess = 1
        return _solve_common(self.ctx, call, self.n, params, x0, history)

    def solve_cg_into(self, params, x, use_pc=False):
        """solve_cg writing into the caller's host array `x` (e.g. pinned memory); x0 = 0 unless use_initial_guess."""
        assert x.dtype == np.float64 and x.flags.c_contiguous and x.shape == (self.n,)
        info = CgInfo()
        self.ctx.check(lib().yfd_system_solve_cg(self.ctx._h, self._h, int(use_pc), C.byref(params), x, C.byref(info), None, 0))
        return dict(iterations=info.iterations, nr0=info.nr0, nr=info.nr, converged=bool(info.converged), ms_solve=info.ms_solve,
                    transport=("one GPU", "NCCL", "peer memory")[info.transport])

    def solve_cg_owned_into(self, params, x_owned, use_pc=False):
        """yfd_system_solve_cg_owned: this rank's rows only (internal cell order, see owned_cells), into the caller's host array."""
        assert x_owned.dtype == np.float64 and x_owned.flags.c_contiguous and x_owned.shape == (self.local_sizes()["n_local"],)
        info = CgInfo()
        self.ctx.check(lib().yfd_system_solve_cg_owned(self.ctx._h, self._h, int(use_pc), C.byref(params), x_owned.ctypes.data_as(C.c_void_p), C.byref(info)))
        return dict(iterations=info.iterations, nr0=info.nr0, nr=info.nr, converged=bool(info.converged), ms_solve=info.ms_solve,
                    transport=("one GPU", "NCCL", "peer memory")[info.transport])

    def owned_cells(self):
        """REFERENCE ids of the cells whose rows this rank owns, in the order solve_cg_owned_into returns them (DG systems)."""
        out = np.empty(self.local_sizes()["n_local"] // self.info.basis_size, np.int64)
        self.ctx.check(lib().yfd_system_owned_cells(self._h, out))
        return out

    def solve_cg_device(self, params=None, use_pc=False, d_x=None):
        """Resident solve: x0 = 0, solution left in the device buffer `d_x` (raw pointer int or None)."""
        params = params or CgParams()
        info = CgInfo()
        self.ctx.check(lib().yfd_system_solve_cg_device(self.ctx._h, self._h, int(use_pc), C.byref(params),
                                                        C.c_void_p(d_x) if d_x else None, C.byref(info)))
        return dict(iterations=info.iterations, nr0=info.nr0, nr=info.nr, converged=bool(info.converged), ms_solve=info.ms_solve,
                    transport=("one GPU", "NCCL", "peer memory")[info.transport])

    def set_pmg_params(self, nu=0, coarse_degree=0, smooth_ratio=0.0):
        """Parameters of the p-multigrid preconditioner (use_pc=PC_PMG); 0 = default."""
        self.ctx.check(lib().yfd_system_set_pmg_params(self._h, nu, coarse_degree, smooth_ratio))

    def pmg_info(self):
        pi = PmgInfo()
        self.ctx.check(lib().yfd_system_pmg_info(self.ctx._h, self._h, C.byref(pi)))
        return dict(nlevels=pi.nlevels, nu=pi.nu, coarse_degree=pi.coarse_degree, smooth_ratio=pi.smooth_ratio,
                    degree=list(pi.degree)[:pi.nlevels], lambda_max=list(pi.lambda_max)[:pi.nlevels])

    def set_operator_mode(self, mode):
        """OPERATOR_AUTO / OPERATOR_STREAMED / OPERATOR_CLASSES (include/yfd.h)."""
        self.ctx.check(lib().yfd_system_set_operator_mode(self._h, mode))

    def operator_info(self):
        oi = OperatorInfo()
        self.ctx.check(lib().yfd_system_operator_info(self._h, C.byref(oi)))
        return {k: getattr(oi, k) for k, _ in OperatorInfo._fields_}

    def apply(self, x, plain=False):
        """y = A x (reference numbering); plain=True: the independent one-thread-per-row kernel."""
        x = np.ascontiguousarray(x, np.float64)
        assert x.shape == (self.n,)
        y = np.empty(self.n, np.float64)
        self.ctx.check(lib().yfd_system_apply(self.ctx._h, self._h, 1 if plain else 0, x, y))
        return y

    def true_residual(self, x=None, d_x=None):
        """||b - A x|| / ||b|| with the independent kernel; x = host array or d_x = raw device pointer (reference numbering)."""
        r = C.c_double()
        if x is not None:
            x = np.ascontiguousarray(x, np.float64)
            assert x.shape == (self.n,)
        self.ctx.check(lib().yfd_system_true_residual(self.ctx._h, self._h, _ptr(x), C.c_void_p(d_x) if d_x else None, C.byref(r)))
        return r.value

    def true_residual_noalloc(self, d_x):
        """yfd_system_true_residual_noalloc: the same number from a call that allocates nothing (static device scratch,
        at most NOALLOC_MAX_ROWS rows, one GPU); d_x = raw device pointer, reference numbering.  No reference counterpart."""
        r = C.c_double()
        self.ctx.check(lib().yfd_system_true_residual_noalloc(self.ctx._h, self._h, C.c_void_p(d_x), C.byref(r)))
        return r.value

    def local_sizes(self):
        a, b, c, d = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
        lib().yfd_system_local_sizes(self._h, C.byref(a), C.byref(b), C.byref(c), C.byref(d))
        return dict(n_local=a.value, nnz_local=b.value, n_halo=c.value, row_begin=d.value)

    def time_spmv(self, reps=20):
        ms, b1, b2 = C.c_double(), C.c_double(), C.c_double()
        self.ctx.check(lib().yfd_system_time_spmv(self.ctx._h, self._h, reps, C.byref(ms), C.byref(b1), C.byref(b2)))
        return dict(ms=ms.value, csr_bytes=b1.value, kernel_bytes=b2.value)

    # CFEM only
    def expand(self, sol):
        sol = np.ascontiguousarray(sol, np.float64)
        a, b = C.c_int64(), C.c_int64()
        out = np.empty(self._npoints(), np.float64)
        self.ctx.check(lib().yfd_cfem_expand(self._h, sol, out))
        return out

    def compress_map(self):
        out = np.empty(self._npoints(), np.int64)
        self.ctx.check(lib().yfd_cfem_get_compress_map(self._h, out))
        return out

    def _npoints(self):
        n = C.c_int64()
        self.ctx.check(lib().yfd_cfem_node_count(self._h, C.byref(n)))
        return n.value


class Maxwell:
    """maxwell_context + assemble / apply_initial_condition / do_timestep / report_errors
    (methods/dg_maxwell_2D.hpp, maxwell_driver.cpp:133-272) on the device (yfd_maxwell)."""

    def __init__(self, ctx, mesh, degree, delta_t=0.1, eta=1.0, upwind=False, euler=False, mu_r=None, eps_r=None):
        self.ctx = ctx
        cfg = MaxwellConfig(degree, delta_t, eta, int(upwind), TI_EULER if euler else TI_RK4)
        h = C.c_void_p()
        self.info = MaxwellInfo()
        mu = None if mu_r is None else np.ascontiguousarray(mu_r, np.float64)
        ep = None if eps_r is None else np.ascontiguousarray(eps_r, np.float64)
        self._h = None
        ctx.check(lib().yfd_maxwell_create(ctx._h, mesh._h, C.byref(cfg), _ptr(mu), _ptr(ep), C.byref(h), C.byref(self.info)))
        self._h = h
        ctx._children.add(self)
        self.ncells, self.bs, self.nf, self.ndofs = (self.info.ncells, self.info.basis_size, self.info.faces_per_elem,
                                                      self.info.ndofs)
        self.delta_t = delta_t

    def close(self):
        if self._h:
            lib().yfd_maxwell_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_operator_mode(self, mode):
        """OPERATOR_AUTO / OPERATOR_STREAMED / OPERATOR_CLASSES: which operator storage the time loop uses."""
        self.ctx.check(lib().yfd_maxwell_set_operator_mode(self._h, mode))
        lib().yfd_maxwell_get_info(self._h, C.byref(self.info))

    def set_time_integration(self, delta_t, euler=False):
        self.ctx.check(lib().yfd_maxwell_set_time_integration(self._h, delta_t, TI_EULER if euler else TI_RK4))
        self.delta_t = delta_t

    def operator(self, blocks=True):
        nc, nf, bs = self.ncells, self.nf, self.bs
        L = 3 * bs
        on = np.empty((nc, L, L)) if blocks else None
        off = np.empty((nc, nf, L, L)) if blocks else None
        nb = np.empty((nc, nf), np.int64)
        has = np.empty((nc, nf), np.int32)
        gM = np.empty((nc, bs, bs))
        self.ctx.check(lib().yfd_maxwell_get_operator(self._h, _ptr(on), _ptr(off), _ptr(nb), _ptr(has), _ptr(gM)))
        return dict(ondiag=on, offdiag=off, neigh=nb, has=has, gM=gM)

    def apply_initial_condition(self, fields=FIELDS_CAVITY_11, t=0.0):
        self.ctx.check(lib().yfd_maxwell_apply_initial_condition(self._h, fields, t))

    def dofs(self):
        d = np.empty(self.ndofs)
        self.ctx.check(lib().yfd_maxwell_get_dofs(self._h, d))
        return d

    def ncells_owned(self):
        nc = C.c_int64()
        self.ctx.check(lib().yfd_maxwell_owned_cells(self._h, C.byref(nc), None))
        return nc.value

    def owned_cells(self):
        """REFERENCE ids of the cells this rank owns, in the order dofs_owned returns them."""
        nc = C.c_int64()
        self.ctx.check(lib().yfd_maxwell_owned_cells(self._h, C.byref(nc), None))
        out = np.empty(nc.value, np.int64)
        self.ctx.check(lib().yfd_maxwell_owned_cells(self._h, C.byref(nc), out.ctypes.data_as(C.c_void_p)))
        return out

    def dofs_owned(self, out=None):
        """yfd_maxwell_get_dofs_owned: [ncells_owned][3][bs], no gather (one GPU: all cells, internal order)."""
        nc = C.c_int64()
        self.ctx.check(lib().yfd_maxwell_owned_cells(self._h, C.byref(nc), None))
        if out is None:
            out = np.empty(nc.value * 3 * self.bs)
        assert out.dtype == np.float64 and out.flags.c_contiguous and out.size == nc.value * 3 * self.bs
        self.ctx.check(lib().yfd_maxwell_get_dofs_owned(self._h, out.ctypes.data_as(C.c_void_p)))
        return out

    def set_dofs(self, d):
        d = np.ascontiguousarray(d, np.float64)
        if d.size != self.ndofs:
            raise YfdError(8, "wrong size of the dof vector")
        self.ctx.check(lib().yfd_maxwell_set_dofs(self._h, d))

    def apply_operator(self, v):
        v = np.ascontiguousarray(v, np.float64)
        if v.size != self.ndofs:
            raise YfdError(8, "wrong size of the dof vector")
        out = np.empty(self.ndofs)
        self.ctx.check(lib().yfd_maxwell_apply_operator(self._h, v, out))
        return out

    def timesteps(self, nsteps=1):
        info = MaxwellStepInfo()
        self.ctx.check(lib().yfd_maxwell_timesteps(self._h, nsteps, C.byref(info)))
        return info

    def report_errors(self, t, fields=FIELDS_CAVITY_11):
        e = np.empty(3)
        self.ctx.check(lib().yfd_maxwell_report_errors(self._h, fields, t, e))
        return e

    def time_apply(self, reps=20):
        ms, rb, kb = C.c_double(), C.c_double(), C.c_double()
        self.ctx.check(lib().yfd_maxwell_time_apply(self._h, reps, C.byref(ms), C.byref(rb), C.byref(kb)))
        return ms.value, rb.value, kb.value

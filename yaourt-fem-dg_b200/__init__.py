"""yaourt-fem-dg_b200: B200 (sm_100a) implementation of Yaourt-FEM-DG's assemble-and-solve hot path.

Everything computes in libyfd_b200.so (hand-written CUDA behind the C ABI of include/yfd.h); this
package is the ctypes binding plus thin host-side objects.  There is no CPU fallback: importing
works without a GPU (so the ABI can be inspected), creating a Context does not.
"""
from .capi import (  # noqa: F401
    TRI, QUAD, FN_ZERO, FN_SINSIN, FN_ONE, FN_SINSIN_SOL, FN_ADVECTION_RHS, FN_SIN_X_SOL, REORDER_NONE, REORDER_MORTON,
    YfdError, lib, lib_path, Context, HostMesh, Mesh, Csr, System, CgParams, CgInfo, AssemblyInfo,
    declared_symbols, partition_plan, partition_plan_from_cuts, Maxwell, MaxwellConfig, MaxwellInfo, MaxwellStepInfo,
    TI_RK4, TI_EULER, FIELDS_CAVITY_11, FIELDS_ZERO, PC_NONE, PC_JACOBI, PC_BLOCK_JACOBI,
    OPERATOR_AUTO, OPERATOR_STREAMED, OPERATOR_CLASSES, OperatorInfo, PC_PMG, PmgInfo, NOALLOC_MAX_ROWS,
)

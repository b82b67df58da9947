#!/usr/bin/env python3
"""Generate tests/golden/struct_*.npz: the UNMODIFIED reference (oracle/_ref/libyfd_ref.so) run on nx x ny structured meshes
it did not make itself -- the mesh comes from yfd_mesher_create_structured (host C++, no GPU needed) and is handed to the
reference through its public points / faces / cells vectors (oracle/ref_harness.cpp :: yfd_ref_mesh_set); connectivity, the
DG-SIP system and the CG solution are the reference's own.  Pins the structured mesher and the oracle on sizes between the
reference mesher's refinement levels (BASELINE configs "64x64", "~20M dofs").

    python tools/make_golden_structured.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import refbind  # noqa: E402
import yaourt_fem_dg_b200 as yfd  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
R = refbind.Ref()
for name, t, nx, ny, k, eta, pc in [("tri_5x7_k2", 0, 5, 7, 2, 2.0, False), ("quad_6x4_k3", 1, 6, 4, 3, 2.0, False), ("tri_8x8_k1_pc", 0, 8, 8, 1, 1.0, True)]:
    hm = yfd.HostMesh.structured(t, nx, ny)
    m = R.set_mesh(t, hm.points, hm.cells, hm.faces, hm.face_boundary, hm.face_boundary_id)
    R.set_cg(rr_tol=1e-12, rr_max=1e4, max_iter=-1)
    o = R.run_dg(k, eta=eta, use_pc=pc)
    np.savez_compressed(os.path.join(OUT, "struct_%s.npz" % name), mtype=t, nx=nx, ny=ny, degree=k, eta=eta, use_pc=int(pc),
                        points=hm.points, cells=hm.cells, faces=hm.faces, face_boundary=hm.face_boundary.astype(np.uint8),
                        face_boundary_id=hm.face_boundary_id, face_owners=m["face_owners"],
                        rowptr=o["rowptr"], colidx=o["colidx"], values=o["values"], rhs=o["rhs"], sol=o["sol"], iters=o["iters"],
                        pc_diag=o["pc_diag"] if pc else np.zeros(0), l2_qp=o["l2_qp"])
    print(name, "cells", len(hm.cells), "n", o["n"], "nnz", o["nnz"], "iters", o["iters"], "L2", o["l2_qp"])

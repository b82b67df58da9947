#!/usr/bin/env python3
"""Generate tests/golden/adv_*.npz by RUNNING the unmodified dg2d_advection.cpp (oracle/_ref/libyfd_ref_adv.so, built from
/root/reference by oracle/Makefile): mesh (shattered points dumped), assembled CSR, rhs, the driver's CG-on-normal-equations
residual history (first 30 iterations) and its L2 errors at the driver's own settings (rr_tol 1e-8, max_iter 2N).

    python tools/make_golden_advection.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import refbind  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
R = refbind.Ref(advection=True)
for name, t, r, k, up, sh in [("tri_r2_k1", 0, 2, 1, 0, 0), ("tri_r1_k2_up_shatter", 0, 1, 2, 1, 1), ("quad_r1_k3_up", 1, 1, 3, 1, 0),
                              ("quad_r2_k1_shatter", 1, 2, 1, 0, 1)]:
    m = R.mesh(t, r, shatter=bool(sh))
    R.set_cg(rr_tol=1e-8, rr_max=1e4, max_iter=30)
    o = R.run_dg(k, eta=1.0, upwinding=bool(up))
    R.set_cg(rr_tol=1e-8, rr_max=1e4, max_iter=-1)
    full = R.run_dg(k, eta=1.0, upwinding=bool(up))
    np.savez_compressed(os.path.join(OUT, "adv_%s.npz" % name), mtype=t, levels=r, degree=k, upwinding=up, shatter=sh, eta=1.0,
                        points=m["points"], cells=m["cells"], faces=m["faces"], face_boundary=m["face_boundary"].astype(np.uint8),
                        rowptr=o["rowptr"], colidx=o["colidx"], values=o["values"], rhs=o["rhs"], history30=o["history"],
                        sol30=o["sol"], iters=full["iters"], l2_qp=full["l2_qp"], l2_mm=full["l2_mm"])
    print(name, o["n"], o["nnz"], full["iters"], full["l2_qp"])

#!/usr/bin/env python3
"""Generate tests/golden/bicgstab.npz by RUNNING the reference's own bicgstab (core/solvers.hpp:246-467, through
oracle/_ref/libyfd_ref_adv.so :: yfd_ref_bicgstab) on the advection-reaction systems already stored in
tests/golden/adv_*.npz (assembled by the unmodified dg2d_advection.cpp) and on two SPD systems of tests/golden/dg_*.npz, without and with the Jacobi iM the drivers build.
Stored per system: the reference's history file (nr/nr0, 6 digits), its iteration count and its solution.

    python tools/make_golden_bicgstab.py
"""
import glob
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import refbind  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
R = refbind.Ref(advection=True)
out = {}
# non-symmetric systems from dg2d_advection.cpp plus two SPD systems from dg2d_diffusion.cpp (tests/golden/dg_*.npz)
PATHS = sorted(glob.glob(os.path.join(GOLD, "adv_*.npz"))) + [os.path.join(GOLD, f) for f in ("dg_quad_r1_k3.npz", "dg_tri_r1_k2_pc.npz")]
for path in PATHS:
    name = os.path.basename(path)[:-4]
    g = np.load(path)
    n = len(g["rhs"])
    diag = np.array([g["values"][g["rowptr"][i]:g["rowptr"][i + 1]][g["colidx"][g["rowptr"][i]:g["rowptr"][i + 1]] == i][0] for i in range(n)])
    for tag, pc in (("plain", None), ("jacobi", 1.0 / diag)):
        res = R.bicgstab(g["rowptr"], g["colidx"], g["values"], g["rhs"], pc_diag=pc, rr_tol=1e-10, rr_max=1e8, max_iter=4 * n)
        out["%s_%s_history" % (name, tag)] = res["history"]
        out["%s_%s_x" % (name, tag)] = res["x"]
        out["%s_%s_iters" % (name, tag)] = res["iterations"]
        print(name, tag, "n", n, "iterations", res["iterations"], "final rr", res["history"][-1])
np.savez_compressed(os.path.join(GOLD, "bicgstab.npz"), **out)

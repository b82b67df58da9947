#!/usr/bin/env python3
"""Generate tests/golden/maxwell_*.npz by RUNNING the unmodified reference DG Maxwell solver
(oracle/_ref/libyfd_ref_maxwell.so = methods/dg_maxwell_2D.hpp + maxwell_driver.cpp built from /root/reference by
oracle/Makefile).  Dev-time tool: needs /root/reference at build time, the committed fixtures do not.  Every array is
an output of the reference's own code: maxwell_context (mesh, shatter), assemble (gM, gOp_ondiag, gOp_offdiag,
offdiag_neigh_offsets), apply_initial_condition (gDofs), do_timestep, report_errors, and run_maxwell_solver's error
log.  LAPACK dgesv comes from oracle/lu_shim.c (no system LAPACK in this image): the fixtures pin the reference's
algorithm; tests/test_oracle_lapack.py re-derives them with the same sources linked against a real LAPACK (OpenBLAS 0.3.15,
`make -C oracle lapack_pin`) and finds them reproduced to <= 1.6e-14, which closes SURVEY.md 8c's "parity unpinned".

    python tools/make_golden_maxwell.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import refbind  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
R = refbind.RefMaxwell()

# (name, mesh type, levels, degree, upwind, shatter, euler, dt, steps)
CASES = [
    ("tri_r1_k1", 0, 1, 1, 0, 0, 0, 0.005, 10),
    ("tri_r1_k2_up_shatter", 0, 1, 2, 1, 1, 0, 0.005, 10),
    ("quad_r1_k3", 1, 1, 3, 0, 0, 0, 0.002, 10),
    ("tri_r0_k4_up", 0, 0, 4, 1, 0, 0, 0.002, 10),
    ("quad_r1_k2_up_shatter", 1, 1, 2, 1, 1, 0, 0.005, 10),
    ("tri_r1_k1_euler", 0, 1, 1, 0, 0, 1, 0.001, 10),
]
for name, t, r, k, up, sh, eu, dt, steps in CASES:
    R.create(t, r, k, dt=dt, eta=1.0, upwind=up, euler=eu, shatter=sh)
    m = R.mesh()
    R.assemble()
    op = R.operator()
    R.init()
    d0 = R.dofs()
    errs = []
    R.step(1)
    d1 = R.dofs()
    errs.append(R.errors(0))
    R.step(steps - 1)
    dN = R.dofs()
    errs.append(R.errors(steps - 1))
    np.savez_compressed(os.path.join(OUT, "maxwell_%s.npz" % name),
                        mtype=t, levels=r, degree=k, upwind=up, shatter=sh, euler=eu, dt=dt, steps=steps, eta=1.0,
                        points=m["points"], cells=m["cells"], faces=m["faces"], face_boundary=m["face_boundary"].astype(np.uint8),
                        ondiag=op["ondiag"], offdiag=op["offdiag"], neigh=op["neigh"], has=op["has"], gM=op["gM"],
                        dofs0=d0, dofs1=d1, dofsN=dN, errors=np.array(errs))
    print(name, "cells", R.ncells, "bs", R.bs, "errors", errs[-1])

# non-uniform material (mu_r, eps_r are public members of maxwell_context, methods/dg_maxwell_2D.hpp:101)
R.create(0, 1, 2, dt=0.005, eta=1.0, upwind=1)
m = R.mesh()
rng = np.random.default_rng(7)
mu = 1.0 + rng.random(R.ncells)
ep = 1.0 + 3.0 * rng.random(R.ncells)
R.set_material(mu, ep)
R.assemble()
op = R.operator()
R.init()
d0 = R.dofs()
R.step(5)
np.savez_compressed(os.path.join(OUT, "maxwell_tri_r1_k2_up_material.npz"),
                    mtype=0, levels=1, degree=2, upwind=1, shatter=0, euler=0, dt=0.005, steps=5, eta=1.0, mu_r=mu, eps_r=ep,
                    points=m["points"], cells=m["cells"], faces=m["faces"], face_boundary=m["face_boundary"].astype(np.uint8),
                    ondiag=op["ondiag"], offdiag=op["offdiag"], neigh=op["neigh"], has=op["has"], gM=op["gM"],
                    dofs0=d0, dofsN=R.dofs())

# the driver exactly as shipped: error log of run_maxwell_solver (maxwell_driver.cpp:194-272)
logs = {}
for name, t, r, k, up, dt, steps in [("tri_r2_k2", 0, 2, 2, 0, 0.01, 20), ("quad_r2_k1_up", 1, 2, 1, 1, 0.01, 20),
                                     ("tri_r3_k3_up", 0, 3, 3, 1, 0.002, 10)]:
    logs[name] = R.run_driver(t, r, k, dt, steps, upwind=up)
    logs[name + "_cfg"] = np.array([t, r, k, up, dt, steps], np.float64)
    print(name, logs[name][-1])
np.savez_compressed(os.path.join(OUT, "maxwell_driver_logs.npz"), **logs)

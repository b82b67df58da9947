#!/usr/bin/env python3
"""Small pass over every hand-rolled synchronisation / bulk-copy path, sized for compute-sanitizer (10-50x slowdown):

    compute-sanitizer --tool memcheck  python tools/sanitize_target.py
    compute-sanitizer --tool racecheck python tools/sanitize_target.py
    compute-sanitizer --tool synccheck python tools/sanitize_target.py

Covers: dg_spmv_tma with deep rings (mbarrier phase flips, cp.async.bulk byte counts rounded up into the padding),
csr_spmv_tma, the class kernels (dg_spmv_class, pmg_sweep, maxwell_stage_class), the radix sort / fold, the element kernels,
block-Jacobi, BiCGStab.  Prints one line per case; exits non-zero if a result is off (the sanitizer reports the rest)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import yaourt_fem_dg_b200 as yfd

ctx = yfd.Context(0)
bad = 0


def check(name, ok, extra=""):
    global bad
    bad += 0 if ok else 1
    print("%-58s %s %s" % (name, "ok" if ok else "FAIL", extra), flush=True)


for cell, r, k in ((yfd.TRI, 3, 1), (yfd.QUAD, 2, 3), (yfd.TRI, 2, 4)):
    mesh = yfd.Mesh(ctx, yfd.HostMesh.create(cell, r))
    sysm = ctx.dg_assemble_sip(mesh, k, build_pc=True)
    p = yfd.CgParams(rr_tol=1e-10, rr_max=1e4, max_iter=2 * sysm.n)
    x = np.random.default_rng(1).standard_normal(sysm.n)
    y_cls = sysm.apply(x)
    a = sysm.solve_cg(p, use_pc=True)        # Jacobi, the reference's -p: plain CG does not converge in 2n iterations at tri k=4
    check("class SpMV + CG cell=%d r=%d k=%d" % (cell, r, k), a["converged"] and sysm.true_residual(a["x"]) <= 1e-9,
          "classes=%d" % sysm.operator_info()["nclasses"])
    if k > 1 or cell == yfd.QUAD:
        b = sysm.solve_cg(p, use_pc=yfd.PC_PMG)
        check("  p-multigrid CG", b["converged"] and np.linalg.norm(b["x"] - a["x"]) <= 1e-7 * np.linalg.norm(a["x"]), "its=%d" % b["iterations"])
    sysm.set_operator_mode(yfd.OPERATOR_STREAMED)
    for stages in (2, 3, 4):
        os.environ.update(YFD_SPMV_STAGES=str(stages), YFD_SPMV_CTAS="2", YFD_SPMV_WARPS="1")
        y = sysm.apply(x)
        c = sysm.solve_cg(p, use_pc=yfd.PC_JACOBI)
        for key in ("YFD_SPMV_STAGES", "YFD_SPMV_CTAS", "YFD_SPMV_WARPS"):
            os.environ.pop(key)
        check("  streamed SpMV, ring of %d, 2 CTAs x 1 warp" % stages,
              np.abs(y - y_cls).max() <= 1e-11 * np.abs(y).max() and c["converged"])
    d = sysm.solve_cg(p, use_pc=yfd.PC_BLOCK_JACOBI)
    check("  block-Jacobi CG", d["converged"])
    sysm.close(); mesh.close()

mesh = yfd.Mesh(ctx, yfd.HostMesh.create(yfd.TRI, 4))
cf = ctx.cfem_assemble_p1(mesh)
os.environ.update(YFD_SPMV_STAGES="2", YFD_SPMV_CTAS="4", YFD_SPMV_WARPS="1")
res = cf.solve_cg(yfd.CgParams(rr_tol=1e-10, rr_max=1e3, max_iter=4 * cf.n))
for key in ("YFD_SPMV_STAGES", "YFD_SPMV_CTAS", "YFD_SPMV_WARPS"):
    os.environ.pop(key)
check("CFEM P1: sort + fold + csr_spmv_tma (deep ring) + CG", res["converged"] and cf.true_residual(res["x"]) <= 1e-9)
p2 = ctx.cfem_assemble_p2(mesh)
res = p2.solve_cg(yfd.CgParams(rr_tol=1e-10, rr_max=1e3, max_iter=4 * p2.n))
check("CFEM P2", res["converged"])
adv = ctx.dg_assemble_advection(mesh, 1, upwinding=True)
res = adv.solve_bicgstab(yfd.CgParams(rr_tol=1e-8, rr_max=1e8, max_iter=400))
check("advection-reaction + BiCGStab", np.isfinite(res["x"]).all())
res = adv.solve_cg(yfd.CgParams(rr_tol=1e-8, rr_max=1e4, max_iter=200, use_normal_eqns=1))
check("advection-reaction + CG on the normal equations", np.isfinite(res["x"]).all())
for cell, r, k, up in ((yfd.TRI, 2, 4, False), (yfd.QUAD, 2, 2, True)):
    mx = yfd.Maxwell(ctx, yfd.Mesh(ctx, yfd.HostMesh.create(cell, r)), k, delta_t=1e-3, upwind=up)
    mx.set_operator_mode(yfd.OPERATOR_CLASSES)
    mx.apply_initial_condition(); mx.timesteps(3); a = mx.dofs()
    mx.set_operator_mode(yfd.OPERATOR_STREAMED)
    mx.apply_initial_condition(); mx.timesteps(3); b = mx.dofs()
    check("Maxwell cell=%d k=%d upwind=%d: class vs streamed stages" % (cell, k, up), np.linalg.norm(a - b) <= 1e-11 * np.linalg.norm(b),
          "classes=%d" % mx.info.nclasses)
ctx.close()
print("sanitize_target: %s" % ("PASS" if bad == 0 else "FAIL"))
sys.exit(1 if bad else 0)

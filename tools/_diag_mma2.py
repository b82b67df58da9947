import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np
import yaourt_fem_dg_b200 as yfd
ctx = yfd.Context(0)
hm = yfd.HostMesh.structured(yfd.QUAD, 1448)
mesh = yfd.Mesh(ctx, hm)
sysm = ctx.dg_assemble_sip(mesh, 3)
p = yfd.CgParams(rr_tol=1e-10, rr_max=1e4, max_iter=2 * sysm.n)
for mode in sys.argv[1:]:
    os.environ["YFD_CLS_MMA"] = mode
    r = sysm.solve_cg(p, history=True)
    h = r["history"]
    print("mode", mode, "its", r["iterations"], "ms %.1f" % r["ms_solve"], "hist", ["%.6e" % v for v in h[[1, 10, 100, 1000, 2000, 3000]]], flush=True)

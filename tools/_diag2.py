import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np
import yaourt_fem_dg_b200 as yfd
ctx = yfd.Context(0)
hm = yfd.HostMesh.structured(yfd.QUAD, 1448)
mesh = yfd.Mesh(ctx, hm)
sysm = ctx.dg_assemble_sip(mesh, 3)
p = yfd.CgParams(rr_tol=1e-10, rr_max=1e4, max_iter=2 * sysm.n)
out = {}
for mode, name in ((yfd.OPERATOR_AUTO, "class"), (yfd.OPERATOR_STREAMED, "streamed")):
    sysm.set_operator_mode(mode)
    r = sysm.solve_cg(p, history=True)
    t = sysm.true_residual(r["x"])
    qp, mm = ctx.dg_postprocess(mesh, 3, r["x"])
    out[name] = r
    h = r["history"]
    print(name, "its", r["iterations"], "true residual %.2e" % t, "l2 %.3e" % np.sqrt(qp), "ms", r["ms_solve"])
    print("   history at 500,1000,2000,3000,3139,4000,5000:", [("%.2e" % h[i]) if i < len(h) else None for i in (500, 1000, 2000, 3000, 3139, 4000, 5000)])
a, b = out["class"], out["streamed"]
m = min(len(a["history"]), len(b["history"]))
rel = np.abs(a["history"][:m] - b["history"][:m]) / b["history"][:m]
print("history rel diff at 10,100,500,1000,2000,3000:", [("%.1e" % rel[i]) for i in (10, 100, 500, 1000, 2000, 3000) if i < m])
print("sol diff %.2e" % (np.linalg.norm(a["x"] - b["x"]) / np.linalg.norm(b["x"])))

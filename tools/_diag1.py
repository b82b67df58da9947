import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np
import yaourt_fem_dg_b200 as yfd
ctx = yfd.Context(0)
for nx in (64, 200, 550):
    hm = yfd.HostMesh.structured(yfd.QUAD, nx)
    mesh = yfd.Mesh(ctx, hm)
    sysm = ctx.dg_assemble_sip(mesh, 3)
    oi = sysm.operator_info()
    x = np.random.default_rng(4).standard_normal(sysm.n)
    y_cls = sysm.apply(x); y_plain = sysm.apply(x, plain=True)
    sysm.set_operator_mode(yfd.OPERATOR_STREAMED)
    y_str = sysm.apply(x); y_str_plain = sysm.apply(x, plain=True)
    sc = np.abs(y_str).max()
    d1 = np.abs(y_cls - y_plain); d2 = np.abs(y_cls - y_str); d3 = np.abs(y_str - y_str_plain)
    print("nx", nx, "classes", oi["nclasses"], "dev", oi["max_deviation"], "tol", oi["tolerance"])
    print("   cls-vs-plain(expanded) %.2e  cls-vs-streamed %.2e  streamed-vs-plain %.2e (rel to max|y|)" % (d1.max()/sc, d2.max()/sc, d3.max()/sc))
    bad = np.flatnonzero(d1 > 1e-11 * sc)
    print("   rows off in cls-vs-plain:", len(bad), bad[:10], (bad[:10] // 10))
    # symmetry of the applied operator through the kernel: x.(A z) - z.(A x)
    sysm.set_operator_mode(yfd.OPERATOR_AUTO)
    z = np.random.default_rng(5).standard_normal(sysm.n)
    a = x @ sysm.apply(z); b = z @ y_cls
    sysm.set_operator_mode(yfd.OPERATOR_STREAMED)
    a2 = x @ sysm.apply(z); b2 = z @ y_str
    nrm = np.linalg.norm(x) * np.linalg.norm(y_str)
    print("   asymmetry class %.2e streamed %.2e" % (abs(a - b) / nrm, abs(a2 - b2) / nrm))
    p = yfd.CgParams(rr_tol=1e-10, rr_max=1e4, max_iter=2 * sysm.n)
    sysm.set_operator_mode(yfd.OPERATOR_AUTO)
    r1 = sysm.solve_cg(p); t1 = sysm.true_residual(r1["x"])
    sysm.set_operator_mode(yfd.OPERATOR_STREAMED)
    r2 = sysm.solve_cg(p); t2 = sysm.true_residual(r2["x"])
    print("   CG class its %d true %.2e | streamed its %d true %.2e | sol diff %.2e" % (r1["iterations"], t1, r2["iterations"], t2,
          np.linalg.norm(r1["x"] - r2["x"]) / np.linalg.norm(r2["x"])))
    sysm.close(); mesh.close()

"""class SpMV: tensor-pipe kernel (dg_spmv_class_mma) against the CUDA-core class kernel and the plain one-thread-per-row product."""
import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np
import yaourt_fem_dg_b200 as yfd
ctx = yfd.Context(0)
cases = [(yfd.QUAD, 3, "s", 64), (yfd.QUAD, 3, "s", 1448), (yfd.QUAD, 2, "s", 700), (yfd.QUAD, 4, "s", 500),
         (yfd.TRI, 3, "r", 7), (yfd.TRI, 2, "r", 9), (yfd.TRI, 4, "r", 6), (yfd.TRI, 4, "r", 8), (yfd.TRI, 3, "r", 9), (yfd.QUAD, 3, "r", 3)]
for ct, k, kind, size in cases:
    hm = yfd.HostMesh.structured(ct, size) if kind == "s" else yfd.HostMesh.create(ct, size)
    mesh = yfd.Mesh(ctx, hm)
    sysm = ctx.dg_assemble_sip(mesh, k)
    oi = sysm.operator_info()
    x = np.random.default_rng(4).standard_normal(sysm.n)
    res = {}
    for mode in ("0", "2"):
        os.environ["YFD_CLS_MMA"] = "2" if mode != "0" else "0"
        y = sysm.apply(x)
        t = sysm.time_spmv(50)["ms"]
        res[mode] = (y, t)
    y_plain = sysm.apply(x, plain=True)
    sc = np.abs(y_plain).max()
    print("ct %d k %d %s%d n %d classes %d uses %d | mma-vs-dfma %.2e mma-vs-plain %.2e dfma-vs-plain %.2e | ms dfma %.4f mma %.4f (x%.2f)" % (
        ct, k, kind, size, sysm.n, oi["nclasses"], oi["uses_classes"], np.abs(res["2"][0] - res["0"][0]).max() / sc,
        np.abs(res["2"][0] - y_plain).max() / sc, np.abs(res["0"][0] - y_plain).max() / sc, res["0"][1], res["2"][1], res["0"][1] / res["2"][1]), flush=True)
    if size in (1448, 64, 500, 8):
        p = yfd.CgParams(rr_tol=1e-10, rr_max=1e4, max_iter=2 * sysm.n)
        for mode in ("0", "2"):
            os.environ["YFD_CLS_MMA"] = mode
            r = sysm.solve_cg(p)
            print("   CG mode %s: its %d ms %.1f true %.2e" % (mode, r["iterations"], r["ms_solve"], sysm.true_residual(r["x"])), flush=True)
    sysm.close(); mesh.close()

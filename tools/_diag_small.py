import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np
import yaourt_fem_dg_b200 as yfd
ctx = yfd.Context(0)
for (t, nx, ny, k) in [(0, 5, 7, 4), (0, 5, 3, 4), (0, 4, 4, 4), (0, 3, 6, 4), (1, 6, 6, 4), (0, 5, 7, 3), (0, 5, 7, 2), (0, 9, 9, 4), (0, 12, 12, 4)]:
    hm = yfd.HostMesh.structured(t, nx, ny)
    mesh = yfd.Mesh(ctx, hm)
    sysm = ctx.dg_assemble_sip(mesh, k, eta=1.0, build_pc=True)
    oi = sysm.operator_info()
    try:
        r = sysm.solve_cg(yfd.CgParams(rr_tol=1e-12, rr_max=1e4, max_iter=2 * sysm.n), use_pc=2)
        print(t, nx, ny, k, "classes", oi["nclasses"], "uses", oi["uses_classes"], "tensor", oi["tensor_pipe"], "table", oi["table_bytes"], "its", r["iterations"], "true %.1e" % sysm.true_residual(r["x"]), flush=True)
    except Exception as e:
        print(t, nx, ny, k, "classes", oi["nclasses"], "table", oi["table_bytes"], "ERROR", str(e)[-120:], flush=True)
    sysm.close(); mesh.close()

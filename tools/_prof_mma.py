"""ncu targets: which == dg: DG class SpMV on the tensor pipe, tri k=4 -r 8; which == mx: Maxwell class stage, tri k=4 -r 9."""
import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np
import yaourt_fem_dg_b200 as yfd
ctx = yfd.Context(0)
which = sys.argv[1]
if which == "dg":
    mesh = yfd.Mesh(ctx, yfd.HostMesh.create(yfd.TRI, 8))
    sysm = ctx.dg_assemble_sip(mesh, 4)
    print(sysm.operator_info(), sysm.time_spmv(20), flush=True)
else:
    mesh = yfd.Mesh(ctx, yfd.HostMesh.create(yfd.TRI, 9))
    os.environ["YFD_MX_CLASSES"] = "2"
    mx = yfd.Maxwell(ctx, mesh, 4, delta_t=1e-4)
    mx.apply_initial_condition()
    print(mx.time_apply(10), flush=True)
    print(mx.timesteps(5).ms_total / 5, flush=True)

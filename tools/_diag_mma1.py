import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np
import yaourt_fem_dg_b200 as yfd
ctx = yfd.Context(0)
hm = yfd.HostMesh.structured(yfd.QUAD, 1448)
mesh = yfd.Mesh(ctx, hm)
sysm = ctx.dg_assemble_sip(mesh, 3)
for mode in sys.argv[1:]:
    os.environ["YFD_CLS_MMA"] = mode
    print(mode, sysm.time_spmv(20)["ms"], flush=True)

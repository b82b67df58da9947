import sys, os
sys.path.insert(0, os.getcwd())
import yaourt_fem_dg_b200 as yfd
ctx = yfd.Context(0)
mesh = yfd.Mesh(ctx, yfd.HostMesh.structured(yfd.QUAD, 724))
for _ in range(2):
    s = ctx.dg_assemble_sip(mesh, 3)
    print("cells", mesh.ncells, "elements ms", s.info.ms_elements, flush=True)
    s.close()

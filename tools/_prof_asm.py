import sys, os
sys.path.insert(0, os.getcwd())
import yaourt_fem_dg_b200 as yfd
ctx = yfd.Context(0)
ct, r, k = (int(v) for v in sys.argv[1:4]) if len(sys.argv) > 3 else (0, 8, 4)
mesh = yfd.Mesh(ctx, yfd.HostMesh.create(ct, r))
for _ in range(2):
    mx = yfd.Maxwell(ctx, mesh, k, delta_t=1e-4)
    print("cells", mx.ncells, "assemble ms", mx.info.ms_assemble, flush=True)
    mx.close()

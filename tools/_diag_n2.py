import os, sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, os.getcwd())
import yaourt_fem_dg_b200 as yfd
rank, world, lrank = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lrank)
dist.init_process_group("nccl", device_id=torch.device("cuda", lrank))
ctx = yfd.Context(lrank, stream=torch.cuda.current_stream().cuda_stream)
ids = [yfd.Context.comm_unique_id() if rank == 0 else None]
dist.broadcast_object_list(ids, src=0)
ctx.attach_comm(rank, world, ids[0])
for (t, r, k, pc) in ((0, (5, 7), 4, 2), (0, (5, 7), 4, 0), (0, (5, 7), 4, 1), (0, 3, 4, 2), (0, (9, 9), 4, 2)):
    hm = yfd.HostMesh.structured(t, r[0], r[1]) if isinstance(r, tuple) else yfd.HostMesh.create(t, r)
    mesh = yfd.Mesh(ctx, hm)
    sysm = ctx.dg_assemble_sip(mesh, k, eta=1.0, build_pc=bool(pc))
    oi = sysm.operator_info()
    try:
        res = sysm.solve_cg(yfd.CgParams(rr_tol=1e-12, rr_max=1e4, max_iter=2 * sysm.n), use_pc=pc)
        print(rank, (t, r, k, pc), oi, "its", res["iterations"], flush=True)
    except Exception as e:
        print(rank, (t, r, k, pc), oi, "ERROR", str(e)[-160:], flush=True)
    sysm.close(); mesh.close()
dist.destroy_process_group()

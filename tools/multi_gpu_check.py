#!/usr/bin/env python3
"""Multi-GPU parity check (run under torchrun on the GPU box; the driver's pytest -m gpu tier has one GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/multi_gpu_check.py

Every rank assembles its block of the replicated mesh, the ranks solve together (NCCL halo exchange + scalar
all-reduce), and rank 0 compares the gathered solution / rhs with the UNMODIFIED reference (oracle/_ref)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import torch
import torch.distributed as dist
import yaourt_fem_dg_b200 as yfd

rank, world, lrank = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lrank)
dist.init_process_group("nccl", device_id=torch.device("cuda", lrank))
ctx = yfd.Context(lrank, stream=torch.cuda.current_stream().cuda_stream)
ids = [yfd.Context.comm_unique_id() if rank == 0 else None]
dist.broadcast_object_list(ids, src=0)
ctx.attach_comm(rank, world, ids[0])

import refbind
ref = refbind.Ref() if refbind.available() else None
fails = 0
# r = refinement level of the reference mesher, or (nx, ny) of the structured one.  The block-Jacobi cases with an odd
# block size (k = 1, 4: bs = 3, 15) and an odd owned-cell count (3 ranks on 1024 / 256 cells; 2 ranks on 5 x 7 x 2
# triangles) are the ones whose halo blocks start at an odd double offset (misaligned bulk-copy source)
CASES = ((0, 4, 1, False), (1, 3, 2, True), (0, 3, 3, False), (1, 4, 3, False), (1, 4, 3, 2), (0, 3, 2, 2),
         (0, 4, 1, 2), (0, 3, 4, 2), (0, (5, 7), 1, 2), (0, (5, 7), 4, 2), (1, (9, 5), 1, 2))
for (t, r, k, pc) in CASES:
    hm = yfd.HostMesh.structured(t, r[0], r[1]) if isinstance(r, tuple) else yfd.HostMesh.create(t, r)
    mesh = yfd.Mesh(ctx, hm)
    sysm = ctx.dg_assemble_sip(mesh, k, eta=1.0, build_pc=bool(pc))      # pc = 2: block-Jacobi (yfd.PC_BLOCK_JACOBI)
    loc = sysm.local_sizes()
    p = yfd.CgParams(rr_tol=1e-12, rr_max=1e4, max_iter=2 * sysm.n)
    res = sysm.solve_cg(p, use_pc=pc, history=True)
    rhs = sysm.rhs()
    tres = sysm.true_residual(res["x"])          # independent kernel, sums combined over the ranks
    fails += 0 if tres <= 1e-10 else 1
    # yfd_system_solve_cg_owned: this rank's rows of the same solution, bit for bit, without the gather
    x_own = np.empty(loc["n_local"])
    sysm.solve_cg_owned_into(p, x_own, use_pc=pc)
    bs_ = sysm.info.basis_size
    own_ok = np.array_equal(x_own.reshape(-1, bs_), res["x"].reshape(-1, bs_)[sysm.owned_cells()])
    fails += 0 if own_ok else 1
    line = "case t=%d r=%s k=%d pc=%d: n=%d local rows %d halo dofs %d iters %d true residual %.1e owned rows %s" % (t, r, k, pc, sysm.n, loc["n_local"], loc["n_halo"], res["iterations"], tres, "OK" if own_ok else "WRONG")
    if ref is not None and rank == 0:
        if isinstance(r, tuple):
            ref.set_mesh(t, hm.points, hm.cells, hm.faces, hm.face_boundary, hm.face_boundary_id)
        else:
            ref.mesh(t, r)
        ref.set_cg(rr_tol=1e-12, rr_max=1e4, max_iter=-1)
        out = ref.run_dg(k, eta=1.0, use_pc=bool(pc))
        e_rhs = np.abs(rhs - out["rhs"]).max() / np.abs(out["rhs"]).max()
        e_sol = np.linalg.norm(res["x"] - out["sol"]) / np.linalg.norm(out["sol"])
        # block-Jacobi against the reference's point-Jacobi run: fewer iterations from k = 2 on, on par at k = 1
        ok = e_rhs <= 1e-12 and e_sol <= 1e-10 and ((res["iterations"] < out["iters"] if k >= 2 else res["iterations"] <= 1.1 * out["iters"] + 2) if pc == 2 else
                                                     abs(res["iterations"] - out["iters"]) <= max(3, 0.05 * out["iters"]))
        fails += 0 if ok else 1
        line += " | vs reference: rhs %.1e sol %.1e ref iters %d %s" % (e_rhs, e_sol, out["iters"], "OK" if ok else "FAIL")
    # every rank must hold the same full solution
    xs = torch.from_numpy(res["x"]).cuda()
    lo, hi = xs.clone(), xs.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    same = bool(torch.equal(lo, hi))
    fails += 0 if same else 1
    if rank == 0:
        print(line + (" | identical on all ranks" if same else " | RANKS DIFFER"), flush=True)
    sysm.close(); mesh.close()
# ---- DG Maxwell: every rank advances its block, one halo exchange per operator application ----
mref = refbind.RefMaxwell() if refbind.maxwell_available() else None
for (t, r, k, up) in ((0, 3, 2, False), (1, 3, 3, True), (0, 2, 4, True), (0, 4, 1, False)):
    dt, steps = 0.002, 10
    mesh = yfd.Mesh(ctx, yfd.HostMesh.create(t, r))
    mx = yfd.Maxwell(ctx, mesh, k, delta_t=dt, upwind=up)
    mx.apply_initial_condition()
    d0 = mx.dofs()
    mx.timesteps(steps)
    d = mx.dofs()
    err = mx.report_errors((steps - 1) * dt)
    line = "maxwell t=%d r=%d k=%d upwind=%d: %d unknowns, %d steps" % (t, r, k, up, mx.ndofs, steps)
    if mref is not None and rank == 0:
        mref.create(t, r, k, dt=dt, upwind=up)
        mref.assemble(); mref.init()
        e0 = np.linalg.norm(d0 - mref.dofs()) / np.linalg.norm(d0)
        mref.step(steps)
        want = mref.dofs()
        e1 = np.linalg.norm(d - want) / np.linalg.norm(want)
        e2 = np.abs(err - mref.errors(steps - 1)).max()
        ok = e0 <= 1e-10 and e1 <= 1e-10 and e2 <= 1e-10
        fails += 0 if ok else 1
        line += " | vs reference: init %.1e dofs %.1e error norms %.1e %s" % (e0, e1, e2, "OK" if ok else "FAIL")
    xs = torch.from_numpy(d).cuda()
    lo, hi = xs.clone(), xs.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    same = bool(torch.equal(lo, hi))
    fails += 0 if same else 1
    # the owned slice (yfd_maxwell_get_dofs_owned / _owned_cells) is this rank's part of the gathered vector, bit for bit,
    # and the ranks' cells cover the mesh exactly once
    ids, own = mx.owned_cells(), mx.dofs_owned()
    L3 = 3 * mx.bs
    owned_ok = np.array_equal(own.reshape(-1, L3), d.reshape(-1, L3)[ids])
    cover = torch.zeros(mx.ncells, device="cuda", dtype=torch.int32)
    cover[torch.from_numpy(ids).cuda()] += 1
    dist.all_reduce(cover)
    owned_ok = owned_ok and bool((cover == 1).all())
    fails += 0 if owned_ok else 1
    if rank == 0:
        print(line + (" | identical on all ranks" if same else " | RANKS DIFFER") + (" | owned slices OK" if owned_ok else " | OWNED SLICES WRONG"), flush=True)
    mx.close(); mesh.close()
t = torch.tensor([fails], device="cuda")
dist.all_reduce(t)
if rank == 0:
    print("multi_gpu_check: %s (%d ranks)" % ("PASS" if t.item() == 0 else "FAIL", world), flush=True)
ctx.close()
dist.destroy_process_group()
sys.exit(0 if t.item() == 0 else 1)

#!/usr/bin/env python3
"""Dev tool: time the Maxwell operator application and a few RK4 steps for some (mesh, r, k) on cuda:0."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import yaourt_fem_dg_b200 as yfd  # noqa: E402

cases = [tuple(int(x) for x in a.split(",")) for a in sys.argv[1:]] or [(0, 8, 4, 0), (0, 9, 4, 0), (0, 9, 4, 1), (1, 9, 3, 0), (0, 10, 1, 0), (0, 10, 2, 0)]
ctx = yfd.Context(0)
var = os.environ.get("YFD_MX_PREFETCH", "1")
for t, r, k, up in cases:
    t0 = time.time()
    hm = yfd.HostMesh.create(t, r)
    mesh = yfd.Mesh(ctx, hm)
    t1 = time.time()
    mx = yfd.Maxwell(ctx, mesh, k, delta_t=1e-5, upwind=bool(up))
    mx.apply_initial_condition()
    ms, rb, kb = mx.time_apply(20)
    info = mx.timesteps(10)
    e = mx.report_errors(10e-5)
    print(json.dumps(dict(variant=var, type=t, r=r, k=k, upwind=up, cells=mx.ncells, dofs=mx.ndofs, ms_assemble=mx.info.ms_assemble,
                          ms_apply=ms, kernel_gbs=kb / ms / 1e6, ref_layout_gbs=rb / ms / 1e6, kernel_bytes=kb,
                          ms_rk4_step=info.ms_per_step, mdofs_per_s=mx.ndofs / info.ms_per_step / 1e3,
                          errors=list(e), host_mesh_s=t1 - t0)), flush=True)
    mx.close(); mesh.close()
    ctx.trim()

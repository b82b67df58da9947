#!/usr/bin/env python3
"""Time the CG SpMV of one workload under different ring configurations (YFD_SPMV_STAGES / YFD_SPMV_WARPS).
Tuning helper, run on the GPU box:  python tools/spmv_sweep.py dg_sip_quad_k3_r9"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import yaourt_fem_dg_b200 as yfd  # noqa: E402

names = sys.argv[1:] or ["dg_sip_quad_k3_r9"]
ctx = yfd.Context(0)
for name in names:
    wl = bench.WORKLOADS[name]
    mesh = yfd.Mesh(ctx, yfd.HostMesh.create(wl["cell"], wl["r"]))
    sysm = ctx.dg_assemble_sip(mesh, wl["k"])
    for stages in (0, 2, 3, 4):
        for warps in (0, 4, 6, 8, 12, 16):
            os.environ["YFD_SPMV_STAGES"] = str(stages)
            os.environ["YFD_SPMV_WARPS"] = str(warps)
            try:
                r = sysm.time_spmv(reps=20)
            except yfd.YfdError as e:
                print(name, stages, warps, "failed", e)
                continue
            print(json.dumps(dict(workload=name, stages=stages, warps=warps, ms=round(r["ms"], 4),
                                  kernel_gbs=round(r["kernel_bytes"] / r["ms"] / 1e6, 1), csr_gbs=round(r["csr_bytes"] / r["ms"] / 1e6, 1))))
    sysm.close(); mesh.close()

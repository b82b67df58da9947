#!/usr/bin/env python3
"""Generate tests/golden/*.npz by RUNNING the unmodified reference (oracle/_ref/libyfd_ref.so, built
from /root/reference by oracle/Makefile).  Dev-time tool: it needs /root/reference at build time, the
committed fixtures do not.  Every array below is an OUTPUT OF THE REFERENCE'S OWN CODE (meshers.hpp,
quadratures.hpp, bases.hpp, dg2d_diffusion.cpp, methods/dg.hpp, blaze_sparse_init.hpp, solvers.hpp,
continuous_fem.cpp, methods/cfem.hpp); nothing here comes from this repository's oracle or CUDA code.

    python tools/make_golden.py            # rewrites tests/golden/
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import refbind  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
os.makedirs(OUT, exist_ok=True)
ref = refbind.Ref()


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def mesh_arrays(m):
    return dict(points=m["points"], cells=m["cells"], faces=m["faces"], face_boundary=m["face_boundary"].astype(np.uint8),
                face_owners=m["face_owners"])


# ---- meshes: full arrays for r <= 2, digests up to r = 6 (mesher<Mesh>::create_mesh + compute_connectivity) ----
digests = {}
for t, name in ((0, "tri"), (1, "quad")):
    for r in range(0, 7):
        m = mesh_arrays(ref.mesh(t, r))
        if r <= 2:
            np.savez_compressed(os.path.join(OUT, "mesh_%s_r%d.npz" % (name, r)), **m)
        for k, v in m.items():
            digests["%s_r%d_%s" % (name, r, k)] = sha(v if k != "points" else np.asarray(v, np.float64))
        digests["%s_r%d_sizes" % (name, r)] = "%d %d %d" % (len(m["points"]), len(m["cells"]), len(m["faces"]))
with open(os.path.join(OUT, "mesh_digests.txt"), "w") as f:
    for k in sorted(digests):
        f.write("%s %s\n" % (k, digests[k]))

# ---- numerics probes: integrate(cell), integrate(face), basis eval / eval_grads, normal, diameter ----
probes = {}
for t, name in ((0, "tri"), (1, "quad")):
    m = ref.mesh(t, 1, shatter=True)              # shattered: generic (non-congruent) cells
    probes["%s_points" % name] = m["points"]; probes["%s_cells" % name] = m["cells"]; probes["%s_faces" % name] = m["faces"]
    for deg in (0, 1, 2, 3, 4, 6, 8, 12):
        probes["%s_cellq_d%d" % (name, deg)] = np.stack([ref.cell_quadrature(c, deg) for c in range(4)])
        probes["%s_faceq_d%d" % (name, deg)] = np.stack([ref.face_quadrature(f, deg) for f in range(6)])
    for k in (0, 1, 2, 3, 4, 6):
        ph, dph = zip(*[ref.basis_eval(c, k, 0.31 + 0.1 * c, 0.47 - 0.05 * c) for c in range(4)])
        probes["%s_phi_k%d" % (name, k)] = np.stack(ph); probes["%s_dphi_k%d" % (name, k)] = np.stack(dph)
    nv = m["cells"].shape[1]
    geo = np.array([[*ref.face_geometry(c, e)[0], ref.face_geometry(c, e)[1]] for c in range(4) for e in range(nv)])
    probes["%s_face_geometry" % name] = geo
np.savez_compressed(os.path.join(OUT, "probes.npz"), **probes)

# ---- DG-SIP systems: raw triplet stream, CSR, rhs, Jacobi diagonal, CG solution + history at rr 1e-12 ----
CASES = [("tri", 0, 1, 1, False, False), ("tri", 0, 1, 2, True, False), ("quad", 1, 1, 3, False, False),
         ("quad", 1, 1, 1, True, False), ("tri", 0, 2, 1, True, True), ("quad", 1, 1, 2, False, True), ("tri", 0, 0, 4, False, False)]
for name, t, r, k, pc, shatter in CASES:
    m = ref.mesh(t, r, shatter=shatter)
    ref.set_cg(rr_tol=1e-12, rr_max=1e4, max_iter=-1)
    out = ref.run_dg(k, eta=1.0, use_pc=pc, raw_triplets=True)
    d = dict(mesh_arrays(m), degree=k, eta=1.0, use_pc=int(pc), rowptr=out["rowptr"], colidx=out["colidx"], values=out["values"],
             rhs=out["rhs"], sol=out["sol"], iters=out["iters"], history=out["history"], raw_i=out["raw"][0], raw_j=out["raw"][1],
             raw_v=out["raw"][2], l2_qp=out["l2_qp"], l2_mm=out["l2_mm"])
    if pc:
        d["pc_diag"] = out["pc_diag"]
    np.savez_compressed(os.path.join(OUT, "dg_%s_r%d_k%d%s%s.npz" % (name, r, k, "_pc" if pc else "", "_shatter" if shatter else "")), **d)

# ---- the driver's printed L2 errors for its default run and one level up (dg2d_diffusion.cpp:225-268) ----
rows = []
for t, r, k in ((0, 4, 1), (0, 3, 2), (1, 3, 2)):
    ref.mesh(t, r)
    ref.set_cg(rr_tol=1e-8, rr_max=1e4, max_iter=-1)
    out = ref.run_dg(k)
    rows.append((t, r, k, out["n"], out["nnz"], out["iters"], out["l2_qp"], out["l2_mm"], out["mesh_h"]))
np.savez_compressed(os.path.join(OUT, "dg_driver_runs.npz"), runs=np.array(rows))

# ---- CFEM P1 (continuous_fem.cpp with its refinement count redirected) ----
for r in (1, 2):
    out = ref.run_cfem(r, raw_triplets=True)
    m = ref.mesh(0, r)
    np.savez_compressed(os.path.join(OUT, "cfem_r%d.npz" % r), **mesh_arrays(m), rowptr=out["rowptr"], colidx=out["colidx"],
                        values=out["values"], rhs=out["rhs"], sol=out["sol"], iters=out["iters"], history=out["history"],
                        raw_i=out["raw"][0], raw_j=out["raw"][1], raw_v=out["raw"][2])

tot = sum(os.path.getsize(os.path.join(OUT, f)) for f in os.listdir(OUT))
print("wrote %d files, %.1f KB" % (len(os.listdir(OUT)), tot / 1024))

#!/usr/bin/env python
"""bench.py -- assembled+solved DOFs/s and CG SpMV HBM GB/s of the DG-SIP assemble-and-solve path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

One "step" = one pass of the hot path over one mesh: per-element DG-SIP blocks -> block CSR ->
CG to rr_tol (reference algorithm, core/solvers.hpp:46-138; no preconditioner, as the reference
driver defaults), i.e. everything run_diffusion_solver does between mesh creation and post-process
(dg2d_diffusion.cpp:131-219).

  value   DOFs / s with the mesh already resident in HBM (device-timed, max over ranks)
  e2e     the same through the C ABI with HOST buffers: mesh arrays pinned on the host are uploaded,
          connectivity built, system assembled and solved, solution copied back -- all in the timed region
  roofline   the dominant kernel (the CG SpMV): algorithmic CSR bytes (12 nnz + 8(n+1)
             + 16 n) / CUDA-event time per launch, against MEASURED_PEAKS.json
  cpu_baseline   the UNMODIFIED reference (oracle/_ref, Blaze, -fopenmp over all host cores) on a
                 bounded sample of the same discretisation, timed on this box's host cores

`--impl reference` times only that CPU reference arm (rank 0; other ranks exit 0).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import numpy as np

# name -> (cell type, refinement, degree, use_pc, rr_tol, cpu sample refinement for one reference step,
#          cpu sample refinement for the in-run cpu_baseline)
WORKLOADS = {
    # BASELINE.json configs[2] at its stated size: "DG diffusion on quadrilaterals degree 3, ~20M dofs".  The reference mesher
    # only offers 10.49M (-r 9) or 41.9M (-r 10) dofs, so the mesh is the nx x nx structured one (yfd_mesher_create_structured,
    # same conventions; the reference consumes it unchanged, tests/test_gpu_structured.py)
    "dg_sip_quad_k3_20m": dict(cell=1, nx=1448, k=3, pc=False, rr_tol=1e-10, ref_r=6, base_r=7,
                               desc="DG-SIP diffusion, quadrangles k=3, 1448 x 1448 structured: 2,096,704 cells, 20,967,040 dofs, nnz 1,047,772,800 (BASELINE configs[2], ~20M dofs)"),
    # configs[0] as written: "64x64 structured triangle mesh, degree 1" = 2 * 64^2 triangles
    "dg_sip_tri_k1_64x64": dict(cell=0, nx=64, k=1, pc=False, rr_tol=1e-8, ref_r=6, base_r=6,
                                desc="DG-SIP diffusion, triangles k=1, 64 x 64 structured (8,192 cells, 24,576 dofs; BASELINE configs[0])"),
    # BASELINE.json configs[2]: "DG diffusion on quadrilaterals degree 3, ~20M dofs" -- the reference
    # mesher only offers 10.49M (r=9) or 41.9M (r=10) dofs (size table)
    "dg_sip_quad_k3_r9": dict(cell=1, r=9, k=3, pc=False, rr_tol=1e-10, ref_r=5, base_r=6,
                              desc="DG-SIP diffusion, quadrangles k=3, -r 9: 1,048,576 cells, 10,485,760 dofs, nnz 523,878,400 (BASELINE configs[2])"),
    "dg_sip_quad_k3_r8": dict(cell=1, r=8, k=3, pc=False, rr_tol=1e-10, ref_r=5, base_r=6,
                              desc="DG-SIP diffusion, quadrangles k=3, -r 8: 262,144 cells, 2,621,440 dofs"),
    # configs[0]: the reference's own CPU-runnable case
    "dg_sip_tri_k1_r6": dict(cell=0, r=6, k=1, pc=False, rr_tol=1e-8, ref_r=6, base_r=6,
                             desc="DG-SIP diffusion, triangles k=1, -r 6 (64 boundary segments per side): 16,384 cells, 49,152 dofs (BASELINE configs[0])"),
    # configs[4] mesh one level down (r=11 is the 1e8-dof point)
    "dg_sip_tri_k2_r9": dict(cell=0, r=9, k=2, pc=False, rr_tol=1e-10, ref_r=5, base_r=6,
                             desc="DG-SIP diffusion, triangles k=2, -r 9: 1,048,576 cells, 6,291,456 dofs"),
    # the same two configurations with the block-Jacobi preconditioner (extension: SURVEY.md 8f rank 3)
    "dg_sip_quad_k3_20m_bj": dict(cell=1, nx=1448, k=3, pc=2, rr_tol=1e-10, ref_r=5, base_r=6,
                                  desc="DG-SIP diffusion, quadrangles k=3, 1448 x 1448 structured, 20,967,040 dofs, block-Jacobi CG (extension)"),
    # ... and with the p-multigrid preconditioner (extension: YFD_PC_PMG, yfd_pmg.cu); reference arm: its best, Jacobi (-p)
    "dg_sip_quad_k3_20m_pmg": dict(cell=1, nx=1448, k=3, pc=3, rr_tol=1e-10, ref_r=5, base_r=6,
                                   desc="DG-SIP diffusion, quadrangles k=3, 1448 x 1448 structured, 20,967,040 dofs, p-multigrid CG (extension)"),
    "dg_sip_tri_k2_r11_pmg": dict(cell=0, r=11, k=2, pc=3, rr_tol=1e-10, ref_r=5, base_r=6,
                                  desc="DG-SIP diffusion, triangles k=2, -r 11: 16,777,216 cells, 100,663,296 dofs (BASELINE configs[4]), p-multigrid CG (extension)"),
    "dg_sip_tri_k2_r9_bj": dict(cell=0, r=9, k=2, pc=2, rr_tol=1e-10, ref_r=5, base_r=6,
                                desc="DG-SIP diffusion, triangles k=2, -r 9: 6,291,456 dofs, block-Jacobi CG (extension)"),
    "dg_sip_tri_k2_r11": dict(cell=0, r=11, k=2, pc=True, rr_tol=1e-10, ref_r=5, base_r=6,
                              desc="DG-SIP diffusion, triangles k=2, -r 11: 16,777,216 cells, 100,663,296 dofs (BASELINE configs[4])"),
    # configs[1] names a P2 continuous FEM the reference does not have (methods/cfem.hpp is P1 only); this is its P1 path
    "cfem_p1_r11": dict(kind="cfem", cell=0, r=11, k=1, pc=False, rr_tol=1e-8, rr_max=1e3, ref_r=7, base_r=8,
                        desc="continuous P1 FEM Poisson, triangles -r 11: 33,554,432 cells, 8,380,417 dofs (BASELINE configs[1] is P2; the reference is P1 only)"),
    "cfem_p1_r10": dict(kind="cfem", cell=0, r=10, k=1, pc=False, rr_tol=1e-8, rr_max=1e3, ref_r=7, base_r=8,
                        desc="continuous P1 FEM Poisson, triangles -r 10: 8,388,608 cells, 2,095,105 dofs"),
    # configs[1] as written: "continuous_fem P2 Poisson on triangles, ~4M dofs" -- an extension (parity unpinned): P2 on the
    # structured mesher, (2 nx - 1)^2 unknowns.  CPU baseline = the oracle's P2 statement + its CG (a scalar port)
    "cfem_p2_4m": dict(kind="cfem", order=2, cell=0, nx=1000, k=2, pc=False, rr_tol=1e-8, rr_max=1e3, ref_r=7, base_r=8,
                       desc="continuous P2 FEM Poisson, triangles, 1000 x 1000 structured: 2,000,000 cells, 3,996,001 dofs (BASELINE configs[1]; "
                            "extension, the reference is P1 only: parity unpinned)"),
    # configs[3]: "dg_maxwell_2D time-domain explicit run, degree 4 triangles, matrix-free element+face operator"
    "maxwell_tri_k4_r9": dict(kind="maxwell", cell=0, r=9, k=4, upwind=False, dt=2e-5, timesteps=50, ref_r=5, base_r=6,
                              desc="DG Maxwell TM, triangles k=4, -r 9, centred fluxes, RK4: 1,048,576 cells, 47,185,920 unknowns (BASELINE configs[3])"),
    "maxwell_tri_k4_r8_upwind": dict(kind="maxwell", cell=0, r=8, k=4, upwind=True, dt=4e-5, timesteps=50, ref_r=5, base_r=6,
                                     desc="DG Maxwell TM, triangles k=4, -r 8, upwind fluxes, RK4: 262,144 cells, 11,796,480 unknowns"),
    "maxwell_quad_k3_r9": dict(kind="maxwell", cell=1, r=9, k=3, upwind=False, dt=2e-5, timesteps=50, ref_r=5, base_r=6,
                               desc="DG Maxwell TM, quadrangles k=3, -r 9, centred fluxes, RK4: 1,048,576 cells, 31,457,280 unknowns"),
}
DEFAULT_WORKLOAD = "dg_sip_quad_k3_20m"


def make_host_mesh(yfd, wl):
    if "nx" in wl:
        return yfd.HostMesh.structured(wl["cell"], wl["nx"])
    return yfd.HostMesh.create(wl["cell"], wl["r"])


def bs_of(k):
    return (k + 2) * (k + 1) // 2


# ------------------------------------------------------------------------------------------------
# clocks: nvidia-smi sampled DURING the timed region (B200_PROFILING.md)
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_id):
        self.gpu_id, self.rows, self.proc, self.thr, self.t0 = gpu_id, [], None, None, None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu_id), "--query-gpu=" + self.FIELDS,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.thr = threading.Thread(target=self._read, daemon=True)
        self.thr.start()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")] + [time.time()])

    def mark(self):
        """The timed region starts now: nvidia-smi was started earlier (before the warm-up) so that even a region of a
        few hundred ms is sampled; only rows that arrive from here on count."""
        self.t0 = time.time()

    def stop(self):
        if not self.proc:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        self.thr.join(timeout=2)
        sm, smax, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = self.rows
        if self.t0 is not None:
            inside = [r for r in rows if r[-1] >= self.t0]
            rows = inside if inside else rows[-1:]          # region shorter than one period: the sample nearest to it
        for r in rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0])); smax.append(float(r[1])); power.append(float(r[2]))
            except ValueError:
                continue
            for nme, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        if not sm:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["no samples"])
        return dict(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(smax)), reasons=sorted(reasons),
                    samples=len(sm), power_w_max=max(power))


# ------------------------------------------------------------------------------------------------
# CPU reference arm (oracle/_ref = the unmodified reference; the only place bench.py touches oracle/)
# ------------------------------------------------------------------------------------------------
def reference_step(ref, wl, r):
    """One full reference pass at refinement r: mesh (untimed) -> cell loop -> finalize -> CG."""
    if wl.get("order") == 2:
        return port_step_p2(wl, r)
    ref.set_cg(rr_tol=wl["rr_tol"], rr_max=wl.get("rr_max", 1e4), max_iter=-1)
    if wl.get("kind") == "cfem":
        out = ref.run_cfem(r)               # continuous_fem.cpp: builds its own mesh (counted in t_blocks)
    else:
        ref.mesh(wl["cell"], r)
        out = ref.run_dg(wl["k"], eta=1.0, use_pc=bool(wl["pc"]))     # pc = 2, 3 (block-Jacobi, p-multigrid) have no reference: its best, Jacobi (-p)
    t = out["t_blocks"] + out["t_csr"] + out["t_cg"]
    return dict(n=out["n"], nnz=out["nnz"], iters=out["iters"], t=t, t_blocks=out["t_blocks"], t_csr=out["t_csr"], t_cg=out["t_cg"])


def port_step_p2(wl, r):
    """P2 has no reference implementation: time the oracle's C statement of it (scalar port), mesh untimed."""
    import oraclebind
    o = oraclebind.Oracle()
    m = o.mesh(0, r)
    t0 = time.perf_counter()
    s = o.cfem_p2(m["points"], m["cells"], m["faces"], m["face_boundary"])
    t1 = time.perf_counter()
    res = o.cg(s["rowptr"], s["colidx"], s["values"], s["rhs"], rr_tol=wl["rr_tol"], rr_max=wl.get("rr_max", 1e4),
               max_iter=4 * s["n"], history=False)
    t2 = time.perf_counter()
    return dict(n=s["n"], nnz=s["nnz"], iters=res["iterations"], t=t2 - t0, t_blocks=t1 - t0, t_csr=0.0, t_cg=t2 - t1)


def load_reference():
    import refbind
    omp = refbind.available(omp=True)
    if not refbind.available(omp=omp):
        return None, 0, None
    cores = os.cpu_count() or 1
    if omp:
        os.environ["OMP_NUM_THREADS"] = str(cores)      # torchrun pins it to 1; the reference arm uses every host core
    ref = refbind.Ref(omp=omp)
    return ref, (ref.num_threads() if omp else 1), "reference"


PORT_P2_LABEL = "oracle/ C statement of the P2 extension + its scalar CG (the reference has no P2: port)"


def load_arm(wl):
    """(handle, cores, kind, label) of the CPU implementation timed for this workload."""
    if wl.get("order") == 2:
        return object(), 1, "port", PORT_P2_LABEL
    ref, cores, kind = load_reference()
    return ref, cores, kind, "unmodified reference (Blaze, -O3 -fopenmp: only its CG is threaded)"


def cpu_baseline(wl, r):
    ref, cores, kind, label = load_arm(wl)
    if ref is None:
        return None
    s = reference_step(ref, wl, r)
    return dict(value=s["n"] / s["t"], unit="DOFs/s", cores=cores, kind=kind, dofs=int(s["n"]), nnz=int(s["nnz"]), cg_iterations=int(s["iters"]),
                sample=label + ", same discretisation at -r %d: "
                       "%d dofs, nnz %d, %d CG iterations to rr %.0e; blocks %.2fs + finalize %.2fs + CG %.2fs; "
                       "%.2f ns per dof per CG iteration" % (r, s["n"], s["nnz"], s["iters"], wl["rr_tol"], s["t_blocks"],
                                                              s["t_csr"], s["t_cg"], 1e9 * s["t_cg"] / max(1, s["n"] * s["iters"])))


def cpu_single_thread(wl, r):
    """the reference AS SHIPPED -- it has no threading -- on one host core, at the reference arm's size."""
    import refbind
    if wl.get("order") == 2 or not refbind.available(omp=False):
        return None
    s = reference_step(refbind.Ref(omp=False), wl, r)
    return dict(value=s["n"] / s["t"], unit="DOFs/s", cores=1, kind="reference", dofs=int(s["n"]), cg_iterations=int(s["iters"]),
                sample="unmodified reference as shipped (no OpenMP), -r %d: blocks %.2fs + finalize %.2fs + CG %.2fs" % (r, s["t_blocks"], s["t_csr"], s["t_cg"]))


def run_reference_arm(args, wl, rank):
    if rank != 0:
        return 0
    ref, cores, kind, label = load_arm(wl)
    if ref is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libyfd_ref.so not built (needs /root/reference at build time)"}))
        return 0
    r = wl["ref_r"]
    for _ in range(args.warmup):
        reference_step(ref, wl, r)
    t0 = time.perf_counter()
    tot = 0.0
    last = None
    for _ in range(args.steps):
        last = reference_step(ref, wl, r)
        tot += last["t"]
    wall = time.perf_counter() - t0
    ms = 1e3 * tot / args.steps
    value = last["n"] / (tot / args.steps)
    sample = ("each step = cell loop + finalize + CG of the %s at -r %d (%d dofs, nnz %d, %d iterations, rr_tol %.0e) "
              "of the workload's discretisation; mesh creation untimed; wall %.1fs" % (label, r, last["n"], last["nnz"], last["iters"], wl["rr_tol"], wall))
    line = {"impl": "reference", "metric": "assembled+solved DOFs/sec", "value": value, "unit": "DOFs/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            # `description` names the workload's discretisation; dofs / nnz / cg_iterations are what THIS arm ran (a bounded
            # sample of it: the GPU line's same_size leg times the GPU path on exactly this system)
            "config": {"workload": args.workload, "description": wl["desc"], "sample": sample, "sample_refinement": r,
                       "dofs": int(last["n"]), "nnz": int(last["nnz"]), "cg_iterations": int(last["iters"]),
                       "phases_s": {"blocks": last["t_blocks"], "finalize": last["t_csr"], "cg": last["t_cg"]}},
            "cpu_baseline": {"value": value, "unit": "DOFs/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": "DOFs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------------
# DG Maxwell workloads (BASELINE configs[3]): one step = assemble + initial condition + T RK4 timesteps + error norms
# ------------------------------------------------------------------------------------------------
MAXWELL_METRIC, MAXWELL_UNIT = "assembled+time-stepped DOF updates/sec", "DOF-steps/s"


def maxwell_reference_step(R, wl, r):
    """The unmodified reference (oracle/_ref/libyfd_ref_maxwell*.so): maxwell_context (mesh: untimed) -> assemble ->
    apply_initial_condition -> T x do_timestep -> report_errors."""
    R.create(wl["cell"], r, wl["k"], dt=wl["dt"], upwind=wl["upwind"])
    t0 = time.perf_counter()
    R.assemble()
    R.init()
    t1 = time.perf_counter()
    R.step(wl["timesteps"])
    t2 = time.perf_counter()
    R.errors(wl["timesteps"] - 1)
    t3 = time.perf_counter()
    return dict(n=3 * R.bs * R.ncells, t=t3 - t0, t_assemble=t1 - t0, t_steps=t2 - t1, t_errors=t3 - t2, cells=R.ncells)


def load_maxwell_reference():
    import refbind
    omp = refbind.maxwell_available(omp=True)
    if not refbind.maxwell_available(omp=omp):
        return None, 0
    cores = os.cpu_count() or 1
    if omp:
        os.environ["OMP_NUM_THREADS"] = str(cores)
    R = refbind.RefMaxwell(omp=omp)
    return R, (R.num_threads() if omp else 1)


def maxwell_sample_text(wl, r, s):
    return ("unmodified reference (Blaze dense blocks, -O3 -fopenmp; its cell loops are serial), same discretisation at -r %d: "
            "%d cells, %d unknowns, %d RK4 timesteps; assemble+IC %.2fs, timesteps %.2fs (%.1f ns per unknown per step), errors %.2fs"
            % (r, s["cells"], s["n"], wl["timesteps"], s["t_assemble"], s["t_steps"], 1e9 * s["t_steps"] / (s["n"] * wl["timesteps"]), s["t_errors"]))


def run_maxwell_reference_arm(args, wl, rank):
    if rank != 0:
        return 0
    R, cores = load_maxwell_reference()
    if R is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libyfd_ref_maxwell.so not built (needs /root/reference at build time)"}))
        return 0
    r = wl["ref_r"]
    for _ in range(args.warmup):
        maxwell_reference_step(R, wl, r)
    tot, last = 0.0, None
    for _ in range(args.steps):
        last = maxwell_reference_step(R, wl, r)
        tot += last["t"]
    value = last["n"] * wl["timesteps"] / (tot / args.steps)
    sample = maxwell_sample_text(wl, r, last)
    print(json.dumps({"impl": "reference", "metric": MAXWELL_METRIC, "value": value, "unit": MAXWELL_UNIT, "n_gpus": args.gpus,
                      "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot / args.steps, "higher_is_better": True,
                      "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                      "config": {"workload": args.workload, "description": wl["desc"], "sample": sample},
                      "cpu_baseline": {"value": value, "unit": MAXWELL_UNIT, "cores": cores, "kind": "reference", "sample": sample},
                      "e2e": {"value": value, "unit": MAXWELL_UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                      "gpu_launches": 0}))
    return 0


def run_maxwell(args, wl, rank, world, local_rank):
    import torch
    import yaourt_fem_dg_b200 as yfd
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path is CUDA only (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d: launch with torch.distributed.run --nproc-per-node %d" % (args.gpus, world, args.gpus))
    stream = torch.cuda.current_stream()
    ctx = yfd.Context(local_rank, stream=stream.cuda_stream)
    if world > 1:
        ids = [yfd.Context.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.attach_comm(rank, world, ids[0])

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if dist is None:
            return ms
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    T = args.profile_iters or wl["timesteps"]
    hm = make_host_mesh(yfd, wl)                             # mesher: input generation, untimed
    info = {}
    e2e_buf = {}

    def one_pass(mesh, x_host=None):
        mx = yfd.Maxwell(ctx, mesh, wl["k"], delta_t=wl["dt"], upwind=wl["upwind"])
        mx.apply_initial_condition()
        st = mx.timesteps(T)
        err = mx.report_errors((T - 1) * wl["dt"])
        if x_host is not None:
            if world > 1:        # several ranks: every rank reads back the unknowns of ITS cells (no gather, 1/N of the bytes per rank)
                nown = len(x_host) // world + 3 * mx.bs * 2
                if "own" not in e2e_buf or e2e_buf["own"].numel() < mx.ncells_owned() * 3 * mx.bs:
                    e2e_buf["own"] = torch.empty(max(nown, mx.ncells_owned() * 3 * mx.bs), dtype=torch.float64).pin_memory()
                mx.dofs_owned(e2e_buf["own"].numpy()[:mx.ncells_owned() * 3 * mx.bs])
                e2e_buf["d2h"] = mx.ncells_owned() * 3 * mx.bs * 8
            else:
                ctx.check(yfd.lib().yfd_maxwell_get_dofs(mx._h, x_host))
        info.update(n=mx.ndofs, cells=mx.ncells, ms_assemble=mx.info.ms_assemble, ms_timesteps=st.ms_total, errors=[float(e) for e in err],
                    operator_bytes=mx.info.operator_bytes, reference_operator_bytes=mx.info.reference_operator_bytes)
        return mx

    mesh = yfd.Mesh(ctx, hm)
    sampler = ClockSampler("GPU-" + str(torch.cuda.get_device_properties(local_rank).uuid).replace("GPU-", ""))
    sampler.start()                                     # nvidia-smi is up before the warm-up; rows count from mark()
    for _ in range(args.warmup):
        one_pass(mesh).close()
    barrier()
    sampler.mark()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = ctx.launches
    e0.record(stream)
    for _ in range(args.steps):
        one_pass(mesh).close()
    e1.record(stream)
    barrier()
    clocks = sampler.stop()
    launches = ctx.launches - l0
    ms_step = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    n = info["n"]
    value = n * T / (ms_step * 1e-3)

    # roofline leg: the operator application alone, CUDA events on the launching stream
    mx = one_pass(mesh)
    ms_apply, ref_bytes, kern_bytes = mx.time_apply(30)
    uses_classes = bool(mx.info.uses_classes)
    cls_info = dict(uses_classes=uses_classes, nclasses=int(mx.info.nclasses), reject_reason=int(mx.info.class_reject_reason),
                    max_deviation=mx.info.class_max_deviation, tolerance=mx.info.class_tolerance, table_bytes=int(mx.info.class_table_bytes))
    streamed, other = None, None
    if cls_info["nclasses"] > 0 and cls_info["reject_reason"] == 0:
        # both kernels on the same operator (AUTO took the faster one when the operator was created)
        mx.set_operator_mode(yfd.OPERATOR_STREAMED)
        ms_s, _, kb_s = mx.time_apply(10)
        mx.set_operator_mode(yfd.OPERATOR_CLASSES)
        ms_c, _, kb_c = mx.time_apply(10)
        mx.set_operator_mode(yfd.OPERATOR_AUTO)
        if uses_classes:
            streamed = dict(ms_per_launch=ms_s, kernel_requested_bytes_per_launch=kb_s)
        else:
            other = dict(kernel="maxwell_stage_class (not chosen: slower than the streamed kernel on this operator)", ms_per_launch=ms_c,
                         kernel_requested_bytes_per_launch=kb_c)
    fp64_peak = ctx.measure_fp64()
    tensor_pipe = bool(getattr(mx.info, "class_tensor_pipe", 0))
    tensor_peak = ctx.measure_fp64_tensor() if tensor_pipe else None
    # block products the kernel EXECUTES per cell: 4 cell products + per face 3 (self) + 3 (neighbour) in the streamed / CUDA-core
    # kernels; the tensor-pipe kernel factors the flux operands and needs 2 + 2 per face (yfd_maxwell_mma.cuh)
    products = (4 + 4 * mx.nf) if tensor_pipe else (4 + 6 * mx.nf)
    flops_apply = 2.0 * mx.bs * mx.bs * products * (info["cells"] / world)
    mx.close()
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "spmv_traffic.json")
    if os.path.exists(tpath) and world == 1:
        traffic = json.load(open(tpath)).get(args.workload)
    achieved = ref_bytes / (ms_apply * 1e-3) / 1e9
    req = kern_bytes / (ms_apply * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": ("maxwell_stage_class (operator application, one block set per congruence class)" if uses_classes else "maxwell_stage (operator application, factored blocks)") + ("" if world == 1 else ", rank 0's share of the cells"),
                "achieved": req, "peak": peak, "unit": "GB/s", "frac": req / peak, "hw_frac": req / peak,
                "traffic": None if uses_classes else traffic,
                "traffic_source": None if (uses_classes or traffic is None) else "static: profiles/spmv_traffic.json (one ncu --set full capture; not a live counter)",
                "peak_source": peak_src, "kernel_requested_bytes_per_launch": kern_bytes, "ms_per_launch": ms_apply,
                "reference_layout": {"bytes_per_launch": ref_bytes, "gbs": achieved, "frac": achieved / peak},
                "fp64": {"achieved_tflops": flops_apply / (ms_apply * 1e-3) / 1e12, "peak_tflops": fp64_peak,
                         "frac": flops_apply / (ms_apply * 1e-3) / 1e12 / fp64_peak, "flops_per_launch": flops_apply,
                         "block_products_per_cell": products,
                         "peak_source": "DFMA microbenchmark run just now on this device (yfd_ctx_measure_fp64)"},
                "operator": cls_info,
                "note": "achieved / frac = the bytes the kernel requests per launch / CUDA-event time against the measured HBM peak; reference_layout = "
                        "the reference's dense layout, (1+nf)(3bs)^2 8 B per cell + 2 vectors (SURVEY.md 8, cfg 4), over the same time"
                        + ("; this mesh has %d congruence classes of cells, the kernel multiplies by one block set per class and streams vectors and "
                           "per-cell coefficients only: it is bound by FP64 issue (see fp64 / bound), its HBM fraction is low BY DESIGN; streamed_kernel is the "
                           "HBM-bound general path on the same operator" % cls_info["nclasses"] if uses_classes else "")}
    if uses_classes and tensor_pipe:
        # the class kernel on the FP64 tensor pipe is bound by that pipe, not by HBM: report it against the DMMA peak measured
        # now on this device (mma.sync m8n8k4 f64 microbenchmark, yfd_ctx_measure_fp64_tensor), with the HBM figures beside it
        tf = flops_apply / (ms_apply * 1e-3) / 1e12
        roofline.update({"bound": "tensor", "kernel": "maxwell_stage_class_mma (operator application, one block set per congruence class, mma.sync m8n8k4 f64)"
                                                       + ("" if world == 1 else ", rank 0's share of the cells"),
                         "achieved": tf, "peak": tensor_peak, "unit": "TFLOP/s", "frac": tf / tensor_peak,
                         "peak_source": "FP64 mma.sync microbenchmark run just now on this device (yfd_ctx_measure_fp64_tensor); useful flops only: "
                                        "2 bs^2 per block product EXECUTED (4 + 4 nf per cell after factoring the flux operands; the streamed kernel needs 4 + 6 nf), "
                                        "the padding of bs to the 8x8x4 tiles is not counted",
                         "hbm": {"requested_gbs": req, "peak": peak, "frac": req / peak, "peak_source": peak_src}})
        roofline.pop("hw_frac", None)
        roofline["fp64"]["tensor_peak_tflops"] = tensor_peak
    if streamed is not None:
        sg = streamed["kernel_requested_bytes_per_launch"] / (streamed["ms_per_launch"] * 1e-3) / 1e9
        streamed.update(kernel="maxwell_stage (factored blocks streamed per cell)", achieved=sg, frac=sg / peak, speedup_of_class_kernel=streamed["ms_per_launch"] / ms_apply)
        roofline["streamed_kernel"] = streamed
    if other is not None:
        roofline["class_kernel"] = other

    e2e = None
    if not args.no_e2e:
        keep, arrs = [], {}
        for nme in ("points", "cells", "faces", "face_boundary"):
            t, a = pinned_like(torch, getattr(hm, nme))
            keep.append(t); arrs[nme] = a
        xt = torch.empty(n, dtype=torch.float64).pin_memory()
        x_host = xt.numpy()
        hm_p = yfd.HostMesh(hm.cell_type, arrs["points"], arrs["cells"], arrs["faces"], arrs["face_boundary"])

        def step_e2e():
            m2 = yfd.Mesh(ctx, hm_p)
            one_pass(m2, x_host).close()
            m2.close()
        mesh.close()
        step_e2e()
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        f0.record(stream)
        for _ in range(args.steps):
            step_e2e()
        f1.record(stream)
        barrier()
        wall = (time.perf_counter() - t0) / args.steps
        ms_e2e = max_over_ranks(f0.elapsed_time(f1)) / args.steps
        e2e = {"value": n * T / (ms_e2e * 1e-3), "unit": MAXWELL_UNIT, "h2d_bytes_per_step": int(sum(a.nbytes for a in arrs.values())),
               "d2h_bytes_per_step": int(e2e_buf.get("d2h", x_host.nbytes)), "ms_per_step": ms_e2e, "wall_ms_per_step": 1e3 * wall,
               "path": "yfd_mesh_create -> yfd_maxwell_create -> apply_initial_condition -> timesteps -> report_errors -> "
                       + ("get_dofs (host) through include/yfd.h" if world == 1 else
                          "get_dofs_owned (every rank reads back the unknowns of its own cells: d2h_bytes_per_step is rank 0's share; "
                          "h2d: every rank uploads the whole mesh) through include/yfd.h")}

    base = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:      # the CPU baseline is an N=1 figure (torchrun pins ranks to one thread)
        R, cores = load_maxwell_reference()
        if R is not None:
            sres = maxwell_reference_step(R, wl, wl["base_r"])
            base = dict(value=sres["n"] * wl["timesteps"] / sres["t"], unit=MAXWELL_UNIT, cores=cores, kind="reference",
                        sample=maxwell_sample_text(wl, wl["base_r"], sres))
    line = {"profiling_run_not_a_bench_value": True} if args.profile_iters else {}
    line.update({"metric": MAXWELL_METRIC, "value": value, "unit": MAXWELL_UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                 "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                 "data": "synthetic",
                 "config": {"workload": args.workload, "description": wl["desc"], "dofs": int(n), "cells": int(info["cells"]),
                            "timesteps_per_step": T, "delta_t": wl["dt"], "time_integrator": "RK4 (4 operator applications per timestep)",
                            "l2_policy": "inputs larger than L2: the factored operator is %.1f GB vs 126 MB L2" % (1e-9 * info["operator_bytes"]),
                            "operator_bytes": int(info["operator_bytes"]), "reference_operator_bytes": int(info["reference_operator_bytes"]),
                            "phases_ms": {"assemble": info["ms_assemble"], "timesteps": info["ms_timesteps"]},
                            "l2_errors_vs_analytic": info["errors"], "parallelism": "1 GPU" if world == 1 else "contiguous blocks of the Z-curve cell order over %d GPUs; one NCCL halo exchange "
                                           "(3 bs doubles per cut-face cell) before each of the 4 operator applications of a timestep" % world},
                 "roofline": roofline, "cpu_baseline": base, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks})
    if rank == 0:
        print(json.dumps(line))
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def pinned_like(torch, a):
    t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    return t, t.numpy()


def same_size_leg(torch, yfd, ctx, stream, wl, r, cpu, assemble, cfem, reps=3):
    """The GPU path on the reference mesher's -r `r` mesh of the workload's discretisation -- exactly the system the CPU
    reference was timed on (`cpu` = cpu_baseline(wl, r)) -- resident and end to end."""
    hm = yfd.HostMesh.create(wl["cell"], r)
    nd = (len(hm.points) + len(hm.faces)) if cfem else bs_of(wl["k"]) * len(hm.cells)
    params = yfd.CgParams(rr_tol=wl["rr_tol"], rr_max=wl.get("rr_max", 1e4), max_iter=2 * nd)
    mesh = yfd.Mesh(ctx, hm)
    out = {}

    def resident():
        sysm = assemble(mesh)
        d_x = torch.empty(sysm.n, dtype=torch.float64, device="cuda")
        res = sysm.solve_cg_device(params, use_pc=wl["pc"], d_x=d_x.data_ptr())
        out.update(n=sysm.n, nnz=sysm.nnz, iters=res["iterations"], converged=res["converged"])
        return sysm, d_x
    sysm, d_x = resident()
    out["true_residual"] = sysm.true_residual(d_x=d_x.data_ptr())
    if not cfem:
        out["l2_error"] = float(np.sqrt(ctx.dg_postprocess(mesh, wl["k"], None, device_ptr=d_x.data_ptr())[0]))
    sysm.close()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        resident()[0].close()
    e1.record(stream)
    torch.cuda.synchronize()
    ms_res = e0.elapsed_time(e1) / reps
    mesh.close()
    arrs, keep = {}, []
    for nme in ("points", "cells", "faces", "face_boundary"):
        t, a = pinned_like(torch, getattr(hm, nme))
        keep.append(t); arrs[nme] = a
    xt = torch.empty(out["n"], dtype=torch.float64).pin_memory()
    hm_p = yfd.HostMesh(hm.cell_type, arrs["points"], arrs["cells"], arrs["faces"], arrs["face_boundary"])

    def e2e():
        m2 = yfd.Mesh(ctx, hm_p)
        s2 = assemble(m2)
        s2.solve_cg_into(params, xt.numpy(), use_pc=wl["pc"])
        s2.close(); m2.close()
    e2e()
    torch.cuda.synchronize()
    e0.record(stream)
    for _ in range(reps):
        e2e()
    e1.record(stream)
    torch.cuda.synchronize()
    ms_e2e = e0.elapsed_time(e1) / reps
    n = out["n"]
    assert n == cpu["dofs"] and out["nnz"] == cpu["nnz"], "same_size: the two arms assembled different systems"
    return {"refinement": r, "dofs": int(n), "nnz": int(out["nnz"]), "gpu_dofs_per_s": n / (ms_res * 1e-3), "gpu_e2e_dofs_per_s": n / (ms_e2e * 1e-3),
            "cpu_dofs_per_s": cpu["value"], "cpu_cores": cpu["cores"], "cpu_kind": cpu["kind"], "gpu_iters": int(out["iters"]), "cpu_iters": int(cpu["cg_iterations"]),
            "gpu_converged": bool(out["converged"]), "gpu_true_residual": out["true_residual"], "gpu_l2_error": out.get("l2_error"),
            "ratio_resident": n / (ms_res * 1e-3) / cpu["value"], "ratio_e2e": n / (ms_e2e * 1e-3) / cpu["value"]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=os.environ.get("YFD_BENCH_WORKLOAD", DEFAULT_WORKLOAD), choices=sorted(WORKLOADS))
    ap.add_argument("--pc", type=int, default=None, choices=[0, 1, 2, 3], help="override the workload's preconditioner: 0 none, 1 Jacobi (-p), 2 block-Jacobi")
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="weak: the structured mesh grows with the number of GPUs (nx = 1122 sqrt(N): 12.6 M dofs per GPU at k=3, 1.0e8 dofs at N=8)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--single-shot", action="store_true",
                    help="one untimed-warm-up-free pass (W=0, K=1, no e2e, no CPU baseline) for solves that take minutes "
                         "(the 1e8-dof target): a recorded run, not the driver's bench contract")
    ap.add_argument("--profile-iters", type=int, default=0,
                    help="ncu helper: cap CG at this many iterations, one step, no warm-up; the printed line is NOT a bench value")
    args = ap.parse_args()
    wl = dict(WORKLOADS[args.workload])
    if args.pc is not None:
        wl["pc"] = args.pc
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.scaling == "weak":
        # fixed load per GPU: the reference mesher only refines x4, the structured one (same conventions, parity-pinned by
        # tests/test_gpu_structured.py) makes any size.  Iteration counts grow with the mesh, so the line also reports the
        # time per CG iteration and DOF-iterations/s, the size-independent figures
        if "nx" not in wl or wl.get("kind"):
            raise SystemExit("--scaling weak is defined for the structured DG workloads (dg_sip_quad_k3_20m*)")
        wl["nx"] = int(round(1122 * world ** 0.5))
        ncell = wl["nx"] ** 2 * (2 if wl["cell"] == 0 else 1)
        wl["desc"] = "DG-SIP diffusion, %s k=%d, %d x %d structured: %d cells, %d dofs (weak scaling: 1122^2 cells per GPU)" % (
            "quadrangles" if wl["cell"] == 1 else "triangles", wl["k"], wl["nx"], wl["nx"], ncell, ncell * bs_of(wl["k"]))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_maxwell_reference_arm(args, wl, rank) if wl.get("kind") == "maxwell" else run_reference_arm(args, wl, rank)
    if args.profile_iters > 0 or args.single_shot:
        args.warmup, args.steps, args.no_e2e, args.no_cpu_baseline = 0, 1, True, True
    elif args.warmup < 3:
        args.warmup = 3

    if wl.get("kind") == "maxwell":
        return run_maxwell(args, wl, rank, world, local_rank)
    import torch
    import yaourt_fem_dg_b200 as yfd
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path is CUDA only (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d: launch with torch.distributed.run --nproc-per-node %d" % (args.gpus, world, args.gpus))

    stream = torch.cuda.current_stream()
    ctx = yfd.Context(local_rank, stream=stream.cuda_stream)
    if world > 1:
        ids = [yfd.Context.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.attach_comm(rank, world, ids[0])

    cfem = wl.get("kind") == "cfem"
    if cfem and world > 1:
        raise SystemExit("the CFEM path is not partitioned (DESIGN.md section 6): run it with --gpus 1")
    k, bs = wl["k"], (1 if cfem else bs_of(wl["k"]))

    def assemble(m):
        if cfem:
            return ctx.cfem_assemble_p2(m) if wl.get("order") == 2 else ctx.cfem_assemble_p1(m)
        return ctx.dg_assemble_sip(m, k, eta=1.0, build_pc=wl["pc"] == 1)
    hm = make_host_mesh(yfd, wl)                            # mesher: input generation, untimed
    params = yfd.CgParams(rr_tol=wl["rr_tol"], rr_max=wl.get("rr_max", 1e4),
                          max_iter=args.profile_iters or 2 * ((len(hm.points) + len(hm.faces)) if wl.get("kind") == "cfem" else bs_of(wl["k"]) * len(hm.cells)))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if dist is None:
            return ms
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------- resident leg: mesh in HBM before the clock starts ----------------
    mesh = yfd.Mesh(ctx, hm)
    n = assemble(mesh).n if cfem else bs * mesh.ncells
    d_x = torch.empty(n if world == 1 else 1, dtype=torch.float64, device="cuda")
    info = {}

    x_shot = np.empty(n, np.float64) if (args.single_shot and world > 1) else None
    kept = []

    def step_resident():
        t0 = time.perf_counter()
        sysm = assemble(mesh)
        t1 = time.perf_counter()
        if x_shot is not None:          # --single-shot on several ranks: the gathered host solution feeds the anchors below
            res = sysm.solve_cg_into(params, x_shot, use_pc=wl["pc"])
        else:
            res = sysm.solve_cg_device(params, use_pc=wl["pc"], d_x=d_x.data_ptr() if world == 1 else None)
        info.update(host_ms_assemble=1e3 * (t1 - t0), host_ms_solve=1e3 * (time.perf_counter() - t1))
        info.update(res, nnz=sysm.nnz, n=sysm.n, ms_elements=sysm.info.ms_elements, ms_structure=sysm.info.ms_structure)
        info.update(sysm.local_sizes())
        return sysm

    gpu_uuid = "GPU-" + str(torch.cuda.get_device_properties(local_rank).uuid).replace("GPU-", "")
    sampler = ClockSampler(gpu_uuid)
    sampler.start()                                     # nvidia-smi is up before the warm-up; rows count from mark()
    for _ in range(args.warmup):
        step_resident().close()
    barrier()
    sampler.mark()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = ctx.launches
    e0.record(stream)
    for _ in range(args.steps):
        sy = step_resident()
        if args.single_shot:
            kept.append(sy)             # a solve that takes minutes is not repeated for the roofline leg
        else:
            sy.close()
    e1.record(stream)
    barrier()
    clocks = sampler.stop()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    launches = ctx.launches - l0
    ms_step = ms_total / args.steps
    value = n / (ms_step * 1e-3)

    # ---------------- roofline leg: the CG SpMV, CUDA events on the launching stream ----------------
    sysm = kept.pop() if kept else step_resident()
    # correctness anchors of the solution just produced (after the timed region, never inside it): the TRUE residual
    # ||b - A x|| / ||b|| through an independent one-thread-per-row kernel (yfd_system_true_residual; the solver itself
    # stops on the recurrence residual, core/solvers.hpp:99,124) and the driver's L2 error against the analytic
    # solution (dg2d_diffusion.cpp:225-268, yfd_dg_postprocess)
    anchors = {}
    if x_shot is not None:
        anchors["true_residual"] = sysm.true_residual(x=x_shot)
        if not cfem:
            qp, mm = ctx.dg_postprocess(mesh, k, x_shot)
            anchors["l2_error"], anchors["l2_error_projection"] = float(np.sqrt(qp)), float(np.sqrt(mm))
    if world == 1:
        anchors["true_residual"] = sysm.true_residual(d_x=d_x.data_ptr())
        if not cfem:
            qp, mm = ctx.dg_postprocess(mesh, k, None, device_ptr=d_x.data_ptr())
            anchors["l2_error"], anchors["l2_error_projection"] = float(np.sqrt(qp)), float(np.sqrt(mm))
    sp = sysm.time_spmv(reps=50)
    oi = None if cfem else sysm.operator_info()
    sp_streamed = None
    if oi and oi["uses_classes"]:            # the general (streamed) kernel on the same system, timed beside the class kernel
        sysm.set_operator_mode(yfd.OPERATOR_STREAMED)
        sp_streamed = sysm.time_spmv(reps=20)
        sysm.set_operator_mode(yfd.OPERATOR_AUTO)
    fp64_peak = ctx.measure_fp64()
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    csr_gbs = sp["csr_bytes"] / (sp["ms"] * 1e-3) / 1e9
    achieved = sp["kernel_bytes"] / (sp["ms"] * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "spmv_traffic.json")
    if os.path.exists(tpath) and world == 1:
        traffic = json.load(open(tpath)).get(args.workload)
    kname = "csr_spmv_tma" if cfem else (("dg_spmv_class_mma" if oi.get("tensor_pipe") else "dg_spmv_class") if oi["uses_classes"] else "dg_spmv_tma")
    roofline = {"bound": "hbm", "kernel": kname + " (y = A d fused with d.y)%s" % ("" if world == 1 else ", rank 0's share of the rows"), "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "hw_frac": achieved / peak, "traffic": traffic,
                "traffic_source": None if traffic is None else "static: profiles/spmv_traffic.json (one ncu --set full capture of this kernel on this workload, dram__bytes_read.sum + dram__bytes_write.sum; not a live counter)",
                "peak_source": peak_src,
                "kernel_requested_bytes_per_launch": sp["kernel_bytes"], "ms_per_launch": sp["ms"],
                "csr_equivalent": {"bytes_per_launch": sp["csr_bytes"], "gbs": csr_gbs, "frac": csr_gbs / peak},
                "note": "achieved / frac = the bytes the kernel requests per launch (DG: 8 B per nonzero -- the block layout stores no "
                        "column indices -- + block lists + x + y) / CUDA-event launch time, against the measured HBM peak: the fraction "
                        "the hardware sustains.  csr_equivalent = SURVEY.md 8d's CSR figure (12 B per nonzero + vectors) over the same "
                        "time: what a CSR kernel would have to sustain to be as fast (can exceed the peak)"}
    flops_spmv = 2.0 * info["nnz"]
    roofline["fp64"] = {"achieved_tflops": flops_spmv / (sp["ms"] * 1e-3) / 1e12, "peak_tflops": fp64_peak,
                        "tensor_peak_tflops": ctx.measure_fp64_tensor(),
                        "frac": flops_spmv / (sp["ms"] * 1e-3) / 1e12 / fp64_peak,
                        "peak_source": "DFMA microbenchmark run just now on this device (yfd_ctx_measure_fp64)"}
    if oi is not None:
        roofline["operator"] = oi
    if sp_streamed is not None:
        sg = sp_streamed["kernel_bytes"] / (sp_streamed["ms"] * 1e-3) / 1e9
        roofline["streamed_kernel"] = {"kernel": "dg_spmv_tma (the general path: per-cell values streamed by TMA; what this system would run without congruent cells)",
                                       "ms_per_launch": sp_streamed["ms"], "kernel_requested_bytes_per_launch": sp_streamed["kernel_bytes"],
                                       "achieved": sg, "frac": sg / peak, "speedup_of_class_kernel": sp_streamed["ms"] / sp["ms"]}
        roofline["note"] += ("; this mesh consists of %d congruence classes of cells, so the solver multiplies by one block row per class held in "
                             "shared memory (dg_spmv_class): the matrix stream is gone, the kernel is bound by FP64 issue / shared-memory broadcasts "
                             "(see fp64) and by the x / y vectors, and its HBM fraction is low BY DESIGN; streamed_kernel is the HBM-bound general path "
                             "on the same system" % oi["nclasses"])
    # element kernel (dg_sip_cells): SURVEY.md 8d's flop count and minimum bytes over its CUDA-event time
    roofline_elements = None
    if not cfem:
        nvv = 3 if wl["cell"] == 0 else 4
        nq = {1: 3, 2: 6, 3: 12, 4: 16, 5: 25, 6: 33}.get(k, 1) if nvv == 3 else (k + 1) ** 2
        ncl = info["n_local"] / bs
        fl = ncl * (nq * (5 * bs * bs + 2 * bs) + nvv * (k + 1) * (12 * bs * bs + 6 * bs))
        by = 8.0 * info["nnz_local"] + 8.0 * info["n_local"]
        tms = info["ms_elements"] * 1e-3
        roofline_elements = {"kernel": "dg_sip_cells (K, Att, Atn, rhs per cell on quadrature)", "ms": info["ms_elements"], "flops": fl, "bytes_written": by,
                             "achieved_tflops": fl / tms / 1e12, "fp64_frac": fl / tms / 1e12 / fp64_peak, "written_gbs": by / tms / 1e9, "hbm_frac": by / tms / 1e9 / peak,
                             "note": "one lane per block entry, 2-5 scalar LDS per FMA: shared-memory latency bound, far from both roofs; 2-3 % of a step"}
    # whole CG iteration (SpMV + 2 fused vector kernels) over the timed solve: B_cg = B_spmv + 72 n,
    # summed over ranks
    it = max(1, info["iterations"])
    nnz_global = info["nnz"]
    if dist is not None:
        tn = torch.tensor([info["nnz"]], device="cuda", dtype=torch.int64)
        dist.all_reduce(tn)
        nnz_global = int(tn.item())
    # bytes the three kernels of an iteration request: matrix (DG: 8 B/nnz + block lists; CSR: 12 B/nnz + offsets) + x, y + 72 n
    mat_bytes = (12.0 * nnz_global + 8.0 * (info["n"] + 1)) if cfem else (8.0 * nnz_global + 12.0 * nnz_global / (bs * bs) + 8.0 * info["n"] / bs)
    if oi is not None and oi["uses_classes"]:
        mat_bytes = (sp["kernel_bytes"] - 16.0 * info["n_local"]) * world       # index lists only
    cg_gbs = (mat_bytes + 16.0 * info["n"] + 72.0 * info["n"]) * it / (info["ms_solve"] * 1e-3) / 1e9
    # the whole CG iteration = this SpMV + cg_update_xr + cg_update_d (72 n bytes, both at the HBM roof): requested bytes / time
    roofline["cg_iteration"] = {"requested_gbs": cg_gbs, "frac": cg_gbs / peak / max(1, world), "ms": info["ms_solve"] / it,
                                "note": "bytes the three kernels of an iteration request (index lists or matrix + x, y + 72 n) over the solve's time per iteration, per GPU"}
    sysm.close()

    # ---------------- end-to-end leg: host buffers in, host solution out ----------------
    e2e = None
    if not args.no_e2e:
        keep = []
        arrs = {}
        for nme in ("points", "cells", "faces", "face_boundary"):
            t, a = pinned_like(torch, getattr(hm, nme))
            keep.append(t); arrs[nme] = a
        xt = torch.empty(n, dtype=torch.float64).pin_memory()
        x_host = xt.numpy()
        hm_p = yfd.HostMesh(hm.cell_type, arrs["points"], arrs["cells"], arrs["faces"], arrs["face_boundary"])
        h2d = sum(a.nbytes for a in arrs.values())
        d2h = x_host.nbytes

        own = {}

        def step_e2e():
            m2 = yfd.Mesh(ctx, hm_p)
            s2 = assemble(m2)
            if world > 1 and not cfem:      # several ranks: every rank reads back ITS rows (no gather, 1/N of the bytes per rank)
                nl = s2.local_sizes()["n_local"]
                if "t" not in own or own["t"].numel() < nl:
                    own["t"] = torch.empty(nl, dtype=torch.float64).pin_memory()
                s2.solve_cg_owned_into(params, own["t"].numpy()[:nl], use_pc=wl["pc"])
                own["d2h"] = nl * 8
            else:
                s2.solve_cg_into(params, x_host, use_pc=wl["pc"])
            s2.close(); m2.close()

        mesh.close()
        step_e2e()
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        f0.record(stream)
        for _ in range(args.steps):
            step_e2e()
        f1.record(stream)
        barrier()
        wall = (time.perf_counter() - t0) / args.steps
        # the events are recorded before the first / after the last call, so they span the host-side work too
        ms_e2e = max_over_ranks(f0.elapsed_time(f1)) / args.steps
        if world > 1:                       # several ranks: the anchors come from the gathered host solution (untimed extra pass)
            m2 = yfd.Mesh(ctx, hm_p)
            s2 = assemble(m2)
            s2.solve_cg_into(params, x_host, use_pc=wl["pc"])
            anchors["true_residual"] = s2.true_residual(x=x_host)
            if not cfem:
                qp, mm = ctx.dg_postprocess(m2, k, x_host)
                anchors["l2_error"], anchors["l2_error_projection"] = float(np.sqrt(qp)), float(np.sqrt(mm))
            s2.close(); m2.close()
        e2e = {"value": n / (ms_e2e * 1e-3), "unit": "DOFs/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(own.get("d2h", d2h)),
               "ms_per_step": ms_e2e, "wall_ms_per_step": 1e3 * wall,
               "path": ("yfd_mesh_create -> %s -> yfd_system_solve_cg (host x) through include/yfd.h" if "d2h" not in own else
                        "yfd_mesh_create -> %s -> yfd_system_solve_cg_owned (every rank reads back its own rows: d2h_bytes_per_step is rank 0's "
                        "share; h2d: every rank uploads the whole mesh) through include/yfd.h") % (("yfd_cfem_assemble_p2" if wl.get("order") == 2 else "yfd_cfem_assemble_p1") if cfem else "yfd_dg_assemble_sip")}

    base = None
    same_size = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:      # the CPU baseline is an N=1 figure (torchrun pins ranks to one thread)
        base = cpu_baseline(wl, wl["base_r"])
        if base is not None:
            st = cpu_single_thread(wl, wl["ref_r"])
            if st is not None:
                base["as_shipped_single_thread"] = st
        # like-for-like: the GPU path on EXACTLY the systems the CPU reference runs (the reference arm's -r ref_r sample and
        # this line's cpu_baseline at -r base_r), resident and end to end, with both sides' iteration counts
        if base is not None:
            same_size = []
            for rr in sorted({wl["ref_r"], wl["base_r"]}):
                cpu = base if rr == wl["base_r"] else cpu_baseline(wl, rr)
                if cpu is None or wl.get("order") == 2:
                    continue
                same_size.append(same_size_leg(torch, yfd, ctx, stream, wl, rr, cpu, assemble, cfem))

    if rank == 0:
        line = {"profiling_run_not_a_bench_value": True} if args.profile_iters else ({"single_shot": True} if args.single_shot else {})
        line.update({"metric": "assembled+solved DOFs/sec", "value": value, "unit": "DOFs/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": args.workload, "description": wl["desc"], "dofs": int(info["n"]), "nnz": int(nnz_global),
                           "cg": "rr_tol %.0e, rr_max per the reference driver, max_iter 2N, x0 = 0, %s" % (wl["rr_tol"], ("no preconditioner (reference default)", "Jacobi (-p)", "block-Jacobi (extension, YFD_PC_BLOCK_JACOBI; reference arm: Jacobi)", "p-multigrid V-cycle (extension, YFD_PC_PMG; reference arm: Jacobi)")[int(wl["pc"])]),
                           "cg_iterations": int(info["iterations"]), "converged": bool(info["converged"]),
                           "true_residual": anchors.get("true_residual"), "l2_error": anchors.get("l2_error"),
                           "l2_error_projection": anchors.get("l2_error_projection"),
                           "anchors_note": "true_residual = ||b - A x|| / ||b|| of the returned solution by an independent one-thread-per-row kernel "
                                           "(yfd_system_true_residual); l2_error = the driver's L2 error against sin(pi x) sin(pi y) at the quadrature "
                                           "points (dg2d_diffusion.cpp:225-268), expected O(h^(k+1))",
                           "l2_policy": "inputs larger than L2: CSR values alone are %.1f GB vs 126 MB L2" % (8e-9 * info["nnz"]),
                           "phases_ms": {"structure": info["ms_structure"], "elements": info["ms_elements"], "cg": info["ms_solve"],
                                         "host_wall_assemble_call": info["host_ms_assemble"], "host_wall_solve_call": info["host_ms_solve"]},
                           "ms_per_cg_iteration": info["ms_solve"] / it, "dof_iterations_per_s": info["n"] * it / (info["ms_solve"] * 1e-3),
                           "cg_iteration_gbs": cg_gbs, "parallelism": "1 GPU" if world == 1 else "contiguous blocks of the Z-curve cell order over %d GPUs: rank 0 holds %d rows + %d halo dofs; "
                                          "transport: %s (%s)" % (world, info["n_local"], info["n_halo"], info.get("transport"),
                                                                  "per CG iteration 2 cross-rank scalar sums through NVLink peer memory; no exchange of d: every rank advances its halo "
                                                                  "copies of r and d from the neighbours' y = A d, stored into its landing zone ahead of the r.r sum"
                                                                  if info.get("transport") == "peer memory" else "per CG iteration one halo exchange and 2 scalar all-reduces")},
                "roofline": roofline, "roofline_elements": roofline_elements, "cpu_baseline": base, "same_size": same_size, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks})
        print(json.dumps(line))
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())

"""GPU: the reference's UNMODIFIED dg2d_diffusion.cpp recompiled against include/yaourt_b200/dropin*.hpp
(oracle/_ref/dropin_dg2d_diffusion, built by oracle/Makefile where /root/reference exists).  Its finalize() ->
blaze::init_from_triplets and its conjugated_gradient calls run on the GPU; the driver's own printed L2 errors must
match what the reference itself produces (golden fixture + live oracle/_ref)."""
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, golden

pytestmark = pytest.mark.gpu
EXE = os.path.join(ROOT, "oracle", "_ref", "dropin_dg2d_diffusion")


def run(args, env=None):
    if not os.path.exists(EXE):
        pytest.skip("oracle/_ref/dropin_dg2d_diffusion not built (needs /root/reference at build time)")
    e = dict(os.environ)
    e.update(env or {})
    res = subprocess.run([EXE] + args, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=e, timeout=600, cwd="/tmp")
    assert res.returncode == 0, res.stdout[-2000:]
    return res.stdout


@pytest.mark.parametrize("row", [0, 1, 2])
def test_unmodified_driver_on_gpu_prints_the_reference_errors(row):
    t, r, k, n, nnz, iters, l2_qp, l2_mm, h = golden("dg_driver_runs.npz")["runs"][row]
    out = run(["-m", "tri" if t == 0 else "quad", "-r", str(int(r)), "-k", str(int(k))])
    qp = float(re.search(r"L2-norm error \(qp\):\s+(\S+)", out).group(1))
    mm = float(re.search(r"L2-norm error \(mm\):\s+(\S+)", out).group(1))
    it = int(re.search(r"\[yaourt_b200\] CG on GPU: iterations (\d+)", out).group(1))
    assert "converged 1" in out
    # the driver prints 6 significant digits; CG stops at rr 1e-8, so errors agree to ~1e-6 relative
    assert abs(qp - l2_qp) <= 2e-5 * l2_qp and abs(mm - l2_mm) <= 2e-5 * l2_mm
    assert abs(it - iters) <= max(3, 0.05 * iters)


def test_preconditioned_driver_flag(ref):
    out = run(["-r", "3", "-k", "2", "-p"])
    ref.mesh(0, 3)
    ref.set_cg(rr_tol=1e-8, rr_max=1e4, max_iter=-1)
    o = ref.run_dg(2, use_pc=True)
    qp = float(re.search(r"L2-norm error \(qp\):\s+(\S+)", out).group(1))
    it = int(re.search(r"iterations (\d+)", out).group(1))
    assert abs(qp - o["l2_qp"]) <= 2e-5 * o["l2_qp"] and abs(it - o["iters"]) <= max(3, 0.05 * o["iters"])


@pytest.mark.parametrize("spec,t,r,k,pc", [("tri 3 2 1", 0, 3, 2, True), ("quad 3 3 0", 1, 3, 3, False)])
def test_batched_entry_from_cpp(ref, spec, t, r, k, pc):
    out = run([], env={"YFD_DROPIN_BATCHED": spec})
    m = re.search(r"batched: n (\d+) nnz (\d+) iterations (\d+) sol_norm (\S+) rhs_norm (\S+)", out)
    n, nnz, it, sn, rn = int(m.group(1)), int(m.group(2)), int(m.group(3)), float(m.group(4)), float(m.group(5))
    ref.mesh(t, r)
    ref.set_cg(rr_tol=1e-12, rr_max=1e4, max_iter=-1)
    o = ref.run_dg(k, use_pc=pc)
    assert (n, nnz) == (o["n"], o["nnz"])
    assert abs(sn - np.linalg.norm(o["sol"])) <= 1e-10 * np.linalg.norm(o["sol"])
    assert abs(rn - np.linalg.norm(o["rhs"])) <= 1e-12 * np.linalg.norm(o["rhs"])
    assert abs(it - o["iters"]) <= max(3, 0.05 * o["iters"])


# ---- the reference's UNMODIFIED maxwell_driver.cpp against include/yaourt_b200/maxwell.hpp ----
MX_EXE = os.path.join(ROOT, "oracle", "_ref", "dropin_maxwell")


@pytest.mark.parametrize("name,extra", [("tri_r2_k2", []), ("quad_r2_k1_up", ["-u"]), ("tri_r3_k3_up", ["-u"])])
def test_unmodified_maxwell_driver_on_gpu_writes_the_reference_error_log(name, extra, tmp_path):
    """run_maxwell_solver as shipped (maxwell_driver.cpp:194-272): assemble(ctx) and do_timestep(ctx) bind to the GPU
    overloads, everything else is the reference's own code; its error log must equal the log of the reference's own run."""
    if not os.path.exists(MX_EXE):
        pytest.skip("oracle/_ref/dropin_maxwell not built (needs /root/reference at build time)")
    logs = golden("maxwell_driver_logs.npz")
    t, r, k, up, dt, steps = logs[name + "_cfg"]
    errfile = str(tmp_path / "err.txt")
    res = subprocess.run([MX_EXE, "-m", "tri" if t == 0 else "quad", "-r", str(int(r)), "-k", str(int(k)), "-d", repr(float(dt)),
                          "-t", str(int(steps)), "-E", errfile] + extra,
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600, cwd="/tmp")
    assert res.returncode == 0, res.stdout[-2000:]
    assert "[yaourt_b200] maxwell contexts on the GPU: 1" in res.stdout
    rows = np.array([list(map(float, ln.split())) for ln in open(errfile) if ln.strip() and not ln.startswith("#")])
    want = logs[name]
    assert rows.shape == want.shape
    assert np.array_equal(rows[:, 0], want[:, 0])
    assert np.allclose(rows[:, 2:5], want[:, 2:5], rtol=2e-5, atol=0)          # 6 printed digits


# ---- the reference's UNMODIFIED dg2d_advection.cpp: finalize() and CG on the normal equations on the GPU ----
ADV_EXE = os.path.join(ROOT, "oracle", "_ref", "dropin_dg2d_advection")


@pytest.mark.parametrize("name", ["tri_r2_k1", "quad_r1_k3_up"])
def test_unmodified_advection_driver_on_gpu(name):
    """cgp.use_normal_eqns = true (dg2d_advection.cpp:225) goes through the drop-in conjugated_gradient -> yfd_cg_solve,
    which builds A^T on the device; the driver's printed L2 error must match the reference's own run."""
    if not os.path.exists(ADV_EXE):
        pytest.skip("oracle/_ref/dropin_dg2d_advection not built (needs /root/reference at build time)")
    g = golden("adv_%s.npz" % name)
    args = ["-m", "tri" if int(g["mtype"]) == 0 else "quad", "-r", str(int(g["levels"])), "-k", str(int(g["degree"]))]
    if int(g["upwinding"]):
        args.append("-u")
    res = subprocess.run([ADV_EXE] + args, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600, cwd="/tmp")
    assert res.returncode == 0, res.stdout[-2000:]
    qp = float(re.search(r"L2-norm error \(qp\):\s+(\S+)", res.stdout).group(1))
    it = int(re.search(r"\[yaourt_b200\] CG on GPU: iterations (\d+)", res.stdout).group(1))
    # CG on A^T A is ill-conditioned: iteration counts of two correct implementations differ by a few percent
    assert abs(qp - float(g["l2_qp"])) <= 0.02 * float(g["l2_qp"])
    assert abs(it - int(g["iters"])) <= max(5, 0.15 * int(g["iters"]))


# ---- the batched C++ entries SURVEY.md 8b asks for beyond the reference's per-block interface ----
@pytest.mark.parametrize("spec,t,r,k,pc", [("tri 3 2 1 fn", 0, 3, 2, True), ("quad 2 3 0 fn", 1, 2, 3, False)])
def test_batched_entry_with_function_ids(ref, spec, t, r, k, pc):
    """assemble_dg_sip(msh, degree, eta, rhs_fn_id, dirichlet_fn_id, ...) with the driver's own data functions"""
    out = run([], env={"YFD_DROPIN_BATCHED": spec})
    m = re.search(r"batched: n (\d+) nnz (\d+) iterations (\d+) sol_norm (\S+) rhs_norm (\S+)", out)
    n, nnz, it, sn, rn = int(m.group(1)), int(m.group(2)), int(m.group(3)), float(m.group(4)), float(m.group(5))
    ref.mesh(t, r)
    ref.set_cg(rr_tol=1e-12, rr_max=1e4, max_iter=-1)
    o = ref.run_dg(k, use_pc=pc)
    assert (n, nnz) == (o["n"], o["nnz"])
    assert abs(sn - np.linalg.norm(o["sol"])) <= 1e-10 * np.linalg.norm(o["sol"])
    assert abs(rn - np.linalg.norm(o["rhs"])) <= 1e-12 * np.linalg.norm(o["rhs"])
    assert abs(it - o["iters"]) <= max(3, 0.05 * o["iters"])


@pytest.mark.parametrize("levels", [2, 4])
def test_batched_cfem_entry(ref, levels):
    """assemble_cfem_p1(msh, rhs_fn_id, &cgp) against the unmodified continuous_fem.cpp"""
    out = run([], env={"YFD_DROPIN_BATCHED": "tri %d 1 0 cfem" % levels})
    m = re.search(r"batched cfem: n (\d+) nnz (\d+) iterations (\d+) sol_norm (\S+) rhs_norm (\S+) expanded_norm (\S+) dirichlet (\d+)", out)
    n, nnz, it = int(m.group(1)), int(m.group(2)), int(m.group(3))
    sn, rn, en, nd = float(m.group(4)), float(m.group(5)), float(m.group(6)), int(m.group(7))
    ref.set_cg(rr_tol=1e-12, rr_max=1e3)
    o = ref.run_cfem(levels)
    assert (n, nnz) == (o["n"], o["nnz"])
    assert abs(sn - np.linalg.norm(o["sol"])) <= 1e-10 * np.linalg.norm(o["sol"])
    assert abs(rn - np.linalg.norm(o["rhs"])) <= 1e-12 * np.linalg.norm(o["rhs"])
    assert abs(en - sn) <= 1e-12 * sn                      # expand() pads the Dirichlet nodes with zeros
    assert nd == 4 * 2 ** levels                           # boundary points of the reference's triangle mesh
    assert abs(it - o["iters"]) <= max(3, 0.05 * o["iters"])


@pytest.mark.parametrize("name", ["tri_r2_k1", "quad_r1_k3_up"])
def test_batched_advection_entry(name):
    """assemble_dg_advection(msh, degree, eta, use_upwinding, &cgp) with cgp.use_normal_eqns as the driver solves
    (dg2d_advection.cpp:225): sizes and rhs of the reference's fixture, solution in the same distance class"""
    if not os.path.exists(ADV_EXE):
        pytest.skip("oracle/_ref/dropin_dg2d_advection not built (needs /root/reference at build time)")
    try:
        g = golden("adv_%s.npz" % name)
    except FileNotFoundError:
        pytest.skip("no fixture adv_%s.npz" % name)
    spec = "%s %d %d %d" % ("tri" if int(g["mtype"]) == 0 else "quad", int(g["levels"]), int(g["degree"]), int(g["upwinding"]))
    e = dict(os.environ)
    e["YFD_DROPIN_BATCHED"] = spec
    res = subprocess.run([ADV_EXE], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=e, timeout=600, cwd="/tmp")
    assert res.returncode == 0, res.stdout[-2000:]
    m = re.search(r"batched: n (\d+) nnz (\d+) iterations (\d+) sol_norm (\S+) rhs_norm (\S+)", res.stdout)
    n, nnz, it, sn, rn = int(m.group(1)), int(m.group(2)), int(m.group(3)), float(m.group(4)), float(m.group(5))
    assert n == len(g["rhs"]) and nnz == len(g["values"])
    assert abs(rn - np.linalg.norm(g["rhs"])) <= 1e-12 * np.linalg.norm(g["rhs"])
    # CG on A^T A: iterates of two correct implementations drift (deviation 3); the solution is pinned by
    # a direct solve of the reference's own matrix
    import scipy.sparse as sp
    import scipy.sparse.linalg as spla
    A = sp.csr_matrix((g["values"], g["colidx"], g["rowptr"]), shape=(n, n))
    want = spla.spsolve(A.tocsc(), g["rhs"])
    assert abs(sn - np.linalg.norm(want)) <= 1e-4 * np.linalg.norm(want)
    assert abs(it - int(g["iters"])) <= max(5, 0.15 * int(g["iters"]))

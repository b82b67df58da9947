"""GPU parity of the DG Maxwell TM path (yfd_maxwell_*) through the C ABI:
  - against the golden vectors written from the UNMODIFIED reference (tests/golden/maxwell_*.npz),
  - against the reference itself (oracle/_ref/libyfd_ref_maxwell.so) and the C restatement on larger meshes,
  - size-independent properties at sizes no CPU oracle run fits (energy conservation of the centred scheme,
    convergence order of the driver's error norms, linearity of the operator).

The reference's operator is dense per cell; yfd_maxwell_get_operator rebuilds that layout from the factored
blocks the kernels keep, so the comparison is entry by entry in the reference's numbering.
Tolerances: offdiag_neigh_offsets bit-exact; mass matrices 1e-13; operator entries 1e-11 relative to the operator's
max-norm (M^-1 is an LU here and LAPACK/closed forms there: cond(M) * eps, cond(M) <= 1.5e5 at k = 4);
dofs after time stepping 1e-10 relative L2 (north_star's solution bar)."""
import glob
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu
OP_TOL, SOL_TOL = 1e-11, 1e-10
FILES = sorted(f for f in glob.glob(os.path.join(GOLDEN, "maxwell_*.npz")) if "driver_logs" not in f)


def _mesh(yfd, ctx, mtype, points, cells, faces, fb, reorder=1):
    hm = yfd.HostMesh(int(mtype), points, cells, faces, np.asarray(fb, np.uint8))
    return yfd.Mesh(ctx, hm, reorder=reorder)


@pytest.mark.parametrize("reorder", [0, 1])
@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f)[8:-4] for f in FILES])
def test_against_reference_fixture(yfd, ctx, path, reorder):
    g = np.load(path)
    k, up, euler, dt, steps = int(g["degree"]), bool(g["upwind"]), bool(g["euler"]), float(g["dt"]), int(g["steps"])
    mesh = _mesh(yfd, ctx, g["mtype"], g["points"], g["cells"], g["faces"], g["face_boundary"], reorder)
    mx = yfd.Maxwell(ctx, mesh, k, delta_t=dt, eta=float(g["eta"]), upwind=up, euler=euler,
                     mu_r=g["mu_r"] if "mu_r" in g else None, eps_r=g["eps_r"] if "eps_r" in g else None)
    op = mx.operator()
    assert np.array_equal(op["neigh"], g["neigh"]) and np.array_equal(op["has"], g["has"])
    assert np.abs(op["gM"] - g["gM"]).max() <= 1e-13 * np.abs(g["gM"]).max()
    sc = np.abs(g["ondiag"]).max()
    assert np.abs(op["ondiag"] - g["ondiag"]).max() <= OP_TOL * sc
    assert np.abs(op["offdiag"] - g["offdiag"]).max() <= OP_TOL * sc
    # apply_initial_condition
    mx.apply_initial_condition()
    d0 = mx.dofs()
    assert np.linalg.norm(d0 - g["dofs0"]) <= SOL_TOL * np.linalg.norm(g["dofs0"])
    # do_timestep from the reference's own initial dofs
    mx.set_dofs(g["dofs0"])
    if "dofs1" in g:
        mx.timesteps(1)
        assert np.linalg.norm(mx.dofs() - g["dofs1"]) <= SOL_TOL * np.linalg.norm(g["dofs1"])
        e = mx.report_errors(0.0)
        assert np.abs(e - g["errors"][0]).max() <= 1e-11
        mx.timesteps(steps - 1)
    else:
        mx.timesteps(steps)
    assert np.linalg.norm(mx.dofs() - g["dofsN"]) <= SOL_TOL * np.linalg.norm(g["dofsN"])
    if "errors" in g:
        e = mx.report_errors((steps - 1) * dt)
        assert np.abs(e - g["errors"][-1]).max() <= 1e-11


def test_apply_operator_matches_dense_reference_blocks(yfd, ctx):
    """v_next = D_c v_c + sum_f O_f v_neigh with the reference's dense blocks from the fixture (apply_operator,
    methods/dg_maxwell_2D.hpp:586-625) against the factored kernel."""
    g = np.load(os.path.join(GOLDEN, "maxwell_tri_r1_k2_up_shatter.npz"))
    k = int(g["degree"])
    mesh = _mesh(yfd, ctx, g["mtype"], g["points"], g["cells"], g["faces"], g["face_boundary"])
    mx = yfd.Maxwell(ctx, mesh, k, upwind=True)
    rng = np.random.default_rng(3)
    nc, nf = g["neigh"].shape
    L = g["ondiag"].shape[1]
    v = rng.standard_normal(nc * L)
    want = np.einsum("cij,cj->ci", g["ondiag"], v.reshape(nc, L))
    for f in range(nf):
        want += g["has"][:, f, None] * np.einsum("cij,cj->ci", g["offdiag"][:, f], v.reshape(nc, L)[g["neigh"][:, f]])
    got = mx.apply_operator(v)
    assert np.abs(got - want.ravel()).max() <= 1e-11 * np.abs(want).max()


@pytest.mark.parametrize("mtype,r,k,up,sh", [(0, 3, 1, 0, 0), (0, 3, 2, 1, 1), (1, 3, 3, 0, 1), (0, 2, 4, 1, 0), (1, 2, 4, 0, 0),
                                             (0, 1, 5, 0, 1), (1, 1, 6, 1, 0), (0, 2, 0, 1, 0), (0, 1, 6, 0, 0), (1, 2, 5, 1, 1)])
def test_against_live_reference(yfd, ctx, mtype, r, k, up, sh):
    import refbind
    if not refbind.maxwell_available():
        pytest.skip("oracle/_ref/libyfd_ref_maxwell.so not built")
    R = refbind.RefMaxwell()
    dt = 0.002
    R.create(mtype, r, k, dt=dt, upwind=up, shatter=sh)
    m = R.mesh(); R.assemble(); rop = R.operator(); R.init()
    mesh = _mesh(yfd, ctx, mtype, m["points"], m["cells"], m["faces"], m["face_boundary"])
    mx = yfd.Maxwell(ctx, mesh, k, delta_t=dt, upwind=bool(up))
    op = mx.operator()
    assert np.array_equal(op["neigh"], rop["neigh"]) and np.array_equal(op["has"], rop["has"])
    sc = np.abs(rop["ondiag"]).max()
    tol = OP_TOL * (1 if k <= 4 else 100)           # cond(M) grows ~30x per degree in the monomial basis
    assert np.abs(op["ondiag"] - rop["ondiag"]).max() <= tol * sc
    assert np.abs(op["offdiag"] - rop["offdiag"]).max() <= tol * sc
    d0 = R.dofs()
    mx.set_dofs(d0)
    R.step(8); mx.timesteps(8)
    want = R.dofs()
    assert np.linalg.norm(mx.dofs() - want) <= SOL_TOL * (1 if k <= 4 else 100) * np.linalg.norm(want)
    assert np.abs(mx.report_errors(7 * dt) - R.errors(7)).max() <= 1e-10


def test_driver_error_log(yfd, ctx):
    """run_maxwell_solver exactly as shipped (maxwell_driver.cpp:194-272): its error log, cycle by cycle, on the
    reference's mesh at -r 3, degree 3, upwind."""
    logs = np.load(os.path.join(GOLDEN, "maxwell_driver_logs.npz"))
    for name in ("tri_r2_k2", "quad_r2_k1_up", "tri_r3_k3_up"):
        t, r, k, up, dt, steps = logs[name + "_cfg"]
        mesh = yfd.Mesh(ctx, yfd.HostMesh.create(int(t), int(r)))
        mx = yfd.Maxwell(ctx, mesh, int(k), delta_t=dt, upwind=bool(up))
        mx.apply_initial_condition()
        for cyc in range(int(steps)):
            mx.timesteps(1)
            e = mx.report_errors(cyc * dt)
            assert np.allclose(e, logs[name][cyc][2:5], rtol=2e-5, atol=0)      # 6 printed digits


def test_centred_scheme_conserves_energy_at_scale(yfd, ctx):
    """Property at a size no CPU oracle run fits (tri -r 7, k = 3: 65 536 cells, 1.97 M unknowns): with centred
    fluxes and PEC walls the semi-discrete operator L satisfies <v, M L v> = 0, so RK4 changes the discrete
    energy sum_c v_c^T M_c v_c only at O(dt^5) per step; upwinding must dissipate it."""
    mesh = yfd.Mesh(ctx, yfd.HostMesh.create(0, 7))
    dt = 2e-4
    mx = yfd.Maxwell(ctx, mesh, 3, delta_t=dt)
    gM = mx.operator(blocks=False)["gM"]

    def energy(d):
        v = d.reshape(mx.ncells, 3, mx.bs)
        return float(np.einsum("cfi,cij,cfj->", v, gM, v))
    mx.apply_initial_condition()
    e0 = energy(mx.dofs())
    mx.timesteps(50)
    e1 = energy(mx.dofs())
    assert abs(e1 - e0) <= 1e-9 * e0
    # a rough (random) state has jumps everywhere: the centred scheme still conserves, upwinding must dissipate
    rng = np.random.default_rng(5)
    rough = rng.standard_normal(mx.ndofs)
    mx.set_time_integration(dt / 10)          # RK4 itself damps modes with lambda dt ~ 1: keep the rough state well resolved in time
    mx.set_dofs(rough)
    r0 = energy(rough)
    mx.timesteps(20)
    r1 = energy(mx.dofs())
    assert abs(r1 - r0) <= 1e-7 * r0
    up = yfd.Maxwell(ctx, mesh, 3, delta_t=dt / 10, upwind=True)
    up.set_dofs(rough)
    up.timesteps(20)
    r2 = energy(up.dofs())
    assert r2 < 0.99 * r0


def test_convergence_order(yfd, ctx):
    """The driver's L2 errors (report_errors) after the same physical time on -r 3,4,5 with k = 2, upwind fluxes:
    order k+1 = 3 within the usual margin."""
    errs = []
    for r in (3, 4, 5):
        mesh = yfd.Mesh(ctx, yfd.HostMesh.create(0, r))
        steps = 40 * 2 ** (r - 3)
        dt = 0.1 / steps
        mx = yfd.Maxwell(ctx, mesh, 2, delta_t=dt, upwind=True)
        mx.apply_initial_condition()
        mx.timesteps(steps)
        errs.append(mx.report_errors(0.1))
    errs = np.array(errs)
    orders = np.log2(errs[:-1] / errs[1:])
    assert (orders > 2.5).all(), orders


def test_linearity_and_zero(yfd, ctx):
    mesh = yfd.Mesh(ctx, yfd.HostMesh.create(1, 4))
    mx = yfd.Maxwell(ctx, mesh, 4, upwind=True)
    rng = np.random.default_rng(11)
    a, b = rng.standard_normal(mx.ndofs), rng.standard_normal(mx.ndofs)
    la, lb, lab = mx.apply_operator(a), mx.apply_operator(b), mx.apply_operator(2.0 * a - 3.0 * b)
    assert np.abs(lab - (2.0 * la - 3.0 * lb)).max() <= 1e-11 * np.abs(lab).max()
    assert not mx.apply_operator(np.zeros(mx.ndofs)).any()
    mx.apply_initial_condition(fields=yfd.FIELDS_ZERO)
    mx.timesteps(3)
    assert not mx.dofs().any()


def test_errors_and_edge_cases(yfd, ctx):
    mesh = yfd.Mesh(ctx, yfd.HostMesh.create(0, 0))
    with pytest.raises(yfd.YfdError):
        yfd.Maxwell(ctx, mesh, 7)                      # beyond the instantiated degrees
    with pytest.raises(yfd.YfdError):
        yfd.Maxwell(ctx, mesh, -1)
    with pytest.raises(yfd.YfdError):
        yfd.Maxwell(ctx, mesh, 1, mu_r=np.ones(mesh.ncells))          # mu_r without eps_r
    mx = yfd.Maxwell(ctx, mesh, 1)
    with pytest.raises(yfd.YfdError):
        mx.set_dofs(np.zeros(5))
    assert mx.timesteps(0).steps == 0
    info = mx.info
    assert info.ndofs == 3 * 3 * mesh.ncells and info.reference_operator_bytes == 4 * 81 * 8 * mesh.ncells

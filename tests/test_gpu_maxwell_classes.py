"""GPU: the congruence-class cache of the DG Maxwell operator (maxwell_stage_class, yfd_maxwell_step.cu) -- BASELINE
configs[3] "matrix-free element+face operator": one set of 2 + 2 nf blocks per class of congruent cells, the time loop
streams vectors only.  The parity suite proper (tests/test_gpu_maxwell.py: dofs against the unmodified reference, 1e-10)
runs through this path wherever the mesh has classes; here the class path is compared with the streamed path on the same
operator, and its accept / decline logic is pinned."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


CASES = [(0, 3, 1, False), (0, 3, 2, True), (0, 2, 3, False), (0, 3, 4, False), (0, 2, 4, True), (1, 3, 1, True), (1, 2, 3, False), (1, 3, 2, True)]


@pytest.mark.parametrize("t,r,k,up", CASES)
def test_class_operator_equals_streamed_operator(yfd, ctx, t, r, k, up):
    mesh = yfd.Mesh(ctx, yfd.HostMesh.create(t, r))
    mx = yfd.Maxwell(ctx, mesh, k, delta_t=0.002, upwind=up)
    # the cache is there (AUTO then follows the timing made at creation; the test pins the class kernel explicitly)
    assert mx.info.class_reject_reason == 0 and 1 <= mx.info.nclasses <= 64, (mx.info.nclasses, mx.info.class_reject_reason)
    assert mx.info.class_max_deviation <= mx.info.class_tolerance <= 1e-10
    mx.set_operator_mode(yfd.OPERATOR_CLASSES)
    assert mx.info.uses_classes
    rng = np.random.default_rng(k + 7 * r)
    v = rng.standard_normal(mx.ndofs)
    y_cls = mx.apply_operator(v)
    mx.set_dofs(v)
    mx.timesteps(6)
    d_cls = mx.dofs()
    mx.set_operator_mode(yfd.OPERATOR_STREAMED)
    assert not mx.info.uses_classes
    y_str = mx.apply_operator(v)
    mx.set_dofs(v)
    mx.timesteps(6)
    d_str = mx.dofs()
    assert np.abs(y_cls - y_str).max() <= 1e-11 * np.abs(y_str).max()
    assert np.linalg.norm(d_cls - d_str) <= 1e-11 * np.linalg.norm(d_str)
    # Euler through the class path
    mx.set_operator_mode(yfd.OPERATOR_CLASSES)
    mx.set_time_integration(0.001, euler=True)
    mx.set_dofs(v); mx.timesteps(3); e_cls = mx.dofs()
    mx.set_operator_mode(yfd.OPERATOR_STREAMED)
    mx.set_dofs(v); mx.timesteps(3); e_str = mx.dofs()
    assert np.linalg.norm(e_cls - e_str) <= 1e-11 * np.linalg.norm(e_str)


def test_nonuniform_material_keeps_classes(yfd, ctx):
    """mu, eps vary per cell: they live in the per-cell coefficients, not in the blocks, so the classes stay."""
    hm = yfd.HostMesh.create(yfd.TRI, 3)
    mesh = yfd.Mesh(ctx, hm)
    rng = np.random.default_rng(1)
    mu, eps = 1.0 + rng.random(len(hm.cells)), 1.0 + 2.0 * rng.random(len(hm.cells))
    mx = yfd.Maxwell(ctx, mesh, 2, delta_t=0.002, upwind=True, mu_r=mu, eps_r=eps)
    mx.set_operator_mode(yfd.OPERATOR_CLASSES)
    assert mx.info.uses_classes
    v = rng.standard_normal(mx.ndofs)
    a = mx.apply_operator(v)
    mx.set_operator_mode(yfd.OPERATOR_STREAMED)
    b = mx.apply_operator(v)
    assert np.abs(a - b).max() <= 1e-11 * np.abs(b).max()


def test_shattered_mesh_declines(yfd, ctx):
    import refbind
    if not refbind.maxwell_available():
        pytest.skip("oracle/_ref not built")
    R = refbind.RefMaxwell()
    R.create(0, 3, 2, dt=0.002, upwind=False, shatter=True)
    m = R.mesh()
    mesh = yfd.Mesh(ctx, yfd.HostMesh(0, m["points"], m["cells"], m["faces"], m["face_boundary"].astype(np.uint8)))
    mx = yfd.Maxwell(ctx, mesh, 2, delta_t=0.002)
    assert not mx.info.uses_classes and mx.info.class_reject_reason in (1, 2)
    with pytest.raises(yfd.YfdError):
        mx.set_operator_mode(yfd.OPERATOR_CLASSES)
    # and the streamed path still reproduces the reference
    R.assemble(); R.init()
    mx.apply_initial_condition()
    R.step(5); mx.timesteps(5)
    want = R.dofs()
    assert np.linalg.norm(mx.dofs() - want) <= 1e-10 * np.linalg.norm(want)


def test_switch_and_large_mesh(yfd, ctx):
    old = os.environ.get("YFD_MX_CLASSES")
    os.environ["YFD_MX_CLASSES"] = "0"
    try:
        mx = yfd.Maxwell(ctx, yfd.Mesh(ctx, yfd.HostMesh.create(yfd.QUAD, 3)), 2)
        assert not mx.info.uses_classes and mx.info.nclasses == 0
    finally:
        if old is None:
            os.environ.pop("YFD_MX_CLASSES", None)
        else:
            os.environ["YFD_MX_CLASSES"] = old
    # 65,536 triangles k=3: the centred scheme conserves the discrete energy through the class path
    mx = yfd.Maxwell(ctx, yfd.Mesh(ctx, yfd.HostMesh.create(yfd.TRI, 7)), 3, delta_t=2e-4, upwind=False)
    assert mx.info.class_reject_reason == 0 and 1 <= mx.info.nclasses <= 64
    mx.set_operator_mode(yfd.OPERATOR_CLASSES)
    mx.apply_initial_condition()
    mx.timesteps(20)
    e1 = mx.report_errors(19 * 2e-4)
    mx.set_operator_mode(yfd.OPERATOR_STREAMED)
    mx.apply_initial_condition()
    mx.timesteps(20)
    e2 = mx.report_errors(19 * 2e-4)
    assert np.abs(e1 - e2).max() <= 1e-11


@pytest.mark.parametrize("t,r,k,up", [(0, 3, 4, False), (0, 3, 2, True), (1, 3, 3, False), (0, 3, 3, True)])
def test_split_rk_equals_fused_rk(yfd, ctx, t, r, k, up):
    """The RK4 step of the tensor-pipe class kernel in its two forms -- vector work fused into the stage kernel's sinks, or plain
    applications + maxwell_rk_vector streaming kernels (YFD_MX_SPLIT_RK) -- performs the same operations in the same order:
    identical bits after several steps, odd vector lengths included."""
    mesh = yfd.Mesh(ctx, yfd.HostMesh.create(t, r))
    res = {}
    for mode in ("0", "2"):
        os.environ["YFD_MX_SPLIT_RK"] = mode
        try:
            mx = yfd.Maxwell(ctx, mesh, k, delta_t=0.002, upwind=up)
            mx.set_operator_mode(yfd.OPERATOR_CLASSES)
            assert mx.info.class_tensor_pipe
            mx.set_dofs(np.random.default_rng(3).standard_normal(mx.ndofs))
            mx.timesteps(5)
            res[mode] = mx.dofs()
            mx.close()
        finally:
            os.environ.pop("YFD_MX_SPLIT_RK", None)
    assert np.array_equal(res["0"], res["2"])
    mesh.close()

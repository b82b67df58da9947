"""GPU: yfd_system_true_residual_noalloc -- the true residual from a call that allocates no device memory (static device
scratch, include/yfd.h).  No counterpart in the reference.

Checked through the C ABI against the existing entry point (yfd_system_true_residual) on the same system and the same
device vector: same arithmetic in the same order, so the two numbers are compared at 1e-12 relative (they are in fact the
same bits), on two calls in a row with different operands -- the second call finds the first one's contents in the
shared scratch.  Collected last on purpose.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _device_vector(torch, a):
    t = torch.from_numpy(np.ascontiguousarray(a, np.float64)).to("cuda:0")
    torch.cuda.synchronize()
    return t


def _systems(yfd, ctx):
    # DG-SIP with the congruence-class cache in use, the same system on the streamed per-cell values, a DG mesh with ragged
    # block rows on triangles, and a continuous-FEM (plain CSR) system
    mq = yfd.Mesh(ctx, yfd.HostMesh.create(yfd.QUAD, 3))
    mt = yfd.Mesh(ctx, yfd.HostMesh.create(yfd.TRI, 3))
    s1 = ctx.dg_assemble_sip(mq, 3)
    s2 = ctx.dg_assemble_sip(mq, 2)
    s2.set_operator_mode(yfd.OPERATOR_STREAMED)
    s3 = ctx.dg_assemble_sip(mt, 1)
    s4 = ctx.cfem_assemble_p1(mt)
    return [("dg quad k=3 classes", s1), ("dg quad k=2 streamed", s2), ("dg tri k=1", s3), ("cfem p1", s4)], [mq, mt]


def test_no_alloc_true_residual_matches_the_allocating_entry_point(yfd, ctx):
    import torch
    systems, meshes = _systems(yfd, ctx)
    assert systems[0][1].operator_info()["uses_classes"] and not systems[1][1].operator_info()["uses_classes"]
    rng = np.random.default_rng(11)
    for name, sysm in systems:
        assert sysm.n <= yfd.NOALLOC_MAX_ROWS
        sol = sysm.solve_cg(yfd.CgParams(rr_tol=1e-10, rr_max=1e4, max_iter=4 * sysm.n))
        # two calls in a row: a converged solution (small residual: cancellation in b - A x), then a random vector
        # (residual O(1)); then the first again, which must not have been disturbed by what the second left behind
        operands = [sol["x"], rng.standard_normal(sysm.n), sol["x"]]
        got, want = [], []
        for x in operands:
            t = _device_vector(torch, x)
            want.append(sysm.true_residual(d_x=t.data_ptr()))
            got.append(sysm.true_residual_noalloc(t.data_ptr()))
        for g, w in zip(got, want):
            assert np.isfinite(g) and abs(g - w) <= 1e-12 * max(abs(w), 1e-300), (name, got, want)
        assert got[0] == got[2], (name, got)
        assert got[1] > 1e3 * got[0], (name, got)             # the two operands really are different problems
        # and against the host route of the existing entry point
        assert abs(got[0] - sysm.true_residual(x=sol["x"])) <= 1e-12 * got[0], name


def test_no_alloc_call_leaves_device_memory_alone(yfd, ctx):
    """cudaMemGetInfo before and after: the call neither takes nor returns device memory (after a first call, which may
    load the module lazily)."""
    import torch
    mesh = yfd.Mesh(ctx, yfd.HostMesh.create(yfd.QUAD, 3))
    sysm = ctx.dg_assemble_sip(mesh, 3)
    t = _device_vector(torch, np.ones(sysm.n))
    first = sysm.true_residual_noalloc(t.data_ptr())
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    launches0 = ctx.launches
    again = sysm.true_residual_noalloc(t.data_ptr())
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(0)[0] == free0
    assert again == first and ctx.launches - launches0 == 3    # rows, partial sums, totals
    sysm.close(); mesh.close()


def test_no_alloc_rejects_what_the_scratch_cannot_hold(yfd, ctx):
    import torch
    # 232 x 232 quadrangles at k=3: 538 240 rows > YFD_NOALLOC_MAX_ROWS = 524 288
    mesh = yfd.Mesh(ctx, yfd.HostMesh.structured(yfd.QUAD, 232))
    sysm = ctx.dg_assemble_sip(mesh, 3)
    assert sysm.n > yfd.NOALLOC_MAX_ROWS
    t = _device_vector(torch, np.zeros(sysm.n))
    with pytest.raises(yfd.YfdError) as e:
        sysm.true_residual_noalloc(t.data_ptr())
    assert "YFD_NOALLOC_MAX_ROWS" in str(e.value) and "524288" in str(e.value)
    assert np.isfinite(sysm.true_residual(d_x=t.data_ptr()))       # the allocating entry point takes it
    with pytest.raises(yfd.YfdError):
        sysm.true_residual_noalloc(0)                                # null operand
    sysm.close(); mesh.close()

"""GPU: device post-processing of a DG solution (yfd_dg_postprocess) against the unmodified reference's
status.L2_errsq_qp / L2_errsq_mm (dg2d_diffusion.cpp:225-268), and the convergence orders it exists to measure."""
import numpy as np
import pytest

from conftest import host_mesh_from_ref

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mtype,r,k,shatter", [(0, 3, 1, 0), (0, 2, 2, 1), (1, 2, 3, 0), (0, 1, 4, 1), (1, 2, 2, 1), (0, 3, 3, 0), (1, 1, 5, 0)])
def test_postprocess_matches_reference(yfd, ref, ctx, mtype, r, k, shatter):
    m = ref.mesh(mtype, r, shatter=bool(shatter))
    ref.set_cg(rr_tol=1e-12, rr_max=1e4, max_iter=-1)
    out = ref.run_dg(k, eta=1.0)
    mesh = yfd.Mesh(ctx, host_mesh_from_ref(yfd, m))
    qp, mm = ctx.dg_postprocess(mesh, k, out["sol"])         # the reference's own solution vector in, its two norms out
    assert abs(np.sqrt(qp) - out["l2_qp"]) <= 1e-11 * max(out["l2_qp"], 1e-300)
    # the mm measure goes through M^-1 (LAPACK there, in-warp LU here): cond(M) * eps
    assert abs(np.sqrt(mm) - out["l2_mm"]) <= 1e-9 * max(out["l2_mm"], 1e-300)


def test_postprocess_device_pointer_and_orders(yfd, ctx):
    """assemble -> CG -> post-process with the solution never leaving the device; L2 error order k+1 on -r 4,5,6."""
    import torch
    # tri k = 1 with the driver's default eta = 1 is under-penalised: the reference itself shows orders 1.48, 1.28 there
    for mtype, k, eta in ((0, 1, 10.0), (1, 2, 1.0)):
        errs = []
        for r in (4, 5, 6):
            mesh = yfd.Mesh(ctx, yfd.HostMesh.create(mtype, r))
            sysm = ctx.dg_assemble_sip(mesh, k, eta=eta)
            d_x = torch.empty(sysm.n, dtype=torch.float64, device="cuda")
            sysm.solve_cg_device(yfd.CgParams(rr_tol=1e-12, rr_max=1e4, max_iter=2 * sysm.n), d_x=d_x.data_ptr())
            ctx.synchronize()
            qp, mm = ctx.dg_postprocess(mesh, k, None, device_ptr=d_x.data_ptr())
            qp2, mm2 = ctx.dg_postprocess(mesh, k, d_x.cpu().numpy())
            assert qp == qp2 and mm == mm2                    # fixed-order reduction: bit-reproducible
            errs.append(np.sqrt([qp, mm]))
        errs = np.array(errs)
        orders = np.log2(errs[:-1] / errs[1:])
        assert (orders > k + 1 - 0.35).all(), orders


def test_postprocess_exact_for_polynomials(yfd, ctx):
    """u_h == reference function when it lies in the space: constant 1 with k = 0 and k = 2 (first dof, rest 0)."""
    mesh = yfd.Mesh(ctx, yfd.HostMesh.create(1, 3))
    for k in (0, 2):
        bs = (k + 1) * (k + 2) // 2
        sol = np.zeros(mesh.ncells * bs)
        sol[::bs] = 1.0
        qp, mm = ctx.dg_postprocess(mesh, k, sol, ref_fn=yfd.FN_ONE)
        assert qp <= 1e-28 and mm <= 1e-26

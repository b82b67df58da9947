"""Maxwell stage kernels on one operator: streamed / class (CUDA cores) / class on the tensor pipe, application alone and RK4 steps."""
import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np
import yaourt_fem_dg_b200 as yfd
ctx = yfd.Context(0)
cases = [(yfd.TRI, 9, 4), (yfd.QUAD, 9, 3), (yfd.TRI, 9, 2), (yfd.TRI, 8, 3)] if len(sys.argv) < 2 else [(int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]))]
for ct, r, k in cases:
    hm = yfd.HostMesh.create(ct, r)
    mesh = yfd.Mesh(ctx, hm)
    out = []
    ref = None
    for name, env in (("streamed", None), ("class dfma", dict(YFD_MX_CLASS_MMA="0")), ("class mma", dict(YFD_MX_CLASS_MMA="1"))):
        os.environ["YFD_MX_CLASSES"] = "2"
        for kk in ("YFD_MX_CLASS_MMA", "YFD_MX_MMA_TILES"):
            os.environ.pop(kk, None)
        if env:
            os.environ.update(env)
        mx = yfd.Maxwell(ctx, mesh, k, delta_t=1e-4)
        mx.set_operator_mode(yfd.OPERATOR_STREAMED if env is None else yfd.OPERATOR_CLASSES)
        mx.apply_initial_condition()
        ms_apply = mx.time_apply(20)[0]
        mx.apply_initial_condition()
        st = mx.timesteps(20)
        u = np.empty(mx.ndofs)
        ctx.check(yfd.lib().yfd_maxwell_get_dofs(mx._h, u))
        if ref is None:
            ref = u
        out.append("%s: apply %.3f ms, RK4 step %.3f ms, dofs vs streamed %.1e" % (name, ms_apply, st.ms_total / 20, np.abs(u - ref).max() / np.abs(ref).max()))
        mx.close()
    print("ct %d r %d k %d | " % (ct, r, k) + " | ".join(out), flush=True)
    mesh.close()

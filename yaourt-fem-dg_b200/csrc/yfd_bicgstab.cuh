/*
 * yfd_bicgstab.cuh -- BiCGStab on the device; replaces yaourt::bicgstab, both overloads
 * (core/solvers.hpp:246-354 plain, :356-467 with iM).  Included by yfd_cg.cu after launch_spmv.
 *
 * The reference's loop, statement by statement (:311-346):
 *     rho_old = rho; rho = r.r0; if |rho| < 1e-9 { r = b - A x; r0 = r; rho = r.r0 }        (ABSOLUTE threshold)
 *     beta = (rho/rho_old)(alpha/omega); p = r + beta (p - omega v); [y = iM p]; v = A p|y;
 *     alpha = rho / v.r0; s = r - alpha v; [z = iM s]; t = A s|z;
 *     omega = t.s / t.t  [iMt = iM t: iMt.z / iMt.iMt];
 *     x = x + alpha p|y + omega s|z; r = s - omega t; nr = |r|; iter++
 * The restart threshold is absolute, and right-hand sides of FEM systems are small (they scale with the cell area), so
 * the restart is the COMMON path once the residual has dropped a few orders: it must not leave the device.  The
 * restart SpMV y = A x is therefore part of every iteration's launch sequence and gated on the device: it is issued
 * through the dot-fused SpMV kernels, which return at once when the `done` field of the cg_scalars they are given is
 * set -- a second cg_scalars object (`gate`) whose `done` means "no restart this iteration" turns the kernel into a
 * no-op without touching the SpMV code.  Everything else is as in the CG: scalars stay on the device, `batch`
 * iterations are one CUDA graph, the host polls one flag per replay.
 *
 * iM is applied as a diagonal (the only preconditioner the reference's drivers build, dg2d_diffusion.cpp:200-212).
 * Not supported: use_normal_eqns in the overload without iM (:289,:318-337; the overload with iM ignores the flag, as here),
 * distributed systems.
 */
#pragma once

struct bicg_scalars {
    double rho, rho_old, alpha, omega;
    double vr0;                 /* v . r0                                          */
    double ts, tt;              /* t.s, t.t  (with iM: iMt.z, iMt.iMt)             */
    double rho_next;            /* r . r0 of the residual the next iteration starts from */
    double nr0, nr;
    long long iter, max_iter;
    double rr_tol, rr_max;
    int    done, restart;
    unsigned int ticket[4];
};

/* what the last block of bicg_init / bicg_update_xr does for the NEXT iteration: loop test (:311), rho_old / rho
 * (:324-326), restart decision (:327), and the two device-side gates */
__device__ __forceinline__ void bicg_next_iteration(bicg_scalars *bs, cg_scalars *g_main, cg_scalars *g_restart)
{
    const double rel = bs->nr / bs->nr0;
    const int done = !(rel > bs->rr_tol && bs->iter < bs->max_iter && rel < bs->rr_max);
    bs->rho_old = bs->rho;
    bs->rho = bs->rho_next;
    const int restart = !done && fabs(bs->rho) < 1e-9;
    bs->restart = restart;
    bs->done = done;
    g_main->done = done;
    g_restart->done = !restart;
}

/* r = b - y (y = A x0 or null); r0 = r; p = v = 0 (:288-295) */
__global__ void __launch_bounds__(CG_THREADS)
bicg_init(const double *__restrict__ b, const double *__restrict__ y, double *__restrict__ r, double *__restrict__ r0,
          double *__restrict__ p, double *__restrict__ v, int64_t n, bicg_scalars *__restrict__ bs,
          cg_scalars *__restrict__ g_main, cg_scalars *__restrict__ g_restart, double *__restrict__ partials,
          double *__restrict__ history, int64_t history_cap)
{
    double rr = 0.0;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const double ri = y ? b[i] - y[i] : b[i];
        r[i] = ri; r0[i] = ri; p[i] = 0.0; v[i] = 0.0;
        rr += ri * ri;
    }
    __shared__ double tot[1];
    const double s[1] = { rr };
    if (grid_sum<1>(s, partials, &bs->ticket[0], tot) && threadIdx.x == 0) {
        const double nr = sqrt(tot[0]);
        bs->nr0 = nr; bs->nr = nr;
        bs->rho = 1.0; bs->alpha = 1.0; bs->omega = 1.0;            /* :286 */
        bs->rho_next = tot[0];                                       /* r . r0 with r0 = r */
        bs->iter = 0;
        if (history && history_cap > 0) history[0] = nr / nr;
        bicg_next_iteration(bs, g_main, g_restart);
    }
}

/* gated: r = b - A x; r0 = r; rho = r . r0 (:327-335) */
__global__ void __launch_bounds__(CG_THREADS)
bicg_restart(const double *__restrict__ b, const double *__restrict__ y, double *__restrict__ r, double *__restrict__ r0,
             int64_t n, bicg_scalars *__restrict__ bs, double *__restrict__ partials)
{
    if (bs->done || !bs->restart) return;
    double rr = 0.0;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const double ri = b[i] - y[i];
        r[i] = ri; r0[i] = ri;
        rr += ri * ri;
    }
    __shared__ double tot[1];
    const double s[1] = { rr };
    if (grid_sum<1>(s, partials, &bs->ticket[1], tot) && threadIdx.x == 0) bs->rho = tot[0];
}

/* p = r + beta (p - omega v); yv = iM p (:337-339 / :437-439) */
__global__ void __launch_bounds__(CG_THREADS)
bicg_update_p(double *__restrict__ p, const double *__restrict__ r, const double *__restrict__ v, const double *__restrict__ pc,
              double *__restrict__ yv, int64_t n, const bicg_scalars *__restrict__ bs)
{
    if (bs->done) return;
    const double omega = bs->omega;
    const double beta = (bs->rho / bs->rho_old) * (bs->alpha / omega);
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const double pi = r[i] + beta * (p[i] - omega * v[i]);
        p[i] = pi;
        if (pc) yv[i] = pc[i] * pi;
    }
}

/* vr0 = v . r0 (:346 / :441) */
__global__ void __launch_bounds__(CG_THREADS)
bicg_dot_vr0(const double *__restrict__ v, const double *__restrict__ r0, int64_t n, bicg_scalars *__restrict__ bs,
             double *__restrict__ partials)
{
    if (bs->done) return;
    double acc = 0.0;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) acc += v[i] * r0[i];
    __shared__ double tot[1];
    const double s[1] = { acc };
    if (grid_sum<1>(s, partials, &bs->ticket[2], tot) && threadIdx.x == 0) { bs->vr0 = tot[0]; bs->alpha = bs->rho / tot[0]; }
}

/* s = r - alpha v; z = iM s (:347 / :442-443).  alpha was stored by bicg_dot_vr0's last block */
__global__ void __launch_bounds__(CG_THREADS)
bicg_update_s(double *__restrict__ s, const double *__restrict__ r, const double *__restrict__ v, const double *__restrict__ pc,
              double *__restrict__ z, int64_t n, const bicg_scalars *__restrict__ bs)
{
    if (bs->done) return;
    const double alpha = bs->alpha;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const double si = r[i] - alpha * v[i];
        s[i] = si;
        if (pc) z[i] = pc[i] * si;
    }
}

/* ts = t.s, tt = t.t; with iM: iMt = iM t, ts = iMt.z, tt = iMt.iMt (:354 / :446-447) */
__global__ void __launch_bounds__(CG_THREADS)
bicg_dots_t(const double *__restrict__ t, const double *__restrict__ sz /* s, or z with iM */, const double *__restrict__ pc,
            int64_t n, bicg_scalars *__restrict__ bs, double *__restrict__ partials)
{
    if (bs->done) return;
    double a = 0.0, c = 0.0;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const double w = pc ? pc[i] * t[i] : t[i];
        a += w * sz[i];
        c += w * w;
    }
    __shared__ double tot[2];
    const double s[2] = { a, c };
    if (grid_sum<2>(s, partials, &bs->ticket[2], tot) && threadIdx.x == 0) { bs->ts = tot[0]; bs->tt = tot[1]; bs->omega = tot[0] / tot[1]; }
}

/* x = x + alpha p|y + omega s|z; r = s - omega t; nr = |r|; iter++ (:356-360 / :449-453); r . r0 for the next pass */
__global__ void __launch_bounds__(CG_THREADS)
bicg_update_xr(double *__restrict__ x, double *__restrict__ r, const double *__restrict__ py, const double *__restrict__ sz,
               const double *__restrict__ s, const double *__restrict__ t, const double *__restrict__ r0, int64_t n,
               bicg_scalars *__restrict__ bs, cg_scalars *__restrict__ g_main, cg_scalars *__restrict__ g_restart,
               double *__restrict__ partials, double *__restrict__ history, int64_t history_cap)
{
    if (bs->done) return;
    const double alpha = bs->alpha, omega = bs->omega;
    double rr = 0.0, rr0 = 0.0;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        x[i] = x[i] + alpha * py[i] + omega * sz[i];
        const double ri = s[i] - omega * t[i];
        r[i] = ri;
        rr += ri * ri;
        rr0 += ri * r0[i];
    }
    __shared__ double tot[2];
    const double v[2] = { rr, rr0 };
    if (grid_sum<2>(v, partials, &bs->ticket[3], tot) && threadIdx.x == 0) {
        const double nr = sqrt(tot[0]);
        bs->nr = nr;
        bs->rho_next = tot[1];
        const long long it = bs->iter + 1;
        bs->iter = it;
        if (history && it < history_cap) history[it] = nr / bs->nr0;
        bicg_next_iteration(bs, g_main, g_restart);
    }
}

static yfd_status bicgstab_device(yfd_ctx *ctx, const spmv_op &op, const double *d_b, double *d_x, const double *d_pc,
                                  const yfd_cg_params *p, yfd_cg_info *info, double *host_history, int64_t history_cap)
{
    const int64_t n = op.nrows;
    /* the reference's overload WITH iM never looks at use_normal_eqns (core/solvers.hpp:356-467 has no such branch): with a
     * preconditioner the flag is ignored, as there; the overload without iM does honour it (:289,:318-337) -- not built */
    if (p->use_normal_eqns && !d_pc) return fail(ctx, YFD_ERR_UNSUPPORTED, "bicgstab without iM on the normal equations (core/solvers.hpp:289,318-337)");
    if (op.plan) return fail(ctx, YFD_ERR_UNSUPPORTED, "bicgstab on a distributed system");
    dbuf<double> r, r0, pv, v, s, t, y, yv, z, partials, hist;
    dbuf<bicg_scalars> bs;
    dbuf<cg_scalars> gates;                 /* [0]: done mirror for the main SpMVs, [1]: restart gate */
    for (dbuf<double> *b : { &r, &r0, &pv, &v, &s, &t, &y }) YFD_TRY(b->alloc(ctx, (size_t) n, 2));
    if (d_pc) { YFD_TRY(yv.alloc(ctx, (size_t) n, 2)); YFD_TRY(z.alloc(ctx, (size_t) n, 2)); }
    YFD_TRY(partials.alloc(ctx, (size_t) 2 * CG_MAX_PARTIALS));
    YFD_TRY(bs.alloc(ctx, 1)); YFD_TRY(gates.alloc(ctx, 2)); YFD_TRY(gates.zero(ctx));
    int64_t hcap = 0;
    if (p->save_iteration_history && host_history && history_cap > 0) {
        hcap = std::min<int64_t>(history_cap, p->max_iter + 1);
        YFD_TRY(hist.alloc(ctx, (size_t) hcap));
    }
    bicg_scalars h0 = {};
    h0.max_iter = p->max_iter; h0.rr_tol = p->rr_tol; h0.rr_max = p->rr_max;
    YFD_CUDA(ctx, cudaMemcpyAsync(bs.p, &h0, sizeof(h0), cudaMemcpyHostToDevice, ctx->stream));
    cg_scalars *g_main = gates.p, *g_restart = gates.p + 1;
    const double *py = d_pc ? yv.p : pv.p, *sz = d_pc ? z.p : s.p;

    gpu_timer timer;
    timer.start(ctx->stream);
    const unsigned vg = vec_grid(ctx, n);
    if (p->use_initial_guess) {
        YFD_TRY(launch_spmv<false>(ctx, op, d_x, y.p, nullptr, nullptr));
        YFD_LAUNCH(ctx, bicg_init, vg, CG_THREADS, 0, d_b, (const double *) y.p, r.p, r0.p, pv.p, v.p, n, bs.p, g_main, g_restart, partials.p, hist.p, hcap);
    } else {
        YFD_CUDA(ctx, cudaMemsetAsync(d_x, 0, (size_t) n * sizeof(double), ctx->stream));
        YFD_LAUNCH(ctx, bicg_init, vg, CG_THREADS, 0, d_b, (const double *) nullptr, r.p, r0.p, pv.p, v.p, n, bs.p, g_main, g_restart, partials.p, hist.p, hcap);
    }

    int batch = p->check_every > 0 ? p->check_every : 32;
    if ((int64_t) batch > p->max_iter) batch = (int) std::max<int64_t>(p->max_iter, 1);
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    const int64_t launches_before = ctx->launches;
    YFD_CUDA(ctx, cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeRelaxed));
    yfd_status cap_st = YFD_OK;
    for (int it = 0; it < batch && cap_st == YFD_OK; it++) {
        cap_st = launch_spmv<true>(ctx, op, d_x, y.p, g_restart, partials.p);            /* y = A x, only when restarting */
        if (cap_st != YFD_OK) break;
        bicg_restart<<<vg, CG_THREADS, 0, ctx->stream>>>(d_b, y.p, r.p, r0.p, n, bs.p, partials.p);
        bicg_update_p<<<vg, CG_THREADS, 0, ctx->stream>>>(pv.p, r.p, v.p, d_pc, yv.p, n, bs.p);
        cap_st = launch_spmv<true>(ctx, op, py, v.p, g_main, partials.p);                /* v = A p|y */
        if (cap_st != YFD_OK) break;
        bicg_dot_vr0<<<vg, CG_THREADS, 0, ctx->stream>>>(v.p, r0.p, n, bs.p, partials.p);
        bicg_update_s<<<vg, CG_THREADS, 0, ctx->stream>>>(s.p, r.p, v.p, d_pc, z.p, n, bs.p);
        cap_st = launch_spmv<true>(ctx, op, sz, t.p, g_main, partials.p);                /* t = A s|z */
        if (cap_st != YFD_OK) break;
        bicg_dots_t<<<vg, CG_THREADS, 0, ctx->stream>>>(t.p, sz, d_pc, n, bs.p, partials.p);
        bicg_update_xr<<<vg, CG_THREADS, 0, ctx->stream>>>(d_x, r.p, py, sz, s.p, t.p, r0.p, n, bs.p, g_main, g_restart, partials.p, hist.p, hcap);
        ctx->launches += 6;
    }
    {
        cudaError_t e = cudaStreamEndCapture(ctx->stream, &graph);
        if (cap_st != YFD_OK) { if (graph) cudaGraphDestroy(graph); return cap_st; }
        if (e != cudaSuccess) return fail(ctx, YFD_ERR_CUDA, "graph capture: %s", cudaGetErrorString(e));
    }
    const int64_t launches_per_graph = ctx->launches - launches_before;
    ctx->launches = launches_before;
    {
        cudaError_t e = cudaGraphInstantiate(&exec, graph, 0);
        if (e != cudaSuccess) { cudaGraphDestroy(graph); return fail(ctx, YFD_ERR_CUDA, "graph instantiate: %s", cudaGetErrorString(e)); }
    }
    bicg_scalars hs = {};
    yfd_status st = YFD_OK;
    for (;;) {
        cudaError_t e = cudaMemcpyAsync(&hs, bs.p, sizeof(hs), cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) { st = fail(ctx, YFD_ERR_CUDA, "BiCGStab poll: %s", cudaGetErrorString(e)); break; }
        if (hs.done) break;
        e = cudaGraphLaunch(exec, ctx->stream);
        if (e != cudaSuccess) { st = fail(ctx, YFD_ERR_CUDA, "graph launch: %s", cudaGetErrorString(e)); break; }
        ctx->launches += launches_per_graph;
    }
    timer.stop(ctx->stream);
    cudaGraphExecDestroy(exec);
    cudaGraphDestroy(graph);
    if (st != YFD_OK) return st;
    if (hcap > 0) {
        const int64_t len = std::min<int64_t>(hcap, hs.iter + 1);
        YFD_TRY(hist.download(ctx, host_history, (size_t) len));
        if (info) info->history_len = len;
    } else if (info) info->history_len = 0;
    if (info) {
        info->iterations = hs.iter;
        info->nr0 = hs.nr0;
        info->nr = hs.nr;
        info->converged = (hs.nr / hs.nr0 <= p->rr_tol) ? 1 : 0;
        info->ms_solve = timer.ms();
        info->ms_spmv = 0.0;
        info->transport = 0;
    }
    return YFD_OK;
}

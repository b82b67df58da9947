/*
 * yfd_maxwell_step.cu -- the DG Maxwell time loop: one kernel per operator application, fused with the
 * Runge-Kutta vector updates that follow it.
 *
 * Replaces apply_operator + the RK4 / Euler combinations of do_timestep (methods/dg_maxwell_2D.hpp:586-657).
 * The reference multiplies, per cell, a dense 3bs x 3bs block with the cell's unknowns and one more per face
 * with the neighbour's (HBM-bound block GEMV over (1+nf)(3bs)^2 doubles per cell).  Only 4 (centred fluxes) or
 * 9 (upwind) of the 9 sub-blocks are non-zero and every one is a scalar multiple of Gx = M^-1 Sx, Gy = M^-1 Sy,
 * T_f = M^-1 tmass_f or N_f = M^-1 nmass_f (:290-295, :356-381, :404-410), so this kernel streams exactly those
 * 2 + 2 nf blocks per cell and applies the 3x3 pattern in registers:
 *     a_f(v) = T_f v_c - N_f v_n                                   (N_f = 0, v_n = 0 on the boundary)
 *     Hx' = 1/mu  [ -Gy Ez + sum_f  ny cE a_f(Ez) - vH ( ny ny a_f(Hx) - nx ny a_f(Hy) ) ]
 *     Hy' = 1/mu  [ +Gx Ez + sum_f -nx cE a_f(Ez) + vH ( nx ny a_f(Hx) - nx nx a_f(Hy) ) ]
 *     Ez' = 1/eps [ -Gy Hx + Gx Hy + sum_f cH ( ny a_f(Hx) - nx a_f(Hy) ) - vE a_f(Ez) ]
 * with (cE, cH, vH, vE) = (kappa_E, kappa_H, nu_H, nu_E) on interior faces and (2 kappa_E, 0, 0, 2 nu_E) on
 * boundary faces (:396-410).
 *
 * Kernel shape (sm_100a): one warp per chunk of CPW = 32/bs consecutive cells, lane = (cell slot, row i).
 * The operator is laid out [chunk][block][j][row] (yfd_maxwell.cuh), so for each (block, j) the warp reads one
 * contiguous line, every entry is used by exactly one lane, and nothing is staged: the loads are fully unrolled
 * per block and marked evict-first (each byte is touched once per application).  Lane 0 opens with one
 * cp.async.bulk.prefetch.L2 of the whole chunk slab, so DRAM works on all of it while the first blocks are consumed.  The cell's own unknowns and
 * its neighbours' are gathered once (coalesced bs-wide runs, L2 hits thanks to the Z-curve order) into a small
 * per-warp shared-memory table [group][field][j][slot] from which every lane reads its cell's v[j] as a
 * conflict-free broadcast.  HBM-bound fp64 FMA work, 0.7 flop/B: no tensor cores.
 *
 * RK4 without extra passes (the reference forms k1..k4 and four temporaries, :636-657): stage s writes the
 * operand of stage s+1 and keeps  k1  and  k2  so that the last stage can form the reference's expression
 * u + ((1/6) dt) (k1 + 2 (k2 + k3) + k4)  in its own order.  Per step: 4 launches.
 */
#include "yfd_maxwell.cuh"
#include "yfd_class_core.cuh"
#include "yfd_mma_core.cuh"
#include <cstdlib>
#include <algorithm>

namespace yfd {

constexpr int MX_WARPS = 8;

__device__ __forceinline__ double ld_stream(const double *p)
{
    return __ldcs(p);          /* ld.global.cs: streamed once, evict first */
}

/* one lane asks the L2 to start fetching `bytes` (multiple of 16) at p: no registers, no shared memory */
__device__ __forceinline__ void l2_prefetch_bulk(const void *p, uint32_t bytes)
{
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}

template <int NV, int K, int MINB>
__global__ void __launch_bounds__(MX_WARPS * 32, MINB)
maxwell_stage(const mx_step_args a, const int mode, const int prefetch)
{
    constexpr int BS = mx_bs(K), CPW = mx_cpw(BS), RWP = mx_rwp(BS), NB = mx_nb(NV), L = 3 * BS;
    constexpr int GRP = 3 * BS * CPW;                 /* one group of the v table: [field][j][slot] */
    extern __shared__ double sm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double *VS = sm + (size_t) warp * ((1 + NV) * GRP);
    const int64_t chunk = (int64_t) blockIdx.x * MX_WARPS + warp;
    if (chunk >= a.nchunks) return;                   /* warps are independent: no CTA-wide barrier below */

    const int slot = lane / BS, i = lane - slot * BS;
    const bool inwarp = slot < CPW;
    const int64_t cell = chunk * CPW + slot;
    const bool valid = inwarp && cell < a.ncells;
    const int r = inwarp ? lane : 0;                  /* lanes past the chunk's rows read row 0 and store nothing */
    const int sl = inwarp ? slot : 0;
    if (prefetch && lane == 0)                        /* the whole chunk's operator slab: DRAM starts on all of it now */
        l2_prefetch_bulk(a.op + (size_t) chunk * (NB * BS * RWP), (uint32_t) (NB * BS * RWP * sizeof(double)));

    /* ---- gather: own unknowns and the neighbours' ---- */
    double own[3] = {0.0, 0.0, 0.0};
    if (valid) {
        const double *v = a.vin + cell * L + i;
        own[0] = __ldg(v); own[1] = __ldg(v + BS); own[2] = __ldg(v + 2 * BS);
    }
    if (inwarp) {
#pragma unroll
        for (int fld = 0; fld < 3; fld++) VS[(fld * BS + i) * CPW + slot] = own[fld];
    }
#pragma unroll
    for (int f = 0; f < NV; f++) {
        const int32_t nb = valid ? __ldg(&a.nbr[cell * NV + f]) : -1;
        double w0 = 0.0, w1 = 0.0, w2 = 0.0;
        if (nb >= 0) {
            const double *v = a.vin + (int64_t) nb * L + i;
            w0 = __ldg(v); w1 = __ldg(v + BS); w2 = __ldg(v + 2 * BS);
        }
        if (inwarp) {
            double *G = VS + (1 + f) * GRP;
            G[(0 * BS + i) * CPW + slot] = w0; G[(1 * BS + i) * CPW + slot] = w1; G[(2 * BS + i) * CPW + slot] = w2;
        }
    }
    __syncwarp();

    /* ---- cell blocks: Gx (b = 0), Gy (b = 1) ---- */
    const double *opc = a.op + (size_t) chunk * (NB * BS * RWP) + r;
    const double *V0 = VS + sl;                       /* own group, this lane's slot */
    double gxe = 0.0, gxh = 0.0, gye = 0.0, gyh = 0.0;
    {
        double wx[BS], wy[BS];
#pragma unroll
        for (int j = 0; j < BS; j++) { wx[j] = ld_stream(opc + j * RWP); wy[j] = ld_stream(opc + (BS + j) * RWP); }
#pragma unroll
        for (int j = 0; j < BS; j++) {
            const double hx = V0[(0 * BS + j) * CPW], hy = V0[(1 * BS + j) * CPW], ez = V0[(2 * BS + j) * CPW];
            gxe += wx[j] * ez; gxh += wx[j] * hy;
            gye += wy[j] * ez; gyh += wy[j] * hx;
        }
    }
    double Hx = -gye, Hy = gxe, Ez = gxh - gyh;

    /* ---- faces: a_f = T_f v_c - N_f v_n ---- */
#pragma unroll
    for (int f = 0; f < NV; f++) {
        const double *Tf = opc + (size_t) (2 + f) * BS * RWP, *Nf = opc + (size_t) (2 + NV + f) * BS * RWP;
        const double *VN = VS + (1 + f) * GRP + sl;
        double wt[BS], wn[BS];
#pragma unroll
        for (int j = 0; j < BS; j++) { wt[j] = ld_stream(Tf + j * RWP); wn[j] = ld_stream(Nf + j * RWP); }
        double p = 0.0, q = 0.0, s = 0.0;
#pragma unroll
        for (int j = 0; j < BS; j++) {
            p += wt[j] * V0[(0 * BS + j) * CPW]; q += wt[j] * V0[(1 * BS + j) * CPW]; s += wt[j] * V0[(2 * BS + j) * CPW];
            p -= wn[j] * VN[(0 * BS + j) * CPW]; q -= wn[j] * VN[(1 * BS + j) * CPW]; s -= wn[j] * VN[(2 * BS + j) * CPW];
        }
        mx_face_coef c = {0, 0, 0, 0, 0, 0};
        if (valid) c = a.fc[cell * NV + f];
        Hx += c.ny * c.cE * s - c.vH * (c.ny * c.ny * p - c.nx * c.ny * q);
        Hy += -c.nx * c.cE * s + c.vH * (c.nx * c.ny * p - c.nx * c.nx * q);
        Ez += c.cH * (c.ny * p - c.nx * q) - c.vE * s;
    }
    if (!valid) return;
    const double inv_mu = __ldg(&a.cs[2 * cell]), inv_eps = __ldg(&a.cs[2 * cell + 1]);
    const double kv[3] = {inv_mu * Hx, inv_mu * Hy, inv_eps * Ez};

    /* ---- sink: plain application or one Runge-Kutta stage (methods/dg_maxwell_2D.hpp:630-657), no FMA
     * contraction so the combinations round as the reference's separate multiply and add ---- */
    const int64_t base = cell * L + i;
#pragma unroll
    for (int fld = 0; fld < 3; fld++) {
        const int64_t idx = base + fld * BS;
        const double k = kv[fld];
        switch (mode) {
        case MX_APPLY: a.out[idx] = k; break;
        case MX_RK1:   a.acc[idx] = k; a.out[idx] = __dadd_rn(own[fld], __dmul_rn(a.dt / 2., k)); break;
        case MX_RK2:   a.kb[idx] = k;  a.out[idx] = __dadd_rn(a.u[idx], __dmul_rn(a.dt / 2., k)); break;
        case MX_RK3:   a.acc[idx] = __dadd_rn(a.acc[idx], __dmul_rn(__dadd_rn(a.kb[idx], k), 2.0));
                       a.out[idx] = __dadd_rn(a.u[idx], __dmul_rn(a.dt, k)); break;
        case MX_RK4:   a.u[idx] = __dadd_rn(a.u[idx], __dmul_rn(__dadd_rn(a.acc[idx], k), (1. / 6.) * a.dt)); break;
        case MX_EULER: a.out[idx] = __dadd_rn(own[fld], __dmul_rn(a.dt, k)); break;
        }
    }
}

/* ---- the same operator application from ONE block set per congruence class (BASELINE configs[3]: "matrix-free
 * element+face operator") ------------------------------------------------------------------------------------------
 * lane <-> cell over the class-sorted cell list, 32 cells of one class per warp; the class's blocks are read as
 * warp-uniform loads (one request feeds every lane), the per-cell state is the vectors, nbr, fc and cs.  The linear
 * maps are applied AFTER combining the fields, one output field at a time (15 accumulators + 15 operands in registers):
 *     Hx' = 1/mu  [ Gy (-Ez) + sum_f T_f w_c - N_f w_n ],  w = ny cE Ez - vH ny ny Hx + vH nx ny Hy
 *     Hy' = 1/mu  [ Gx ( Ez) + sum_f T_f w_c - N_f w_n ],  w = -nx cE Ez + vH nx ny Hx - vH nx nx Hy
 *     Ez' = 1/eps [ Gx Hy + Gy (-Hx) + sum_f T_f w_c - N_f w_n ],  w = cH ny Hx - cH nx Hy - vE Ez
 * which is the streamed kernel's formula with the scalar factors moved inside the block products (22 products of bs x bs
 * per triangle either way).  Same sinks (plain application / RK stages) as maxwell_stage. */
template <int BS>
__device__ __forceinline__ void mx_combo(const double *__restrict__ v /* [3][BS] of one cell */, double ahx, double ahy, double aez,
                                         double sign, double (&w)[BS])
{
#pragma unroll
    for (int i = 0; i < BS; i++) w[i] = 0.0;
    double t[BS];
    if (ahx != 0.0) { load_vec<BS>(v, t);
#pragma unroll
        for (int i = 0; i < BS; i++) w[i] = fma(sign * ahx, t[i], w[i]); }
    if (ahy != 0.0) { load_vec<BS>(v + BS, t);
#pragma unroll
        for (int i = 0; i < BS; i++) w[i] = fma(sign * ahy, t[i], w[i]); }
    if (aez != 0.0) { load_vec<BS>(v + 2 * BS, t);
#pragma unroll
        for (int i = 0; i < BS; i++) w[i] = fma(sign * aez, t[i], w[i]); }
}

template <int NV, int K, int C>
__global__ void __launch_bounds__(CLS_THREADS, C == 1 ? 2 : 1)
maxwell_stage_class(const mx_step_args a, const int mode)
{
    constexpr int BS = mx_bs(K), NB = mx_nb(NV), SS = cls_slot_stride(BS), L = 3 * BS;
    const int lane = threadIdx.x & 31;
    const int64_t gw = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
    /* C cells per lane: every class-block entry fetched is used C times (yfd_class_core.cuh) */
    const int64_t nunits = a.ntiles * (CLS_TILE / (32 * C));
    for (int64_t u = gw; u < nunits; u += nw) {
        const int k = __ldg(&a.tile_cls[(u * (32 * C)) / CLS_TILE]);
        int32_t c[C];
#pragma unroll
        for (int q = 0; q < C; q++) c[q] = __ldg(&a.work_cell[u * (32 * C) + lane + 32 * q]);
        if (c[0] < 0) continue;
        bool live[C];
#pragma unroll
        for (int q = 0; q < C; q++) { live[q] = c[q] >= 0; if (!live[q]) c[q] = c[0]; }      /* padding: cell 0's operands, no sink */
        const double *Tk = a.table + (size_t) k * (NB * SS);
#pragma unroll 1
        for (int fld = 0; fld < 3; fld++) {
            double acc[C][BS], w[C][BS];
#pragma unroll
            for (int q = 0; q < C; q++)
#pragma unroll
                for (int i = 0; i < BS; i++) acc[q][i] = 0.0;
            /* cell blocks (Gx = block 0, Gy = block 1): Hx: Gy (-Ez) ; Hy: Gx Ez ; Ez: Gx Hy + Gy (-Hx) */
#pragma unroll
            for (int q = 0; q < C; q++)
                mx_combo<BS>(a.vin + (int64_t) c[q] * L, 0.0, fld == 2 ? 1.0 : 0.0, fld == 2 ? 0.0 : 1.0, fld == 0 ? -1.0 : 1.0, w[q]);
            block_times_n<BS, C, true>(Tk + (fld == 0 ? SS : 0), w, acc);
            if (fld == 2) {
#pragma unroll
                for (int q = 0; q < C; q++) mx_combo<BS>(a.vin + (int64_t) c[q] * L, 1.0, 0.0, 0.0, -1.0, w[q]);
                block_times_n<BS, C, true>(Tk + SS, w, acc);
            }
#pragma unroll 1
            for (int f = 0; f < NV; f++) {
                double ahx[C], ahy[C], aez[C];
#pragma unroll
                for (int q = 0; q < C; q++) {
                    const mx_face_coef fq = a.fc[(int64_t) c[q] * NV + f];
                    if (fld == 0)      { aez[q] = fq.ny * fq.cE;  ahx[q] = -fq.vH * fq.ny * fq.ny; ahy[q] = fq.vH * fq.nx * fq.ny; }
                    else if (fld == 1) { aez[q] = -fq.nx * fq.cE; ahx[q] = fq.vH * fq.nx * fq.ny;  ahy[q] = -fq.vH * fq.nx * fq.nx; }
                    else               { ahx[q] = fq.cH * fq.ny;  ahy[q] = -fq.cH * fq.nx;         aez[q] = -fq.vE; }
                    mx_combo<BS>(a.vin + (int64_t) c[q] * L, ahx[q], ahy[q], aez[q], 1.0, w[q]);
                }
                block_times_n<BS, C, true>(Tk + (size_t) (2 + f) * SS, w, acc);
                /* whether face f has a neighbour is part of the class key: the same answer for every cell of the tile */
                if (__ldg(&a.nbr[(int64_t) c[0] * NV + f]) >= 0) {
#pragma unroll
                    for (int q = 0; q < C; q++) {
                        const int32_t nb = __ldg(&a.nbr[(int64_t) c[q] * NV + f]);
                        mx_combo<BS>(a.vin + (int64_t) nb * L, ahx[q], ahy[q], aez[q], -1.0, w[q]);
                    }
                    block_times_n<BS, C, true>(Tk + (size_t) (2 + NV + f) * SS, w, acc);
                }
            }
#pragma unroll
            for (int q = 0; q < C; q++) {
                if (!live[q]) continue;
                const double sc = fld == 2 ? __ldg(&a.cs[2 * c[q] + 1]) : __ldg(&a.cs[2 * c[q]]);
                const int64_t base = (int64_t) c[q] * L + fld * BS;
#pragma unroll
                for (int i = 0; i < BS; i++) {
                    const int64_t idx = base + i;
                    const double kk = sc * acc[q][i];
                    switch (mode) {
                    case MX_APPLY: a.out[idx] = kk; break;
                    case MX_RK1:   a.acc[idx] = kk; a.out[idx] = __dadd_rn(__ldg(&a.vin[idx]), __dmul_rn(a.dt / 2., kk)); break;
                    case MX_RK2:   a.kb[idx] = kk;  a.out[idx] = __dadd_rn(a.u[idx], __dmul_rn(a.dt / 2., kk)); break;
                    case MX_RK3:   a.acc[idx] = __dadd_rn(a.acc[idx], __dmul_rn(__dadd_rn(a.kb[idx], kk), 2.0));
                                   a.out[idx] = __dadd_rn(a.u[idx], __dmul_rn(a.dt, kk)); break;
                    case MX_RK4:   a.u[idx] = __dadd_rn(a.u[idx], __dmul_rn(__dadd_rn(a.acc[idx], kk), (1. / 6.) * a.dt)); break;
                    case MX_EULER: a.out[idx] = __dadd_rn(__ldg(&a.vin[idx]), __dmul_rn(a.dt, kk)); break;
                    }
                }
            }
        }
    }
}

/* ---- class build: table from the representatives' blocks, verification of every cell against it ---- */
__global__ void mx_cls_build_table(const int64_t *__restrict__ start, const uint32_t *__restrict__ sorted_cell, const double *__restrict__ op,
                                   int bs, int nb, int cpw, int rwp, double *__restrict__ table, double *__restrict__ scale)
{
    __shared__ double sm[32];
    const int k = blockIdx.x, bs2 = bs * bs, ss = cls_slot_stride(bs);
    const int64_t c = sorted_cell[start[k]];
    const int64_t chunk = c / cpw;
    const int slot = (int) (c - chunk * cpw);
    const double *src = op + (size_t) chunk * ((size_t) nb * bs * rwp) + slot * bs;
    double mx = 0.0;
    for (int t = threadIdx.x; t < nb * ss; t += blockDim.x) {
        const int b = t / ss, m = t - b * ss, j = m / bs, i = m - j * bs;
        const double v = m < bs2 ? src[(size_t) (b * bs + j) * rwp + i] : 0.0;
        table[(size_t) k * nb * ss + t] = v;
        mx = fmax(mx, fabs(v));
    }
    for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int) (blockDim.x >> 5); w++) mx = fmax(mx, sm[w]);
        scale[k] = mx;
    }
}

__global__ void mx_cls_verify(const int32_t *__restrict__ cell_cls, const double *__restrict__ op, const double *__restrict__ table,
                              const double *__restrict__ scale, int64_t ncells, int bs, int nb, int cpw, int rwp,
                              unsigned long long *__restrict__ worst_bits)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((int64_t) gridDim.x * blockDim.x) >> 5;
    const int bs2 = bs * bs, ss = cls_slot_stride(bs);
    double worst = 0.0;
    for (int64_t c = warp; c < ncells; c += nwarps) {
        const int k = cell_cls[c];
        const int64_t chunk = c / cpw;
        const int slot = (int) (c - chunk * cpw);
        const double *src = op + (size_t) chunk * ((size_t) nb * bs * rwp) + slot * bs;
        const double *T = table + (size_t) k * nb * ss;
        double d = 0.0;
        for (int t = lane; t < nb * bs2; t += 32) {
            const int b = t / bs2, m = t - b * bs2, j = m / bs, i = m - j * bs;
            d = fmax(d, fabs(src[(size_t) (b * bs + j) * rwp + i] - T[b * ss + m]));
        }
        worst = fmax(worst, d / scale[k]);
    }
    for (int o = 16; o > 0; o >>= 1) worst = fmax(worst, __shfl_xor_sync(0xffffffffu, worst, o));
    if (lane == 0 && worst > 0.0) atomicMax(worst_bits, (unsigned long long) __double_as_longlong(worst));
}

yfd_status maxwell_build_classes(yfd_ctx *ctx, yfd_maxwell *m, uint64_t *d_keys, int key_bits, double tol)
{
    mx_classes &cl = m->classes;
    cl = mx_classes();
    const int bs = m->bs, nb = mx_nb(m->nv), ss = cls_slot_stride(bs), cpw = mx_cpw(bs), rwp = mx_rwp(bs);
    const int64_t nc = m->ncells;
    if (nc < 1 || m->degree < 1 || m->degree > 4) return YFD_OK;
    class_groups g;
    /* the table is read from global memory (L1 / L2 hits): 2 MB keeps it L2-resident beside the vectors' working set */
    YFD_TRY(group_cells_by_key(ctx, nc, d_keys, key_bits, (size_t) nb * ss * sizeof(double), (size_t) 2 << 20, g));
    cl.ncls = (int) std::min<int64_t>(g.ncls, 0x7fffffff);
    if (g.reject) { cl.reject = g.reject; return YFD_OK; }
    dbuf<double> scale;
    dbuf<unsigned long long> worst;
    YFD_TRY(cl.table.alloc(ctx, (size_t) g.ncls * nb * ss, 2)); YFD_TRY(scale.alloc(ctx, (size_t) g.ncls));
    YFD_TRY(worst.alloc(ctx, 1)); YFD_TRY(worst.zero(ctx));
    YFD_LAUNCH(ctx, mx_cls_build_table, (unsigned) g.ncls, 128, 0, (const int64_t *) g.start.p, (const uint32_t *) g.sorted_cell.p,
               (const double *) m->op.p, bs, nb, cpw, rwp, cl.table.p, scale.p);
    const int64_t want = ceil_div(nc, 8), cap = (int64_t) ctx->sm_count * 16;
    YFD_LAUNCH(ctx, mx_cls_verify, (unsigned) std::max<int64_t>(1, std::min(want, cap)), 256, 0, (const int32_t *) g.cell_cls.p, (const double *) m->op.p,
               (const double *) cl.table.p, (const double *) scale.p, nc, bs, nb, cpw, rwp, worst.p);
    unsigned long long wb = 0;
    YFD_TRY(worst.download(ctx, &wb, 1));
    double dev;
    memcpy(&dev, &wb, sizeof(dev));
    cl.max_dev = dev; cl.tol = tol; cl.npos = g.npos; cl.ntiles = g.ntiles;
    cl.work_cell = std::move(g.work_cell); cl.tile_cls = std::move(g.tile_cls); cl.cell_cls = std::move(g.cell_cls);
    if (!(dev <= tol)) { cl.reject = 3; return YFD_OK; }
    YFD_TRY(mma_fragment_table(ctx, cl.table.p, cl.ncls, bs, nb, cl.table_mma));
    cl.ready = true;
    return YFD_OK;
}

#include "yfd_maxwell_mma.cuh"

template <int NV, int K>
static yfd_status launch_stage_class(yfd_ctx *ctx, int mode, const mx_step_args &a)
{
    /* block sizes the tensor-pipe kernel exists for (YFD_MX_CLASS_MMA=0: the CUDA-core class kernel) */
    if constexpr (mx_bs(K) == 6 || mx_bs(K) == 10 || mx_bs(K) == 15) {
        if (a.table_mma && env_int("YFD_MX_CLASS_MMA", 1) != 0) {
            int64_t grid = (int64_t) ctx->sm_count * 2;
            const int64_t want = ceil_div(a.ntiles, MMA_THREADS / 32);
            if (grid > want) grid = want;
            if (grid < 1) grid = 1;
            if (a.upwind) YFD_LAUNCH(ctx, (maxwell_stage_class_mma<NV, K, true>), (unsigned) grid, MMA_THREADS, 0, a, mode);
            else          YFD_LAUNCH(ctx, (maxwell_stage_class_mma<NV, K, false>), (unsigned) grid, MMA_THREADS, 0, a, mode);
            return YFD_OK;
        }
    }
    /* two cells per lane (255 registers, one CTA per SM) unless YFD_MX_CLASS_CELLS=1 asks for the one-cell variant */
    const int cells = env_int("YFD_MX_CLASS_CELLS", 2) == 1 ? 1 : 2;
    int64_t grid = (int64_t) ctx->sm_count * (cells == 1 ? 2 : 1);
    const int64_t want = ceil_div(a.ntiles * (CLS_TILE / (32 * cells)), CLS_THREADS / 32);
    if (grid > want) grid = want;
    if (grid < 1) grid = 1;
    if (cells == 1) YFD_LAUNCH(ctx, (maxwell_stage_class<NV, K, 1>), (unsigned) grid, CLS_THREADS, 0, a, mode);
    else            YFD_LAUNCH(ctx, (maxwell_stage_class<NV, K, 2>), (unsigned) grid, CLS_THREADS, 0, a, mode);
    return YFD_OK;
}

template <int NV>
static yfd_status dispatch_stage_class(yfd_ctx *ctx, int degree, int mode, const mx_step_args &a)
{
    switch (degree) {
    case 1: return launch_stage_class<NV, 1>(ctx, mode, a);
    case 2: return launch_stage_class<NV, 2>(ctx, mode, a);
    case 3: return launch_stage_class<NV, 3>(ctx, mode, a);
    case 4: return launch_stage_class<NV, 4>(ctx, mode, a);
    }
    return fail(ctx, YFD_ERR_UNSUPPORTED, "Maxwell class operator: degrees 1..4");
}

/* Measured on B200 (profiles/README.md, r01 Maxwell sweep): the L2 bulk prefetch is worth +8-10 % at every degree;
 * 2 CTAs per SM (128 registers: all loads of a face in flight at once) is best for bs >= 6, 4 CTAs for bs <= 3. */
template <int K>
struct mx_tune { static constexpr int MINB = K <= 1 ? 4 : 2; };

static int mx_prefetch()
{
    static const int on = [] { const char *e = getenv("YFD_MX_PREFETCH"); return e ? atoi(e) : 1; }();
    return on;
}

template <int NV, int K>
static yfd_status launch_stage(yfd_ctx *ctx, int mode, const mx_step_args &a)
{
    constexpr int BS = mx_bs(K), CPW = mx_cpw(BS);
    const size_t smem = (size_t) MX_WARPS * (1 + NV) * 3 * BS * CPW * sizeof(double);
    const unsigned grid = (unsigned) ceil_div(a.nchunks, MX_WARPS);
    YFD_LAUNCH(ctx, (maxwell_stage<NV, K, mx_tune<K>::MINB>), grid, MX_WARPS * 32, smem, a, mode, mx_prefetch());
    return YFD_OK;
}

template <int NV>
static yfd_status dispatch_stage(yfd_ctx *ctx, int degree, int mode, const mx_step_args &a)
{
    switch (degree) {
    case 0: return launch_stage<NV, 0>(ctx, mode, a);
    case 1: return launch_stage<NV, 1>(ctx, mode, a);
    case 2: return launch_stage<NV, 2>(ctx, mode, a);
    case 3: return launch_stage<NV, 3>(ctx, mode, a);
    case 4: return launch_stage<NV, 4>(ctx, mode, a);
    case 5: return launch_stage<NV, 5>(ctx, mode, a);
    case 6: return launch_stage<NV, 6>(ctx, mode, a);
    }
    return fail(ctx, YFD_ERR_UNSUPPORTED, "Maxwell degree %d: kernels are instantiated for 0..6", degree);
}

/* The RK4 vector work of one stage as its own streaming kernel (same operations, same order as the fused sinks above).  With the
 * tensor-pipe class kernel the fused sinks are scattered 8 / 16-byte accesses of 4 lanes per cell and cost 0.31 ms per stage on
 * top of the 0.49 ms application (tri k=4 -r 9); written by the application as one plain store per entry and combined here
 * with fully coalesced accesses at the HBM rate, the same vectors cost 0.17-0.35 ms.
 *   1: o = u + (dt/2) k          (k = acc)          2: o = u + (dt/2) k          (k = kb)
 *   3: acc += 2 (kb + k); k = u + dt k  (in place)  4: u += (dt/6) (acc + k) */
__global__ void __launch_bounds__(256)
maxwell_rk_vector(int stage, int64_t n, double dt, double *__restrict__ u, double *__restrict__ acc, const double *__restrict__ kb,
                  double *__restrict__ k, double *__restrict__ o)
{
    const int64_t stride = (int64_t) gridDim.x * blockDim.x, n2 = n >> 1;
    auto one = [&](double uu, double ac, double b, double kk, double &r0, double &r1) {
        switch (stage) {
        case 1: case 2: r0 = __dadd_rn(uu, __dmul_rn(dt / 2., kk)); break;
        case 3: r0 = __dadd_rn(ac, __dmul_rn(__dadd_rn(b, kk), 2.0)); r1 = __dadd_rn(uu, __dmul_rn(dt, kk)); break;
        default: r0 = __dadd_rn(uu, __dmul_rn(__dadd_rn(ac, kk), (1. / 6.) * dt)); break;
        }
    };
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
        const double2 uu = reinterpret_cast<const double2 *>(u)[i], kk = reinterpret_cast<const double2 *>(k)[i];
        double2 ac = make_double2(0.0, 0.0), b = make_double2(0.0, 0.0);
        if (stage >= 3) ac = reinterpret_cast<const double2 *>(acc)[i];
        if (stage == 3) b = reinterpret_cast<const double2 *>(kb)[i];
        double a0 = 0, a1 = 0, b0 = 0, b1 = 0;
        one(uu.x, ac.x, b.x, kk.x, a0, a1);
        one(uu.y, ac.y, b.y, kk.y, b0, b1);
        if (stage <= 2) reinterpret_cast<double2 *>(o)[i] = make_double2(a0, b0);
        else if (stage == 3) { reinterpret_cast<double2 *>(acc)[i] = make_double2(a0, b0); reinterpret_cast<double2 *>(k)[i] = make_double2(a1, b1); }
        else reinterpret_cast<double2 *>(u)[i] = make_double2(a0, b0);
    }
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
        const int64_t i = n - 1;
        double r0 = 0, r1 = 0;
        one(u[i], stage >= 3 ? acc[i] : 0.0, stage == 3 ? kb[i] : 0.0, k[i], r0, r1);
        if (stage <= 2) o[i] = r0; else if (stage == 3) { acc[i] = r0; k[i] = r1; } else u[i] = r0;
    }
}

yfd_status maxwell_launch_rk_vector(yfd_ctx *ctx, int stage, int64_t n, double dt, double *u, double *acc, const double *kb, double *k, double *o)
{
    if (n <= 0) return YFD_OK;
    const int64_t want = ceil_div(n / 2 + 1, 256), cap = (int64_t) ctx->sm_count * 16;
    YFD_LAUNCH(ctx, maxwell_rk_vector, (unsigned) (want < cap ? want : cap), 256, 0, stage, n, dt, u, acc, kb, k, o);
    return YFD_OK;
}

yfd_status maxwell_launch_stage(yfd_ctx *ctx, int nv, int degree, int mode, const mx_step_args &a)
{
    if (a.nchunks <= 0) return YFD_OK;
    if (a.table) return nv == 3 ? dispatch_stage_class<3>(ctx, degree, mode, a) : dispatch_stage_class<4>(ctx, degree, mode, a);
    return nv == 3 ? dispatch_stage<3>(ctx, degree, mode, a) : dispatch_stage<4>(ctx, degree, mode, a);
}

} // namespace yfd

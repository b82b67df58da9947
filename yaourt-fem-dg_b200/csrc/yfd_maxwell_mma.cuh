/*
 * yfd_maxwell_mma.cuh -- maxwell_stage_class on the FP64 tensor pipe (included by yfd_maxwell_step.cu inside namespace yfd).
 *
 * The class-cache form of the Maxwell operator is 22 (triangles) / 28 (quadrangles) products of a bs x bs class block with a
 * per-cell combination of the three fields: GEMM-shaped once 8 cells of one class are taken together (M = 8 cells, N = 8
 * rows, K = 4 columns per mma.sync.m8n8k4; roles and column numbering as in yfd_class_mma.cuh / yfd_mma_core.cuh).  The
 * CUDA-core kernel needs one block entry delivered per 2-4 DFMA and sits at 15 % of the FP64 peak; here a block fragment
 * (one double per lane, read through L1) feeds 256 multiply-adds.
 *
 * Order of work per m-tile (8 consecutive entries of the class-sorted list): SOURCE-major -- the three fields of a source
 * cell (the cell itself, then the neighbour across each face) are loaded once as operand fragments, combined in registers
 * into the operands of the products, and all three output fields accumulate side by side:
 *     self:  Hx += Gy (-Ez);  Hy += Gx Ez;  Ez += Gx Hy + Gy (-Hx)
 *     face f, self:  Hx += ny (T_f q1),  Hy += -nx (T_f q1),  Ez += T_f q2       (q1, q2: see the kernel)
 *     face f, nbr:   the same with N_f and the opposite sign
 * The sink (plain application, RK4 stages, Euler) is maxwell_stage's.
 *
 * Measured and not kept (profiles/r02_class_mma.md): two m-tiles per warp sharing every block fragment (236-255 registers, one
 * CTA per SM) is 30-45 % slower than one m-tile with two CTAs per SM; sending the results through a per-warp shared-memory
 * buffer so that they leave as flat runs instead of 8-byte pieces of 8 cells is 5-10 % slower as well; loading the next
 * neighbour's fragments while the current one's three products run (24 more registers, spills at 128) changes nothing
 * (0.62 vs 0.61 ms): 16 warps per SM already cover the operand latency that the tensor pipe leaves visible.
 */
#pragma once

template <int BS, int NB>
struct mx_mma {
    using sh = mma_shape<BS>;
    static constexpr int KS = sh::KS, NTM = sh::NTM, LEFT = sh::LEFT, QS = sh::QS, CS = NB * KS * QS;
    static constexpr int NT1 = NTM > 0 ? NTM : 1, LF1 = LEFT > 0 ? LEFT : 1;

    /* acc += (block b of the class) x w, for the 8 cells of the m-tile */
    static __device__ __forceinline__ void product(const double *__restrict__ Fk, int b, int lane, int t, const double (&w)[KS],
                                                   double (&acc)[NT1][2], double (&accl)[LF1])
    {
#pragma unroll
        for (int q = 0; q < KS; q++) {
            const double *Fq = Fk + (b * KS + q) * QS;
#pragma unroll
            for (int h = 0; h < NTM; h++) dmma884(acc[h][0], acc[h][1], w[q], __ldg(Fq + h * 32 + lane));
            if (LEFT == 2) {
                const double2 wl = __ldg(reinterpret_cast<const double2 *>(Fq + NTM * 32 + 2 * t));
                accl[0] = fma(w[q], wl.x, accl[0]); accl[LEFT - 1] = fma(w[q], wl.y, accl[LEFT - 1]);
            }
            if (LEFT == 1) accl[0] = fma(w[q], __ldg(Fq + NTM * 32 + 2 * t), accl[0]);
        }
    }
};

/* flux factors of output field `fld` on one face: w = ahx Hx + ahy Hy + aez Ez (maxwell_stage_class) */
__device__ __forceinline__ void mx_face_factors(const mx_face_coef &fq, int fld, double &ahx, double &ahy, double &aez)
{
    if (fld == 0)      { aez = fq.ny * fq.cE;  ahx = -fq.vH * fq.ny * fq.ny; ahy = fq.vH * fq.nx * fq.ny; }
    else if (fld == 1) { aez = -fq.nx * fq.cE; ahx = fq.vH * fq.nx * fq.ny;  ahy = -fq.vH * fq.nx * fq.nx; }
    else               { ahx = fq.cH * fq.ny;  ahy = -fq.cH * fq.nx;         aez = -fq.vE; }
}

/* UPW = false: centred fluxes, vH = vE = 0 in every face coefficient: the terms they multiply are skipped (exact zeros) */
template <int NV, int K, bool UPW>
__global__ void __launch_bounds__(MMA_THREADS, 2)
maxwell_stage_class_mma(const mx_step_args a, const int mode)
{
    constexpr int BS = mx_bs(K), NB = mx_nb(NV), L = 3 * BS;
    using mm = mx_mma<BS, NB>;
    constexpr int KS = mm::KS, NTM = mm::NTM, LEFT = mm::LEFT, NT1 = mm::NT1, LF1 = mm::LF1;
    static_assert(CLS_TILE == 64, "a warp takes 8 m-tiles of 8 cells");
    const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int64_t gw = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
    for (int64_t tile = gw; tile < a.ntiles; tile += nw) {
        const int64_t pos0 = tile * CLS_TILE;
        const int32_t c_lo = __ldg(&a.work_cell[pos0 + lane]), c_hi = __ldg(&a.work_cell[pos0 + 32 + lane]);
        const int32_t c0 = __shfl_sync(0xffffffffu, c_lo, 0);
        if (c0 < 0) continue;
        const double *Fk = a.table_mma + (size_t) __ldg(&a.tile_cls[tile]) * mm::CS;
        /* whether face f has a neighbour is part of the class key: the same answer for every cell of the tile */
        bool has_nb[NV];
#pragma unroll
        for (int f = 0; f < NV; f++) has_nb[f] = __ldg(&a.nbr[(int64_t) c0 * NV + f]) >= 0;
#pragma unroll 1
        for (int m = 0; m < 8; m++) {
            const int32_t cm = __shfl_sync(0xffffffffu, m < 4 ? c_lo : c_hi, 8 * (m & 3) + g);
            if (__shfl_sync(0xffffffffu, cm, 0) < 0) break;     /* padding from here to the end of the tile */
            const int32_t cc = cm < 0 ? c0 : cm;                /* padding inside the m-tile: cell 0's operands, no sink */
            double acc[3][NT1][2], accl[3][LF1];
#pragma unroll
            for (int fld = 0; fld < 3; fld++) {
#pragma unroll
                for (int h = 0; h < NT1; h++) acc[fld][h][0] = acc[fld][h][1] = 0.0;
#pragma unroll
                for (int r = 0; r < LF1; r++) accl[fld][r] = 0.0;
            }
            double v[3][KS], w[KS];
            /* ---- the cell itself ---- */
#pragma unroll
            for (int fld = 0; fld < 3; fld++) mma_operand<BS>(a.vin + (int64_t) cc * L + fld * BS, t, v[fld]);
#pragma unroll
            for (int q = 0; q < KS; q++) w[q] = -v[2][q];
            mm::product(Fk, 1, lane, t, w, acc[0], accl[0]);                      /* Hx: Gy (-Ez) */
            mm::product(Fk, 0, lane, t, v[2], acc[1], accl[1]);                   /* Hy: Gx Ez    */
            mm::product(Fk, 0, lane, t, v[1], acc[2], accl[2]);                   /* Ez: Gx Hy    */
#pragma unroll
            for (int q = 0; q < KS; q++) w[q] = -v[0][q];
            mm::product(Fk, 1, lane, t, w, acc[2], accl[2]);                      /*     + Gy (-Hx) */
            /* faces.  With z = ny Hx - nx Hy the three flux operands of a source are
             *     Hx: ny q1,   Hy: -nx q1,   Ez: q2,      q1 = cE Ez - vH z,   q2 = cH z - vE Ez
             * (maxwell_stage_class's w_field,f, factored): ONE product B q1 serves both H fields -- its result is scaled by the
             * cell's own ny / -nx afterwards, a lane's accumulators all belong to its cell g -- and B q2 goes straight into Ez:
             * two block products per face and source instead of three (16 instead of 22 per triangle). */
            auto face_source = [&](const mx_face_coef &fq, int b, double sign) {
                double p1[NT1][2], p1l[LF1];
#pragma unroll
                for (int h = 0; h < NT1; h++) p1[h][0] = p1[h][1] = 0.0;
#pragma unroll
                for (int r = 0; r < LF1; r++) p1l[r] = 0.0;
#pragma unroll
                for (int q = 0; q < KS; q++) {
                    const double z = fma(fq.ny, v[0][q], -fq.nx * v[1][q]);
                    w[q] = UPW ? fma(fq.cE, v[2][q], -fq.vH * z) : fq.cE * v[2][q];
                    v[0][q] = sign * (UPW ? fma(fq.cH, z, -fq.vE * v[2][q]) : fq.cH * z);      /* q2 (v[0] is not needed any more) */
                }
                mm::product(Fk, b, lane, t, w, p1, p1l);
                const double sx = sign * fq.ny, sy = -sign * fq.nx;
#pragma unroll
                for (int h = 0; h < NTM; h++) {
                    acc[0][h][0] = fma(sx, p1[h][0], acc[0][h][0]); acc[0][h][1] = fma(sx, p1[h][1], acc[0][h][1]);
                    acc[1][h][0] = fma(sy, p1[h][0], acc[1][h][0]); acc[1][h][1] = fma(sy, p1[h][1], acc[1][h][1]);
                }
#pragma unroll
                for (int r = 0; r < LEFT; r++) { accl[0][r] = fma(sx, p1l[r], accl[0][r]); accl[1][r] = fma(sy, p1l[r], accl[1][r]); }
                mm::product(Fk, b, lane, t, v[0], acc[2], accl[2]);
            };
            /* the cell's own faces need its fields three times: keep Hx aside (face_source overwrites v[0]) */
            double hx[KS];
#pragma unroll
            for (int q = 0; q < KS; q++) hx[q] = v[0][q];
#pragma unroll
            for (int f = 0; f < NV; f++) {
#pragma unroll
                for (int q = 0; q < KS; q++) v[0][q] = hx[q];
                face_source(a.fc[(int64_t) cc * NV + f], 2 + f, 1.0);
            }
            /* ---- the neighbours ---- */
#pragma unroll
            for (int f = 0; f < NV; f++) {
                if (!has_nb[f]) continue;
                const int32_t nb = __ldg(&a.nbr[(int64_t) cc * NV + f]);
#pragma unroll
                for (int fld = 0; fld < 3; fld++) mma_operand<BS>(a.vin + (int64_t) nb * L + fld * BS, t, v[fld]);
                face_source(a.fc[(int64_t) cc * NV + f], 2 + NV + f, -1.0);
            }
            /* ---- sink: rows 8 h + 2 t, + 1 of cell g from the tensor tiles; LEFT rows summed over the cell's 4 lanes,
             * lane t = r keeps row 8 NTM + r ---- */
            auto sink = [&](int64_t idx, double kk) {
                switch (mode) {
                case MX_APPLY: a.out[idx] = kk; break;
                case MX_RK1:   a.acc[idx] = kk; a.out[idx] = __dadd_rn(__ldg(&a.vin[idx]), __dmul_rn(a.dt / 2., kk)); break;
                case MX_RK2:   a.kb[idx] = kk;  a.out[idx] = __dadd_rn(a.u[idx], __dmul_rn(a.dt / 2., kk)); break;
                case MX_RK3:   a.acc[idx] = __dadd_rn(a.acc[idx], __dmul_rn(__dadd_rn(a.kb[idx], kk), 2.0));
                               a.out[idx] = __dadd_rn(a.u[idx], __dmul_rn(a.dt, kk)); break;
                case MX_RK4:   a.u[idx] = __dadd_rn(a.u[idx], __dmul_rn(__dadd_rn(a.acc[idx], kk), (1. / 6.) * a.dt)); break;
                case MX_EULER: a.out[idx] = __dadd_rn(__ldg(&a.vin[idx]), __dmul_rn(a.dt, kk)); break;
                }
            };
#pragma unroll
            for (int fld = 0; fld < 3; fld++) {
#pragma unroll
                for (int r = 0; r < LEFT; r++) {
                    accl[fld][r] += __shfl_xor_sync(0xffffffffu, accl[fld][r], 1);
                    accl[fld][r] += __shfl_xor_sync(0xffffffffu, accl[fld][r], 2);
                }
            }
            /* the same sink on two adjacent rows at once (16-byte accesses: half the requests of the scattered 8-byte ones, which is
             * what the RK stages spend their time on: up to five accesses per entry) */
            auto sink2 = [&](int64_t idx, double kx, double ky) {
                auto ld2 = [](const double *p) { return *reinterpret_cast<const double2 *>(p); };
                auto st2 = [](double *p, double x, double y) { *reinterpret_cast<double2 *>(p) = make_double2(x, y); };
                switch (mode) {
                case MX_APPLY: st2(a.out + idx, kx, ky); break;
                case MX_RK1: { st2(a.acc + idx, kx, ky); const double2 v = __ldg(reinterpret_cast<const double2 *>(a.vin + idx));
                               st2(a.out + idx, __dadd_rn(v.x, __dmul_rn(a.dt / 2., kx)), __dadd_rn(v.y, __dmul_rn(a.dt / 2., ky))); break; }
                case MX_RK2: { st2(a.kb + idx, kx, ky); const double2 v = ld2(a.u + idx);
                               st2(a.out + idx, __dadd_rn(v.x, __dmul_rn(a.dt / 2., kx)), __dadd_rn(v.y, __dmul_rn(a.dt / 2., ky))); break; }
                case MX_RK3: { const double2 ac = ld2(a.acc + idx), kb = ld2(a.kb + idx), v = ld2(a.u + idx);
                               st2(a.acc + idx, __dadd_rn(ac.x, __dmul_rn(__dadd_rn(kb.x, kx), 2.0)), __dadd_rn(ac.y, __dmul_rn(__dadd_rn(kb.y, ky), 2.0)));
                               st2(a.out + idx, __dadd_rn(v.x, __dmul_rn(a.dt, kx)), __dadd_rn(v.y, __dmul_rn(a.dt, ky))); break; }
                case MX_RK4: { const double2 v = ld2(a.u + idx), ac = ld2(a.acc + idx);
                               st2(a.u + idx, __dadd_rn(v.x, __dmul_rn(__dadd_rn(ac.x, kx), (1. / 6.) * a.dt)),
                                   __dadd_rn(v.y, __dmul_rn(__dadd_rn(ac.y, ky), (1. / 6.) * a.dt))); break; }
                case MX_EULER: { const double2 v = __ldg(reinterpret_cast<const double2 *>(a.vin + idx));
                                 st2(a.out + idx, __dadd_rn(v.x, __dmul_rn(a.dt, kx)), __dadd_rn(v.y, __dmul_rn(a.dt, ky))); break; }
                }
            };
            const int64_t cq = cm < 0 ? 0 : cm;
            const double s_mu = __ldg(&a.cs[2 * cq]), s_eps = __ldg(&a.cs[2 * cq + 1]);
            if constexpr (BS == 15) {
                /* a field's 15 rows start at cm * 45 + fld * 15: even or odd with cm + fld.  Even: the lane's own pairs (2t, 2t+1),
                 * (8+2t, 9+2t) are aligned, row 14 is single.  Odd: pairs start at odd rows -- (2t+1, 2t+2), (9+2t, 10+2t), row 0 single --
                 * and the second row of each comes from the next lane of the cell (lane 3 takes row 8 from lane 0). */
                const int src = (lane & ~3) | ((t + 1) & 3);
#pragma unroll
                for (int fld = 0; fld < 3; fld++) {
                    const double sc = fld == 2 ? s_eps : s_mu;
                    const double d0 = sc * acc[fld][0][0], d1 = sc * acc[fld][0][1], d2 = sc * acc[fld][1][0], d3 = sc * acc[fld][1][1];
                    const double nA = __shfl_sync(0xffffffffu, t == 0 ? d2 : d0, src);
                    const double nB = __shfl_sync(0xffffffffu, d2, src);
                    if (cm < 0) continue;
                    const bool odd = ((cm + fld) & 1) != 0;
                    const int64_t base = (int64_t) cm * L + fld * BS;
                    sink2(base + (odd ? 2 * t + 1 : 2 * t), odd ? d1 : d0, odd ? nA : d1);
                    if (t < 3) sink2(base + (odd ? 9 + 2 * t : 8 + 2 * t), odd ? d3 : d2, odd ? nB : d3);
                    if (odd ? t == 0 : t == 3) sink(base + (odd ? 0 : 14), odd ? d0 : d2);
                }
            } else if constexpr (BS == 10 || BS == 6) {
                /* 3 bs even: every pair (2t, 2t+1) is aligned; bs = 10: rows 8, 9 (the DFMA rows, summed over the 4 lanes) leave from lane 0 */
                if (cm >= 0) {
#pragma unroll
                    for (int fld = 0; fld < 3; fld++) {
                        const double sc = fld == 2 ? s_eps : s_mu;
                        const int64_t base = (int64_t) cm * L + fld * BS;
                        if (2 * t + 1 < BS) sink2(base + 2 * t, sc * acc[fld][0][0], sc * acc[fld][0][1]);
                        if (LEFT == 2 && t == 0) sink2(base + 8, sc * accl[fld][0], sc * accl[fld][LEFT - 1]);
                    }
                }
            } else if (cm >= 0) {
#pragma unroll
                for (int fld = 0; fld < 3; fld++) {
                    const double sc = fld == 2 ? s_eps : s_mu;
                    const int64_t base = (int64_t) cm * L + fld * BS;
#pragma unroll
                    for (int h = 0; h < NTM; h++) {
                        const int i = 8 * h + 2 * t;
                        if (i < BS)     sink(base + i, sc * acc[fld][h][0]);
                        if (i + 1 < BS) sink(base + i + 1, sc * acc[fld][h][1]);
                    }
                    if (LEFT > 0 && t < LEFT) sink(base + 8 * NTM + t, sc * (t == 0 ? accl[fld][0] : accl[fld][LEFT - 1]));
                }
            }
        }
    }
}

/*
 * yfd_elem.cuh -- device pieces shared by the element kernels (yfd_dg.cu, yfd_maxwell.cu): the reference's
 * quadrature rules mapped to physical cells, the scaled-monomial basis and the cell frame.
 */
#pragma once
#include "yfd_internal.hpp"

namespace yfd {

constexpr int DG_MAX_DEGREE = 6;          /* bs = 28 */

__host__ __device__ constexpr int dg_nq_cell(int nv, int k)
{
    /* Dunavant point counts for degree 2k (degree 0 uses rule 1) */
    return nv == 4 ? (k + 1) * (k + 1)
                   : (k == 0 ? 1 : k == 1 ? 3 : k == 2 ? 6 : k == 3 ? 12 : k == 4 ? 16 : k == 5 ? 25 : 33);
}

/* cell_basis_bones::eval / eval_grads (core/bases.hpp:96-155): phi = xt^(d-i) yt^i, xt = (x-xb)/(h/2) */
template <int K>
struct basis_eval {
    static constexpr int BS = (K + 2) * (K + 1) / 2;
    /* phi[i], gx[i], gy[i] for all i at physical point (x,y); frame (xb,yb,h) */
    __device__ static __forceinline__ void run(double x, double y, double xb, double yb, double h,
                                               double *phi, double *gx, double *gy, int stride)
    {
        const double half_h = 0.5 * h;
        const double bx = (x - xb) / half_h, by = (y - yb) / half_h;
        const double ih = 2.0 / h;
        double px[K + 1], py[K + 1];
        px[0] = 1.0; py[0] = 1.0;
#pragma unroll
        for (int p = 1; p <= K; p++) { px[p] = px[p - 1] * bx; py[p] = py[p - 1] * by; }
        int pos = 0;
#pragma unroll
        for (int d = 0; d <= K; d++) {
#pragma unroll
            for (int i = 0; i <= d; i++) {
                const int ex = d - i, ey = i;
                phi[pos * stride] = px[ex] * py[ey];
                const double dx = ex == 0 ? 0.0 : (double) ex * ih * px[ex == 0 ? 0 : ex - 1];
                const double dy = ey == 0 ? 0.0 : (double) ey * ih * py[ey == 0 ? 0 : ey - 1];
                gx[pos * stride] = dx * py[ey];
                gy[pos * stride] = px[ex] * dy;
                pos++;
            }
        }
    }
    /* values only */
    __device__ static __forceinline__ void values(double x, double y, double xb, double yb, double h,
                                                  double *phi, int stride)
    {
        const double half_h = 0.5 * h;
        const double bx = (x - xb) / half_h, by = (y - yb) / half_h;
        double px[K + 1], py[K + 1];
        px[0] = 1.0; py[0] = 1.0;
#pragma unroll
        for (int p = 1; p <= K; p++) { px[p] = px[p - 1] * bx; py[p] = py[p - 1] * by; }
        int pos = 0;
#pragma unroll
        for (int d = 0; d <= K; d++)
#pragma unroll
            for (int i = 0; i <= d; i++) { phi[pos * stride] = px[d - i] * py[i]; pos++; }
    }
};

template <int NV>
__device__ __forceinline__ void cell_frame(const double2 *P, double &xb, double &yb, double &h)
{
    /* barycenter (core/mesh.hpp:424-441), diameter = max pairwise vertex distance (:474-487) */
    double sx = P[0].x, sy = P[0].y;
#pragma unroll
    for (int v = 1; v < NV; v++) { sx += P[v].x; sy += P[v].y; }
    xb = sx / (double) NV; yb = sy / (double) NV;
    double d = 0.0;
#pragma unroll
    for (int a = 0; a < NV; a++)
#pragma unroll
        for (int b = a + 1; b < NV; b++) {
            const double dx = P[a].x - P[b].x, dy = P[a].y - P[b].y;
            d = fmax(sqrt(dx * dx + dy * dy), d);
        }
    h = d;
}

/* q-th point of integrate(msh, cell, 2K): triangles -> triangle_quadrature_dunavant (core/quadratures.hpp:583-624,
 * rule = (c0,c1,c2,w)[nq]); quadrangles -> quadrangle_quadrature + bilinear map (:775-861), eta outer / xi inner */
template <int NV, int K>
__device__ __forceinline__ void cell_quad_point(const double2 *P, int q, const double *rule, const double *gx,
                                                const double *gw, double &x, double &y, double &w)
{
    if (NV == 3) {
        const double c0 = __ldg(&rule[4 * q]), c1 = __ldg(&rule[4 * q + 1]);
        const double c2 = __ldg(&rule[4 * q + 2]), cw = __ldg(&rule[4 * q + 3]);
        const double v0x = P[1].x - P[0].x, v0y = P[1].y - P[0].y;
        const double v1x = P[2].x - P[0].x, v1y = P[2].y - P[0].y;
        const double area = fabs((v0x * v1y - v0y * v1x) / 2.0);
        x = P[0].x * c0 + P[1].x * c1 + P[2].x * c2;
        y = P[0].y * c0 + P[1].y * c1 + P[2].y * c2;
        w = cw * area;
    } else {
        const int iy = q / (K + 1), ix = q - iy * (K + 1);
        const double xi = __ldg(&gx[ix]), et = __ldg(&gx[iy]);
        const double w2 = __ldg(&gw[ix]) * __ldg(&gw[iy]);
        x = 0.25 * P[0].x * (1 - xi) * (1 - et) + 0.25 * P[1].x * (1 + xi) * (1 - et) +
            0.25 * P[2].x * (1 + xi) * (1 + et) + 0.25 * P[NV - 1].x * (1 - xi) * (1 + et);
        y = 0.25 * P[0].y * (1 - xi) * (1 - et) + 0.25 * P[1].y * (1 + xi) * (1 - et) +
            0.25 * P[2].y * (1 + xi) * (1 + et) + 0.25 * P[NV - 1].y * (1 - xi) * (1 + et);
        const double j11 = 0.25 * ((P[1].x - P[0].x) * (1 - et) + (P[2].x - P[NV - 1].x) * (1 + et));
        const double j12 = 0.25 * ((P[1].y - P[0].y) * (1 - et) + (P[2].y - P[NV - 1].y) * (1 + et));
        const double j21 = 0.25 * ((P[NV - 1].x - P[0].x) * (1 - xi) + (P[2].x - P[1].x) * (1 + xi));
        const double j22 = 0.25 * ((P[NV - 1].y - P[0].y) * (1 - xi) + (P[2].y - P[1].y) * (1 + xi));
        w = w2 * fabs(j11 * j22 - j12 * j21);
    }
}

/* normal(msh, cl, fc) (core/mesh.hpp:505-530): rotate the edge, flip by the barycentre test, normalise */
__device__ __forceinline__ void face_normal(double2 A, double2 B, double xb, double yb, double &nx, double &ny)
{
    const double ex = B.x - A.x, ey = B.y - A.y;
    nx = -ey; ny = ex;
    const double fbx = (A.x + B.x) / 2.0, fby = (A.y + B.y) / 2.0;
    if (nx * (fbx - xb) + ny * (fby - yb) < 0.0) { nx = -nx; ny = -ny; }
    const double inv = 1.0 / sqrt(nx * nx + ny * ny);
    nx *= inv; ny *= inv;
}

} // namespace yfd

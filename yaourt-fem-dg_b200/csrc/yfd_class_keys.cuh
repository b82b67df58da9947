/*
 * yfd_class_keys.cuh -- the hash that PROPOSES congruence classes of cells (yfd_spmv_class.cu, yfd_maxwell.cu).
 */
#pragma once
#include "yfd_internal.hpp"
#include "yfd_elem.cuh"

namespace yfd {

__device__ __forceinline__ uint64_t cls_mix(uint64_t h, int64_t v)
{
    h ^= (uint64_t) v + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2);
    h ^= h >> 30; h *= 0xbf58476d1ce4e5b9ull;
    h ^= h >> 27; h *= 0x94d049bb133111ebull;
    return h ^ (h >> 31);
}

/* What the blocks of a cell depend on with constant coefficients (dg2d_diffusion.cpp:133-193; methods/dg_maxwell_2D.hpp:
 * 220-461): the cell's vertices relative to its barycentre, and per face whether there is a neighbour and that neighbour's
 * frame (barycentre, diameter) -- the neighbour basis is evaluated from its frame alone (core/bases.hpp:96-155).
 * Everything in units of the cell diameter, rounded to 2^-16; with_size also mixes in the diameter itself (2^-16 relative):
 * the SIP blocks are scale-invariant in 2D, the Maxwell blocks M^-1 S scale like 1/h.  The hash only PROPOSES classes,
 * a verification pass over the assembled blocks decides.  Also reduces the smallest diameter and the largest coordinate
 * (the rounding-noise scale eps |x| / h of the blocks).  Args: pts, cells, cell_faces, face_owners, perm, ncells, cell_begin. */
template <int NV, typename Args>
__global__ void cell_class_keys(const Args a, int key_bits, int with_size, uint64_t *__restrict__ keys,
                                unsigned long long *__restrict__ hmin_bits, unsigned long long *__restrict__ xmax_bits)
{
    const int64_t cell = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    double hloc = INFINITY, xloc = 0.0;
    if (cell < a.ncells) {
        const int32_t rc = a.perm[cell + a.cell_begin];
        double2 P[NV];
#pragma unroll
        for (int v = 0; v < NV; v++) P[v] = __ldg(&a.pts[a.cells[(int64_t) NV * rc + v]]);
        double xb, yb, h;
        cell_frame<NV>(P, xb, yb, h);
        const double q = 65536.0 / h;
        uint64_t key = (uint64_t) NV;
        if (with_size) key = cls_mix(key, llrint(log2(h) * 65536.0));
#pragma unroll
        for (int v = 0; v < NV; v++) {
            key = cls_mix(key, llrint((P[v].x - xb) * q));
            key = cls_mix(key, llrint((P[v].y - yb) * q));
            xloc = fmax(xloc, fmax(fabs(P[v].x), fabs(P[v].y)));
        }
#pragma unroll
        for (int e = 0; e < NV; e++) {
            const int32_t f = a.cell_faces[(int64_t) NV * rc + e];
            const int2 own = __ldg(&a.face_owners[f]);
            const int32_t nrc = own.x == rc ? own.y : own.x;
            if (nrc < 0) { key = cls_mix(key, 0x7fffffffffffll); continue; }
            double2 Q[NV];
#pragma unroll
            for (int v = 0; v < NV; v++) Q[v] = __ldg(&a.pts[a.cells[(int64_t) NV * nrc + v]]);
            double nxb, nyb, nh;
            cell_frame<NV>(Q, nxb, nyb, nh);
            key = cls_mix(key, llrint((nxb - xb) * q));
            key = cls_mix(key, llrint((nyb - yb) * q));
            key = cls_mix(key, llrint(nh * q));
        }
        keys[cell] = key_bits >= 64 ? key : (key & ((1ull << key_bits) - 1));
        hloc = h;
    }
    for (int o = 16; o > 0; o >>= 1) {
        hloc = fmin(hloc, __shfl_xor_sync(0xffffffffu, hloc, o));
        xloc = fmax(xloc, __shfl_xor_sync(0xffffffffu, xloc, o));
    }
    if ((threadIdx.x & 31) == 0 && isfinite(hloc)) {      /* positive doubles: the bit patterns order as the values */
        atomicMin(hmin_bits, (unsigned long long) __double_as_longlong(hloc));
        atomicMax(xmax_bits, (unsigned long long) __double_as_longlong(xloc));
    }
}

} // namespace yfd

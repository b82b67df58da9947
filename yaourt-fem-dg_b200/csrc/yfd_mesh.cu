/*
 * yfd_mesh.cu -- device-resident mesh, connectivity and the internal cell order.
 *
 * Replaces, for the hot path, what the reference recomputes with binary searches inside
 * its cell loop: face_ids()/offset(face) (core/mesh.hpp:287-296, :329-353) and
 * mesh::compute_connectivity (core/mesh.hpp:176-204).  Here both are one pass on the GPU:
 * a thread per (cell, local face) searches the sorted face list, and the two owners of a
 * face are the min and max owning cell id (the reference fills fo[0] then fo[1] while
 * walking cells in ascending id, so fo[0] < fo[1]).
 *
 * New (no reference counterpart): an internal Z-curve renumbering of the cells
 * (perm / iperm).  Reference cell ids follow refinement history, so neighbours are far
 * apart in index space (hard part 1); the element kernel, the CSR and
 * the solver all work in internal order, and every accessor maps back.
 */
#include "yfd_internal.hpp"
#include <cmath>

namespace yfd {

__global__ void narrow_i64(const int64_t *__restrict__ in, int32_t *__restrict__ out, int64_t n,
                           int64_t lo, int64_t hi, int *__restrict__ err)
{
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const int64_t v = in[i];
        if (v < lo || v >= hi) atomicExch(err, 1);
        out[i] = (int32_t) v;
    }
}

/* faces must be strictly increasing in (p0,p1) with p0 < p1; cells strictly increasing lexicographically */
template <int NV>
__global__ void check_sorted(const int32_t *__restrict__ cells, int64_t ncells,
                             const int32_t *__restrict__ faces, int64_t nfaces, int *__restrict__ err)
{
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nfaces) {
        const int32_t a = faces[2 * i], b = faces[2 * i + 1];
        if (a >= b) atomicExch(err, 2);
        if (i > 0) {
            const int32_t pa = faces[2 * i - 2], pb = faces[2 * i - 1];
            if (!(pa < a || (pa == a && pb < b))) atomicExch(err, 2);
        }
    }
    if (i > 0 && i < ncells) {
        bool less = false;
#pragma unroll
        for (int v = 0; v < NV; v++) {
            const int32_t x = cells[NV * (i - 1) + v], y = cells[NV * i + v];
            if (x < y) { less = true; break; }
            if (x > y) break;
        }
        if (!less) atomicExch(err, 3);
    }
}

__device__ __forceinline__ int64_t face_lower_bound(const int2 *__restrict__ faces, int64_t nfaces, int32_t a, int32_t b)
{
    int64_t lo = 0, hi = nfaces;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        const int2 f = __ldg(&faces[mid]);
        if (f.x < a || (f.x == a && f.y < b)) lo = mid + 1; else hi = mid;
    }
    return lo;
}

/* one thread per (cell, local face): face id + owner bookkeeping */
template <int NV>
__global__ void connect_cells(const int32_t *__restrict__ cells, int64_t ncells,
                              const int2 *__restrict__ faces, int64_t nfaces,
                              int32_t *__restrict__ cell_faces,
                              int32_t *__restrict__ own_min, int32_t *__restrict__ own_max, int32_t *__restrict__ own_cnt,
                              int *__restrict__ err)
{
    const int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ncells * NV) return;
    const int64_t c = t / NV;
    const int e = (int) (t - c * NV);
    /* local face order: tri p0p1,p1p2,p0p2 ; quad p0p1,p1p2,p2p3,p0p3 (core/mesh.hpp:299-325) */
    const int ia = (e == NV - 1) ? 0 : e;
    const int ib = (e == NV - 1) ? NV - 1 : e + 1;
    int32_t a = cells[NV * c + ia], b = cells[NV * c + ib];
    if (a > b) { int32_t s = a; a = b; b = s; }
    const int64_t f = face_lower_bound(faces, nfaces, a, b);
    bool found = f < nfaces;
    if (found) { const int2 ff = faces[f]; found = (ff.x == a && ff.y == b); }
    if (!found) {
        /* the reference throws only past-the-end (core/mesh.hpp:292-293) and would otherwise
         * silently use a wrong face; a missing edge is always an error here */
        atomicExch(err, 4);
        cell_faces[t] = -1;
        return;
    }
    cell_faces[t] = (int32_t) f;
    atomicMin(&own_min[f], (int32_t) c);
    atomicMax(&own_max[f], (int32_t) c);
    if (atomicAdd(&own_cnt[f], 1) >= 2) atomicExch(err, 5);
}

__global__ void finish_owners(const int32_t *__restrict__ own_min, const int32_t *__restrict__ own_max,
                              const int32_t *__restrict__ own_cnt, int32_t *__restrict__ owners, int64_t nfaces)
{
    const int64_t f = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= nfaces) return;
    const int cnt = own_cnt[f];
    owners[2 * f]     = cnt >= 1 ? own_min[f] : -1;
    owners[2 * f + 1] = cnt >= 2 ? own_max[f] : -1;
}

__device__ __forceinline__ uint32_t spread16(uint32_t v)
{
    v &= 0xffffu;
    v = (v | (v << 8)) & 0x00ff00ffu;
    v = (v | (v << 4)) & 0x0f0f0f0fu;
    v = (v | (v << 2)) & 0x33333333u;
    v = (v | (v << 1)) & 0x55555555u;
    return v;
}

template <int NV>
__global__ void morton_keys(const int32_t *__restrict__ cells, int64_t ncells, const double2 *__restrict__ pts,
                            double x0, double y0, double sx, double sy, uint64_t *__restrict__ keys)
{
    const int64_t c = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncells) return;
    double bx = 0, by = 0;
#pragma unroll
    for (int v = 0; v < NV; v++) { const double2 p = pts[cells[NV * c + v]]; bx += p.x; by += p.y; }
    bx /= NV; by /= NV;
    int ix = (int) ((bx - x0) * sx), iy = (int) ((by - y0) * sy);
    ix = min(max(ix, 0), 65535); iy = min(max(iy, 0), 65535);
    keys[c] = (uint64_t) (spread16((uint32_t) ix) | (spread16((uint32_t) iy) << 1));
}

__global__ void invert_perm(const uint32_t *__restrict__ perm_u, int32_t *__restrict__ perm, int32_t *__restrict__ iperm, int64_t n)
{
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t r = perm_u[i];
    perm[i] = (int32_t) r;
    iperm[r] = (int32_t) i;
}

__global__ void identity_perm(int32_t *__restrict__ perm, int32_t *__restrict__ iperm, int64_t n)
{
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    perm[i] = (int32_t) i;
    iperm[i] = (int32_t) i;
}

__global__ void widen_i32(const int32_t *__restrict__ in, int64_t *__restrict__ out, int64_t n)
{
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x)
        out[i] = (int64_t) in[i];
}

static inline unsigned grid_for(int64_t n, int block = 256) { return (unsigned) (n > 0 ? ceil_div(n, block) : 1); }

yfd_status upload_narrow(yfd_ctx *ctx, const int64_t *host, int64_t n, int64_t lo, int64_t hi, dbuf<int32_t> &out, int *d_err)
{
    YFD_TRY(out.alloc(ctx, (size_t) n));
    if (n == 0) return YFD_OK;
    dbuf<int64_t> tmp;
    YFD_TRY(tmp.alloc(ctx, (size_t) n));
    YFD_TRY(tmp.upload(ctx, host, (size_t) n));
    const unsigned g = (unsigned) std::min<int64_t>(ceil_div(n, 256), (int64_t) ctx->sm_count * 16);
    YFD_LAUNCH(ctx, narrow_i64, g, 256, 0, tmp.p, out.p, n, lo, hi, d_err);
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return YFD_OK;
}

yfd_status download_widen(yfd_ctx *ctx, const int32_t *d_in, int64_t n, int64_t *host)
{
    if (n == 0) return YFD_OK;
    dbuf<int64_t> tmp;
    YFD_TRY(tmp.alloc(ctx, (size_t) n));
    const unsigned g = (unsigned) std::min<int64_t>(ceil_div(n, 256), (int64_t) ctx->sm_count * 16);
    YFD_LAUNCH(ctx, widen_i32, g, 256, 0, d_in, tmp.p, n);
    return tmp.download(ctx, host, (size_t) n);
}

} // namespace yfd

using namespace yfd;

template <int NV>
static yfd_status build_connectivity(yfd_ctx *ctx, yfd_mesh *m, int *d_err)
{
    const int64_t nc = m->ncells, nf = m->nfaces;
    YFD_LAUNCH(ctx, check_sorted<NV>, grid_for(std::max(nc, nf)), 256, 0, m->cells.p, nc, m->faces.p, nf, d_err);
    dbuf<int32_t> omin, omax, ocnt;
    YFD_TRY(omin.alloc(ctx, (size_t) nf)); YFD_TRY(omax.alloc(ctx, (size_t) nf)); YFD_TRY(ocnt.alloc(ctx, (size_t) nf));
    YFD_CUDA(ctx, cudaMemsetAsync(omin.p, 0x7f, (size_t) nf * 4, ctx->stream));
    YFD_CUDA(ctx, cudaMemsetAsync(omax.p, 0xff, (size_t) nf * 4, ctx->stream));
    YFD_TRY(ocnt.zero(ctx));
    YFD_TRY(m->cell_faces.alloc(ctx, (size_t) nc * NV));
    YFD_TRY(m->face_owners.alloc(ctx, (size_t) nf * 2));
    YFD_LAUNCH(ctx, connect_cells<NV>, grid_for(nc * NV), 256, 0, m->cells.p, nc, (const int2 *) m->faces.p, nf,
               m->cell_faces.p, omin.p, omax.p, ocnt.p, d_err);
    YFD_LAUNCH(ctx, finish_owners, grid_for(nf), 256, 0, omin.p, omax.p, ocnt.p, m->face_owners.p, nf);
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return YFD_OK;
}

template <int NV>
static yfd_status build_order(yfd_ctx *ctx, yfd_mesh *m, const double *host_pts)
{
    const int64_t nc = m->ncells;
    YFD_TRY(m->perm.alloc(ctx, (size_t) nc));
    YFD_TRY(m->iperm.alloc(ctx, (size_t) nc));
    if (m->reorder == YFD_REORDER_NONE || nc < 2) {
        YFD_LAUNCH(ctx, identity_perm, grid_for(nc), 256, 0, m->perm.p, m->iperm.p, nc);
        return YFD_OK;
    }
    double x0 = 1e300, x1 = -1e300, y0 = 1e300, y1 = -1e300;
    for (int64_t i = 0; i < m->npoints; i++) {
        const double x = host_pts[2 * i], y = host_pts[2 * i + 1];
        x0 = x < x0 ? x : x0; x1 = x > x1 ? x : x1;
        y0 = y < y0 ? y : y0; y1 = y > y1 ? y : y1;
    }
    const double sx = x1 > x0 ? 65536.0 / (x1 - x0) : 0.0, sy = y1 > y0 ? 65536.0 / (y1 - y0) : 0.0;
    dbuf<uint64_t> keys, keys2;
    dbuf<uint32_t> vals, vals2;
    YFD_TRY(keys.alloc(ctx, (size_t) nc)); YFD_TRY(keys2.alloc(ctx, (size_t) nc));
    YFD_TRY(vals.alloc(ctx, (size_t) nc)); YFD_TRY(vals2.alloc(ctx, (size_t) nc));
    YFD_LAUNCH(ctx, morton_keys<NV>, grid_for(nc), 256, 0, m->cells.p, nc, (const double2 *) m->points.p, x0, y0, sx, sy, keys.p);
    YFD_TRY(fill_iota_u32(ctx, vals.p, nc));
    /* stable: cells with the same Z-key keep reference order */
    YFD_TRY(radix_sort_pairs(ctx, keys.p, vals.p, keys2.p, vals2.p, nc, 32));
    YFD_LAUNCH(ctx, invert_perm, grid_for(nc), 256, 0, vals.p, m->perm.p, m->iperm.p, nc);
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return YFD_OK;
}

extern "C" {

yfd_status yfd_mesh_create(yfd_ctx *ctx, int cell_type, int64_t npoints, const double *points,
                           int64_t ncells, const int64_t *cells, int64_t nfaces, const int64_t *faces,
                           const uint8_t *face_boundary, int reorder, yfd_mesh **out)
{
    if (!ctx || !out) return YFD_ERR_INVALID_ARGUMENT;
    *out = nullptr;
    yfd::nvtx_range nr_mesh("yfd: mesh upload + connectivity + Z-curve order");
    if (cell_type != YFD_CELL_TRIANGLE && cell_type != YFD_CELL_QUADRANGLE)
        return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "unknown cell type %d", cell_type);
    if (npoints < 0 || ncells < 0 || nfaces < 0 || (npoints && !points) || (ncells && !cells) || (nfaces && (!faces || !face_boundary)))
        return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "null or negative mesh array");
    if (npoints >= 0x7fffffff || ncells >= 0x7fffffff / 4 * 4 || nfaces >= 0x7fffffff)
        return fail(ctx, YFD_ERR_UNSUPPORTED, "mesh too large for 32-bit entity ids");
    if (reorder != YFD_REORDER_NONE && reorder != YFD_REORDER_MORTON)
        return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "unknown reorder mode %d", reorder);
    scoped_device dev(ctx->device);

    yfd_mesh *m = new yfd_mesh();
    m->ctx = ctx;
    m->cell_type = cell_type;
    m->nv = m->nf = (cell_type == YFD_CELL_TRIANGLE) ? 3 : 4;
    m->npoints = npoints; m->ncells = ncells; m->nfaces = nfaces;
    m->reorder = reorder;
    yfd_status st = YFD_OK;
    dbuf<int> derr;
    int herr = 0;
    do {
        if ((st = derr.alloc(ctx, 1)) != YFD_OK) break;
        if ((st = derr.zero(ctx)) != YFD_OK) break;
        if ((st = m->points.alloc(ctx, (size_t) npoints * 2)) != YFD_OK) break;
        if ((st = m->points.upload(ctx, points, (size_t) npoints * 2)) != YFD_OK) break;
        if ((st = upload_narrow(ctx, cells, ncells * m->nv, 0, npoints, m->cells, derr.p)) != YFD_OK) break;
        if ((st = upload_narrow(ctx, faces, nfaces * 2, 0, npoints, m->faces, derr.p)) != YFD_OK) break;
        if ((st = m->face_bnd.alloc(ctx, (size_t) nfaces)) != YFD_OK) break;
        if ((st = m->face_bnd.upload(ctx, face_boundary, (size_t) nfaces)) != YFD_OK) break;
        if ((st = derr.download(ctx, &herr, 1)) != YFD_OK) break;
        if (herr) { st = fail(ctx, YFD_ERR_INVALID_ARGUMENT, "mesh refers to a point id outside [0, npoints)"); break; }
        st = (m->nv == 3) ? build_connectivity<3>(ctx, m, derr.p) : build_connectivity<4>(ctx, m, derr.p);
        if (st != YFD_OK) break;
        if ((st = derr.download(ctx, &herr, 1)) != YFD_OK) break;
        if (herr == 2) { st = fail(ctx, YFD_ERR_INVALID_ARGUMENT, "faces are not sorted by (p0,p1) with p0<p1 (core/mesh.hpp:77-81)"); break; }
        if (herr == 3) { st = fail(ctx, YFD_ERR_INVALID_ARGUMENT, "cells are not sorted lexicographically (core/meshers.hpp:178)"); break; }
        if (herr == 4) { st = fail(ctx, YFD_ERR_FACE_NOT_FOUND, "Mesh face not found"); break; }
        if (herr == 5) { st = fail(ctx, YFD_ERR_TOO_MANY_OWNERS, "BUG: a face has max 2 owners"); break; }
        st = (m->nv == 3) ? build_order<3>(ctx, m, points) : build_order<4>(ctx, m, points);
        if (st != YFD_OK) break;
        cudaError_t e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) { st = fail(ctx, YFD_ERR_CUDA, "mesh upload: %s", cudaGetErrorString(e)); break; }
    } while (0);
    if (st != YFD_OK) { delete m; return st; }
    *out = m;
    return YFD_OK;
}

void yfd_mesh_destroy(yfd_mesh *m)
{
    if (!m) return;
    scoped_device dev(m->ctx->device);
    cudaStreamSynchronize(m->ctx->stream);
    delete m;
}

void yfd_mesh_sizes(const yfd_mesh *m, int64_t *npoints, int64_t *ncells, int64_t *nfaces, int *nv)
{
    if (npoints) *npoints = m->npoints;
    if (ncells) *ncells = m->ncells;
    if (nfaces) *nfaces = m->nfaces;
    if (nv) *nv = m->nv;
}

yfd_status yfd_mesh_get_face_owners(yfd_mesh *m, int64_t *owners)
{
    if (!m || !owners) return YFD_ERR_INVALID_ARGUMENT;
    scoped_device dev(m->ctx->device);
    return download_widen(m->ctx, m->face_owners.p, m->nfaces * 2, owners);
}

yfd_status yfd_mesh_get_cell_faces(yfd_mesh *m, int64_t *cell_faces)
{
    if (!m || !cell_faces) return YFD_ERR_INVALID_ARGUMENT;
    scoped_device dev(m->ctx->device);
    return download_widen(m->ctx, m->cell_faces.p, m->ncells * m->nf, cell_faces);
}

yfd_status yfd_mesh_get_permutation(yfd_mesh *m, int64_t *perm)
{
    if (!m || !perm) return YFD_ERR_INVALID_ARGUMENT;
    scoped_device dev(m->ctx->device);
    return download_widen(m->ctx, m->perm.p, m->ncells, perm);
}

yfd_status yfd_mesh_set_points(yfd_mesh *m, const double *points)
{
    if (!m || !points) return YFD_ERR_INVALID_ARGUMENT;
    scoped_device dev(m->ctx->device);
    YFD_TRY(m->points.upload(m->ctx, points, (size_t) m->npoints * 2));
    YFD_CUDA(m->ctx, cudaStreamSynchronize(m->ctx->stream));
    return YFD_OK;
}

} /* extern "C" */

/*
 * include/yaourt_b200/core.hpp -- shared pieces of the drop-in headers: status -> exception mapping, the process-wide
 * context, mesh upload and CompressedMatrix <-> CSR array conversion.  See dropin.hpp.
 */
#pragma once
#include <cstdint>
#include <fstream>
#include <iostream>
#include <stdexcept>
#include <string>
#include <vector>
#include "yfd.h"

namespace yaourt {
namespace b200 {

/* status -> the exception type the reference throws for the same condition */
inline void check(yfd_ctx *ctx, yfd_status st)
{
    if (st == YFD_OK) return;
    std::string msg = ctx && yfd_last_error(ctx)[0] ? yfd_last_error(ctx) : yfd_status_string(st);
    switch (st) {
    case YFD_ERR_INVALID_ARGUMENT: case YFD_ERR_INVALID_STATE: case YFD_ERR_CELL_NOT_FOUND:
    case YFD_ERR_FACE_NOT_FOUND: case YFD_ERR_DEGREE_TOO_HIGH:
        throw std::invalid_argument(msg);              /* methods/dg.hpp:72-73, core/mesh.hpp:281,293, core/quadratures.hpp:579 */
    case YFD_ERR_NO_NEIGHBOURS: case YFD_ERR_TOO_MANY_OWNERS:
        throw std::logic_error(msg);                   /* core/mesh.hpp:199,256 */
    default:
        throw std::runtime_error(msg);
    }
}

/* One context per process (the reference is single-threaded and has no notion of a device). */
class context {
    yfd_ctx *h_ = nullptr;
public:
    explicit context(int device = 0)
    {
        const yfd_status st = yfd_ctx_create(device, nullptr, &h_);
        if (st != YFD_OK) throw std::runtime_error(std::string("yaourt_b200: ") + yfd_status_string(st));
    }
    context(const context &) = delete;
    context &operator=(const context &) = delete;
    ~context() { yfd_ctx_destroy(h_); }
    yfd_ctx *get() const { return h_; }
    static context &global()
    {
        static context c(0);
        return c;
    }
};

/* mesh<T,2,CellT,FaceT> (core/mesh.hpp:158-171) -> flat arrays -> device mesh */
template <typename Mesh>
yfd_mesh *upload_mesh(context &ctx, const Mesh &msh)
{
    const size_t nv = msh.cells.empty() ? 3 : msh.cells[0].p.size();
    std::vector<double> pts(2 * msh.points.size());
    for (size_t i = 0; i < msh.points.size(); i++) { pts[2 * i] = msh.points[i].x(); pts[2 * i + 1] = msh.points[i].y(); }
    std::vector<int64_t> cells(nv * msh.cells.size()), faces(2 * msh.faces.size());
    std::vector<uint8_t> fb(msh.faces.size());
    for (size_t i = 0; i < msh.cells.size(); i++)
        for (size_t v = 0; v < nv; v++) cells[nv * i + v] = (int64_t) msh.cells[i].p[v];
    for (size_t i = 0; i < msh.faces.size(); i++) {
        faces[2 * i] = (int64_t) msh.faces[i].p0; faces[2 * i + 1] = (int64_t) msh.faces[i].p1;
        fb[i] = msh.faces[i].is_boundary ? 1 : 0;
    }
    yfd_mesh *m = nullptr;
    check(ctx.get(), yfd_mesh_create(ctx.get(), nv == 3 ? YFD_CELL_TRIANGLE : YFD_CELL_QUADRANGLE, (int64_t) msh.points.size(), pts.data(),
                                     (int64_t) msh.cells.size(), cells.data(), (int64_t) msh.faces.size(), faces.data(), fb.data(),
                                     YFD_REORDER_MORTON, &m));
    return m;
}

/* CSR arrays -> blaze::CompressedMatrix through reserve / append / finalize (core/blaze_sparse_init.hpp:134-154) */
template <typename Matrix>
void fill_compressed(Matrix &M, size_t rows, size_t cols, const std::vector<int64_t> &rowptr, const std::vector<int64_t> &colidx,
                     const std::vector<double> &values)
{
    M.resize(rows, cols, false);
    M.reset();
    M.reserve(values.size());
    for (size_t r = 0; r < rows; r++) {
        for (int64_t q = rowptr[r]; q < rowptr[r + 1]; q++) M.append(r, (size_t) colidx[(size_t) q], values[(size_t) q]);
        M.finalize(r);
    }
}

template <typename Matrix>
void flatten_compressed(const Matrix &A, std::vector<int64_t> &rowptr, std::vector<int64_t> &colidx, std::vector<double> &values)
{
    rowptr.assign(A.rows() + 1, 0);
    colidx.clear(); values.clear();
    colidx.reserve(A.nonZeros()); values.reserve(A.nonZeros());
    for (size_t r = 0; r < A.rows(); r++) {
        for (auto it = A.cbegin(r); it != A.cend(r); ++it) { colidx.push_back((int64_t) it->index()); values.push_back(it->value()); }
        rowptr[r + 1] = (int64_t) colidx.size();
    }
}

} // namespace b200
} // namespace yaourt


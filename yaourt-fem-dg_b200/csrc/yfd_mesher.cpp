/*
 * yfd_mesher.cpp -- host mesh generator with the reference's numbering.
 *
 * Replaces mesher<simplicial_mesh<T>> / mesher<quad_mesh<T>> (core/meshers.hpp:130-179,
 * :302-361): unit-square base mesh + uniform red refinement.  The numbering contract
 *  that makes dof maps bit-exact:
 *   - a refinement level numbers the midpoint of face f (f = rank in the (p0,p1)-sorted
 *     face list) as  npoints + f           (core/meshers.hpp:47-55, :203-211);
 *   - quadrangles then append one barycentre per cell in the level's cell order (:242-244);
 *   - children are created in the order t0..t3 / q0..q3 with the vertex tuples of
 *     :90-109 / :256-281, tuples are NOT sorted inside a cell;
 *   - faces are re-sorted + uniqued every level (:117, :289), cells are sorted
 *     lexicographically once at the end (:178, :360).
 * Design differs from the reference (which keeps an array of edge structs and re-sorts
 * them): faces are 64-bit keys, boundary faces (O(sqrt(N)) of them) are carried in a side
 * list, and the per-cell work is an OpenMP loop.
 */
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <new>
#include <vector>
#include "yfd.h"

struct yfd_host_mesh {
    int cell_type = 0, nv = 3;
    std::vector<double>  points;      /* [np][2] */
    std::vector<int64_t> cells;       /* [nc][nv] */
    std::vector<int64_t> faces;       /* [nf][2] */
    std::vector<uint8_t> fbnd;
    std::vector<int64_t> fbid;
};

namespace {

inline uint64_t edge_key(uint32_t a, uint32_t b)
{
    return a < b ? ((uint64_t) a << 32) | b : ((uint64_t) b << 32) | a;
}

struct bnd_edge { uint64_t key; int64_t id; };

struct level_state {
    int nv;
    std::vector<double>   pts;     /* x,y interleaved */
    std::vector<uint64_t> fkey;    /* sorted, unique */
    std::vector<bnd_edge> bnd;     /* sorted by key */
    std::vector<uint32_t> cells;   /* nv per cell, creation order */
};

inline int64_t find_face(const std::vector<uint64_t> &fkey, uint64_t k)
{
    return (int64_t) (std::lower_bound(fkey.begin(), fkey.end(), k) - fkey.begin());
}

inline const bnd_edge *find_bnd(const std::vector<bnd_edge> &bnd, uint64_t k)
{
    auto it = std::lower_bound(bnd.begin(), bnd.end(), k, [](const bnd_edge &e, uint64_t key) { return e.key < key; });
    return (it != bnd.end() && it->key == k) ? &*it : nullptr;
}

void refine_once(level_state &L)
{
    const int nv = L.nv;
    const int64_t np0 = (int64_t) L.pts.size() / 2;
    const int64_t nf0 = (int64_t) L.fkey.size();
    const int64_t nc0 = (int64_t) L.cells.size() / nv;
    const int64_t np1 = np0 + nf0 + (nv == 4 ? nc0 : 0);
    L.pts.resize((size_t) np1 * 2);
    double *P = L.pts.data();

    /* midpoints, numbered by sorted face rank */
#pragma omp parallel for schedule(static) if (nf0 > 100000)
    for (int64_t f = 0; f < nf0; f++) {
        const uint32_t a = (uint32_t) (L.fkey[f] >> 32), b = (uint32_t) L.fkey[f];
        P[2 * (np0 + f)]     = (P[2 * a] + P[2 * b]) / 2.0;
        P[2 * (np0 + f) + 1] = (P[2 * a + 1] + P[2 * b + 1]) / 2.0;
    }

    const int epc = nv == 3 ? 9 : 16;           /* new face records per parent */
    std::vector<uint64_t> nkey((size_t) nc0 * epc);
    std::vector<uint32_t> ncell((size_t) nc0 * 4 * nv);
    /* children of boundary faces: each boundary face has exactly one owner, so writing its two
     * halves at 2*rank(face in bnd list) is race free */
    std::vector<bnd_edge> nbnd(L.bnd.size() * 2);

#pragma omp parallel for schedule(static) if (nc0 > 50000)
    for (int64_t c = 0; c < nc0; c++) {
        const uint32_t *v = &L.cells[(size_t) c * nv];
        uint32_t mid[4];
        for (int e = 0; e < nv; e++) {
            /* local face e: tri p0p1,p1p2,p0p2 ; quad p0p1,p1p2,p2p3,p0p3 (core/mesh.hpp:299-325) */
            uint32_t a, b;
            if (nv == 3) { a = v[e == 2 ? 0 : e]; b = v[e == 2 ? 2 : e + 1]; }
            else         { a = v[e == 3 ? 0 : e]; b = v[e == 3 ? 3 : e + 1]; }
            const uint64_t k = edge_key(a, b);
            const int64_t f = find_face(L.fkey, k);
            mid[e] = (uint32_t) (np0 + f);
            if (const bnd_edge *be = find_bnd(L.bnd, k)) {
                const size_t r = (size_t) (be - L.bnd.data());
                nbnd[2 * r]     = { edge_key(a, mid[e]), be->id };
                nbnd[2 * r + 1] = { edge_key(mid[e], b), be->id };
            }
        }
        uint64_t *K = &nkey[(size_t) c * epc];
        uint32_t *C = &ncell[(size_t) c * 4 * nv];
        if (nv == 3) {
            const uint32_t p0 = v[0], p1 = v[1], p2 = v[2], p0b = mid[0], p1b = mid[1], p2b = mid[2];
            K[0] = edge_key(p0, p0b);  K[1] = edge_key(p0b, p2b); K[2] = edge_key(p0, p2b);
            K[3] = edge_key(p0b, p1);  K[4] = edge_key(p0b, p1b); K[5] = edge_key(p1, p1b);
            K[6] = edge_key(p1b, p2);  K[7] = edge_key(p2b, p1b); K[8] = edge_key(p2, p2b);
            const uint32_t ch[12] = { p0, p0b, p2b,  p0b, p1, p1b,  p1b, p2, p2b,  p0b, p1b, p2b };
            std::memcpy(C, ch, sizeof(ch));
        } else {
            const uint32_t p0 = v[0], p1 = v[1], p2 = v[2], p3 = v[3];
            const uint32_t p0b = mid[0], p1b = mid[1], p2b = mid[2], p3b = mid[3];
            const uint32_t bar = (uint32_t) (np0 + nf0 + c);
            P[2 * bar]     = (P[2 * p0] + P[2 * p1] + P[2 * p2] + P[2 * p3]) / 4.0;
            P[2 * bar + 1] = (P[2 * p0 + 1] + P[2 * p1 + 1] + P[2 * p2 + 1] + P[2 * p3 + 1]) / 4.0;
            K[0]  = edge_key(p0, p0b);  K[1]  = edge_key(p0b, bar); K[2]  = edge_key(p3b, bar); K[3]  = edge_key(p0, p3b);
            K[4]  = edge_key(p0b, p1);  K[5]  = edge_key(p1, p1b);  K[6]  = edge_key(p1b, bar); K[7]  = K[1];
            K[8]  = K[6];               K[9]  = edge_key(p1b, p2);  K[10] = edge_key(p2, p2b);  K[11] = edge_key(p2b, bar);
            K[12] = K[2];               K[13] = K[11];              K[14] = edge_key(p2b, p3);  K[15] = edge_key(p3, p3b);
            const uint32_t ch[16] = { p0, p0b, bar, p3b,  p0b, p1, p1b, bar,  bar, p1b, p2, p2b,  p3b, bar, p2b, p3 };
            std::memcpy(C, ch, sizeof(ch));
        }
    }

    std::sort(nkey.begin(), nkey.end());
    nkey.erase(std::unique(nkey.begin(), nkey.end()), nkey.end());
    std::sort(nbnd.begin(), nbnd.end(), [](const bnd_edge &a, const bnd_edge &b) { return a.key < b.key; });
    L.fkey.swap(nkey);
    L.bnd.swap(nbnd);
    L.cells.swap(ncell);
}

template <int NV>
void sort_cells(std::vector<uint32_t> &cells)
{
    struct tup { uint32_t p[NV]; };
    static_assert(sizeof(tup) == NV * sizeof(uint32_t), "packed");
    tup *b = reinterpret_cast<tup *>(cells.data());
    tup *e = b + cells.size() / NV;
    std::sort(b, e, [](const tup &x, const tup &y) {
        for (int i = 0; i < NV; i++) {
            if (x.p[i] < y.p[i]) return true;
            if (x.p[i] > y.p[i]) return false;
        }
        return false;
    });
}

} // namespace

extern "C" {

yfd_status yfd_mesher_create(int cell_type, int levels, yfd_host_mesh **out)
{
    if (!out) return YFD_ERR_INVALID_ARGUMENT;
    *out = nullptr;
    if ((cell_type != YFD_CELL_TRIANGLE && cell_type != YFD_CELL_QUADRANGLE) || levels < 0 || levels > 13)
        return YFD_ERR_INVALID_ARGUMENT;
    try {
        level_state L;
        auto add_bnd = [&](uint32_t a, uint32_t b, int64_t id) { L.bnd.push_back({ edge_key(a, b), id }); };
        if (cell_type == YFD_CELL_TRIANGLE) {
            /* core/meshers.hpp:133-151 */
            L.nv = 3;
            L.pts = { 0.0, 0.0,  1.0, 0.0,  1.0, 1.0,  0.0, 1.0,  0.5, 0.5 };
            const uint32_t e[8][2] = { {0,1}, {0,3}, {0,4}, {1,2}, {1,4}, {2,3}, {2,4}, {3,4} };
            for (auto &x : e) L.fkey.push_back(edge_key(x[0], x[1]));
            add_bnd(0, 1, 0); add_bnd(0, 3, 3); add_bnd(1, 2, 1); add_bnd(2, 3, 2);
            L.cells = { 0,1,4,  0,4,3,  1,2,4,  2,3,4 };
        } else {
            /* core/meshers.hpp:305-333 */
            L.nv = 4;
            L.pts = { 0.0, 0.0,  1.0, 0.0,  1.0, 1.0,  0.0, 1.0,  0.5, 0.5,  0.5, 0.0,  1.0, 0.5,  0.5, 1.0,  0.0, 0.5 };
            const uint32_t e[12][2] = { {0,5}, {0,8}, {1,5}, {1,6}, {2,6}, {2,7}, {3,7}, {3,8}, {4,5}, {4,6}, {4,7}, {4,8} };
            for (auto &x : e) L.fkey.push_back(edge_key(x[0], x[1]));
            add_bnd(0, 5, 0); add_bnd(0, 8, 3); add_bnd(1, 5, 0); add_bnd(1, 6, 1);
            add_bnd(2, 6, 1); add_bnd(2, 7, 2); add_bnd(3, 7, 2); add_bnd(3, 8, 3);
            L.cells = { 0,5,4,8,  1,6,4,5,  2,7,4,6,  3,8,4,7 };
        }
        std::sort(L.fkey.begin(), L.fkey.end());
        std::sort(L.bnd.begin(), L.bnd.end(), [](const bnd_edge &a, const bnd_edge &b) { return a.key < b.key; });
        for (int l = 0; l < levels; l++) refine_once(L);
        if (L.nv == 3) sort_cells<3>(L.cells); else sort_cells<4>(L.cells);

        yfd_host_mesh *m = new yfd_host_mesh();
        m->cell_type = cell_type;
        m->nv = L.nv;
        m->points.swap(L.pts);
        m->cells.assign(L.cells.begin(), L.cells.end());
        const size_t nf = L.fkey.size();
        m->faces.resize(nf * 2);
        m->fbnd.assign(nf, 0);
        m->fbid.assign(nf, 0);
        for (size_t f = 0; f < nf; f++) {
            m->faces[2 * f] = (int64_t) (L.fkey[f] >> 32);
            m->faces[2 * f + 1] = (int64_t) (uint32_t) L.fkey[f];
        }
        for (const bnd_edge &be : L.bnd) {
            const int64_t f = find_face(L.fkey, be.key);
            m->fbnd[(size_t) f] = 1;
            m->fbid[(size_t) f] = be.id;
        }
        *out = m;
    } catch (const std::bad_alloc &) {
        return YFD_ERR_OUT_OF_MEMORY;
    }
    return YFD_OK;
}

/* nx x ny structured mesh of the unit square in the reference's conventions (new: the reference mesher only offers
 * 4^(r+1) cells; BASELINE's "64x64 triangles", "~20M dofs", "~4M dofs" sizes fall between its levels).
 * Points: id = j (nx+1) + i at (i/nx, j/ny).  Quadrangles: (p00, p10, p11, p01), counter-clockwise like the reference's
 * base cells (core/meshers.hpp:327-330).  Triangles: every quadrangle split along p00-p11 into (p00, p10, p11) and
 * (p00, p11, p01), 2 nx ny cells.  Cells sorted lexicographically on the vertex tuple (core/meshers.hpp:178,:360), faces
 * sorted by (p0 < p1) (core/mesh.hpp:77-81), boundary ids 0 bottom, 1 right, 2 top, 3 left (core/meshers.hpp:139-144). */
yfd_status yfd_mesher_create_structured(int cell_type, int64_t nx, int64_t ny, yfd_host_mesh **out)
{
    if (!out) return YFD_ERR_INVALID_ARGUMENT;
    *out = nullptr;
    if ((cell_type != YFD_CELL_TRIANGLE && cell_type != YFD_CELL_QUADRANGLE) || nx < 1 || ny < 1 ||
        (nx + 1) * (ny + 1) >= 0x7fffffffLL)
        return YFD_ERR_INVALID_ARGUMENT;
    try {
        yfd_host_mesh *m = new yfd_host_mesh();
        m->cell_type = cell_type;
        const int nv = m->nv = cell_type == YFD_CELL_TRIANGLE ? 3 : 4;
        const int64_t npx = nx + 1, np = npx * (ny + 1);
        m->points.resize((size_t) np * 2);
#pragma omp parallel for
        for (int64_t j = 0; j <= ny; j++)
            for (int64_t i = 0; i <= nx; i++) {
                m->points[(size_t) (j * npx + i) * 2] = (double) i / (double) nx;
                m->points[(size_t) (j * npx + i) * 2 + 1] = (double) j / (double) ny;
            }
        const bool tri = nv == 3;
        const int64_t nc = (tri ? 2 : 1) * nx * ny;
        m->cells.resize((size_t) nc * nv);
        /* generation order (j, i[, half]) is already lexicographic on the first vertex p00 = j npx + i, and for the two
         * triangles of a quadrangle (p00, p10, ..) < (p00, p11, ..) since p10 < p11: no sort needed */
#pragma omp parallel for
        for (int64_t j = 0; j < ny; j++)
            for (int64_t i = 0; i < nx; i++) {
                const int64_t p00 = j * npx + i, p10 = p00 + 1, p01 = p00 + npx, p11 = p01 + 1;
                if (tri) {
                    int64_t *c = &m->cells[(size_t) (2 * (j * nx + i)) * 3];
                    c[0] = p00; c[1] = p10; c[2] = p11; c[3] = p00; c[4] = p11; c[5] = p01;
                } else {
                    int64_t *c = &m->cells[(size_t) (j * nx + i) * 4];
                    c[0] = p00; c[1] = p10; c[2] = p11; c[3] = p01;
                }
            }
        /* faces in (p0, p1) order: from point p = (i, j) the edges with p as the smaller end are, by increasing p1,
         * p -> p+1 (horizontal), p -> p+npx (vertical), p -> p+npx+1 (diagonal, triangles only) */
        const int64_t nf = nx * (ny + 1) + ny * (nx + 1) + (tri ? nx * ny : 0);
        m->faces.resize((size_t) nf * 2);
        m->fbnd.assign((size_t) nf, 0);
        m->fbid.assign((size_t) nf, 0);
        int64_t f = 0;
        for (int64_t j = 0; j <= ny; j++)
            for (int64_t i = 0; i <= nx; i++) {
                const int64_t p = j * npx + i;
                if (i < nx) {
                    m->faces[(size_t) f * 2] = p; m->faces[(size_t) f * 2 + 1] = p + 1;
                    if (j == 0) { m->fbnd[(size_t) f] = 1; m->fbid[(size_t) f] = 0; }
                    if (j == ny) { m->fbnd[(size_t) f] = 1; m->fbid[(size_t) f] = 2; }
                    f++;
                }
                if (j < ny) {
                    m->faces[(size_t) f * 2] = p; m->faces[(size_t) f * 2 + 1] = p + npx;
                    if (i == 0) { m->fbnd[(size_t) f] = 1; m->fbid[(size_t) f] = 3; }
                    if (i == nx) { m->fbnd[(size_t) f] = 1; m->fbid[(size_t) f] = 1; }
                    f++;
                }
                if (tri && i < nx && j < ny) {
                    m->faces[(size_t) f * 2] = p; m->faces[(size_t) f * 2 + 1] = p + npx + 1;
                    f++;
                }
            }
        if (f != nf) { delete m; return YFD_ERR_INVALID_STATE; }
        *out = m;
    } catch (const std::bad_alloc &) {
        return YFD_ERR_OUT_OF_MEMORY;
    }
    return YFD_OK;
}

void yfd_host_mesh_destroy(yfd_host_mesh *m) { delete m; }

void yfd_host_mesh_sizes(const yfd_host_mesh *m, int64_t *npoints, int64_t *ncells, int64_t *nfaces, int *nv)
{
    if (npoints) *npoints = (int64_t) m->points.size() / 2;
    if (ncells)  *ncells = (int64_t) m->cells.size() / m->nv;
    if (nfaces)  *nfaces = (int64_t) m->faces.size() / 2;
    if (nv)      *nv = m->nv;
}

const double  *yfd_host_mesh_points(const yfd_host_mesh *m) { return m->points.data(); }
const int64_t *yfd_host_mesh_cells(const yfd_host_mesh *m) { return m->cells.data(); }
const int64_t *yfd_host_mesh_faces(const yfd_host_mesh *m) { return m->faces.data(); }
const uint8_t *yfd_host_mesh_face_boundary(const yfd_host_mesh *m) { return m->fbnd.data(); }
const int64_t *yfd_host_mesh_face_boundary_id(const yfd_host_mesh *m) { return m->fbid.data(); }

} /* extern "C" */

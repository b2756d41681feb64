// ref_shim.cpp -- TEST INFRASTRUCTURE.  C entry points around the UNMODIFIED reference functions.
//
// oracle/build_ref.sh compiles the reference's own translation units where they lie under /root/reference
// (grid/grid.cpp, distance_fields/dfields.cpp, narrow_band/narrowBand.cpp, optimize/primeqp.cpp, optimize/quadprog.cpp,
// marching_cubes/hasher.cpp, marching_cubes/mcubes.cpp) against the stand-in headers of oracle/ref_stubs/ (Eigen, Boost
// and PCL are absent) together with this file into oracle/_ref/libmrb_ref.so.  Nothing here computes anything: every
// function converts flat [x][y][z] fp32 arrays to the reference's `grid`, calls the reference function named in its
// comment, and converts back.  Only tests/ (and the fixture generator tests/golden/make_ref_golden.py) load the library.
// C++03 like the reference (built with -std=gnu++03).
#include <string.h>

#include "grid.h"
#include "dfields.h"
#include "narrowBand.h"
#include "primeqp.h"
#include "quadprog.h"
// mcubes.h DEFINES edgeTable / triTable (non-const globals in a header, mcubes.h:9,43); mcubes.cpp already carries them,
// so this translation unit's unused copies get other names.
#define edgeTable ref_shim_unused_edgeTable
#define triTable ref_shim_unused_triTable
#include "mcubes.h"
#undef edgeTable
#undef triTable

namespace {

gridPtr to_grid(const float* a, const int* dims, const float* scale = 0, const float* shift = 0, int pad = 6) {
    Eigen::Vector3i d; d[0] = dims[0]; d[1] = dims[1]; d[2] = dims[2];
    Eigen::Vector3f sc, sh;
    for (int i = 0; i < 3; i++) { sc[i] = scale ? scale[i] : 1.0f; sh[i] = shift ? shift[i] : 0.0f; }
    gridPtr g(new grid(d, sc, sh, pad));
    if (a) {
        size_t n = 0;
        for (int i = 0; i < dims[0]; i++) for (int j = 0; j < dims[1]; j++) for (int k = 0; k < dims[2]; k++) (*g)[i][j][k] = a[n++];
    }
    return g;
}
void from_grid(gridPtr g, float* out) {
    size_t n = 0;
    for (int i = 0; i < g->dims[0]; i++) for (int j = 0; j < g->dims[1]; j++) for (int k = 0; k < g->dims[2]; k++) out[n++] = (*g)[i][j][k];
}
vector<Eigen::Vector3f> to_normals(const float* nrm, int n) {
    vector<Eigen::Vector3f> v(n);
    for (int i = 0; i < n; i++) { v[i][0] = nrm[3 * i]; v[i][1] = nrm[3 * i + 1]; v[i][2] = nrm[3 * i + 2]; }
    return v;
}
int copy_ints(const vector<int>& v, int* out, int cap, int* n_out) {
    *n_out = (int)v.size();
    if (!out) return 0;
    if ((int)v.size() > cap) return -5;
    for (size_t i = 0; i < v.size(); i++) out[i] = v[i];
    return 0;
}

}  // namespace

extern "C" {

const char* ref_build_info(void) {
#ifdef REF_STANDIN_DOUBLE
    return "reference sources + stand-in headers; eigen/SVD stand-in: independent double-precision Jacobi";
#else
    return "reference sources + stand-in headers; eigen/SVD stand-in: include/mrb_eig3.h arithmetic";
#endif
}

// fastPerim, dfields.cpp:7-37
int ref_fast_perim(const float* vol, const int* dims, float* out) { from_grid(fastPerim(to_grid(vol, dims)), out); return 0; }
// getsqdist, dfields.cpp:41-99
int ref_sqdist(const float* g, const int* dims, float* out) { from_grid(getsqdist(to_grid(g, dims)), out); return 0; }
// getsqrt, dfields.cpp:182-192
int ref_sqrt(const float* g, const int* dims, float* out) { from_grid(getsqrt(to_grid(g, dims)), out); return 0; }
// erode_grid / dilate_grid, narrowBand.cpp:10-79
int ref_erode(const float* g, const int* dims, float* out) { from_grid(erode_grid(to_grid(g, dims)), out); return 0; }
int ref_dilate(const float* g, const int* dims, float* out) { from_grid(dilate_grid(to_grid(g, dims)), out); return 0; }
// createBands, narrowBand.cpp:82-102
int ref_create_bands(const float* margin, const int* dims, float band_size, float* band, float* tight) {
    bandsPtr b = createBands(to_grid(margin, dims), band_size);
    from_grid(b->band, band);
    from_grid(b->tight_band, tight);
    return 0;
}
// findIndexes, grid.cpp:350-371
int ref_find_indexes(const float* g, const int* dims, int* out, int cap, int* n_out) {
    return copy_ints(findIndexes(to_grid(g, dims)), out, cap, n_out);
}
// getIndexMap, grid.cpp:374-389
int ref_index_map(const int* dims, const int* indexes, int n, float* out) {
    vector<int> idx(indexes, indexes + n);
    from_grid(getIndexMap(to_grid(0, dims), idx), out);
    return 0;
}
// getFeatureIndexes, primeqp.cpp:231-244
int ref_feature_indexes(const float* fmap, const float* imap, const int* dims, int* out, int cap, int* n_out) {
    return copy_ints(getFeatureIndexes(to_grid(fmap, dims), to_grid(imap, dims)), out, cap, n_out);
}
// primeQP (primeqp.cpp:168-223) then getJc/getIr/getPr (quadprog.cpp:8-42).  Any output may be NULL.
int ref_prime_qp(const float* volume, const float* margin, const float* band, const float* tight, const int* dims, int n,
                 float* lb, float* ub, float* x0, float* invdg, int* jc, int* ir, float* pr, int nnz_cap, int* nnz_out) {
    bandsPtr b(new bands());
    b->band = to_grid(band, dims);
    b->tight_band = to_grid(tight, dims);
    qp_argsPtr q = primeQP(to_grid(volume, dims), to_grid(margin, dims), b);
    if ((int)q->x.size() != n) return -1;
    for (int i = 0; i < n; i++) {
        if (lb) lb[i] = q->lb[i];
        if (ub) ub[i] = q->ub[i];
        if (x0) x0[i] = q->x[i];
        if (invdg) invdg[i] = q->invdg[i];
    }
    vector<int> vjc = getJc(q->R), vir = getIr(q->R);
    vector<float> vpr = getPr(q->R);
    if (nnz_out) *nnz_out = (int)vir.size();
    if (jc) for (size_t i = 0; i < vjc.size(); i++) jc[i] = vjc[i];
    if (ir || pr) {
        if ((int)vir.size() > nnz_cap) return -5;
        for (size_t i = 0; i < vir.size(); i++) { if (ir) ir[i] = vir[i]; if (pr) pr[i] = vpr[i]; }
    }
    return 0;
}
// doStep (quadprog.cpp:45-61) for every row, `iters` sweeps with the ping-pong of runQP (quadprog.cpp:72-91), on caller CSC arrays.
int ref_do_steps(int n, const int* jc, const int* ir, const float* pr, const float* invdg, const float* lb, const float* ub,
                 const float* x0, int iters, float* x_out) {
    const int nnz = jc[n];
    vector<int> vjc(jc, jc + n + 1), vir(ir, ir + nnz);
    vector<float> vpr(pr, pr + nnz), vinv(invdg, invdg + n), vlb(lb, lb + n), vub(ub, ub + n);
    vector<float> a(x0, x0 + n), b(n, 0.0f);
    vector<float>*in = &a, *out = &b;
    for (int it = 0; it < iters; it++) {
        if (it % 2) { in = &b; out = &a; } else { in = &a; out = &b; }
        for (int r = 0; r < n; r++) doStep(r, *in, *out, vir, vjc, vpr, vinv, vlb, vub);
    }
    for (int i = 0; i < n; i++) x_out[i] = (*out)[i];
    return 0;
}
// optimize, quadprog.cpp:111-147 (the CPU twin of quadprog.cu:236-280): stages 1-3.  feature_map may be NULL.
int ref_optimize(const float* volume, const float* feature_map, const int* dims, int using_features, float* F_out) {
    gridPtr vol = to_grid(volume, dims);
    gridPtr fm = to_grid(feature_map, dims);
    if (!feature_map)
        for (int i = 0; i < dims[0]; i++) for (int j = 0; j < dims[1]; j++) for (int k = 0; k < dims[2]; k++) (*fm)[i][j][k] = 0.0f;
    from_grid(optimize(vol, fm, using_features != 0, false), F_out);
    return 0;
}
// getSurface, dfields.cpp:322-338
int ref_get_surface(const float* vol, const int* dims, int* out, int cap, int* n_out) {
    return copy_ints(getSurface(to_grid(vol, dims)), out, cap, n_out);
}
// getSurfaceNormals, dfields.cpp:341-349
int ref_surface_normals(const float* vol, const int* dims, const int* surface, int n, float* out) {
    vector<int> s(surface, surface + n);
    vector<Eigen::Vector3f> nr = getSurfaceNormals(to_grid(vol, dims), s);
    for (int i = 0; i < n; i++) { out[3 * i] = nr[i][0]; out[3 * i + 1] = nr[i][1]; out[3 * i + 2] = nr[i][2]; }
    return 0;
}
// getFeatureMap, dfields.cpp:447-555
int ref_feature_map(const float* vol, const float* smap, const float* normals, int nn, const int* dims, float thr, float* out) {
    from_grid(getFeatureMap(to_grid(vol, dims), to_grid(smap, dims), to_normals(normals, nn), thr), out);
    return 0;
}
// VertexInterp, mcubes.cpp:369-392
int ref_vertex_interp(float iso, const float* p1, const float* p2, float v1, float v2, float* out) {
    XYZ p = VertexInterp(iso, XYZ(p1[0], p1[1], p1[2]), XYZ(p2[0], p2[1], p2[2]), v1, v2);
    out[0] = p.x; out[1] = p.y; out[2] = p.z;
    return 0;
}
// mcubes, mcubes.cpp:592-662: the whole stage 4 including the PLY file.  surface_map / normals may be NULL when !using_features.
int ref_mcubes_ply(const float* F, const float* smap, const float* normals, int nn, const int* dims, float iso, float ft, float ct,
                   const char* path, int using_features, const float* scale, const float* shift, int pad) {
    gridPtr in = to_grid(F, dims, scale, shift, pad);
    gridPtr sm = to_grid(smap, dims);
    if (!smap)
        for (int i = 0; i < dims[0]; i++) for (int j = 0; j < dims[1]; j++) for (int k = 0; k < dims[2]; k++) (*sm)[i][j][k] = -1.0f;
    mcubes(in, sm, to_normals(normals, normals ? nn : 0), iso, ft, ct, path, using_features != 0);
    return 0;
}
// The body of mcubes (mcubes.cpp:594-622) up to the hash map, returning raw fp32 vertices and int faces instead of text:
// getGridCells -> Polygonise per cell -> flipEdges -> concatenation -> pointHashmap -> index().  Two calls: verts/faces
// NULL to size, then to fill.
int ref_mcubes_mesh(const float* F, const float* smap, const float* normals, int nn, const int* dims, float iso, float ft, float ct,
                    int using_features, float* verts, int* faces, int* nverts, int* nfaces) {
    gridPtr in = to_grid(F, dims);
    gridPtr sm = to_grid(smap, dims);
    if (!smap)
        for (int i = 0; i < dims[0]; i++) for (int j = 0; j < dims[1]; j++) for (int k = 0; k < dims[2]; k++) (*sm)[i][j][k] = -1.0f;
    vector<Eigen::Vector3f> nr = to_normals(normals, normals ? nn : 0);
    const bool uf = using_features != 0;
    vector<GRIDCELL> cells = getGridCells(in, sm, nr, ft, ct, uf);
    for (size_t i = 0; i < cells.size(); i++) Polygonise(cells[i], iso, nr, uf);
    vector<TRIANGLE> new_triangles;
    if (uf) new_triangles = flipEdges(cells, in);
    vector<TRIANGLE> all;
    for (size_t i = 0; i < cells.size(); i++)
        for (size_t j = 0; j < cells[i].triangles.size(); j++) all.push_back(cells[i].triangles[j]);
    for (size_t i = 0; i < new_triangles.size(); i++) all.push_back(new_triangles[i]);
    pointHashmap indexer(all, in);
    *nverts = indexer.numPoints;
    *nfaces = (int)all.size();
    if (verts)
        for (int i = 0; i < indexer.numPoints; i++) { verts[3 * i] = indexer.points[i].x; verts[3 * i + 1] = indexer.points[i].y; verts[3 * i + 2] = indexer.points[i].z; }
    if (faces)
        for (size_t i = 0; i < all.size(); i++)
            for (int c = 0; c < 3; c++) faces[3 * i + c] = indexer.index(all[i].p[c]);
    return 0;
}
// per-cell classification of getGridCells + Polygonise (mcubes.cpp:157-312): feature value after Polygonise and the feature
// point of every surface-crossing cell, in cell order (used to compare classifications with a tolerance on the points).
int ref_feature_cells(const float* F, const float* smap, const float* normals, int nn, const int* dims, float iso, float ft, float ct,
                      int* cell_id, int* feature, float* fpoint, int cap, int* n_out) {
    gridPtr in = to_grid(F, dims);
    gridPtr sm = to_grid(smap, dims);
    vector<Eigen::Vector3f> nr = to_normals(normals, nn);
    vector<GRIDCELL> cells = getGridCells(in, sm, nr, ft, ct, true);
    int n = 0;
    for (size_t i = 0; i < cells.size(); i++) {
        Polygonise(cells[i], iso, nr, true);
        if (!cells[i].isSurface) continue;
        if (cell_id && n < cap) {
            cell_id[n] = (int)i; feature[n] = cells[i].feature;
            fpoint[3 * n] = cells[i].feature_point.x; fpoint[3 * n + 1] = cells[i].feature_point.y; fpoint[3 * n + 2] = cells[i].feature_point.z;
        }
        n++;
    }
    *n_out = n;
    return 0;
}
// grid::getCloudPoint, grid.cpp:204-210
int ref_cloud_point(const float* p, const float* scale, const float* shift, int pad, float* out) {
    int dims[3] = {1, 1, 1};
    gridPtr g = to_grid(0, dims, scale, shift, pad);
    Eigen::Vector3f v; v[0] = p[0]; v[1] = p[1]; v[2] = p[2];
    Eigen::Vector3f r = g->getCloudPoint(v);
    out[0] = r[0]; out[1] = r[1]; out[2] = r[2];
    return 0;
}

}  // extern "C"

// mrb_oracle.cpp -- CPU restatement of the reference hot path.  TEST INFRASTRUCTURE ONLY.
//
// This file is the parity oracle for mesh_reconstruction_b200.  It restates, function by
// function and in the same loop / summation order, the CPU path of CRLab/Mesh_Reconstruction
// (each function cites the reference file:line it follows).  It is NOT part of the product:
// only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
// may load it.  libmrb.so never links or calls it.
//
// Parity pin: the reference ships no tests, golden vectors or fixtures for this path (SURVEY.md
// section 4 / 8c).  The oracle is pinned against the REFERENCE'S OWN CODE: oracle/build_ref.sh
// compiles the reference's hot-path translation units unmodified from /root/reference (stand-in
// headers for the absent Eigen / Boost / PCL, oracle/ref_stubs/) into oracle/_ref/libmrb_ref.so, and
// tests/test_ref_pin.py requires this file to equal it bit for bit on every stage, the whole default
// path (F, mesh, PLY bytes), the feature path and the edge cases, live and through committed
// fixtures generated from it (tests/golden/ref_golden.*).  NOT pinned by that build, and bounded
// by tolerance tests instead: the floating-point path inside Eigen's eigen-solver / JacobiSVD
// (feature sub-path only; our definition is include/mrb_eig3.h) and the assumption that Eigen's
// sparse product returns row indexes ascending inside each column (values are exact integers).
// Also kept: hand-derived known answers (tests/test_oracle_*.py) and tests/golden/ (oracle-made).
//
// Build: g++ -O2 -ffp-contract=off -fno-fast-math -fopenmp -shared -fPIC (oracle/Makefile).
// fp32 everywhere, no FMA contraction (the reference is a Debug x86-64 build, CMakeLists.txt:3).
// OpenMP is used only across independent output elements (bit-identical to the serial loops);
// set OMP_NUM_THREADS=1 for the single-threaded reference timing.
//
// Layout: flat arrays in the reference storage order voxels[x][y][z], idx = (x*ny + y)*nz + z.

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <vector>
#include <algorithm>
#include <unordered_map>
#include <string>

#include "../mesh_reconstruction_b200/csrc/mc_tables.h"   // packed Bourke case table (data)

namespace {

struct Dims { int nx, ny, nz; };
inline size_t at(const Dims& d, int x, int y, int z) { return ((size_t)x * d.ny + y) * d.nz + z; }
inline size_t count(const Dims& d) { return (size_t)d.nx * d.ny * d.nz; }
inline Dims mk(const int* dims) { return Dims{dims[0], dims[1], dims[2]}; }

// grid::sub2ind (grid.cpp:172-174): x-fastest linear index.
inline int sub2ind(const Dims& d, int x, int y, int z) { return z * d.nx * d.ny + y * d.nx + x; }
// grid::ind2sub (grid.cpp:163-170).
inline void ind2sub(const Dims& d, int lin, int& x, int& y, int& z) {
    z = lin / (d.nx * d.ny);
    lin = lin % (d.nx * d.ny);
    y = lin / d.nx;
    x = lin % d.nx;
}

const uint64_t kTri[256] = MRB_MC_TRI_WORDS;
const int kEdgeCorner[12][2] = {{0,1},{1,2},{2,3},{3,0},{4,5},{5,6},{6,7},{7,4},{0,4},{1,5},{2,6},{3,7}};

thread_local std::string g_err;

}  // namespace

extern "C" {

const char* mrbo_last_error() { return g_err.c_str(); }

// ---------------------------------------------------------------------------------------------
// Stage 1
// ---------------------------------------------------------------------------------------------

// fastPerim, dfields.cpp:7-37.
int mrbo_fast_perim(const float* vol, const int* dims, float* out) {
    Dims d = mk(dims);
#pragma omp parallel for collapse(2) schedule(static)
    for (int i = 0; i < d.nx; i++)
        for (int j = 0; j < d.ny; j++)
            for (int k = 0; k < d.nz; k++) {
                float r = 1.0f;
                int ni = 1, nj = 1, nk = 1, pi = 1, pj = 1, pk = 1;
                if (i == 0) ni = 0; else if (i == d.nx - 1) pi = 0;
                if (j == 0) nj = 0; else if (j == d.ny - 1) pj = 0;
                if (k == 0) nk = 0; else if (k == d.nz - 1) pk = 0;
                float c = vol[at(d, i, j, k)];
                for (int a = i - ni; a <= i + pi; a++)
                    for (int b = j - nj; b <= j + pj; b++)
                        for (int e = k - nk; e <= k + pk; e++)
                            if (c != vol[at(d, a, b, e)]) r = 0.0f;
                out[at(d, i, j, k)] = r;
            }
    return 0;
}

// getsqdist, dfields.cpp:41-99 -- the definition: three brute-force full-line passes.
int mrbo_sqdist_brute(const float* in, const int* dims, float* out) {
    Dims d = mk(dims);
    size_t n = count(d);
    std::vector<float> a(in, in + n), b(n);
    // first transformation (along x), float min with sentinel 100000 (dfields.cpp:46-63)
#pragma omp parallel for collapse(2) schedule(static)
    for (int i = 0; i < d.nx; i++)
        for (int j = 0; j < d.ny; j++)
            for (int k = 0; k < d.nz; k++) {
                float mn = 100000;
                for (int q = 0; q < d.nx; q++) {
                    float dist = 100000;
                    if (a[at(d, q, j, k)] == 0.0f) dist = (float)((q - i) * (q - i));
                    if (dist < mn) mn = dist;
                }
                b[at(d, i, j, k)] = mn;
            }
    a = b;
    // second transformation (along y), int accumulate (dfields.cpp:67-79)
#pragma omp parallel for collapse(2) schedule(static)
    for (int i = 0; i < d.nx; i++)
        for (int j = 0; j < d.ny; j++)
            for (int k = 0; k < d.nz; k++) {
                int mn = 100000;
                for (int q = 0; q < d.ny; q++) {
                    int dist = (int)(a[at(d, i, q, k)] + (float)((q - j) * (q - j)));
                    if (dist < mn) mn = dist;
                }
                b[at(d, i, j, k)] = (float)mn;
            }
    a = b;
    // third transformation (along z) (dfields.cpp:83-95)
#pragma omp parallel for collapse(2) schedule(static)
    for (int i = 0; i < d.nx; i++)
        for (int j = 0; j < d.ny; j++)
            for (int k = 0; k < d.nz; k++) {
                int mn = 100000;
                for (int q = 0; q < d.nz; q++) {
                    int dist = (int)(a[at(d, i, j, q)] + (float)((q - k) * (q - k)));
                    if (dist < mn) mn = dist;
                }
                b[at(d, i, j, k)] = (float)mn;
            }
    std::memcpy(out, b.data(), n * sizeof(float));
    return 0;
}

// Exact fast mode (additional, NOT the definition): same result as mrbo_sqdist_brute by
// construction -- every intermediate is an integer < 2^24, so min(100000, exact EDT) can be
// computed with any exact algorithm.  Per line: two-sided scan (pass 1) and an early-exit
// outward search (passes 2, 3).  tests/test_oracle_dfield.py checks both modes agree.
static void line_minplus(const int* g, int n, int* outv) {
    for (int j = 0; j < n; j++) {
        int best = g[j] < 100000 ? g[j] : 100000;
        for (int r = 1; r < n; r++) {
            int r2 = r * r;
            if (r2 >= best) break;
            if (j - r >= 0) { int c = g[j - r] + r2; if (c < best) best = c; }
            if (j + r < n)  { int c = g[j + r] + r2; if (c < best) best = c; }
        }
        outv[j] = best;
    }
}
int mrbo_sqdist_fast(const float* in, const int* dims, float* out) {
    Dims d = mk(dims);
    size_t n = count(d);
    std::vector<int> a(n);
    // pass 1 along x
#pragma omp parallel for collapse(2) schedule(static)
    for (int j = 0; j < d.ny; j++)
        for (int k = 0; k < d.nz; k++) {
            int last = -1;
            for (int i = 0; i < d.nx; i++) {
                if (in[at(d, i, j, k)] == 0.0f) last = i;
                long long v = 100000;
                if (last >= 0) { long long dd = i - last; v = std::min<long long>(100000, dd * dd); }
                a[at(d, i, j, k)] = (int)v;
            }
            last = -1;
            for (int i = d.nx - 1; i >= 0; i--) {
                if (in[at(d, i, j, k)] == 0.0f) last = i;
                if (last >= 0) { long long dd = last - i; long long v = std::min<long long>(100000, dd * dd);
                                 if (v < a[at(d, i, j, k)]) a[at(d, i, j, k)] = (int)v; }
            }
        }
    // pass 2 along y
#pragma omp parallel for schedule(static)
    for (int i = 0; i < d.nx; i++) {
        std::vector<int> g(d.ny), o(d.ny);
        for (int k = 0; k < d.nz; k++) {
            for (int j = 0; j < d.ny; j++) g[j] = a[at(d, i, j, k)];
            line_minplus(g.data(), d.ny, o.data());
            for (int j = 0; j < d.ny; j++) a[at(d, i, j, k)] = o[j];
        }
    }
    // pass 3 along z
#pragma omp parallel for collapse(2) schedule(static)
    for (int i = 0; i < d.nx; i++)
        for (int j = 0; j < d.ny; j++) {
            std::vector<int> o(d.nz);
            line_minplus(&a[at(d, i, j, 0)], d.nz, o.data());
            for (int k = 0; k < d.nz; k++) out[at(d, i, j, k)] = (float)o[k];
        }
    return 0;
}

// getsqrt, dfields.cpp:182-192: (float)sqrt((double)v).
int mrbo_sqrt(const float* in, const int* dims, float* out) {
    size_t n = count(mk(dims));
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; i++) out[i] = (float)std::sqrt((double)in[i]);
    return 0;
}

// getsqrt(getsqdist(fastPerim(volume))), quadprog.cpp:115.  brute != 0 selects the definition.
int mrbo_dfield(const float* vol, const int* dims, float* margin, int brute) {
    size_t n = count(mk(dims));
    std::vector<float> p(n), s(n);
    mrbo_fast_perim(vol, dims, p.data());
    if (brute) mrbo_sqdist_brute(p.data(), dims, s.data());
    else mrbo_sqdist_fast(p.data(), dims, s.data());
    return mrbo_sqrt(s.data(), dims, margin);
}

// ---------------------------------------------------------------------------------------------
// Stage 2
// ---------------------------------------------------------------------------------------------

// erode_grid, narrowBand.cpp:10-43.
int mrbo_erode(const float* g, const int* dims, float* out) {
    Dims d = mk(dims);
    std::memcpy(out, g, count(d) * sizeof(float));
    for (int i = 1; i < d.nx - 1; i++)
        for (int j = 1; j < d.ny - 1; j++)
            for (int k = 1; k < d.nz - 1; k++) {
                float r;
                if (g[at(d, i, j, k)] == 0.0f) r = 0.0f;
                else if (g[at(d, i - 1, j, k)] == 0.0f) r = 0.0f;
                else if (g[at(d, i, j - 1, k)] == 0.0f) r = 0.0f;
                else if (g[at(d, i, j, k - 1)] == 0.0f) r = 0.0f;
                else if (g[at(d, i + 1, j, k)] == 0.0f) r = 0.0f;
                else if (g[at(d, i, j + 1, k)] == 0.0f) r = 0.0f;
                else if (g[at(d, i, j, k + 1)] == 0.0f) r = 0.0f;
                else r = 1.0f;
                out[at(d, i, j, k)] = r;
            }
    return 0;
}

// dilate_grid, narrowBand.cpp:46-79.
int mrbo_dilate(const float* g, const int* dims, float* out) {
    Dims d = mk(dims);
    std::memcpy(out, g, count(d) * sizeof(float));
    for (int i = 1; i < d.nx - 1; i++)
        for (int j = 1; j < d.ny - 1; j++)
            for (int k = 1; k < d.nz - 1; k++) {
                float r;
                if (g[at(d, i, j, k)] == 1.0f) r = 1.0f;
                else if (g[at(d, i - 1, j, k)] == 1.0f) r = 1.0f;
                else if (g[at(d, i, j - 1, k)] == 1.0f) r = 1.0f;
                else if (g[at(d, i, j, k - 1)] == 1.0f) r = 1.0f;
                else if (g[at(d, i + 1, j, k)] == 1.0f) r = 1.0f;
                else if (g[at(d, i, j + 1, k)] == 1.0f) r = 1.0f;
                else if (g[at(d, i, j, k + 1)] == 1.0f) r = 1.0f;
                else r = 0.0f;
                out[at(d, i, j, k)] = r;
            }
    return 0;
}

// createBands, narrowBand.cpp:82-102.
int mrbo_create_bands(const float* margin, const int* dims, float band_size, float* band, float* tight) {
    size_t n = count(mk(dims));
    std::vector<float> b0(n);
    for (size_t i = 0; i < n; i++) b0[i] = (margin[i] <= band_size) ? 1.0f : 0.0f;
    mrbo_erode(b0.data(), dims, tight);
    mrbo_dilate(tight, dims, band);
    return 0;
}

// findIndexes, grid.cpp:350-371: (i,j,k) lexicographic order, x-fastest linear index values.
int mrbo_find_indexes(const float* g, const int* dims, int* idx_out, int capacity, int* n_out) {
    Dims d = mk(dims);
    int n = 0;
    for (int i = 0; i < d.nx; i++)
        for (int j = 0; j < d.ny; j++)
            for (int k = 0; k < d.nz; k++)
                if (g[at(d, i, j, k)] != 0.0f) {
                    if (idx_out) {
                        if (n >= capacity) { g_err = "find_indexes: capacity"; return -5; }
                        idx_out[n] = sub2ind(d, i, j, k);
                    }
                    n++;
                }
    *n_out = n;
    return 0;
}

// getIndexMap, grid.cpp:374-389 (ids stored as float).
int mrbo_index_map(const int* dims, const int* indexes, int n, float* map_out) {
    Dims d = mk(dims);
    size_t tot = count(d);
    for (size_t i = 0; i < tot; i++) map_out[i] = -1.0f;
    for (int i = 0; i < n; i++) {
        int x, y, z; ind2sub(d, indexes[i], x, y, z);
        map_out[at(d, x, y, z)] = (float)i;
    }
    return 0;
}

// getFeatureIndexes, primeqp.cpp:231-244.
int mrbo_feature_indexes(const float* feature_map, const float* index_map, const int* dims,
                         int* out, int capacity, int* n_out) {
    size_t tot = count(mk(dims));
    std::vector<int> v;
    for (size_t i = 0; i < tot; i++)
        if (feature_map[i] == 1.0f) v.push_back((int)index_map[i]);
    std::sort(v.begin(), v.end());
    *n_out = (int)v.size();
    if (out) { if ((int)v.size() > capacity) return -5; std::copy(v.begin(), v.end(), out); }
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Stage 3
// ---------------------------------------------------------------------------------------------

// primeQP (primeqp.cpp:168-223) with getHMat (:7-112), getlb/getub (:115-165) and the CSC
// export getIr/getJc/getPr (quadprog.cpp:8-42).
//
// getHMat builds D (3*ntight rows; row (t,axis) = -2 at voxel t, +1 at t-e, +1 at t+e) from
// triplets and H = D^T * D with Eigen's sparse product.  Here H is assembled as the sum of the
// outer products d_r d_r^T of the rows of D -- the definition of D^T D -- into per-column
// sorted entry lists; all values are small integers, exact in fp32 in any summation order.
// Eigen call sites (setFromTriplets, transpose()*, diagonal().sparseView(), operator-, prune(0,0),
// InnerIterator) contribute exact integer arithmetic and a sorted-CSC ordering contract only.
int mrbo_prime_qp(const float* volume, const float* margin, const float* band, const float* tight,
                  const int* dims, int n,
                  float* lb, float* ub, float* x0, float* invdg,
                  int* jc, int* ir, float* pr, int nnz_capacity, int* nnz_out) {
    Dims d = mk(dims);
    size_t tot = count(d);
    // findIndexes(band) + getIndexMap (primeqp.cpp:170-172)
    std::vector<int> idmap(tot, -1);
    std::vector<int> bx, by, bz;
    bx.reserve(n); by.reserve(n); bz.reserve(n);
    int nb = 0;
    for (int i = 0; i < d.nx; i++)
        for (int j = 0; j < d.ny; j++)
            for (int k = 0; k < d.nz; k++)
                if (band[at(d, i, j, k)] != 0.0f) {
                    idmap[at(d, i, j, k)] = nb++;
                    bx.push_back(i); by.push_back(j); bz.push_back(k);
                }
    if (nb != n) { g_err = "prime_qp: n != band count"; return -1; }

    // H columns: up to 13 (row,value) entries each.
    struct Col { int cnt; int row[13]; float val[13]; };
    std::vector<Col> H(n);
    for (auto& c : H) c.cnt = 0;
    auto add = [&](int r, int c, float v) {
        Col& col = H[c];
        for (int q = 0; q < col.cnt; q++) if (col.row[q] == r) { col.val[q] += v; return; }
        col.row[col.cnt] = r; col.val[col.cnt] = v; col.cnt++;
    };
    // rows of D: one per tight voxel and axis (primeqp.cpp:29-91)
    for (int i = 0; i < d.nx; i++)
        for (int j = 0; j < d.ny; j++)
            for (int k = 0; k < d.nz; k++) {
                if (tight[at(d, i, j, k)] == 0.0f) continue;
                if (i < 1 || j < 1 || k < 1 || i > d.nx - 2 || j > d.ny - 2 || k > d.nz - 2) {
                    g_err = "prime_qp: tight band touches the grid border (reference exits, grid.cpp:20-37)";
                    return -4;
                }
                const int off[3][3] = {{1,0,0},{0,1,0},{0,0,1}};
                for (int a = 0; a < 3; a++) {
                    int c0 = idmap[at(d, i, j, k)];
                    int c1 = idmap[at(d, i - off[a][0], j - off[a][1], k - off[a][2])];
                    int c2 = idmap[at(d, i + off[a][0], j + off[a][1], k + off[a][2])];
                    if (c0 < 0 || c1 < 0 || c2 < 0) {
                        g_err = "prime_qp: tight voxel has a neighbour outside the band (reference: column -1)";
                        return -4;
                    }
                    int cols[3] = {c0, c1, c2};
                    float vals[3] = {-2.0f, 1.0f, 1.0f};
                    for (int p = 0; p < 3; p++)
                        for (int q = 0; q < 3; q++) add(cols[p], cols[q], vals[p] * vals[q]);
                }
            }
    // bounds (primeqp.cpp:115-165) and x0 (:178-187)
    for (int v = 0; v < n; v++) {
        size_t a = at(d, bx[v], by[v], bz[v]);
        float l = (volume[a] == 0.0f) ? -1000.0f : margin[a];
        float u = (volume[a] == 1.0f) ? 1000.0f : -margin[a];
        if (lb) lb[v] = l;
        if (ub) ub[v] = u;
        if (x0) { float x = 0.0f; if (x < l) x = l; if (x > u) x = u; x0[v] = x; }
    }
    // invdg (:189-197) and R = H - diag(H), prune (:199-210); CSC with ascending row ids.
    int nnz = 0;
    if (jc) jc[0] = 0;
    for (int c = 0; c < n; c++) {
        Col& col = H[c];
        int order[13];
        for (int q = 0; q < col.cnt; q++) order[q] = q;
        std::sort(order, order + col.cnt, [&](int a, int b) { return col.row[a] < col.row[b]; });
        float dg = 0.0f;
        for (int q = 0; q < col.cnt; q++) if (col.row[q] == c) dg = col.val[q];
        if (invdg) invdg[c] = (dg != 0.0f) ? (float)(1.0 / (double)dg) : 0.0f;
        for (int q = 0; q < col.cnt; q++) {
            int e = order[q];
            if (col.row[e] == c) continue;          // H - d
            if (col.val[e] == 0.0f) continue;       // prune(0,0)
            if (ir || pr) {
                if (nnz >= nnz_capacity) { g_err = "prime_qp: nnz capacity"; return -5; }
                if (ir) ir[nnz] = col.row[e];
                if (pr) pr[nnz] = col.val[e];
            }
            nnz++;
        }
        if (jc) jc[c + 1] = nnz;
    }
    if (nnz_out) *nnz_out = nnz;
    return 0;
}

// doStep, quadprog.cpp:45-61.
static inline void do_step(int row, const float* in, float* out, const int* ir, const int* jc,
                           const float* pr, const float* invdg, const float* lb, const float* ub) {
    float res = 0;
    int start = jc[row], end = jc[row + 1];
    for (int i = start; i < end; i++) res += pr[i] * in[ir[i]];
    res = (float)((double)(in[row] - res * invdg[row]) * 0.5);
    if (res < lb[row]) res = lb[row];
    if (res > ub[row]) res = ub[row];
    out[row] = res;
}

// runQP, quadprog.cpp:64-105.  Rows of one sweep are independent (Jacobi) -> OpenMP over rows
// is bit-identical to the serial loop.
int mrbo_run_qp(int n, const int* jc, const int* ir, const float* pr, const float* invdg,
                const float* lb, const float* ub, const float* x0, int iters,
                const int* feature_idx, int nfeat, int using_features, float* x_out) {
    std::vector<float> buf1(x0, x0 + n), buf2(n, 0.0f);
    float* in = buf1.data();
    float* out = buf2.data();
    for (int it = 0; it < iters; it++) {
        if (it % 2) { in = buf2.data(); out = buf1.data(); }
        else        { in = buf1.data(); out = buf2.data(); }
#pragma omp parallel for schedule(static)
        for (int r = 0; r < n; r++) do_step(r, in, out, ir, jc, pr, invdg, lb, ub);
    }
    if (iters == 0) out = buf1.data();   // reference would return an uninitialised pointer; define as x0
    if (using_features) {
        // quadprog.cpp:92-101; reading featureIndexes[count] past the end is UB in the reference,
        // the oracle treats it as "no match" (SURVEY Appendix A item 10).
        int cnt = 0;
        for (int i = 0; i < n; i++) {
            if (cnt < nfeat && i == feature_idx[cnt]) { out[i] = 0.2f; cnt++; }
        }
    }
    std::memcpy(x_out, out, (size_t)n * sizeof(float));
    return 0;
}

// optimize, quadprog.cpp:111-147 (quadprog.cu:236-280).  band_size/iters are 4/500 there.
// stage_ms (may be NULL): 4 floats, wall ms of dfield / bands+indexes / primeQP+runQP / (unused).
int mrbo_optimize(const float* volume, const float* feature_map, const int* dims, int using_features,
                  int band_size, int iters, int brute_edt, float* F_out, int* n_band_out, double* stage_ms);

// ---------------------------------------------------------------------------------------------
// Stage 4
// ---------------------------------------------------------------------------------------------
}  // extern "C"

namespace {

struct XYZ { float x, y, z; };
// XYZ::operator<, hasher.h:41-58
inline bool xyz_less(const XYZ& a, const XYZ& b) {
    if (a.x < b.x) return true; else if (a.x > b.x) return false;
    if (a.y < b.y) return true; else if (a.y > b.y) return false;
    if (a.z < b.z) return true; else if (a.z > b.z) return false;
    return false;
}
// XYZ::dist, hasher.h:28-33 (std::sqrt(float) overload: <cmath> + using namespace std).
inline float xyz_dist(const XYZ& l, const XYZ& r) {
    float d = (r.x - l.x) * (r.x - l.x);
    d += (r.y - l.y) * (r.y - l.y);
    d += (r.z - l.z) * (r.z - l.z);
    return std::sqrt(d);
}
// VertexInterp, mcubes.cpp:369-392.
inline XYZ vertex_interp(float iso, XYZ p1, XYZ p2, float v1, float v2) {
    if (xyz_less(p2, p1)) { std::swap(p1, p2); std::swap(v1, v2); }
    XYZ p;
    if (std::fabs(v1 - v2) > 0.00001) {
        p.x = p1.x + (p2.x - p1.x) / (v2 - v1) * (iso - v1);
        p.y = p1.y + (p2.y - p1.y) / (v2 - v1) * (iso - v1);
        p.z = p1.z + (p2.z - p1.z) / (v2 - v1) * (iso - v1);
    } else {
        p = p1;
    }
    return p;
}
// pointHashmap::hashXYZ, hasher.h:99-104.
inline uint64_t hash_xyz(const XYZ& s, const Dims& d) {
    uint64_t val = (uint64_t)((double)s.x * 1000000000.0 * (double)(d.ny * d.nz));
    val += (uint64_t)((double)s.y * 1000000.0 * (double)d.nz);
    val += (uint64_t)((double)s.z * 1000.0);
    return val;
}

}  // namespace

extern "C" {

int mrbo_hash_points(const float* xyz, int n, const int* dims, uint64_t* keys) {
    Dims d = mk(dims);
    for (int i = 0; i < n; i++) keys[i] = hash_xyz(XYZ{xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]}, d);
    return 0;
}

struct mrbo_mesh { float* verts; int* faces; int nverts; int nfaces; };

// mcubes (mcubes.cpp:592-662) default path without the PLY write: getGridCells (:157-191),
// Polygonise (:194-232), getSamplePoints (:314-367), pointHashmap (hasher.cpp:10-32).
// Cells are streamed in (i,j,k) order instead of materialising vector<GRIDCELL>.
int mrbo_mcubes(const float* F, const int* dims, float isolevel, mrbo_mesh* out) {
    Dims d = mk(dims);
    std::vector<XYZ> tri_pts;   // 3 per triangle, reference order
    for (int i = 0; i < d.nx - 1; i++)
        for (int j = 0; j < d.ny - 1; j++)
            for (int k = 0; k < d.nz - 1; k++) {
                XYZ p[8]; float val[8];
                int cubeindex = 0;
                for (int n = 0; n < 8; n++) {
                    int x = i + (int)((n % 4 == 1) || (n % 4) == 2);
                    int y = j + ((n % 4) / 2);
                    int z = k + (n / 4);
                    p[n] = XYZ{(float)x, (float)y, (float)z};
                    val[n] = F[at(d, x, y, z)];
                    if (val[n] < isolevel) cubeindex |= (1 << n);
                }
                uint64_t w = kTri[cubeindex];
                int ntri = (int)(w & 15);
                if (ntri == 0) continue;      // edgeTable[cubeindex]==0  <=>  no triangles
                int edges = 0;
                for (int e = 0; e < 12; e++)
                    if (((cubeindex >> kEdgeCorner[e][0]) & 1) != ((cubeindex >> kEdgeCorner[e][1]) & 1)) edges |= 1 << e;
                XYZ vert[12];
                for (int e = 0; e < 12; e++) {
                    if (edges & (1 << e)) vert[e] = vertex_interp(isolevel, p[kEdgeCorner[e][0]], p[kEdgeCorner[e][1]],
                                                                  val[kEdgeCorner[e][0]], val[kEdgeCorner[e][1]]);
                    else vert[e] = XYZ{-1.0f, -1.0f, -1.0f};
                }
                for (int t = 0; t < ntri; t++) {
                    XYZ a = vert[(w >> (4 + 12 * t)) & 15];
                    XYZ b = vert[(w >> (8 + 12 * t)) & 15];
                    XYZ c = vert[(w >> (12 + 12 * t)) & 15];
                    if (xyz_dist(a, b) < 0.00001) continue;
                    if (xyz_dist(b, c) < 0.00001) continue;
                    if (xyz_dist(c, a) < 0.00001) continue;
                    tri_pts.push_back(a); tri_pts.push_back(b); tri_pts.push_back(c);
                }
            }
    // pointHashmap ctor (hasher.cpp:10-24) + index() (hasher.cpp:28-32)
    std::unordered_map<uint64_t, int> map;
    map.reserve(tri_pts.size());
    std::vector<XYZ> points;
    for (size_t i = 0; i < tri_pts.size(); i++) {
        uint64_t id = hash_xyz(tri_pts[i], d);
        if (map.find(id) == map.end()) { map[id] = (int)points.size(); points.push_back(tri_pts[i]); }
    }
    out->nverts = (int)points.size();
    out->nfaces = (int)(tri_pts.size() / 3);
    out->verts = (float*)std::malloc(sizeof(float) * 3 * std::max<size_t>(1, points.size()));
    out->faces = (int*)std::malloc(sizeof(int) * std::max<size_t>(1, tri_pts.size()));
    for (size_t i = 0; i < points.size(); i++) {
        out->verts[3 * i] = points[i].x; out->verts[3 * i + 1] = points[i].y; out->verts[3 * i + 2] = points[i].z;
    }
    for (size_t i = 0; i < tri_pts.size(); i++) out->faces[i] = map.at(hash_xyz(tri_pts[i], d));
    return 0;
}

void mrbo_mesh_free(mrbo_mesh* m) { std::free(m->verts); std::free(m->faces); m->verts = nullptr; m->faces = nullptr; }

// PLY writer, mcubes.cpp:626-660: default ostream float formatting == "%g" (6 significant digits);
// vertices through grid::getCloudPoint (grid.cpp:204-210).
int mrbo_write_ply(const char* path, const float* verts, int nverts, const int* faces, int nfaces,
                   const float* scale, const float* shift, int pad) {
    FILE* f = std::fopen(path, "w");
    if (!f) { g_err = "write_ply: cannot open"; return -1; }
    std::fprintf(f, "ply\nformat ascii 1.0\nelement vertex %d\nproperty float x\nproperty float y\nproperty float z\n"
                    "element face %d\nproperty list uchar int vertex_indices\nend_header\n", nverts, nfaces);
    for (int i = 0; i < nverts; i++) {
        float p[3];
        for (int a = 0; a < 3; a++) p[a] = ((verts[3 * i + a] - (float)pad) * scale[a]) + shift[a];
        std::fprintf(f, "%g %g %g\n", p[0], p[1], p[2]);
    }
    for (int i = 0; i < nfaces; i++) std::fprintf(f, "3 %d %d %d\n", faces[3 * i], faces[3 * i + 1], faces[3 * i + 2]);
    std::fclose(f);
    return 0;
}

}  // extern "C"

#include <chrono>

extern "C" {

int mrbo_optimize(const float* volume, const float* feature_map, const int* dims, int using_features,
                  int band_size, int iters, int brute_edt, float* F_out, int* n_band_out, double* stage_ms) {
    using clk = std::chrono::steady_clock;
    auto ms = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
    Dims d = mk(dims);
    size_t tot = count(d);
    auto t0 = clk::now();
    std::vector<float> margin(tot), band(tot), tight(tot);
    mrbo_dfield(volume, dims, margin.data(), brute_edt);
    auto t1 = clk::now();
    mrbo_create_bands(margin.data(), dims, (float)band_size, band.data(), tight.data());
    int n = 0;
    mrbo_find_indexes(band.data(), dims, nullptr, 0, &n);
    std::vector<int> indexes(std::max(n, 1));
    mrbo_find_indexes(band.data(), dims, indexes.data(), n, &n);
    std::vector<int> feat;
    int nfeat = 0;
    if (feature_map) {
        std::vector<float> imap(tot);
        mrbo_index_map(dims, indexes.data(), n, imap.data());
        mrbo_feature_indexes(feature_map, imap.data(), dims, nullptr, 0, &nfeat);
        feat.resize(std::max(nfeat, 1));
        mrbo_feature_indexes(feature_map, imap.data(), dims, feat.data(), nfeat, &nfeat);
    }
    auto t2 = clk::now();
    std::vector<float> lb(n), ub(n), x0(n), invdg(n), x(n);
    std::vector<int> jc(n + 1);
    int nnz = 0;
    int rc = mrbo_prime_qp(volume, margin.data(), band.data(), tight.data(), dims, n, lb.data(), ub.data(), x0.data(),
                           invdg.data(), jc.data(), nullptr, nullptr, 0, &nnz);
    if (rc) return rc;
    std::vector<int> ir(std::max(nnz, 1));
    std::vector<float> pr(std::max(nnz, 1));
    rc = mrbo_prime_qp(volume, margin.data(), band.data(), tight.data(), dims, n, nullptr, nullptr, nullptr, nullptr,
                       jc.data(), ir.data(), pr.data(), nnz, &nnz);
    if (rc) return rc;
    auto t3 = clk::now();
    mrbo_run_qp(n, jc.data(), ir.data(), pr.data(), invdg.data(), lb.data(), ub.data(), x0.data(), iters,
                feat.data(), nfeat, using_features, x.data());
    auto t4 = clk::now();
    // F = (2*volume - 1) * (BAND_SIZE + 1), quadprog.cpp:131-142
    for (size_t i = 0; i < tot; i++) F_out[i] = (float)((2.0 * volume[i] - 1.0) * ((double)band_size + 1.0));
    for (int i = 0; i < n; i++) {
        int xx, yy, zz; ind2sub(d, indexes[i], xx, yy, zz);
        F_out[at(d, xx, yy, zz)] = x[i];
    }
    if (n_band_out) *n_band_out = n;
    if (stage_ms) { stage_ms[0] = ms(t0, t1); stage_ms[1] = ms(t1, t2); stage_ms[2] = ms(t2, t3); stage_ms[3] = ms(t3, t4); }
    return 0;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// Feature pre-pass (dfields.cpp:199-349, 447-555).  PARITY UNPINNED against the reference: the
// normals come from Eigen::SelfAdjointEigenSolver there; here from the deterministic Jacobi solver
// of include/mrb_eig3.h (the definition both the oracle and the CUDA kernels use).
// ---------------------------------------------------------------------------------------------
#include "../include/mrb_eig3.h"

namespace {

// isSurface, dfields.cpp:216-232.  Returns -1 if the 3x3x3 window leaves the grid (reference: exit(1)).
inline int is_surface(const float* g, const Dims& d, int i, int j, int k) {
    if (g[at(d, i, j, k)] != 1.0f) return 0;
    if (i < 1 || j < 1 || k < 1 || i > d.nx - 2 || j > d.ny - 2 || k > d.nz - 2) return -1;
    for (int a = i - 1; a <= i + 1; a++)
        for (int b = j - 1; b <= j + 1; b++)
            for (int c = k - 1; c <= k + 1; c++)
                if (g[at(d, a, b, c)] == 0.0f) return 1;
    return 0;
}

}  // namespace

extern "C" {

// getSurface, dfields.cpp:322-338: x-fastest linear indexes of the surface voxels in (i,j,k) order.
int mrbo_get_surface(const float* vol, const int* dims, int* out, int capacity, int* n_out) {
    Dims d = mk(dims);
    int n = 0;
    for (int i = 0; i < d.nx; i++)
        for (int j = 0; j < d.ny; j++)
            for (int k = 0; k < d.nz; k++) {
                int s = is_surface(vol, d, i, j, k);
                if (s < 0) { g_err = "get_surface: occupied voxel on the grid border (reference exits, grid.cpp:20-37)"; return -4; }
                if (s) {
                    if (out) { if (n >= capacity) return -5; out[n] = sub2ind(d, i, j, k); }
                    n++;
                }
            }
    *n_out = n;
    return 0;
}

// getSurfaceNormals, dfields.cpp:341-349 with getCovariance (:270-290), getSurfaceNeighbors (:235-251),
// getCentroid (:254-267), getNormalVector (:293-309), orientNormal (:312-319), getNeighbors (:199-213).
int mrbo_surface_normals(const float* vol, const int* dims, const int* surface, int n, float* normals) {
    Dims d = mk(dims);
    int bad = 0;
#pragma omp parallel for schedule(static)
    for (int s = 0; s < n; s++) {
        int x, y, z;
        ind2sub(d, surface[s], x, y, z);
        if (x < 2 || y < 2 || z < 2 || x > d.nx - 3 || y > d.ny - 3 || z > d.nz - 3) { bad = 1; continue; }
        // surface neighbours in (i,j,k) order
        int px[27], py[27], pz[27], cnt = 0;
        for (int i = x - 1; i <= x + 1; i++)
            for (int j = y - 1; j <= y + 1; j++)
                for (int k = z - 1; k <= z + 1; k++)
                    if (vol[at(d, i, j, k)] > 0 && is_surface(vol, d, i, j, k) == 1) { px[cnt] = i; py[cnt] = j; pz[cnt] = k; cnt++; }
        float c[3] = {0.0f, 0.0f, 0.0f};
        for (int q = 0; q < cnt; q++) { c[0] += (float)px[q]; c[1] += (float)py[q]; c[2] += (float)pz[q]; }
        c[0] = c[0] / (float)cnt; c[1] = c[1] / (float)cnt; c[2] = c[2] / (float)cnt;
        float cov[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        for (int q = 0; q < cnt; q++) {
            float v[3] = {(float)px[q] - c[0], (float)py[q] - c[1], (float)pz[q] - c[2]};
            for (int r = 0; r < 3; r++)
                for (int cc = 0; cc < 3; cc++) cov[r * 3 + cc] = cov[r * 3 + cc] + v[r] * v[cc];
        }
        for (int e = 0; e < 9; e++) cov[e] = cov[e] / (float)cnt;
        float nrm[3];
        mrb_normal_from_covariance(cov, nrm);
        // orientNormal: centroid of all occupied (>0) voxels of the 3x3x3 window, relative to the voxel
        float oc[3] = {0.0f, 0.0f, 0.0f};
        int on = 0;
        for (int i = x - 1; i <= x + 1; i++)
            for (int j = y - 1; j <= y + 1; j++)
                for (int k = z - 1; k <= z + 1; k++)
                    if (vol[at(d, i, j, k)] > 0) { oc[0] += (float)i; oc[1] += (float)j; oc[2] += (float)k; on++; }
        oc[0] = oc[0] / (float)on; oc[1] = oc[1] / (float)on; oc[2] = oc[2] / (float)on;
        oc[0] -= (float)x; oc[1] -= (float)y; oc[2] -= (float)z;
        float prod = nrm[0] * oc[0] + nrm[1] * oc[1] + nrm[2] * oc[2];
        if (prod > 0.0f) { nrm[0] = -nrm[0]; nrm[1] = -nrm[1]; nrm[2] = -nrm[2]; }
        normals[3 * s] = nrm[0]; normals[3 * s + 1] = nrm[1]; normals[3 * s + 2] = nrm[2];
    }
    if (bad) { g_err = "surface_normals: surface voxel within 2 voxels of the border (reference exits)"; return -4; }
    return 0;
}

// getFeatureMap, dfields.cpp:447-555.
int mrbo_feature_map(const float* vol, const float* surface_map, const float* normals, int nnormals, const int* dims,
                     float threshold, float* out) {
    Dims d = mk(dims);
    size_t tot = count(d);
    for (size_t i = 0; i < tot; i++) out[i] = 0.0f;
    for (int i = 0; i < d.nx - 1; i++)
        for (int j = 0; j < d.ny - 1; j++)
            for (int k = 0; k < d.nz - 1; k++) {
                int cubeindex = 0;
                int cx[8], cy[8], cz[8];
                for (int n = 0; n < 8; n++) {
                    cx[n] = i + (((n % 4) == 1 || (n % 4) == 2) ? 1 : 0);
                    cy[n] = j + (n % 4) / 2;
                    cz[n] = k + n / 4;
                    int gv = (int)vol[at(d, cx[n], cy[n], cz[n])];
                    if (gv < 1) cubeindex |= (1 << n);
                }
                if (cubeindex == 0 || cubeindex == 255) continue;      // edgeTable[cubeindex] == 0
                float mn = 5.0f;
                for (int n = 0; n < 7; n++)
                    for (int m = n + 1; m < 8; m++) {
                        int i1 = (int)surface_map[at(d, cx[n], cy[n], cz[n])];
                        int i2 = (int)surface_map[at(d, cx[m], cy[m], cz[m])];
                        if (i1 == -1 || i2 == -1) continue;
                        if (i1 < 0 || i2 < 0 || i1 >= nnormals || i2 >= nnormals) { g_err = "feature_map: normal index out of range"; return -1; }
                        const float* a = normals + 3 * i1;
                        const float* b = normals + 3 * i2;
                        float angle = a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
                        if (angle < mn) mn = angle;
                    }
                if (mn >= threshold) continue;
                for (int a = 0; a < 2; a++)
                    for (int b = 0; b < 2; b++)
                        for (int c = 0; c < 2; c++)
                            if (surface_map[at(d, i + a, j + b, k + c)] > -1.0f) out[at(d, i + a, j + b, k + c)] = 1.0f;
            }
    return 0;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// Feature-preserving marching cubes (mcubes with USING_FEATURES, mcubes.cpp:592-662): getGridCells + isFeature,
// Polygonise with feature points, flipEdges, concatenation, pointHashmap.  The per-cell arithmetic is the shared
// definition in include/mrb_feature_mc.h (PARITY UNPINNED against the reference's Eigen::JacobiSVD); the loops --
// cell order, neighbour order (x-1, x+1, y-1, y+1, z-1, z+1), triangle concatenation -- are restated here.
// ---------------------------------------------------------------------------------------------
#include "../include/mrb_feature_mc.h"

extern "C" int mrbo_mcubes_features(const float* F, const float* surface_map, const float* normals, int nnormals,
                                    const int* dims, float isolevel, float feat_thresh, float corner_thresh, mrbo_mesh* out) {
    (void)nnormals;
    Dims d = mk(dims);
    const int cx = d.nx - 1, cy = d.ny - 1, cz = d.nz - 1;
    int ea[12], eb[12];
    for (int e = 0; e < 12; e++) { ea[e] = kEdgeCorner[e][0]; eb[e] = kEdgeCorner[e][1]; }
    std::vector<MrbFeatCell> cells;                  // surface-crossing cells only, in (i,j,k) order
    std::vector<long long> cell_id;                  // their linear cell index
    std::unordered_map<long long, int> slot;
    for (int i = 0; i < cx; i++)
        for (int j = 0; j < cy; j++)
            for (int k = 0; k < cz; k++) {
                float val[8];
                int nidx[8], cubeindex = 0;
                for (int n = 0; n < 8; n++) {
                    int x = i + (int)((n % 4 == 1) || (n % 4) == 2), y = j + ((n % 4) / 2), z = k + (n / 4);
                    val[n] = F[at(d, x, y, z)];
                    nidx[n] = (int)surface_map[at(d, x, y, z)];
                    if (val[n] < isolevel) cubeindex |= (1 << n);
                }
                if (cubeindex == 0 || cubeindex == 255) continue;     // edgeTable == 0: isSurface = false, no triangles
                MrbFeatCell c;
                mrb_polygonise_feature(i, j, k, val, nidx, normals, isolevel, feat_thresh, corner_thresh, kTri[cubeindex],
                                       cubeindex, ea, eb, &c);
                long long id = ((long long)i * cy + j) * cz + k;
                slot[id] = (int)cells.size();
                cell_id.push_back(id);
                cells.push_back(c);
            }
    // flipEdges, mcubes.cpp:434-588 (cells without triangles cannot take part, so only crossing cells are visited)
    std::vector<MrbTri> new_tris;
    for (size_t a = 0; a < cells.size(); a++) {
        if (cells[a].feature == 0) continue;
        long long id = cell_id[a];
        int z = (int)(id % cz), y = (int)((id / cz) % cy), x = (int)(id / ((long long)cz * cy));
        long long nb[6];
        int nn = 0;
        if (x != 0) nb[nn++] = id - (long long)cy * cz;
        if (x != cx - 1) nb[nn++] = id + (long long)cy * cz;
        if (y != 0) nb[nn++] = id - cz;
        if (y != cy - 1) nb[nn++] = id + cz;
        if (z != 0) nb[nn++] = id - 1;
        if (z != cz - 1) nb[nn++] = id + 1;
        for (int q = 0; q < nn; q++) {
            auto it = slot.find(nb[q]);
            if (it == slot.end() || cells[it->second].feature == 0) continue;
            MrbTri buf[2 * MRB_FEAT_MAX_TRI * MRB_FEAT_MAX_TRI];
            int m = mrb_flip_pair(&cells[a], &cells[it->second], buf, 2 * MRB_FEAT_MAX_TRI * MRB_FEAT_MAX_TRI);
            for (int t = 0; t < m; t++) new_tris.push_back(buf[t]);
        }
    }
    // all_triangles = per-cell lists in cell order ++ new_triangles (mcubes.cpp:606-616); pointHashmap
    std::vector<XYZ> tri_pts;
    for (auto& c : cells)
        for (int t = 0; t < c.ntri; t++)
            for (int v = 0; v < 3; v++) tri_pts.push_back(XYZ{c.tri[t].p[v].x, c.tri[t].p[v].y, c.tri[t].p[v].z});
    for (auto& t : new_tris)
        for (int v = 0; v < 3; v++) tri_pts.push_back(XYZ{t.p[v].x, t.p[v].y, t.p[v].z});
    std::unordered_map<uint64_t, int> map;
    std::vector<XYZ> points;
    for (size_t i = 0; i < tri_pts.size(); i++) {
        uint64_t id = hash_xyz(tri_pts[i], d);
        if (map.find(id) == map.end()) { map[id] = (int)points.size(); points.push_back(tri_pts[i]); }
    }
    out->nverts = (int)points.size();
    out->nfaces = (int)(tri_pts.size() / 3);
    out->verts = (float*)std::malloc(sizeof(float) * 3 * std::max<size_t>(1, points.size()));
    out->faces = (int*)std::malloc(sizeof(int) * std::max<size_t>(1, tri_pts.size()));
    for (size_t i = 0; i < points.size(); i++) { out->verts[3 * i] = points[i].x; out->verts[3 * i + 1] = points[i].y; out->verts[3 * i + 2] = points[i].z; }
    for (size_t i = 0; i < tri_pts.size(); i++) out->faces[i] = map.at(hash_xyz(tri_pts[i], d));
    return 0;
}

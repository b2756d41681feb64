// mrb_frontend.cu -- the front end on the GPU: binvox completion + observed points -> occupancy grid.
//
// Replaces, for mesh_reconstruction.cpp:99-120,
//   binvoxToPCL            binvoxToPCL/binvoxToPcl.cpp:105-169   one thread per sample point of the up-sampled lattice, kept
//                                                                 samples listed by the ordered compaction of mrb_compact.cuh
//   getDensity/getResolution  binvoxToPcl.cpp:172-201, voxelize/voxelize.cpp:6-32   exact nearest neighbour on a counting-sorted
//                                                                 uniform grid, one thread per query; the float running sum of
//                                                                 the reference is order dependent and is taken by one thread
//   getResolutionFactor    binvoxToPcl.cpp:203-219               host arithmetic on the two statistics
//   voxelizeData           voxelize.cpp:34-51,87-101             pcl::VoxelGrid restated: leaf index per point, stable radix sort
//                                                                 (CUB, library code outside the hot path), one thread per leaf
//                                                                 sums its points in cloud order
//   grid(cloud, vox)       grid/grid.cpp:84-121                   occupancy bytes + min / max of the leaf centroids (shift, scale)
//   fillGrid               grid.cpp:213-265                       one thread per z column: first run, first gap, next voxel
//   getBinaryVolume        grid.cpp:322-334
// PARITY: the arithmetic the reference itself spells out (the vote, the point transform, the density sums, the grid
// constructor, fillGrid) is pinned by oracle/_ref (tests/test_ref_front.py); what PCL does inside (KdTreeFLANN, VoxelGrid,
// octree) is not in /root/reference and is restated from its documented behaviour: exact NN; leaf = floor(p / leaf size) - floor(min / leaf size),
// one centroid per occupied leaf.  With ConstConf(1) every strength is 1, so the octree strength averaging is a no-op.
#include <cub/device/device_radix_sort.cuh>
#include <math.h>
#include <algorithm>
#include <stdlib.h>
#include <string.h>

#include "mrb_compact.cuh"

namespace mrb {

// ---- order-preserving float <-> uint (for atomicMin / atomicMax on floats) ------------------------------------
__host__ __device__ inline unsigned fenc(float f) {
    unsigned b;
    memcpy(&b, &f, 4);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ inline float fdec(unsigned u) {
    unsigned b = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
    float f;
    memcpy(&f, &b, 4);
    return f;
}

// ---- binvox up-sampling -------------------------------------------------------------------------------------
struct VoteGeom {
    int D, W, H, res;       // depth (x), width (y), height (z) of the binvox grid; samples per cell edge
    float fres;
};

struct Sample { int x, y, z, i, j, k; };

__device__ __forceinline__ Sample sample_of(const VoteGeom& g, size_t idx) {
    const int r3 = g.res * g.res * g.res;
    size_t cell = idx / (size_t)r3;
    int s = (int)(idx - cell * (size_t)r3);
    Sample o;
    o.k = s % g.res; s /= g.res;
    o.j = s % g.res;
    o.i = s / g.res;
    o.z = (int)(cell % (size_t)(g.H - 1)); cell /= (size_t)(g.H - 1);
    o.y = (int)(cell % (size_t)(g.W - 1));
    o.x = (int)(cell / (size_t)(g.W - 1));
    return o;
}

// The sample's lattice position (binvoxToPcl.cpp:133-135): cell corner + i / res, in float.
__device__ __forceinline__ void sample_pos(const VoteGeom& g, const Sample& s, float p[3]) {
    p[0] = __fadd_rn((float)s.x, __fdiv_rn((float)s.i, g.fres));
    p[1] = __fadd_rn((float)s.y, __fdiv_rn((float)s.j, g.fres));
    p[2] = __fadd_rn((float)s.z, __fdiv_rn((float)s.k, g.fres));
}

// The vote of the eight cell corners, each weighted +-1 / (L1 distance + 1e-5) (binvoxToPcl.cpp:137-150).  The L1
// distance separates per axis, so the six |p - corner| terms are formed once; the sum over the corners keeps the
// reference's order n = 4a + 2b + c and its roundings (float adds, the 1e-5 added in double, float division).
struct VotePred {
    const unsigned char* vox;
    VoteGeom g;
    __device__ bool operator()(size_t idx) const {
        const Sample s = sample_of(g, idx);
        float p[3];
        sample_pos(g, s, p);
        float ax[2], ay[2], az[2];
#pragma unroll
        for (int a = 0; a < 2; a++) {
            ax[a] = fabsf(__fsub_rn(p[0], (float)(s.x + a)));
            ay[a] = fabsf(__fsub_rn(p[1], (float)(s.y + a)));
            az[a] = fabsf(__fsub_rn(p[2], (float)(s.z + a)));
        }
        const size_t WH = (size_t)g.W * g.H;
        float val = 0.0f;
#pragma unroll
        for (int n = 0; n < 8; n++) {
            const int a = n >> 2, b = (n >> 1) & 1, c = n & 1;
            float dist = __fadd_rn(__fadd_rn(ax[a], ay[b]), az[c]);
            dist = (float)__dadd_rn((double)dist, 0.00001);
            const int v = vox[(size_t)(s.x + a) * WH + (size_t)(s.z + c) * g.W + (size_t)(s.y + b)];     // get_index, binvoxToPcl.cpp:9-13
            val = __fadd_rn(val, __fdiv_rn((float)(2 * v - 1), dist));
        }
        return val > 0.0f;
    }
};

struct VoteXform { double scale, tx, ty, tz, fD, fW, fH; };

// Kept samples -> points in the binvox file's frame (binvoxToPcl.cpp:152-156; the expression is evaluated in double there).
__global__ void k_vote_points(const int* __restrict__ pos, int n, VoteGeom g, VoteXform t, float* __restrict__ xyz) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    const Sample s = sample_of(g, (size_t)pos[q]);
    float p[3];
    sample_pos(g, s, p);
    xyz[3 * (size_t)q + 0] = (float)__dadd_rn(__ddiv_rn(__dmul_rn(__dadd_rn((double)p[0], 0.5), t.scale), t.fD), t.tx);
    xyz[3 * (size_t)q + 1] = (float)__dadd_rn(__ddiv_rn(__dmul_rn(__dadd_rn((double)p[1], 0.5), t.scale), t.fW), t.ty);
    xyz[3 * (size_t)q + 2] = (float)__dadd_rn(__ddiv_rn(__dmul_rn(__dadd_rn((double)p[2], 0.5), t.scale), t.fH), t.tz);
}

int dev_binvox_points(mrb_ctx* ctx, const unsigned char* vox_dev, const int dwh[3], const float xform[4], int res, float** xyz_out,
                      int* n_out) {
    *xyz_out = nullptr; *n_out = 0;
    VoteGeom g;
    g.D = dwh[0]; g.W = dwh[1]; g.H = dwh[2]; g.res = res; g.fres = (float)res;
    if (g.D < 2 || g.W < 2 || g.H < 2) return MRB_OK;         // no cell: an empty cloud, as the reference's loops
    const double nsamp = (double)(g.D - 1) * (g.W - 1) * (g.H - 1) * (double)res * res * res;
    if (nsamp > 2147483647.0) { set_error("mrb_binvox_to_points: more than 2^31 - 1 sample points"); return MRB_EINVAL; }
    VotePred pred;
    pred.vox = vox_dev; pred.g = g;
    int* pos = nullptr;
    int n = 0;
    MRB_TRY(compact(ctx, pred, (size_t)nsamp, &pos, &n));
    float* xyz = nullptr;
    MRB_TRY(arena_alloc(ctx, (size_t)(n > 0 ? n : 1) * 3, &xyz));
    VoteXform t;
    t.scale = (double)xform[0]; t.tx = (double)xform[1]; t.ty = (double)xform[2]; t.tz = (double)xform[3];
    t.fD = (double)(float)g.D; t.fW = (double)(float)g.W; t.fH = (double)(float)g.H;
    if (n > 0) MRB_LAUNCH(ctx, k_vote_points, cdiv(n, 256), 256, 0, pos, n, g, t, xyz);
    *xyz_out = xyz; *n_out = n;
    return MRB_OK;
}

// ---- min / max of a point set -------------------------------------------------------------------------------
// mm[0..2] = min x,y,z, mm[3..5] = max, as fenc() words.  skip10: 0 all points, 1 only those with (index % 10) > 0.
__global__ void __launch_bounds__(256) k_points_minmax(const float* __restrict__ xyz, int n, int skip10, unsigned* __restrict__ mm) {
    unsigned lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0u, 0u, 0u};
    for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < n; q += gridDim.x * blockDim.x) {
        if (skip10 && (q % 10) == 0) continue;
#pragma unroll
        for (int a = 0; a < 3; a++) {
            const unsigned e = fenc(xyz[3 * (size_t)q + a]);
            lo[a] = min(lo[a], e); hi[a] = max(hi[a], e);
        }
    }
#pragma unroll
    for (int a = 0; a < 3; a++) {
        lo[a] = __reduce_min_sync(0xffffffffu, lo[a]);
        hi[a] = __reduce_max_sync(0xffffffffu, hi[a]);
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; a++) { atomicMin(&mm[a], lo[a]); atomicMax(&mm[3 + a], hi[a]); }
    }
}

static int points_minmax(mrb_ctx* ctx, const float* xyz, int n, int skip10, float mn[3], float mx[3]) {
    unsigned* mm = nullptr;
    MRB_TRY(arena_alloc(ctx, 8, &mm));
    MRB_CUDA(cudaMemsetAsync(mm, 0xFF, 3 * sizeof(unsigned), ctx->stream));
    MRB_CUDA(cudaMemsetAsync(mm + 3, 0, 3 * sizeof(unsigned), ctx->stream));
    const unsigned grid = std::min<unsigned>(cdiv(n, 256), (unsigned)ctx->sm_count * 8);
    MRB_LAUNCH(ctx, k_points_minmax, grid, 256, 0, xyz, n, skip10, mm);
    unsigned h[6];
    MRB_CUDA(cudaMemcpyAsync(h, mm, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    MRB_TRY(sync(ctx));
    for (int a = 0; a < 3; a++) { mn[a] = fdec(h[a]); mx[a] = fdec(h[3 + a]); }
    return MRB_OK;
}

// ---- nearest-neighbour statistic (getDensity / getResolution) -------------------------------------------------
struct NNGrid {
    float mn[3], cell;
    int dim[3];
};
__device__ __forceinline__ int nn_cell(float v, float mn, float cell) { return (int)floorf(__fdiv_rn(__fsub_rn(v, mn), cell)); }
__device__ __forceinline__ int nn_clamp(int c, int d) { return c < 0 ? 0 : (c >= d ? d - 1 : c); }

__global__ void k_nn_count(const float* __restrict__ xyz, int n, NNGrid g, int* __restrict__ count, int* __restrict__ cell_of) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n || (q % 10) == 0) return;
    const int a = nn_clamp(nn_cell(xyz[3 * (size_t)q], g.mn[0], g.cell), g.dim[0]);
    const int b = nn_clamp(nn_cell(xyz[3 * (size_t)q + 1], g.mn[1], g.cell), g.dim[1]);
    const int c = nn_clamp(nn_cell(xyz[3 * (size_t)q + 2], g.mn[2], g.cell), g.dim[2]);
    const int id = (a * g.dim[1] + b) * g.dim[2] + c;
    cell_of[q] = id;
    atomicAdd(&count[id], 1);
}
// Reference points grouped by cell (the order inside a cell is arbitrary: only a minimum is taken over them).
__global__ void k_nn_fill(const float* __restrict__ xyz, int n, const int* __restrict__ cell_of, const int* __restrict__ start,
                          int* __restrict__ cursor, float* __restrict__ sorted) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n || (q % 10) == 0) return;
    const int id = cell_of[q];
    const int slot = start[id] + atomicAdd(&cursor[id], 1);
    sorted[3 * (size_t)slot] = xyz[3 * (size_t)q];
    sorted[3 * (size_t)slot + 1] = xyz[3 * (size_t)q + 1];
    sorted[3 * (size_t)slot + 2] = xyz[3 * (size_t)q + 2];
}
// One thread per query point (every tenth point): cubes of cells of growing radius r around the query's cell until the
// best squared distance is provably final (everything outside the cube is at least (r - 1/4) cells away; the quarter
// cell absorbs the rounding of the cell coordinates).  Squared distance as FLANN's L2 functor forms it: ((dx^2 + dy^2) + dz^2) in float.
__global__ void k_nn_query(const float* __restrict__ xyz, int n, int nref, NNGrid g, const int* __restrict__ start,
                           const int* __restrict__ count, const float* __restrict__ sorted, float* __restrict__ dist_out) {
    const int qi = blockIdx.x * blockDim.x + threadIdx.x;
    if ((long long)qi * 10 >= n) return;
    const float qx = xyz[30 * (size_t)qi], qy = xyz[30 * (size_t)qi + 1], qz = xyz[30 * (size_t)qi + 2];
    float best = 1e30f;
    auto scan = [&](int s, int e) {
        for (int t = s; t < e; t++) {
            const float dx = __fsub_rn(sorted[3 * (size_t)t], qx), dy = __fsub_rn(sorted[3 * (size_t)t + 1], qy),
                        dz = __fsub_rn(sorted[3 * (size_t)t + 2], qz);
            const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
            if (d2 < best) best = d2;
        }
    };
    const float fc0 = floorf(__fdiv_rn(__fsub_rn(qx, g.mn[0]), g.cell)), fc1 = floorf(__fdiv_rn(__fsub_rn(qy, g.mn[1]), g.cell)),
                fc2 = floorf(__fdiv_rn(__fsub_rn(qz, g.mn[2]), g.cell));
    if (!(fabsf(fc0) < 1048576.0f && fabsf(fc1) < 1048576.0f && fabsf(fc2) < 1048576.0f)) {
        scan(0, nref);          // a query absurdly far from the reference points: every point is a candidate
    } else {
        const int c0 = (int)fc0, c1 = (int)fc1, c2 = (int)fc2;
        // the largest radius that can still meet a cell of the grid
        int rmax = 1;
        rmax = max(rmax, max(c0 + 1, g.dim[0] - c0)); rmax = max(rmax, max(c1 + 1, g.dim[1] - c1)); rmax = max(rmax, max(c2 + 1, g.dim[2] - c2));
        // a query outside the grid: no cube of radius below its Chebyshev distance to the grid meets a cell
        int r0 = 1;
        r0 = max(r0, c0 < 0 ? -c0 : c0 - g.dim[0] + 1); r0 = max(r0, c1 < 0 ? -c1 : c1 - g.dim[1] + 1); r0 = max(r0, c2 < 0 ? -c2 : c2 - g.dim[2] + 1);
        for (int r = r0; r <= rmax; r++) {
            const int a_lo = max(c0 - r, 0), a_hi = min(c0 + r, g.dim[0] - 1);
            const int b_lo = max(c1 - r, 0), b_hi = min(c1 + r, g.dim[1] - 1);
            const int c_lo = max(c2 - r, 0), c_hi = min(c2 + r, g.dim[2] - 1);
            for (int a = a_lo; a <= a_hi; a++)
                for (int b = b_lo; b <= b_hi; b++) {
                    const int row = (a * g.dim[1] + b) * g.dim[2];
                    if (r == r0 || abs(a - c0) == r || abs(b - c1) == r) {
                        // a whole z run of cells is one contiguous range of the sorted points
                        if (c_lo <= c_hi) scan(start[row + c_lo], start[row + c_hi] + count[row + c_hi]);
                    } else {                                   // inside the cube's x / y extent: only its two z faces
                        if (c2 - r >= 0 && c2 - r < g.dim[2]) scan(start[row + c2 - r], start[row + c2 - r] + count[row + c2 - r]);
                        if (c2 + r >= 0 && c2 + r < g.dim[2]) scan(start[row + c2 + r], start[row + c2 + r] + count[row + c2 + r]);
                    }
                }
            if (best < 1e30f && sqrtf(best) <= ((float)r - 0.25f) * g.cell) break;
        }
    }
    dist_out[qi] = (float)sqrt((double)best);          // binvoxToPcl.cpp:191
}
// distTotal += dist in query order, in float (binvoxToPcl.cpp:191, voxelize.cpp:27): one thread, the order is the result.
__global__ void k_nn_sum(const float* __restrict__ dist, int nq, float* __restrict__ total) {
    float t = 0.0f;
    int q = 0;
    for (; q + 8 <= nq; q += 8) {
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; u++) v[u] = dist[q + u];
#pragma unroll
        for (int u = 0; u < 8; u++) t = __fadd_rn(t, v[u]);
    }
    for (; q < nq; q++) t = __fadd_rn(t, dist[q]);
    *total = t;
}

int dev_nn_stats(mrb_ctx* ctx, const float* xyz, int n, float* dist_total, int* count_out) {
    *dist_total = 0.0f; *count_out = 0;
    const int nref = n - (n + 9) / 10, nq = (n + 9) / 10;
    if (nref <= 0) return MRB_OK;
    float mn[3], mx[3];
    MRB_TRY(points_minmax(ctx, xyz, n, 1, mn, mx));
    // about eight reference points per cell; at most 2^24 cells
    NNGrid g;
    double ext[3], vol = 1.0;
    for (int a = 0; a < 3; a++) { ext[a] = (double)mx[a] - (double)mn[a]; if (!(ext[a] > 0)) ext[a] = 0; }
    double emax = std::max(ext[0], std::max(ext[1], ext[2]));
    if (!(emax > 0)) emax = 1.0;
    for (int a = 0; a < 3; a++) vol *= std::max(ext[a], emax * 1e-3);
    double cell = cbrt(vol * 8.0 / (double)nref);
    for (;;) {
        double cells = 1.0;
        for (int a = 0; a < 3; a++) cells *= floor(ext[a] / cell) + 1.0;
        if (cells <= 16777216.0) break;
        cell *= 1.26;
    }
    g.cell = (float)cell;
    for (int a = 0; a < 3; a++) { g.mn[a] = mn[a]; g.dim[a] = (int)floor(ext[a] / (double)g.cell) + 2; }
    const size_t ncell = (size_t)g.dim[0] * g.dim[1] * g.dim[2];
    int *cnt = nullptr, *start = nullptr, *cursor = nullptr, *cell_of = nullptr;
    float *sorted = nullptr, *dist = nullptr, *total = nullptr;
    MRB_TRY(arena_alloc(ctx, ncell, &cnt));
    MRB_TRY(arena_alloc(ctx, ncell + 1, &start));
    MRB_TRY(arena_alloc(ctx, ncell, &cursor));
    MRB_TRY(arena_alloc(ctx, (size_t)n, &cell_of));
    MRB_TRY(arena_alloc(ctx, (size_t)nref * 3, &sorted));
    MRB_TRY(arena_alloc(ctx, (size_t)nq, &dist));
    MRB_TRY(arena_alloc(ctx, 1, &total));
    MRB_CUDA(cudaMemsetAsync(cnt, 0, ncell * sizeof(int), ctx->stream));
    MRB_CUDA(cudaMemsetAsync(cursor, 0, ncell * sizeof(int), ctx->stream));
    MRB_LAUNCH(ctx, k_nn_count, cdiv(n, 256), 256, 0, xyz, n, g, cnt, cell_of);
    MRB_TRY(scan_exclusive(ctx, cnt, start, ncell, start + ncell));
    MRB_LAUNCH(ctx, k_nn_fill, cdiv(n, 256), 256, 0, xyz, n, cell_of, start, cursor, sorted);
    MRB_LAUNCH(ctx, k_nn_query, cdiv(nq, 128), 128, 0, xyz, n, nref, g, start, cnt, sorted, dist);
    MRB_LAUNCH(ctx, k_nn_sum, 1, 1, 0, dist, nq, total);
    MRB_CUDA(cudaMemcpyAsync(ctx->h_mail + 40, total, sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    MRB_TRY(sync(ctx));
    memcpy(dist_total, ctx->h_mail + 40, sizeof(float));
    *count_out = nq;
    return MRB_OK;
}

// ---- voxel grid, occupancy, gap fill ---------------------------------------------------------------------------
struct LeafGeom {
    float inv;          // 1 / leaf size (pcl::VoxelGrid::inverse_leaf_size_)
    int minb[3], div[3];
    int pad, dims[3];
};
__global__ void k_leaf_keys(const float* __restrict__ xyz, int n, LeafGeom g, unsigned* __restrict__ key, int* __restrict__ val) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    const int i = (int)floorf(__fmul_rn(xyz[3 * (size_t)q], g.inv)) - g.minb[0];
    const int j = (int)floorf(__fmul_rn(xyz[3 * (size_t)q + 1], g.inv)) - g.minb[1];
    const int k = (int)floorf(__fmul_rn(xyz[3 * (size_t)q + 2], g.inv)) - g.minb[2];
    key[q] = (unsigned)i + (unsigned)g.div[0] * ((unsigned)j + (unsigned)g.div[1] * (unsigned)k);     // VoxelGrid's leaf index (x fastest)
    val[q] = q;
}
// One thread per leaf (the first position of its run in the sorted list): centroid = float sum of its points in cloud
// order / count; the leaf's voxel is marked occupied; min / max of all centroids (grid.cpp:104-109).
__global__ void __launch_bounds__(256) k_leaf_centroids(const float* __restrict__ xyz, const unsigned* __restrict__ key,
                                                        const int* __restrict__ val, int n, LeafGeom g, unsigned char* __restrict__ occ,
                                                        unsigned* __restrict__ mm, int* __restrict__ nleaf) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0u, 0u, 0u};
    bool head = false;
    if (s < n) {
        const unsigned kk = key[s];
        head = s == 0 || key[s - 1] != kk;
        if (head) {
            float sx = 0.0f, sy = 0.0f, sz = 0.0f;
            int m = 0;
            for (int t = s; t < n && key[t] == kk; t++, m++) {
                const size_t p = (size_t)val[t];
                sx = __fadd_rn(sx, xyz[3 * p]); sy = __fadd_rn(sy, xyz[3 * p + 1]); sz = __fadd_rn(sz, xyz[3 * p + 2]);
            }
            const float fm = (float)m;
            const float c[3] = {__fdiv_rn(sx, fm), __fdiv_rn(sy, fm), __fdiv_rn(sz, fm)};
#pragma unroll
            for (int a = 0; a < 3; a++) lo[a] = hi[a] = fenc(c[a]);
            const int i = (int)(kk % (unsigned)g.div[0]), j = (int)((kk / (unsigned)g.div[0]) % (unsigned)g.div[1]),
                      k = (int)(kk / ((unsigned)g.div[0] * (unsigned)g.div[1]));
            occ[((size_t)(i + g.pad) * g.dims[1] + (size_t)(j + g.pad)) * g.dims[2] + (size_t)(k + g.pad)] = 1;
        }
    }
    const unsigned heads = __ballot_sync(0xffffffffu, head);
    if (heads == 0) return;
#pragma unroll
    for (int a = 0; a < 3; a++) {
        lo[a] = __reduce_min_sync(0xffffffffu, lo[a]);
        hi[a] = __reduce_max_sync(0xffffffffu, hi[a]);
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; a++) { atomicMin(&mm[a], lo[a]); atomicMax(&mm[3 + a], hi[a]); }
        atomicAdd(nleaf, __popc(heads));
    }
}
// fillGrid (grid.cpp:213-265): in every z column the gap between the first run of occupied voxels and the next occupied
// voxel is closed when it is narrower than res_factor + 1 (the confidences written there are all > 0, i.e. occupied).
__global__ void k_fill_columns(unsigned char* __restrict__ occ, int ncol, int nz, int max_width) {
    const int col = blockIdx.x * blockDim.x + threadIdx.x;
    if (col >= ncol) return;
    unsigned char* c = occ + (size_t)col * nz;
    int k = 0;
    while (k < nz && !c[k]) k++;            // first occupied voxel
    while (k < nz && c[k]) k++;             // end of the first run = start of the gap
    const int g0 = k;
    while (k < nz && !c[k]) k++;            // next occupied voxel
    if (k >= nz || g0 >= nz || k - g0 >= max_width) return;
    for (int t = g0; t < k; t++) c[t] = 1;
}
// getBinaryVolume (grid.cpp:322-334)
__global__ void k_occ_to_volume(const unsigned char* __restrict__ occ, size_t n, float* __restrict__ vol) {
    for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (size_t)gridDim.x * blockDim.x) vol[q] = occ[q] ? 1.0f : 0.0f;
}

struct VolumeMeta { int dims[3]; float scale[3], shift[3]; int pad, nleaf; };

int dev_voxelize_leaf(mrb_ctx* ctx, const float* xyz, int n, float leaf, int res_factor, float** vol_out, VolumeMeta* meta) {
    *vol_out = nullptr;
    if (n <= 0 || !(leaf > 0.0f)) { set_error("mrb_voxelize_leaf: empty cloud or leaf size <= 0"); return MRB_EINVAL; }
    float mn[3], mx[3];
    MRB_TRY(points_minmax(ctx, xyz, n, 0, mn, mx));
    LeafGeom g;
    g.inv = 1.0f / leaf;
    g.pad = 6;                                                        // grid.cpp:91
    double total = 1.0, cells = 1.0;
    for (int a = 0; a < 3; a++) {
        const double lo = floor((double)(mn[a] * g.inv)), hi = floor((double)(mx[a] * g.inv));
        if (fabs(lo) > 2e9 || fabs(hi) > 2e9) { set_error("mrb_voxelize_leaf: leaf size too small for the cloud's extent"); return MRB_EINVAL; }
        g.minb[a] = (int)lo;
        g.div[a] = (int)hi - g.minb[a] + 1;
        g.dims[a] = g.div[a] + 2 * g.pad;
        cells *= (double)g.div[a];
        total *= (double)g.dims[a];
    }
    if (cells > 4294967295.0 || total > 2147483647.0) { set_error("mrb_voxelize_leaf: more than 2^31 - 1 voxels"); return MRB_EINVAL; }
    const size_t nvox = (size_t)total;
    unsigned *key = nullptr, *key2 = nullptr, *mm = nullptr;
    int *val = nullptr, *val2 = nullptr, *nleaf = nullptr;
    unsigned char* occ = nullptr;
    float* vol = nullptr;
    MRB_TRY(arena_alloc(ctx, (size_t)n, &key));
    MRB_TRY(arena_alloc(ctx, (size_t)n, &key2));
    MRB_TRY(arena_alloc(ctx, (size_t)n, &val));
    MRB_TRY(arena_alloc(ctx, (size_t)n, &val2));
    MRB_TRY(arena_alloc(ctx, 8, &mm));
    MRB_TRY(arena_alloc(ctx, 1, &nleaf));
    MRB_TRY(arena_alloc(ctx, nvox, &occ));
    MRB_TRY(arena_alloc(ctx, nvox, &vol));
    MRB_LAUNCH(ctx, k_leaf_keys, cdiv(n, 256), 256, 0, xyz, n, g, key, val);
    // stable sort by leaf index: inside a leaf the points stay in cloud order
    int bits = 1;
    while (bits < 32 && (double)(1ull << bits) < cells) bits++;
    size_t tmp_bytes = 0;
    MRB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, key, key2, val, val2, n, 0, bits, ctx->stream));
    char* tmp = nullptr;
    MRB_TRY(arena_alloc(ctx, tmp_bytes + 16, &tmp));
    MRB_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, key, key2, val, val2, n, 0, bits, ctx->stream));
    MRB_CUDA(cudaMemsetAsync(mm, 0xFF, 3 * sizeof(unsigned), ctx->stream));
    MRB_CUDA(cudaMemsetAsync(mm + 3, 0, 3 * sizeof(unsigned), ctx->stream));
    MRB_CUDA(cudaMemsetAsync(nleaf, 0, sizeof(int), ctx->stream));
    MRB_CUDA(cudaMemsetAsync(occ, 0, nvox, ctx->stream));
    MRB_LAUNCH(ctx, k_leaf_centroids, cdiv(n, 256), 256, 0, xyz, key2, val2, n, g, occ, mm, nleaf);
    if (res_factor > 0)
        MRB_LAUNCH(ctx, k_fill_columns, cdiv((size_t)g.dims[0] * g.dims[1], 128), 128, 0, occ, g.dims[0] * g.dims[1], g.dims[2], res_factor + 1);
    MRB_LAUNCH(ctx, k_occ_to_volume, std::min<unsigned>(cdiv(nvox, 256), (unsigned)ctx->sm_count * 16), 256, 0, occ, nvox, vol);
    unsigned h[6];
    MRB_CUDA(cudaMemcpyAsync(h, mm, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    MRB_CUDA(cudaMemcpyAsync(ctx->h_mail + 41, nleaf, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    MRB_TRY(sync(ctx));
    for (int a = 0; a < 3; a++) {
        meta->dims[a] = g.dims[a];
        // grid.cpp:94-96,104-109: shift / high_vals start at +-10000 and follow the centroids
        const float lo = std::min(10000.0f, fdec(h[a])), hi = std::max(-10000.0f, fdec(h[3 + a]));
        meta->shift[a] = lo;
        meta->scale[a] = (hi - lo) / ((float)g.div[a]);           // grid.cpp:118-120
    }
    meta->pad = g.pad;
    meta->nleaf = ctx->h_mail[41];
    *vol_out = vol;
    return MRB_OK;
}

// device -> caller: 0 malloc'd host array, 1 the arena (valid until the next call), 2 the context's pinned staging
static int export_floats(mrb_ctx* ctx, const float* dev, size_t count, int device_out, float** out) {
    if (device_out == 1) { *out = (float*)dev; return sync(ctx); }
    char* pin = nullptr;
    MRB_TRY(pinned_reserve(ctx, (count > 0 ? count : 1) * sizeof(float), &pin));
    if (count > 0) MRB_CUDA(cudaMemcpyAsync(pin, dev, count * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    MRB_TRY(sync(ctx));
    if (device_out == 2) { *out = (float*)pin; return MRB_OK; }
    float* h = (float*)malloc((count > 0 ? count : 1) * sizeof(float));
    if (!h) { set_error("front end: host malloc failed"); return MRB_ENOMEM; }
    if (count > 0) memcpy(h, pin, count * sizeof(float));
    *out = h;
    return MRB_OK;
}

static int export_volume(mrb_ctx* ctx, const float* vol, const VolumeMeta& m, int device_out, mrb_volume* out) {
    memset(out, 0, sizeof(*out));
    for (int a = 0; a < 3; a++) { out->dims[a] = m.dims[a]; out->scale[a] = m.scale[a]; out->shift[a] = m.shift[a]; }
    out->pad = m.pad;
    out->occupied_leaves = m.nleaf;
    out->on_device = device_out == 2 ? 2 : (device_out ? 1 : 0);
    return export_floats(ctx, vol, (size_t)m.dims[0] * m.dims[1] * m.dims[2], out->on_device, &out->voxels);
}

}  // namespace mrb

using namespace mrb;

extern "C" {

int mrb_binvox_to_points(mrb_ctx* ctx, const unsigned char* voxels, const int dwh[3], const float xform[4], int res_factor,
                         int device_out, mrb_points* out) {
    MRB_REQUIRE(ctx && voxels && dwh && xform && out, "bad argument");
    MRB_REQUIRE(res_factor >= 1, "Resolution factor must be at least 1");        // binvoxToPcl.cpp:106-109
    MRB_REQUIRE(dwh[0] >= 1 && dwh[1] >= 1 && dwh[2] >= 1, "bad binvox dimensions");
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    const unsigned char* dv = nullptr;
    MRB_TRY(stage_in(ctx, voxels, (size_t)dwh[0] * dwh[1] * dwh[2], &dv));
    float* xyz = nullptr;
    int n = 0;
    MRB_TRY(dev_binvox_points(ctx, dv, dwh, xform, res_factor, &xyz, &n));
    out->n = n;
    out->on_device = device_out == 2 ? 2 : (device_out ? 1 : 0);
    return export_floats(ctx, xyz, (size_t)n * 3, out->on_device, &out->xyz);
}

void mrb_points_free(mrb_points* p) {
    if (!p) return;
    if (p->on_device == 0) free(p->xyz);
    p->xyz = nullptr; p->n = 0;
}

int mrb_nn_stats(mrb_ctx* ctx, const float* xyz, int n, float* dist_total, int* count) {
    MRB_REQUIRE(ctx && dist_total && count && n >= 0 && (n == 0 || xyz), "bad argument");
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    const float* d = nullptr;
    MRB_TRY(stage_in(ctx, xyz, (size_t)n * 3, &d));
    return dev_nn_stats(ctx, d, n, dist_total, count);
}

int mrb_voxelize_leaf(mrb_ctx* ctx, const float* xyz, int n, float leaf, int res_factor, int device_out, mrb_volume* out) {
    MRB_REQUIRE(ctx && xyz && out && n > 0, "bad argument");
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    const float* d = nullptr;
    MRB_TRY(stage_in(ctx, xyz, (size_t)n * 3, &d));
    float* vol = nullptr;
    VolumeMeta m;
    MRB_TRY(dev_voxelize_leaf(ctx, d, n, leaf, res_factor, &vol, &m));
    return export_volume(ctx, vol, m, device_out, out);
}

void mrb_volume_free(mrb_volume* v) {
    if (!v) return;
    if (v->on_device == 0) free(v->voxels);
    v->voxels = nullptr;
}

int mrb_front_end(mrb_ctx* ctx, const unsigned char* voxels, const int dwh[3], const float xform[4], const float* observed, int n_observed,
                  int res_factor, int device_out, mrb_volume* out, mrb_front_info* info) {
    MRB_REQUIRE(ctx && voxels && dwh && xform && out && n_observed >= 0 && (n_observed == 0 || observed), "bad argument");
    MRB_REQUIRE(dwh[0] >= 1 && dwh[1] >= 1 && dwh[2] >= 1, "bad binvox dimensions");
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    mrb_front_info fi;
    memset(&fi, 0, sizeof(fi));
    const unsigned char* dv = nullptr;
    const float* dobs = nullptr;
    MRB_TRY(stage_in(ctx, voxels, (size_t)dwh[0] * dwh[1] * dwh[2], &dv));
    MRB_TRY(stage_in(ctx, observed, (size_t)n_observed * 3, &dobs));
    int res = res_factor;
    if (res < 1) {
        // getResolutionFactor (binvoxToPcl.cpp:203-219): 3/8 of the observed density against the density of the completion at res 1
        MRB_TRY(dev_nn_stats(ctx, dobs, n_observed, &fi.observed_dist_total, &fi.observed_count));
        const float other = 3.0 * ((float)fi.observed_count / fi.observed_dist_total) / 8.0;
        float* init = nullptr;
        int ninit = 0;
        MRB_TRY(dev_binvox_points(ctx, dv, dwh, xform, 1, &init, &ninit));
        MRB_TRY(dev_nn_stats(ctx, init, ninit, &fi.completed_dist_total, &fi.completed_count));
        const float density = (float)fi.completed_count / fi.completed_dist_total;
        fi.observed_density = other;
        fi.completed_density = density;
        res = density > other ? 1 : (int)ceil(other / density);
        if (res < 1) { set_error("mrb_front_end: the density ratio gives no resolution factor (empty cloud?)"); return MRB_EINVAL; }
    }
    fi.res_factor = res;
    float* pred = nullptr;
    int npred = 0;
    MRB_TRY(dev_binvox_points(ctx, dv, dwh, xform, res, &pred, &npred));
    fi.n_predicted = npred;
    // full_confidence + assign_confidence with ConstConf(1) (mesh_reconstruction.cpp:107-112): observed points first, then the completion
    const long long nall = (long long)n_observed + npred;
    if (nall <= 0 || nall > 2147483647LL / 3) { set_error("mrb_front_end: empty or too large cloud"); return MRB_EINVAL; }
    float* all = nullptr;
    MRB_TRY(arena_alloc(ctx, (size_t)nall * 3, &all));
    if (n_observed > 0) MRB_CUDA(cudaMemcpyAsync(all, dobs, (size_t)n_observed * 3 * sizeof(float), cudaMemcpyDeviceToDevice, ctx->stream));
    if (npred > 0) MRB_CUDA(cudaMemcpyAsync(all + (size_t)n_observed * 3, pred, (size_t)npred * 3 * sizeof(float), cudaMemcpyDeviceToDevice, ctx->stream));
    fi.n_cloud = (int)nall;
    // getResolution (voxelize.cpp:6-32): leaf size = twice the mean nearest-neighbour distance
    float dt = 0.0f;
    int cnt = 0;
    MRB_TRY(dev_nn_stats(ctx, all, (int)nall, &dt, &cnt));
    fi.leaf = 2.0 * dt / (float)cnt;
    float* vol = nullptr;
    VolumeMeta m;
    MRB_TRY(dev_voxelize_leaf(ctx, all, (int)nall, fi.leaf, res, &vol, &m));
    if (info) *info = fi;
    return export_volume(ctx, vol, m, device_out, out);
}

}  // extern "C"

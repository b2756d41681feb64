// mrb_features.cu -- feature pre-pass of --feature-detection: surface voxels, surface normals, feature map.
//
// Replaces getSurface / isSurface (distance_fields/dfields.cpp:216-232, 322-338), getSurfaceNormals with
// getCovariance / getSurfaceNeighbors / getCentroid / getNormalVector / orientNormal (dfields.cpp:199-349)
// and getFeatureMap (dfields.cpp:447-555).  One thread per surface voxel / per cell; all arithmetic in
// the reference's loop order with separately rounded fp32 operations (-fmad=false).  The 3x3 eigen-solve
// is the deterministic Jacobi iteration of include/mrb_eig3.h (the reference calls Eigen there: the
// feature sub-path is parity-unpinned against the reference and bit-exact against our oracle).
#include "mrb_internal.cuh"
#include "mrb_compact.cuh"
#include "../../include/mrb_eig3.h"

namespace mrb {

// isSurface (dfields.cpp:216-232): value == 1 and a 0 in the 3x3x3 window.  -1: the window leaves the grid.
__device__ __forceinline__ int is_surface(const float* __restrict__ g, const Dims& d, int i, int j, int k) {
    if (g[d.at(i, j, k)] != 1.0f) return 0;
    if (i < 1 || j < 1 || k < 1 || i > d.nx - 2 || j > d.ny - 2 || k > d.nz - 2) return -1;
    for (int a = i - 1; a <= i + 1; a++)
        for (int b = j - 1; b <= j + 1; b++)
            for (int c = k - 1; c <= k + 1; c++)
                if (g[d.at(a, b, c)] == 0.0f) return 1;
    return 0;
}

struct SurfacePred {
    const float* g;
    Dims d;
    int* err;
    __device__ bool operator()(size_t idx) const {
        int z = (int)(idx % d.nz);
        size_t t = idx / d.nz;
        int y = (int)(t % d.ny);
        int x = (int)(t / d.ny);
        int s = is_surface(g, d, x, y, z);
        if (s < 0) { *err = 1; return false; }
        return s == 1;
    }
};

__global__ void k_pos_to_linear_f(const int* __restrict__ pos, int n, Dims d, int* __restrict__ lin) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int p = pos[i];
    int z = p % d.nz;
    int t = p / d.nz;
    int y = t % d.ny;
    int x = t / d.ny;
    lin[i] = z * d.nx * d.ny + y * d.nx + x;
}

__global__ void __launch_bounds__(128) k_surface_normals(const float* __restrict__ vol, Dims d, const int* __restrict__ surface,
                                                         int n, float* __restrict__ normals, int* __restrict__ err) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    int lin = surface[s];
    int z = lin / (d.nx * d.ny);
    int r = lin % (d.nx * d.ny);
    int y = r / d.nx;
    int x = r % d.nx;
    if (lin < 0 || (size_t)lin >= d.count() || x < 2 || y < 2 || z < 2 || x > d.nx - 3 || y > d.ny - 3 || z > d.nz - 3) {
        *err = 1;
        return;
    }
    // getSurfaceNeighbors + getCentroid
    unsigned mask = 0;          // bit (a*9+b*3+c) set for surface neighbours, (i,j,k) order
    int cnt = 0;
    float c0 = 0.0f, c1 = 0.0f, c2 = 0.0f;
    for (int a = 0; a < 3; a++)
        for (int b = 0; b < 3; b++)
            for (int c = 0; c < 3; c++) {
                int i = x - 1 + a, j = y - 1 + b, k = z - 1 + c;
                if (vol[d.at(i, j, k)] > 0 && is_surface(vol, d, i, j, k) == 1) {
                    mask |= 1u << (a * 9 + b * 3 + c);
                    cnt++;
                    c0 += (float)i; c1 += (float)j; c2 += (float)k;
                }
            }
    c0 = c0 / (float)cnt; c1 = c1 / (float)cnt; c2 = c2 / (float)cnt;
    // getCovariance
    float cov[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int a = 0; a < 3; a++)
        for (int b = 0; b < 3; b++)
            for (int c = 0; c < 3; c++) {
                if (!(mask & (1u << (a * 9 + b * 3 + c)))) continue;
                float v[3] = {(float)(x - 1 + a) - c0, (float)(y - 1 + b) - c1, (float)(z - 1 + c) - c2};
                for (int rr = 0; rr < 3; rr++)
                    for (int cc = 0; cc < 3; cc++) cov[rr * 3 + cc] = cov[rr * 3 + cc] + v[rr] * v[cc];
            }
    for (int e = 0; e < 9; e++) cov[e] = cov[e] / (float)cnt;
    float nrm[3];
    mrb_normal_from_covariance(cov, nrm);
    // orientNormal
    float o0 = 0.0f, o1 = 0.0f, o2 = 0.0f;
    int on = 0;
    for (int i = x - 1; i <= x + 1; i++)
        for (int j = y - 1; j <= y + 1; j++)
            for (int k = z - 1; k <= z + 1; k++)
                if (vol[d.at(i, j, k)] > 0) { o0 += (float)i; o1 += (float)j; o2 += (float)k; on++; }
    o0 = o0 / (float)on; o1 = o1 / (float)on; o2 = o2 / (float)on;
    o0 -= (float)x; o1 -= (float)y; o2 -= (float)z;
    float prod = nrm[0] * o0 + nrm[1] * o1 + nrm[2] * o2;
    if (prod > 0.0f) { nrm[0] = -nrm[0]; nrm[1] = -nrm[1]; nrm[2] = -nrm[2]; }
    normals[3 * (size_t)s] = nrm[0]; normals[3 * (size_t)s + 1] = nrm[1]; normals[3 * (size_t)s + 2] = nrm[2];
}

__global__ void k_fill_f(float* __restrict__ p, size_t n, float v) {
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (size_t)gridDim.x * blockDim.x) p[idx] = v;
}

// getFeatureMap (dfields.cpp:491-552): one thread per cell; marking is idempotent, so the write order is free.
__global__ void __launch_bounds__(256) k_feature_map(const float* __restrict__ vol, const float* __restrict__ smap,
                                                     const float* __restrict__ normals, int nnormals, Dims d, float threshold,
                                                     float* __restrict__ out, int* __restrict__ err) {
    const int cx = d.nx - 1, cy = d.ny - 1, cz = d.nz - 1;
    const size_t ncell = (size_t)cx * cy * cz;
    for (size_t cell = (size_t)blockIdx.x * blockDim.x + threadIdx.x; cell < ncell; cell += (size_t)gridDim.x * blockDim.x) {
        int k = (int)(cell % cz);
        size_t t = cell / cz;
        int j = (int)(t % cy);
        int i = (int)(t / cy);
        int cubeindex = 0;
        size_t a8[8];
        for (int n = 0; n < 8; n++) {
            a8[n] = d.at(i + ((((n % 4) == 1) || ((n % 4) == 2)) ? 1 : 0), j + (n % 4) / 2, k + n / 4);
            if ((int)vol[a8[n]] < 1) cubeindex |= (1 << n);
        }
        if (cubeindex == 0 || cubeindex == 255) continue;
        float mn = 5.0f;
        int idx8[8];
        for (int n = 0; n < 8; n++) idx8[n] = (int)smap[a8[n]];
        for (int n = 0; n < 7; n++)
            for (int m = n + 1; m < 8; m++) {
                int i1 = idx8[n], i2 = idx8[m];
                if (i1 == -1 || i2 == -1) continue;
                if (i1 < 0 || i2 < 0 || i1 >= nnormals || i2 >= nnormals) { *err = 1; continue; }
                const float* a = normals + 3 * (size_t)i1;
                const float* b = normals + 3 * (size_t)i2;
                float angle = a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
                if (angle < mn) mn = angle;
            }
        if (mn >= threshold) continue;
        for (int a = 0; a < 2; a++)
            for (int b = 0; b < 2; b++)
                for (int c = 0; c < 2; c++) {
                    size_t q = d.at(i + a, j + b, k + c);
                    if (smap[q] > -1.0f) out[q] = 1.0f;
                }
    }
}

}  // namespace mrb

using namespace mrb;

extern "C" {

int mrb_get_surface(mrb_ctx* ctx, const float* volume, const int dims[3], int* idx_out, int capacity, int* n_out) {
    MRB_REQUIRE(ctx && volume && dims && n_out, "NULL argument");
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    Dims d = mkdims(dims);
    const float* dv = nullptr;
    MRB_TRY(stage_in(ctx, volume, d.count(), &dv));
    int* err = ctx->d_mail + 48;
    MRB_CUDA(cudaMemsetAsync(err, 0, sizeof(int), ctx->stream));
    int* pos = nullptr;
    int n = 0;
    MRB_TRY(compact(ctx, SurfacePred{dv, d, err}, d.count(), &pos, &n));
    int e = 0;
    MRB_TRY(read_int(ctx, err, &e));
    if (e) { set_error("mrb_get_surface: occupied voxel on the grid border (the reference exits, grid.cpp:20-37)"); return MRB_EBORDER; }
    *n_out = n;
    if (!idx_out || n == 0) return sync(ctx);
    if (n > capacity) { set_error("mrb_get_surface: capacity %d < %d", capacity, n); return MRB_ECAPACITY; }
    Out<int> o;
    MRB_TRY(o.init(ctx, idx_out, (size_t)n));
    MRB_LAUNCH(ctx, k_pos_to_linear_f, cdiv(n, 256), 256, 0, pos, n, d, o.dev);
    MRB_TRY(o.commit(ctx));
    return sync(ctx);
}

int mrb_surface_normals(mrb_ctx* ctx, const float* volume, const int dims[3], const int* surface, int n, float* normals_out) {
    MRB_REQUIRE(ctx && volume && dims && n >= 0 && (n == 0 || (surface && normals_out)), "bad argument");
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    if (n == 0) return MRB_OK;
    Dims d = mkdims(dims);
    const float* dv = nullptr;
    const int* ds = nullptr;
    MRB_TRY(stage_in(ctx, volume, d.count(), &dv));
    MRB_TRY(stage_in(ctx, surface, (size_t)n, &ds));
    Out<float> o;
    MRB_TRY(o.init(ctx, normals_out, (size_t)n * 3));
    int* err = ctx->d_mail + 48;
    MRB_CUDA(cudaMemsetAsync(err, 0, sizeof(int), ctx->stream));
    MRB_LAUNCH(ctx, k_surface_normals, cdiv(n, 128), 128, 0, dv, d, ds, n, o.dev, err);
    int e = 0;
    MRB_TRY(read_int(ctx, err, &e));
    if (e) { set_error("mrb_surface_normals: surface voxel within 2 voxels of the grid border (the reference exits)"); return MRB_EBORDER; }
    MRB_TRY(o.commit(ctx));
    return sync(ctx);
}

int mrb_feature_map(mrb_ctx* ctx, const float* volume, const float* surface_map, const float* normals, int nnormals,
                    const int dims[3], float threshold, float* feature_map_out) {
    MRB_REQUIRE(ctx && volume && surface_map && dims && feature_map_out && nnormals >= 0 && (nnormals == 0 || normals), "bad argument");
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    Dims d = mkdims(dims);
    const float *dv = nullptr, *dm = nullptr, *dn = nullptr;
    MRB_TRY(stage_in(ctx, volume, d.count(), &dv));
    MRB_TRY(stage_in(ctx, surface_map, d.count(), &dm));
    MRB_TRY(stage_in(ctx, normals, (size_t)(nnormals > 0 ? nnormals : 1) * 3, &dn));
    Out<float> o;
    MRB_TRY(o.init(ctx, feature_map_out, d.count()));
    size_t want = (d.count() + 255) / 256, cap = (size_t)ctx->sm_count * 16;
    unsigned g = (unsigned)(want < cap ? want : cap);
    MRB_LAUNCH(ctx, k_fill_f, g, 256, 0, o.dev, d.count(), 0.0f);
    int* err = ctx->d_mail + 48;
    MRB_CUDA(cudaMemsetAsync(err, 0, sizeof(int), ctx->stream));
    if (d.nx > 1 && d.ny > 1 && d.nz > 1) MRB_LAUNCH(ctx, k_feature_map, g, 256, 0, dv, dm, dn, nnormals, d, threshold, o.dev, err);
    int e = 0;
    MRB_TRY(read_int(ctx, err, &e));
    if (e) { set_error("mrb_feature_map: surface map refers to a normal index outside [0, nnormals)"); return MRB_EINVAL; }
    MRB_TRY(o.commit(ctx));
    return sync(ctx);
}

}  // extern "C"

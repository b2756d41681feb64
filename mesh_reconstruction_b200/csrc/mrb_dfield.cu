// mrb_dfield.cu -- stage 1: perimeter mask, exact separable squared EDT, sqrt.
//
// Replaces fastPerim / getsqdist / getsqrt (distance_fields/dfields.cpp:7-99,182-192) and the
// reference's naive kernels fastPerimGPU / sqdistpt1-3 / sqrtgpu (dfields.cu:62-159).
//
// Data layout: dense [x][y][z] arrays, z contiguous.  Every kernel maps consecutive lanes to
// consecutive z, so every global access of a warp is one or two contiguous 128-byte lines.
//
// EDT: result = min(100000, exact squared Euclidean distance to the nearest zero voxel); all
// intermediates are integers (< 2^24, so the reference's fp32/int mix is exact):
//   pass X  one thread per (y,z) line, forward + backward running distance -> u16 d1 (clamped
//           at 317: 317^2 > 100000, so a clamped site can never win below the 100000 cap);
//   pass Y  one thread per voxel, exact outward search over y' = y +- r while r^2 < best;
//   pass Z  same along z, optionally fused with the correctly rounded sqrt.
// The outward search is exact (a candidate at offset r costs at least r^2) and its cost is
// proportional to the distance actually found instead of the reference's full-line scan.
#include "mrb_internal.cuh"

namespace mrb {

static const int kClamp = MRB_SQDIST_CLAMP;
static const int kD1Clamp = 317;

// ---- perimeter --------------------------------------------------------------------------------
// out = 1 unless a voxel of the border-clamped 3x3x3 window differs from the centre.
// MaskOut: write a u8 "is zero site" mask (perim == 0) instead of the float grid.
template <bool MaskOut>
__global__ void __launch_bounds__(256) k_perim(const float* __restrict__ vol, Dims d, float* __restrict__ out,
                                               uint8_t* __restrict__ zmask) {
    size_t n = d.count();
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (size_t)gridDim.x * blockDim.x) {
        int z = (int)(idx % d.nz);
        size_t t = idx / d.nz;
        int y = (int)(t % d.ny);
        int x = (int)(t / d.ny);
        float c = __ldg(vol + idx);
        int x0 = x > 0 ? x - 1 : x, x1 = x < d.nx - 1 ? x + 1 : x;
        int y0 = y > 0 ? y - 1 : y, y1 = y < d.ny - 1 ? y + 1 : y;
        int z0 = z > 0 ? z - 1 : z, z1 = z < d.nz - 1 ? z + 1 : z;
        bool same = true;
        for (int a = x0; a <= x1; a++)
            for (int b = y0; b <= y1; b++) {
                const float* row = vol + ((size_t)a * d.ny + b) * d.nz;
                for (int e = z0; e <= z1; e++) same = same && (__ldg(row + e) == c);
            }
        if (MaskOut) zmask[idx] = same ? 0 : 1;
        else out[idx] = same ? 1.0f : 0.0f;
    }
}

// ---- EDT pass X -------------------------------------------------------------------------------
template <typename SrcT>
__device__ __forceinline__ bool is_site(const SrcT* p, size_t i);
template <>
__device__ __forceinline__ bool is_site<float>(const float* p, size_t i) { return __ldg(p + i) == 0.0f; }
template <>
__device__ __forceinline__ bool is_site<uint8_t>(const uint8_t* p, size_t i) { return __ldg(p + i) != 0; }

template <typename SrcT>
__global__ void __launch_bounds__(128) k_edt_x(const SrcT* __restrict__ src, Dims d, uint16_t* __restrict__ d1) {
    size_t plane = (size_t)d.ny * d.nz;
    size_t line = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (line >= plane) return;
    int dist = kD1Clamp;
    for (int x = 0; x < d.nx; x++) {
        size_t i = (size_t)x * plane + line;
        dist = is_site<SrcT>(src, i) ? 0 : min(dist + 1, kD1Clamp);
        d1[i] = (uint16_t)dist;
    }
    dist = kD1Clamp;
    for (int x = d.nx - 1; x >= 0; x--) {
        size_t i = (size_t)x * plane + line;
        int f = d1[i];
        dist = (f == 0) ? 0 : min(dist + 1, kD1Clamp);
        if (dist < f) d1[i] = (uint16_t)dist;
    }
}

// ---- EDT pass Y -------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_edt_y(const uint16_t* __restrict__ d1, Dims d, int* __restrict__ g2) {
    size_t n = d.count();
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (size_t)gridDim.x * blockDim.x) {
        int y = (int)((idx / d.nz) % d.ny);
        int c = __ldg(d1 + idx);
        int best = min(c * c, kClamp);
        int rmax = max(y, d.ny - 1 - y);
        for (int r = 1; r <= rmax; r++) {
            int r2 = r * r;
            if (r2 >= best) break;
            if (y - r >= 0) { int v = __ldg(d1 + idx - (size_t)r * d.nz); best = min(best, v * v + r2); }
            if (y + r < d.ny) { int v = __ldg(d1 + idx + (size_t)r * d.nz); best = min(best, v * v + r2); }
        }
        g2[idx] = best;
    }
}

// ---- EDT pass Z (+ optional sqrt) ---------------------------------------------------------------
template <bool Sqrt>
__global__ void __launch_bounds__(256) k_edt_z(const int* __restrict__ g2, Dims d, float* __restrict__ out) {
    size_t n = d.count();
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (size_t)gridDim.x * blockDim.x) {
        int z = (int)(idx % d.nz);
        int best = min(__ldg(g2 + idx), kClamp);
        int rmax = max(z, d.nz - 1 - z);
        for (int r = 1; r <= rmax; r++) {
            int r2 = r * r;
            if (r2 >= best) break;
            if (z - r >= 0) best = min(best, __ldg(g2 + idx - r) + r2);
            if (z + r < d.nz) best = min(best, __ldg(g2 + idx + r) + r2);
        }
        float f = (float)best;
        out[idx] = Sqrt ? __fsqrt_rn(f) : f;
    }
}

__global__ void __launch_bounds__(256) k_sqrt(const float* __restrict__ in, size_t n, float* __restrict__ out) {
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (size_t)gridDim.x * blockDim.x)
        out[idx] = __fsqrt_rn(in[idx]);   // == (float)sqrt((double)v): double rounding is innocuous for sqrt
}

static unsigned grid_for(mrb_ctx* ctx, size_t n, int block) {
    size_t want = (n + block - 1) / block;
    size_t cap = (size_t)ctx->sm_count * 16;
    return (unsigned)(want < cap ? want : cap);
}

int dev_fast_perim(mrb_ctx* ctx, const float* vol, Dims d, float* out) {
    MRB_LAUNCH(ctx, (k_perim<false>), grid_for(ctx, d.count(), 256), 256, 0, vol, d, out, (uint8_t*)nullptr);
    return MRB_OK;
}

template <typename SrcT>
static int edt_from(mrb_ctx* ctx, const SrcT* src, Dims d, float* out, bool fuse_sqrt) {
    size_t n = d.count();
    uint16_t* d1 = nullptr;
    int* g2 = nullptr;
    MRB_TRY(arena_alloc(ctx, n, &d1));
    MRB_TRY(arena_alloc(ctx, n, &g2));
    size_t plane = (size_t)d.ny * d.nz;
    MRB_LAUNCH(ctx, (k_edt_x<SrcT>), cdiv(plane, 128), 128, 0, src, d, d1);
    MRB_LAUNCH(ctx, k_edt_y, grid_for(ctx, n, 256), 256, 0, d1, d, g2);
    if (fuse_sqrt) MRB_LAUNCH(ctx, (k_edt_z<true>), grid_for(ctx, n, 256), 256, 0, g2, d, out);
    else MRB_LAUNCH(ctx, (k_edt_z<false>), grid_for(ctx, n, 256), 256, 0, g2, d, out);
    return MRB_OK;
}

int dev_sqdist(mrb_ctx* ctx, const float* in, Dims d, float* out, bool fuse_sqrt) {
    return edt_from<float>(ctx, in, d, out, fuse_sqrt);
}

int dev_sqrt(mrb_ctx* ctx, const float* in, size_t n, float* out) {
    MRB_LAUNCH(ctx, k_sqrt, grid_for(ctx, n, 256), 256, 0, in, n, out);
    return MRB_OK;
}

int dev_dfield(mrb_ctx* ctx, const float* vol, Dims d, float* margin) {
    uint8_t* zmask = nullptr;
    MRB_TRY(arena_alloc(ctx, d.count(), &zmask));
    MRB_LAUNCH(ctx, (k_perim<true>), grid_for(ctx, d.count(), 256), 256, 0, vol, d, (float*)nullptr, zmask);
    return edt_from<uint8_t>(ctx, zmask, d, margin, true);
}

}  // namespace mrb

// ---- C-ABI ------------------------------------------------------------------------------------
using namespace mrb;

static int check_dims(const int* dims) {
    if (!dims || dims[0] < 1 || dims[1] < 1 || dims[2] < 1) { set_error("bad dims"); return MRB_EINVAL; }
    if ((size_t)dims[0] * dims[1] * dims[2] > (size_t)0x7fffffff) { set_error("grid too large (> 2^31-1 voxels)"); return MRB_EINVAL; }
    return MRB_OK;
}

template <typename F>
static int grid_to_grid(mrb_ctx* ctx, const float* in, const int* dims, float* out, F body) {
    MRB_REQUIRE(ctx && in && out, "NULL argument");
    MRB_TRY(check_dims(dims));
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    Dims d = mkdims(dims);
    const float* din = nullptr;
    MRB_TRY(stage_in(ctx, in, d.count(), &din));
    Out<float> o;
    MRB_TRY(o.init(ctx, out, d.count()));
    MRB_TRY(body(din, d, o.dev));
    MRB_TRY(o.commit(ctx));
    return sync(ctx);
}

extern "C" {

int mrb_fast_perim(mrb_ctx* ctx, const float* volume, const int dims[3], float* out) {
    return grid_to_grid(ctx, volume, dims, out, [&](const float* i, Dims d, float* o) { return dev_fast_perim(ctx, i, d, o); });
}
int mrb_sqdist(mrb_ctx* ctx, const float* in, const int dims[3], float* out) {
    return grid_to_grid(ctx, in, dims, out, [&](const float* i, Dims d, float* o) { return dev_sqdist(ctx, i, d, o, false); });
}
int mrb_sqrt(mrb_ctx* ctx, const float* in, const int dims[3], float* out) {
    return grid_to_grid(ctx, in, dims, out, [&](const float* i, Dims d, float* o) { return dev_sqrt(ctx, i, d.count(), o); });
}
int mrb_dfield(mrb_ctx* ctx, const float* volume, const int dims[3], float* margin) {
    return grid_to_grid(ctx, volume, dims, margin, [&](const float* i, Dims d, float* o) { return dev_dfield(ctx, i, d, o); });
}

}  // extern "C"

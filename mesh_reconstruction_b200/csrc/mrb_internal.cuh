// mrb_internal.cuh -- shared internals of libmrb.so (context, arena, staging, launch helpers).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>

#include "../../include/mrb.h"

namespace mrb {

void set_error(const char* fmt, ...);

#define MRB_CUDA(call)                                                                         \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess) {                                                               \
            ::mrb::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            return MRB_ECUDA;                                                                  \
        }                                                                                      \
    } while (0)

#define MRB_TRY(call)                  \
    do {                               \
        int rc_ = (call);              \
        if (rc_ != MRB_OK) return rc_; \
    } while (0)

#define MRB_REQUIRE(cond, msg)                  \
    do {                                        \
        if (!(cond)) {                          \
            ::mrb::set_error("%s: %s", __func__, msg); \
            return MRB_EINVAL;                  \
        }                                       \
    } while (0)

struct Dims {
    int nx, ny, nz;
    __host__ __device__ size_t count() const { return (size_t)nx * ny * nz; }
    __host__ __device__ size_t at(int x, int y, int z) const { return ((size_t)x * ny + y) * nz + z; }
};
inline Dims mkdims(const int* d) { return Dims{d[0], d[1], d[2]}; }

// Grow-only device arena.  Steady state: one cudaMalloc'd block, bump allocation, reset per call.
struct Arena {
    struct Block { char* base; size_t cap; size_t used; };
    std::vector<Block> blocks;
    size_t high_water = 0;      // bytes requested during the current call
    int alloc(size_t bytes, void** out);
    int reset();                // start of a top-level call; consolidates to one block if needed
    void release();
};

}  // namespace mrb

struct mrb_ctx {
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    mrb::Arena arena;
    uint64_t launches = 0;
    // small pinned host mailbox for scalar read-backs
    int* h_mail = nullptr;      // 64 ints, pinned
    int* d_mail = nullptr;      // 64 ints, device
    int* d_bar = nullptr;       // 64 ints, device: grid-barrier words (two 128-byte lines)
    char* h_pinned = nullptr;   // grow-only pinned staging buffer for device->host results
    size_t h_pinned_cap = 0;
    // mrb_reconstruct: the kernel that writes F also packs the "F < isolevel" bits marching cubes starts from
    // (one read of F less).  want_sign_bits / sign_iso are set by the caller of stages 1-3; sign_bits / sign_F by the
    // writer; consumed (and cleared) by classify_cells when F and the isolevel match.
    // grow-only device buffers outside the arena (which every entry point resets): intermediates of calls that are
    // composed from several entry points (mrb_reconstruct with features)
    char* side[2] = {nullptr, nullptr};
    size_t side_cap[2] = {0, 0};
    // arenas of destroyed plans (mrb_qp_plan_destroy), reused by the next mrb_qp_plan_create on this context: a plan per
    // call would otherwise pay cudaMalloc + cudaFree (about 2 ms for a 256^3 grid) every time
    std::vector<mrb::Arena> plan_arenas;
    bool want_sign_bits = false;
    float sign_iso = 0.0f;
    const float* sign_F = nullptr;
    uint32_t* sign_bits = nullptr;
};

namespace mrb {

// Pinned host staging of at least `bytes` (grow-only; contents valid until the next reserve).
int pinned_reserve(mrb_ctx* ctx, size_t bytes, char** out);
// Device side buffer `which` (0 / 1) of at least `bytes` (grow-only; contents valid until the next reserve of the same buffer).
int side_reserve(mrb_ctx* ctx, int which, size_t bytes, char** out);

// True when p is device-accessible memory (device or managed).
bool is_device_ptr(const void* p);

template <typename T>
int arena_alloc(mrb_ctx* ctx, size_t count, T** out) {
    void* p = nullptr;
    int rc = ctx->arena.alloc(count * sizeof(T), &p);
    *out = (T*)p;
    return rc;
}

// Input staging: device pointer for `src` (copied to the arena when src is a host pointer).
template <typename T>
int stage_in(mrb_ctx* ctx, const T* src, size_t count, const T** dev) {
    if (src == nullptr) { *dev = nullptr; return MRB_OK; }
    if (is_device_ptr(src)) { *dev = src; return MRB_OK; }
    T* d = nullptr;
    MRB_TRY(arena_alloc(ctx, count, &d));
    MRB_CUDA(cudaMemcpyAsync(d, src, count * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
    *dev = d;
    return MRB_OK;
}

// Output staging: device pointer to write to; commit() copies back when dst is a host pointer.
template <typename T>
struct Out {
    T* dst = nullptr;
    T* dev = nullptr;
    size_t count = 0;
    bool host = false;
    int init(mrb_ctx* ctx, T* dst_, size_t count_) {
        dst = dst_; count = count_;
        if (dst == nullptr) { dev = nullptr; return MRB_OK; }
        if (is_device_ptr(dst)) { dev = dst; host = false; return MRB_OK; }
        host = true;
        return arena_alloc(ctx, count, &dev);
    }
    int commit(mrb_ctx* ctx, size_t n = (size_t)-1) {
        if (!host || dst == nullptr) return MRB_OK;
        if (n == (size_t)-1) n = count;
        if (n == 0) return MRB_OK;
        MRB_CUDA(cudaMemcpyAsync(dst, dev, n * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
        return MRB_OK;
    }
};

inline int sync(mrb_ctx* ctx) {
    MRB_CUDA(cudaStreamSynchronize(ctx->stream));
    return MRB_OK;
}

// Launch helper: counts launches (bench gpu_launches) and checks the launch error.
#define MRB_LAUNCH(ctx, kernel, grid, block, smem, ...)                             \
    do {                                                                            \
        kernel<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__);            \
        (ctx)->launches++;                                                          \
        MRB_CUDA(cudaGetLastError());                                               \
    } while (0)

// Opt-in to more than 48 KB of dynamic shared memory.  The attribute belongs to the function, not to the launch, and contexts on
// several host threads launch the same kernels with different sizes: always ask for the largest size any call uses (220 KB), so
// that one thread can never lower it under another thread's launch.
inline cudaError_t allow_dyn_smem(const void* kern, size_t smem) {
    if (smem <= 48 * 1024) return cudaSuccess;
    const size_t most = (size_t)220 * 1024;
    return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(smem > most ? smem : most));
}

inline unsigned cdiv(size_t a, size_t b) { return (unsigned)((a + b - 1) / b); }

// ---- device-wide primitives (mrb_scan.cu) ----
// Exclusive prefix sum of n int32 (in may alias out).  total_dev (optional) receives the sum.
int scan_exclusive(mrb_ctx* ctx, const int* in, int* out, size_t n, int* total_dev);
// Reads one int from device memory (synchronises the stream).
int read_int(mrb_ctx* ctx, const int* dev, int* host);

// ---- stage entry points on device pointers (used by the fused path) ----
struct BandDev {          // stage-2 products of the fused path, all device-resident
    int n_band = 0;
    int n_tight = 0;
    int* band_id = nullptr;     // N^3 int32: band id or -1   (z-fastest dense grid)
    int* band_pos = nullptr;    // B   int32: dense offset (x*ny+y)*nz+z of band voxel i
    uint32_t* code = nullptr;   // B   packed per-voxel word, see mrb_qp.cu
    float* margin = nullptr;    // B   margin (distance to the perimeter) of band voxel i
    int* nbr = nullptr;         // 8*B band ids of the x/y neighbours, SoA: nbr[k*B + i],
                                //     k: 0:x-2 1:x-1 2:y-2 3:y-1 4:y+1 5:y+2 6:x+1 7:x+2 (self where absent)
};

int dev_fast_perim(mrb_ctx* ctx, const float* vol, Dims d, float* out);
int dev_sqdist(mrb_ctx* ctx, const float* in, Dims d, float* out, bool fuse_sqrt);
int dev_sqrt(mrb_ctx* ctx, const float* in, size_t n, float* out);
int dev_dfield(mrb_ctx* ctx, const float* vol, Dims d, float* margin);
int dev_erode(mrb_ctx* ctx, const float* in, Dims d, float* out);
int dev_dilate(mrb_ctx* ctx, const float* in, Dims d, float* out);
int dev_threshold(mrb_ctx* ctx, const float* margin, size_t n, float band_size, float* out);
// Ordered compaction of non-zero voxels: pos_out[i] = dense offset; returns count (device scalar + host).
int dev_compact_nonzero(mrb_ctx* ctx, const float* grid, Dims d, int** pos_out, int* n_out);

int dev_qp_build(mrb_ctx* ctx, const float* volume, const float* margin, const float* band, const float* tight,
                 Dims d, BandDev* bd);
int dev_qp_solve(mrb_ctx* ctx, const BandDev& bd, Dims d, int band_size, int iters, float* x_out /*B*/,
                 float* last_delta_host);
int dev_feature_reset(mrb_ctx* ctx, const int* feat_dev, int nfeat, int n, float* x);
int dev_assemble_F(mrb_ctx* ctx, const float* volume, const BandDev& bd, const float* x, Dims d, int band_size,
                   float* F);
// ordered list of surface-crossing cells (linear cell ids over (nx-1, ny-1, nz-1), k fastest)
int classify_cells(mrb_ctx* ctx, const float* F, Dims d, float iso, int** cells_out, int* ncells_out);
// triangle soup (3 floats per corner) + hashXYZ keys -> per-corner vertex ids and unique vertices (first-appearance order)
int dev_dedup(mrb_ctx* ctx, const float* soup, const unsigned long long* keys, int ncorner, int* faces, float** verts_out,
              int* nverts_out, int** vid_out = nullptr);
int export_mesh(mrb_ctx* ctx, float* verts_dev, int* faces_dev, int nverts, int nfaces, int device_out, mrb_mesh* out);
// Stacked batch: d is nobj objects of xpitch x-planes each, stacked along x.  Faces come back with per-object vertex
// ids; face_off_host / vert_off_host (nobj + 1 entries) delimit the objects.  *seam_err_host = 1 (and no mesh) if a
// surface-crossing cell spans two objects.
struct McStack {
    int nobj, xpitch;
    unsigned long long kstride;
    int *face_off_host, *vert_off_host, *seam_err_host;
};
int dev_mcubes(mrb_ctx* ctx, const float* F, Dims d, float isolevel, float** verts_dev, int** faces_dev,
               int* nverts, int* nfaces, const McStack* stack = nullptr, int xorg = 0);

}  // namespace mrb

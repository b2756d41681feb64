// mrb_band.cu -- stage 2: narrow band (threshold, 6-neighbour erode / dilate) and ordered compaction.
//
// Replaces createBands / erode_grid / dilate_grid (narrow_band/narrowBand.cpp:10-102) and
// findIndexes / getIndexMap (grid/grid.cpp:350-389).
//
// Compaction is a warp-ballot + prefix-scan: the dense grid is z-fastest, so ascending dense
// offset == the reference's (i,j,k) loop order, and an exclusive scan of per-tile counts followed
// by ballot/popc ranking inside each warp emits the list in exactly the reference order without
// atomics.
#include "mrb_internal.cuh"
#include "mrb_compact.cuh"

namespace mrb {

static unsigned grid_for(mrb_ctx* ctx, size_t n, int block) {
    size_t want = (n + block - 1) / block;
    size_t cap = (size_t)ctx->sm_count * 16;
    return (unsigned)(want < cap ? want : cap);
}

// Erode==true : narrowBand.cpp:10-43 (tests == 0.0);  false: dilate, narrowBand.cpp:46-79 (tests == 1.0).
template <bool Erode>
__global__ void __launch_bounds__(256) k_morph(const float* __restrict__ g, Dims d, float* __restrict__ out) {
    size_t n = d.count();
    const size_t sx = (size_t)d.ny * d.nz, sy = d.nz;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (size_t)gridDim.x * blockDim.x) {
        int z = (int)(idx % d.nz);
        size_t t = idx / d.nz;
        int y = (int)(t % d.ny);
        int x = (int)(t / d.ny);
        float c = __ldg(g + idx);
        bool interior = x >= 1 && x <= d.nx - 2 && y >= 1 && y <= d.ny - 2 && z >= 1 && z <= d.nz - 2;
        if (!interior) { out[idx] = c; continue; }
        const float key = Erode ? 0.0f : 1.0f;
        bool hit = (c == key) || (__ldg(g + idx - sx) == key) || (__ldg(g + idx - sy) == key) ||
                   (__ldg(g + idx - 1) == key) || (__ldg(g + idx + sx) == key) || (__ldg(g + idx + sy) == key) ||
                   (__ldg(g + idx + 1) == key);
        out[idx] = Erode ? (hit ? 0.0f : 1.0f) : (hit ? 1.0f : 0.0f);
    }
}

__global__ void __launch_bounds__(256) k_threshold(const float* __restrict__ m, size_t n, float band_size,
                                                   float* __restrict__ out) {
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (size_t)gridDim.x * blockDim.x)
        out[idx] = (m[idx] <= band_size) ? 1.0f : 0.0f;
}

int dev_erode(mrb_ctx* ctx, const float* in, Dims d, float* out) {
    MRB_LAUNCH(ctx, (k_morph<true>), grid_for(ctx, d.count(), 256), 256, 0, in, d, out);
    return MRB_OK;
}
int dev_dilate(mrb_ctx* ctx, const float* in, Dims d, float* out) {
    MRB_LAUNCH(ctx, (k_morph<false>), grid_for(ctx, d.count(), 256), 256, 0, in, d, out);
    return MRB_OK;
}
int dev_threshold(mrb_ctx* ctx, const float* margin, size_t n, float band_size, float* out) {
    MRB_LAUNCH(ctx, k_threshold, grid_for(ctx, n, 256), 256, 0, margin, n, band_size, out);
    return MRB_OK;
}

// ---- ordered compaction of non-zero voxels (mrb_compact.cuh) ---------------------------------------
struct NonZeroPred {
    const float* g;
    __device__ bool operator()(size_t idx) const { return g[idx] != 0.0f; }
};

int dev_compact_nonzero(mrb_ctx* ctx, const float* grid, Dims d, int** pos_out, int* n_out) {
    return compact(ctx, NonZeroPred{grid}, d.count(), pos_out, n_out);
}

// dense offset (x*ny+y)*nz+z  ->  reference linear index z*nx*ny + y*nx + x (grid.cpp:172-174)
__global__ void k_pos_to_linear(const int* __restrict__ pos, int n, Dims d, int* __restrict__ lin) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int p = pos[i];
    int z = p % d.nz;
    int t = p / d.nz;
    int y = t % d.ny;
    int x = t / d.ny;
    lin[i] = z * d.nx * d.ny + y * d.nx + x;
}

__global__ void k_fill(float* __restrict__ p, size_t n, float v) {
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (size_t)gridDim.x * blockDim.x) p[idx] = v;
}

// getIndexMap: map(indexes[i]) = (float)i through grid::operator()(int) (grid.cpp:176-183).
// Duplicate indexes: the reference keeps the last write; serialise that case with atomicMax on the id.
__global__ void k_index_scatter(const int* __restrict__ lin, int n, Dims d, float* __restrict__ map, int* __restrict__ err) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int l = lin[i];
    if (l < 0 || (size_t)l >= d.count()) { atomicExch(err, 1); return; }
    int z = l / (d.nx * d.ny);
    int r = l % (d.nx * d.ny);
    int y = r / d.nx;
    int x = r % d.nx;
    // float ids are exact below 2^24 only -- same limitation as the reference (grid.cpp:386)
    atomicMax((int*)(map + d.at(x, y, z)), __float_as_int((float)i));
}

}  // namespace mrb

using namespace mrb;

extern "C" {

static int morph_api(mrb_ctx* ctx, const float* in, const int* dims, float* out, bool erode) {
    MRB_REQUIRE(ctx && in && out && dims, "NULL argument");
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    Dims d = mkdims(dims);
    const float* din = nullptr;
    MRB_TRY(stage_in(ctx, in, d.count(), &din));
    Out<float> o;
    MRB_TRY(o.init(ctx, out, d.count()));
    MRB_TRY(erode ? dev_erode(ctx, din, d, o.dev) : dev_dilate(ctx, din, d, o.dev));
    MRB_TRY(o.commit(ctx));
    return sync(ctx);
}

int mrb_erode(mrb_ctx* ctx, const float* in, const int dims[3], float* out) { return morph_api(ctx, in, dims, out, true); }
int mrb_dilate(mrb_ctx* ctx, const float* in, const int dims[3], float* out) { return morph_api(ctx, in, dims, out, false); }

int mrb_create_bands(mrb_ctx* ctx, const float* margin, const int dims[3], float band_size, float* band, float* tight) {
    MRB_REQUIRE(ctx && margin && band && tight && dims, "NULL argument");
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    Dims d = mkdims(dims);
    const float* dm = nullptr;
    MRB_TRY(stage_in(ctx, margin, d.count(), &dm));
    Out<float> ob, ot;
    MRB_TRY(ob.init(ctx, band, d.count()));
    MRB_TRY(ot.init(ctx, tight, d.count()));
    float* b0 = nullptr;
    MRB_TRY(arena_alloc(ctx, d.count(), &b0));
    MRB_TRY(dev_threshold(ctx, dm, d.count(), band_size, b0));
    MRB_TRY(dev_erode(ctx, b0, d, ot.dev));
    MRB_TRY(dev_dilate(ctx, ot.dev, d, ob.dev));
    MRB_TRY(ob.commit(ctx));
    MRB_TRY(ot.commit(ctx));
    return sync(ctx);
}

int mrb_find_indexes(mrb_ctx* ctx, const float* grid, const int dims[3], int* idx_out, int capacity, int* n_out) {
    MRB_REQUIRE(ctx && grid && dims && n_out, "NULL argument");
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    Dims d = mkdims(dims);
    const float* dg = nullptr;
    MRB_TRY(stage_in(ctx, grid, d.count(), &dg));
    int* pos = nullptr;
    int n = 0;
    MRB_TRY(dev_compact_nonzero(ctx, dg, d, &pos, &n));
    *n_out = n;
    if (!idx_out || n == 0) return sync(ctx);
    if (n > capacity) { set_error("mrb_find_indexes: capacity %d < %d", capacity, n); return MRB_ECAPACITY; }
    Out<int> o;
    MRB_TRY(o.init(ctx, idx_out, (size_t)n));
    MRB_LAUNCH(ctx, k_pos_to_linear, cdiv(n, 256), 256, 0, pos, n, d, o.dev);
    MRB_TRY(o.commit(ctx));
    return sync(ctx);
}

int mrb_index_map(mrb_ctx* ctx, const int dims[3], const int* indexes, int n, float* map_out) {
    MRB_REQUIRE(ctx && dims && map_out && (indexes || n == 0) && n >= 0, "bad argument");
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    Dims d = mkdims(dims);
    const int* di = nullptr;
    MRB_TRY(stage_in(ctx, indexes, (size_t)n, &di));
    Out<float> o;
    MRB_TRY(o.init(ctx, map_out, d.count()));
    MRB_LAUNCH(ctx, k_fill, cdiv(d.count(), 256 * 8), 256, 0, o.dev, d.count(), -1.0f);
    MRB_CUDA(cudaMemsetAsync(ctx->d_mail, 0, sizeof(int), ctx->stream));
    if (n > 0) MRB_LAUNCH(ctx, k_index_scatter, cdiv(n, 256), 256, 0, di, n, d, o.dev, ctx->d_mail);
    int err = 0;
    MRB_TRY(read_int(ctx, ctx->d_mail, &err));
    if (err) { set_error("mrb_index_map: linear index out of range (reference: grid.cpp:177-180 exits)"); return MRB_EINVAL; }
    MRB_TRY(o.commit(ctx));
    return sync(ctx);
}

}  // extern "C"

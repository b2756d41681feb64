// mrb_stage12.cu -- fused-path stages 1+2 for binary occupancy grids: bit-packed perimeter, tile-sparse
// windowed EDT and band flags.
//
// Same results as k_site -> k_tedt_x/y/z -> k_tight -> k_band_flags in mrb_strips.cu (i.e. as fastPerim ->
// getsqdist -> createBands, dfields.cpp:7-99 / narrowBand.cpp:10-102, wherever the band can reach), but
//   * the occupancy grid (values exactly 0 or 1, which is what getBinaryVolume produces, grid.cpp:322-334)
//     is read ONCE and packed to one bit per voxel (32 z per word); the 27-neighbour "window is uniform"
//     test of fastPerim becomes AND-reductions of 27 words per 32 voxels;
//   * perimeter sites exist only near the surface, so the grid is cut into 8x8x32 tiles and the EDT /
//     morphology passes run only on tiles within band_size+1 voxels of a tile that holds a site; every
//     other voxel is "far" / not in the band by construction (filled by memset).
// Grids with any value other than 0 and 1 use the general byte-grid kernels instead (reported through
// *binary_out): fastPerim compares floats with !=, which a bit plane cannot represent.
#include "mrb_internal.cuh"
#include "mrb_compact.cuh"

namespace mrb {

static const int kFar = 255;
static const int kTX = 8, kTY = 8;      // tile = 8 x 8 x 32 voxels = one bit word column block

struct BitDims {
    int nx, ny, nz, nzw;                // nzw = ceil(nz / 32) words per (x,y) row
    int tx, ty, tz;                     // tiles
    __host__ __device__ size_t words() const { return (size_t)nx * ny * nzw; }
    __host__ __device__ size_t tiles() const { return (size_t)tx * ty * tz; }
};

// one warp per (x,y) row: occ bit = (v == 1); flags any value that is neither 0 nor 1
__global__ void __launch_bounds__(256) k_pack_occ(const float* __restrict__ vol, BitDims b, uint32_t* __restrict__ occ,
                                                  int* __restrict__ nonbinary) {
    const int lane = threadIdx.x & 31;
    const size_t rows = (size_t)b.nx * b.ny;
    bool bad = false;
    for (size_t row = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < rows; row += (size_t)gridDim.x * (blockDim.x >> 5)) {
        const float* p = vol + row * b.nz;
        for (int w = 0; w < b.nzw; w++) {
            int z = w * 32 + lane;
            float v = z < b.nz ? __ldg(p + z) : 0.0f;
            bad = bad || (v != 0.0f && v != 1.0f);
            unsigned m = __ballot_sync(0xffffffffu, v == 1.0f);
            if (lane == 0) occ[row * b.nzw + w] = m;
        }
    }
    if (__any_sync(0xffffffffu, bad) && lane == 0) *nonbinary = 1;
}

// site bit = fastPerim would write 0: the border-clamped 3x3x3 window is not uniform (dfields.cpp:13-30).
// Clamping == ignoring out-of-range neighbours.  One thread per word.
__global__ void __launch_bounds__(256) k_site_bits(const uint32_t* __restrict__ occ, BitDims b, uint32_t* __restrict__ site,
                                                   uint8_t* __restrict__ tile_site) {
    const size_t n = b.words();
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (size_t)gridDim.x * blockDim.x) {
        const int w = (int)(idx % b.nzw);
        const size_t t = idx / b.nzw;
        const int y = (int)(t % b.ny), x = (int)(t / b.ny);
        const int nvalid = min(32, b.nz - 32 * w);
        const uint32_t valid = nvalid == 32 ? 0xffffffffu : ((1u << nvalid) - 1u);
        // AND over the (x,y) window of: all-ones test (A*) and all-zeros test (N*) for words w-1, w, w+1
        uint32_t A0 = 0xffffffffu, A1 = 0xffffffffu, A2 = 0xffffffffu, N0 = 0xffffffffu, N1 = 0xffffffffu, N2 = 0xffffffffu;
        for (int a = max(x - 1, 0); a <= min(x + 1, b.nx - 1); a++)
            for (int c = max(y - 1, 0); c <= min(y + 1, b.ny - 1); c++) {
                const uint32_t* row = occ + ((size_t)a * b.ny + c) * b.nzw;
                const uint32_t o1 = __ldg(row + w);
                A1 &= o1; N1 &= ~o1;
                if (w > 0) { const uint32_t o0 = __ldg(row + w - 1); A0 &= o0; N0 &= ~o0; }
                if (w < b.nzw - 1) { const uint32_t o2 = __ldg(row + w + 1); A2 &= o2; N2 &= ~o2; }
            }
        // z direction inside the word with carries; at z = 0 and z = nz-1 the missing neighbour is ignored
        auto z3 = [&](uint32_t m0, uint32_t m1, uint32_t m2) {
            uint32_t lo = (m1 << 1) | (w > 0 ? (m0 >> 31) : (m1 & 1u));                 // neighbour z-1 (replicated at z = 0)
            uint32_t hi = (m1 >> 1);                                                     // neighbour z+1
            if (nvalid == 32) hi |= (w < b.nzw - 1 ? (m2 << 31) : (m1 & 0x80000000u));   // replicated at the last z
            else hi |= (m1 & (1u << (nvalid - 1)));                                      // last valid z inside this word
            return m1 & lo & hi;
        };
        const uint32_t uniform = z3(A0, A1, A2) | z3(N0, N1 & valid, N2);
        const uint32_t s = ~uniform & valid;
        site[idx] = s;
        if (s) tile_site[((size_t)(x / kTX) * b.ty + (y / kTY)) * b.tz + w] = 1;
    }
}

struct ActiveTilePred {      // a tile is active when a tile within (r, r, 1) tiles holds a site
    const uint8_t* tile_site;
    BitDims b;
    int r;
    __device__ bool operator()(size_t t) const {
        const int tz = (int)(t % b.tz);
        const size_t q = t / b.tz;
        const int ty = (int)(q % b.ty), tx = (int)(q / b.ty);
        for (int a = max(tx - r, 0); a <= min(tx + r, b.tx - 1); a++)
            for (int c = max(ty - r, 0); c <= min(ty + r, b.ty - 1); c++)
                for (int e = max(tz - 1, 0); e <= min(tz + 1, b.tz - 1); e++)
                    if (tile_site[((size_t)a * b.ty + c) * b.tz + e]) return true;
        return false;
    }
};

// Voxel loop of one tile: block = one tile, 256 threads, thread -> (lx = tid/32, z lane), loop over ly.
#define MRB_TILE_VOXELS(tiles, b, x, y, z, idx)                                              \
    const int tile_ = tiles[blockIdx.x];                                                       \
    const int tz_ = tile_ % b.tz, ty_ = (tile_ / b.tz) % b.ty, tx_ = tile_ / (b.tz * b.ty);  \
    const int x = tx_ * kTX + (threadIdx.x >> 5), z = tz_ * 32 + (threadIdx.x & 31);           \
    if (x < b.nx && z < b.nz)                                                                  \
        for (int y = ty_ * kTY, idx_base_ = 0; y < min(ty_ * kTY + kTY, b.ny) && idx_base_ == 0; y++)  \
            for (size_t idx = ((size_t)x * b.ny + y) * b.nz + z, once_ = 1; once_; once_ = 0)

__device__ __forceinline__ bool site_at(const uint32_t* __restrict__ site, const BitDims& b, int x, int y, int z) {
    return (__ldg(site + ((size_t)x * b.ny + y) * b.nzw + (z >> 5)) >> (z & 31)) & 1u;
}

// pass X from the site bits: |dx| of the nearest site within the window, else far
__global__ void __launch_bounds__(256) k_t_edt_x(const int* __restrict__ tiles, const uint32_t* __restrict__ site, BitDims b, int W,
                                                 uint8_t* __restrict__ d1) {
    MRB_TILE_VOXELS(tiles, b, x, y, z, idx) {
        int best = kFar;
        if (site_at(site, b, x, y, z)) best = 0;
        else
            for (int r = 1; r <= W; r++) {
                bool hit = (x - r >= 0 && site_at(site, b, x - r, y, z)) || (x + r < b.nx && site_at(site, b, x + r, y, z));
                if (hit) { best = r; break; }
            }
        d1[idx] = (uint8_t)best;
    }
}

__global__ void __launch_bounds__(256) k_t_edt_y(const int* __restrict__ tiles, const uint8_t* __restrict__ d1, BitDims b, int W,
                                                 uint8_t* __restrict__ g2) {
    const int W2 = W * W;
    MRB_TILE_VOXELS(tiles, b, x, y, z, idx) {
        int c = __ldg(d1 + idx);
        int best = c == kFar ? 0x7fff : c * c;
        for (int r = 1; r <= W; r++) {
            int r2 = r * r;
            if (r2 >= best) break;
            if (y - r >= 0) { int v = __ldg(d1 + idx - (size_t)r * b.nz); if (v != kFar) best = min(best, v * v + r2); }
            if (y + r < b.ny) { int v = __ldg(d1 + idx + (size_t)r * b.nz); if (v != kFar) best = min(best, v * v + r2); }
        }
        g2[idx] = (uint8_t)(best <= W2 ? best : kFar);
    }
}

__global__ void __launch_bounds__(256) k_t_edt_z(const int* __restrict__ tiles, const uint8_t* __restrict__ g2, BitDims b, int W,
                                                 uint8_t* __restrict__ sq) {
    const int W2 = W * W;
    MRB_TILE_VOXELS(tiles, b, x, y, z, idx) {
        int c = __ldg(g2 + idx);
        int best = c == kFar ? 0x7fff : c;
        for (int r = 1; r <= W; r++) {
            int r2 = r * r;
            if (r2 >= best) break;
            if (z - r >= 0) { int v = __ldg(g2 + idx - r); if (v != kFar) best = min(best, v + r2); }
            if (z + r < b.nz) { int v = __ldg(g2 + idx + r); if (v != kFar) best = min(best, v + r2); }
        }
        sq[idx] = (uint8_t)(best <= W2 ? best : kFar);
    }
}

// tight = erode6(band0), band0 = (sq != far); border voxels copied (narrowBand.cpp:10-43)
__global__ void __launch_bounds__(256) k_t_tight(const int* __restrict__ tiles, const uint8_t* __restrict__ sq, BitDims b,
                                                 uint8_t* __restrict__ tight) {
    const size_t sx = (size_t)b.ny * b.nz, sy = b.nz;
    MRB_TILE_VOXELS(tiles, b, x, y, z, idx) {
        bool t = __ldg(sq + idx) != kFar;
        bool interior = x >= 1 && x <= b.nx - 2 && y >= 1 && y <= b.ny - 2 && z >= 1 && z <= b.nz - 2;
        if (interior && t)
            t = __ldg(sq + idx - sx) != kFar && __ldg(sq + idx + sx) != kFar && __ldg(sq + idx - sy) != kFar &&
                __ldg(sq + idx + sy) != kFar && __ldg(sq + idx - 1) != kFar && __ldg(sq + idx + 1) != kFar;
        tight[idx] = t ? 1 : 0;
    }
}

// flags: bit0 band = dilate6(tight) (narrowBand.cpp:46-79), bit1 tight; counts[0] band, [1] tight, [2] tight voxel on the border
__global__ void __launch_bounds__(256) k_t_flags(const int* __restrict__ tiles, const uint8_t* __restrict__ tight, BitDims b,
                                                 uint8_t* __restrict__ flags, int* __restrict__ counts) {
    const size_t sx = (size_t)b.ny * b.nz, sy = b.nz;
    int nb = 0, nt = 0;
    {
        MRB_TILE_VOXELS(tiles, b, x, y, z, idx) {
            bool tg = __ldg(tight + idx) != 0;
            bool interior = x >= 1 && x <= b.nx - 2 && y >= 1 && y <= b.ny - 2 && z >= 1 && z <= b.nz - 2;
            bool bd = tg;
            if (interior && !bd)
                bd = __ldg(tight + idx - sx) || __ldg(tight + idx + sx) || __ldg(tight + idx - sy) || __ldg(tight + idx + sy) ||
                     __ldg(tight + idx - 1) || __ldg(tight + idx + 1);
            if (tg && !interior) counts[2] = 1;
            flags[idx] = (uint8_t)((bd ? 1 : 0) | (tg ? 2 : 0));
            nb += bd; nt += tg;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { nb += __shfl_xor_sync(0xffffffffu, nb, o); nt += __shfl_xor_sync(0xffffffffu, nt, o); }
    if ((threadIdx.x & 31) == 0) { if (nb) atomicAdd(&counts[0], nb); if (nt) atomicAdd(&counts[1], nt); }
}

// Stages 1+2 of the fused path for a binary volume.  sq / flags: N^3 bytes each (outputs).  counts: device int[4]
// (band, tight, border error), zeroed by the caller.  *binary_out = 0 when the volume holds another value than
// 0 / 1 (nothing usable was produced: the caller runs the general kernels).
int stage12_binary(mrb_ctx* ctx, const float* vol, Dims d, int W, uint8_t* sq, uint8_t* flags, int* counts, int* binary_out) {
    BitDims b{d.nx, d.ny, d.nz, (d.nz + 31) / 32, (d.nx + kTX - 1) / kTX, (d.ny + kTY - 1) / kTY, (d.nz + 31) / 32};
    const size_t n = d.count(), nw = b.words(), nt = b.tiles();
    uint32_t *occ = nullptr, *site = nullptr;
    uint8_t *tile_site = nullptr, *d1 = nullptr, *g2 = nullptr;
    MRB_TRY(arena_alloc(ctx, nw, &occ));
    MRB_TRY(arena_alloc(ctx, nw, &site));
    MRB_TRY(arena_alloc(ctx, nt, &tile_site));
    MRB_TRY(arena_alloc(ctx, n, &d1));
    MRB_TRY(arena_alloc(ctx, n, &g2));
    int* nonbinary = ctx->d_mail + 52;
    MRB_CUDA(cudaMemsetAsync(nonbinary, 0, sizeof(int), ctx->stream));
    MRB_CUDA(cudaMemsetAsync(tile_site, 0, nt, ctx->stream));
    const size_t rows = (size_t)d.nx * d.ny;
    size_t want = (rows + 7) / 8, cap = (size_t)ctx->sm_count * 32;
    MRB_LAUNCH(ctx, k_pack_occ, (unsigned)(want < cap ? want : cap), 256, 0, vol, b, occ, nonbinary);
    want = (nw + 255) / 256;
    MRB_LAUNCH(ctx, k_site_bits, (unsigned)(want < cap ? want : cap), 256, 0, occ, b, site, tile_site);
    MRB_CUDA(cudaMemcpyAsync(ctx->h_mail + 52, nonbinary, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    // far / empty everywhere; the active tiles overwrite their voxels
    MRB_CUDA(cudaMemsetAsync(d1, kFar, n, ctx->stream));
    MRB_CUDA(cudaMemsetAsync(g2, kFar, n, ctx->stream));
    MRB_CUDA(cudaMemsetAsync(sq, kFar, n, ctx->stream));
    MRB_CUDA(cudaMemsetAsync(flags, 0, n, ctx->stream));
    int* tiles = nullptr;
    int ntiles = 0;
    const int r = (W + 1 + kTX - 1) / kTX;          // tiles within band_size + 1 voxels in x / y (z: one 32-voxel tile)
    MRB_TRY(compact(ctx, ActiveTilePred{tile_site, b, r}, nt, &tiles, &ntiles));    // synchronises the stream
    *binary_out = ctx->h_mail[52] ? 0 : 1;
    if (!*binary_out || ntiles == 0) return MRB_OK;
    uint8_t* tight = d1;                              // d1 is dead after pass Y; reuse it for the tight mask
    MRB_LAUNCH(ctx, k_t_edt_x, ntiles, 256, 0, tiles, site, b, W, d1);
    MRB_LAUNCH(ctx, k_t_edt_y, ntiles, 256, 0, tiles, d1, b, W, g2);
    MRB_LAUNCH(ctx, k_t_edt_z, ntiles, 256, 0, tiles, g2, b, W, sq);
    MRB_CUDA(cudaMemsetAsync(tight, 0, n, ctx->stream));
    MRB_LAUNCH(ctx, k_t_tight, ntiles, 256, 0, tiles, sq, b, tight);
    MRB_LAUNCH(ctx, k_t_flags, ntiles, 256, 0, tiles, tight, b, flags, counts);
    return MRB_OK;
}

}  // namespace mrb

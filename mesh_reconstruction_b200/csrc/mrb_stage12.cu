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

// The same for rows of a multiple of 128 voxels at a 16-byte aligned address: one LDG.128 per lane (a warp reads 512 contiguous
// bytes = four bit words), two rows' worth of loads in flight per warp before the first is used (the scalar kernel above
// reads the 64 MiB grid of the 256^3 workload at 2.2 TB/s, a third of what a plain stream reaches).  Lane L holds voxels
// 4L..4L+3 of a 128-voxel segment: a nibble; the four words of the segment are OR-reductions over groups of eight lanes.
__global__ void __launch_bounds__(256) k_pack_occ_v4(const float4* __restrict__ vol, BitDims b, uint32_t* __restrict__ occ,
                                                     int* __restrict__ nonbinary) {
    const int lane = threadIdx.x & 31;
    const size_t segs = (size_t)b.nx * b.ny * (b.nz / 128);               // 128-voxel segments; occ word index = seg * 4 + w
    const size_t wstep = (size_t)gridDim.x * (blockDim.x >> 5);
    bool bad = false;
    for (size_t seg = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); seg < segs; seg += 4 * wstep) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; u++)
            v[u] = seg + u * wstep < segs ? __ldg(vol + (seg + u * wstep) * 32 + lane) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int u = 0; u < 4; u++) {
            if (seg + u * wstep >= segs) break;                           // warp-uniform
            const float4 a = v[u];
            bad = bad || (a.x != 0.0f && a.x != 1.0f) || (a.y != 0.0f && a.y != 1.0f) || (a.z != 0.0f && a.z != 1.0f) ||
                  (a.w != 0.0f && a.w != 1.0f);
            uint32_t m = ((a.x == 1.0f) ? 1u : 0u) | ((a.y == 1.0f) ? 2u : 0u) | ((a.z == 1.0f) ? 4u : 0u) | ((a.w == 1.0f) ? 8u : 0u);
            m <<= 4 * (lane & 7);
            m |= __shfl_xor_sync(0xffffffffu, m, 1);
            m |= __shfl_xor_sync(0xffffffffu, m, 2);
            m |= __shfl_xor_sync(0xffffffffu, m, 4);
            if ((lane & 7) == 0) occ[(seg + u * wstep) * 4 + (lane >> 3)] = m;
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

// ---- one kernel per active tile: windowed EDT (x, y, z) + erode + dilate in shared memory ------------------------------
// The five tile kernels above re-read their inputs from global memory for every probe.  For band_size <= 8 the whole
// chain fits in shared memory: a block loads the site bits of its tile plus a halo of H = W + 2 voxels (W for the EDT
// window, 1 for the erosion, 1 for the dilation) as one 64-bit word per (x,y) row, runs pass X word-parallel over z,
// passes Y and Z, the 6-neighbour erosion and dilation on byte arrays, and writes sq and flags of its 8x8x32 voxels.
// Same arithmetic as k_t_edt_x/y/z, k_t_tight, k_t_flags (voxels of inactive tiles are "far": see ActiveTilePred).
template <int W>
struct TileGeom {
    static constexpr int H = W + 2;
    static constexpr int RX = kTX + 2 * H, RY = kTY + 2 * H, RZ = 32 + 2 * H;   // site region
    static constexpr int RZP = (RZ + 3) / 4 * 4;                               // byte row pitch
    static constexpr int CX = kTX + 4, CY = kTY + 4, CZ = 32 + 4;               // sq region (tile + 2)
    static constexpr int CZP = (CZ + 3) / 4 * 4;
    static constexpr int UX = kTX + 2, UY = kTY + 2, UZ = 32 + 2;               // tight region (tile + 1)
    static constexpr int UZP = (UZ + 3) / 4 * 4;
    static constexpr size_t bytes = (size_t)RX * RY * 8 + (size_t)CX * RY * RZP + (size_t)CX * CY * RZP + (size_t)CX * CY * CZP +
                                    (size_t)UX * UY * UZP;
};

template <int W>
__global__ void __launch_bounds__(256) k_tile_stage12(const uint8_t* __restrict__ tile_site, int tile_r, const uint32_t* __restrict__ site, BitDims b,
                                                      uint8_t* __restrict__ sq_out, uint8_t* __restrict__ flags_out,
                                                      int* __restrict__ counts) {
    using G = TileGeom<W>;
    constexpr int H = G::H, W2 = W * W;
    extern __shared__ unsigned long long s_raw[];
    unsigned long long* s_site = s_raw;                                   // [RX][RY]
    uint8_t* s_d1 = (uint8_t*)(s_site + G::RX * G::RY);                   // [CX][RY][RZP]
    uint8_t* s_g2 = s_d1 + G::CX * G::RY * G::RZP;                        // [CX][CY][RZP]
    uint8_t* s_sq = s_g2 + G::CX * G::CY * G::RZP;                        // [CX][CY][CZP]
    uint8_t* s_tg = s_sq + G::CX * G::CY * G::CZP;                        // [UX][UY][UZP]
    // one block per tile of the grid; a tile with no perimeter site within reach is left at the memset values
    const int tile = blockIdx.x;
    if (!ActiveTilePred{tile_site, b, tile_r}((size_t)tile)) return;
    const int tz = tile % b.tz, ty = (tile / b.tz) % b.ty, tx = tile / (b.tz * b.ty);
    const int x0 = tx * kTX, y0 = ty * kTY, z0 = tz * 32;
    const int tid = threadIdx.x;
    // site bits: bit i of s_site[rx][ry] is voxel (x0 - H + rx, y0 - H + ry, z0 - H + i); nothing outside the grid
    for (int i = tid; i < G::RX * G::RY; i += 256) {
        const int rx = i / G::RY, ry = i % G::RY;
        const int x = x0 - H + rx, y = y0 - H + ry;
        unsigned long long v = 0;
        if (x >= 0 && x < b.nx && y >= 0 && y < b.ny) {
            const uint32_t* row = site + ((size_t)x * b.ny + y) * b.nzw;
            const unsigned long long w1 = __ldg(row + tz);
            const unsigned long long w0 = tz > 0 ? __ldg(row + tz - 1) : 0u;
            const unsigned long long w2 = tz < b.nzw - 1 ? __ldg(row + tz + 1) : 0u;
            v = (w0 >> (32 - H)) | (w1 << H) | (w2 << (32 + H));
        }
        s_site[i] = v;
    }
    __syncthreads();
    // pass X: column (cx, ry) <-> x = x0 - 2 + cx (site row rx = cx + W); d1 = |dx| of the nearest site within W, word-parallel in z
    for (int i = tid; i < G::CX * G::RY; i += 256) {
        const int cx = i / G::RY, ry = i % G::RY;
        const int rx = cx + W;
        unsigned long long m[W + 1];
        m[0] = s_site[rx * G::RY + ry];
#pragma unroll
        for (int r = 1; r <= W; r++) m[r] = m[r - 1] | s_site[(rx - r) * G::RY + ry] | s_site[(rx + r) * G::RY + ry];
        uint32_t* out = (uint32_t*)(s_d1 + (size_t)i * G::RZP);
        for (int zw = 0; zw < G::RZP / 4; zw++) {
            uint32_t word = 0;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int z = zw * 4 + k;
                uint32_t dd = kFar;
#pragma unroll
                for (int r = W; r >= 0; r--) dd = ((m[r] >> z) & 1ull) ? (uint32_t)r : dd;
                word |= dd << (8 * k);
            }
            out[zw] = word;
        }
    }
    __syncthreads();
    // pass Y: (cx, cy, z) <-> y = y0 - 2 + cy (d1 row ry = cy + W); g2 = min(d1^2 + dy^2) if <= W^2
    for (int i = tid; i < G::CX * G::CY * G::RZP; i += 256) {
        const int z = i % G::RZP, c = i / G::RZP;
        const int cy = c % G::CY, cx = c / G::CY;
        const uint8_t* col = s_d1 + ((size_t)cx * G::RY + cy + W) * G::RZP + z;
        int best = 0x7fff;
#pragma unroll
        for (int dy = -W; dy <= W; dy++) {
            const int v = col[dy * G::RZP];
            if (v != kFar) best = min(best, v * v + dy * dy);
        }
        s_g2[i] = (uint8_t)(best <= W2 ? best : kFar);
    }
    __syncthreads();
    // pass Z: (cx, cy, cz) <-> z = z0 - 2 + cz (g2 index cz + W); sq = min(g2 + dz^2) if <= W^2
    for (int i = tid; i < G::CX * G::CY * G::CZP; i += 256) {
        const int cz = i % G::CZP, c = i / G::CZP;
        int best = 0x7fff;
        if (cz < G::CZ) {
            const uint8_t* row = s_g2 + (size_t)c * G::RZP + cz + W;
#pragma unroll
            for (int dz = -W; dz <= W; dz++) {
                const int v = row[dz];
                if (v != kFar) best = min(best, v + dz * dz);
            }
        }
        s_sq[i] = (uint8_t)(best <= W2 ? best : kFar);
    }
    __syncthreads();
    // tight = erode6(sq != far) on interior voxels, copy on the grid border (narrowBand.cpp:10-43); 0 outside the grid
    for (int i = tid; i < G::UX * G::UY * G::UZP; i += 256) {
        const int uz = i % G::UZP, c = i / G::UZP;
        const int uy = c % G::UY, ux = c / G::UY;
        const int x = x0 - 1 + ux, y = y0 - 1 + uy, z = z0 - 1 + uz;
        bool t = false;
        if (uz < G::UZ && x >= 0 && x < b.nx && y >= 0 && y < b.ny && z >= 0 && z < b.nz) {
            const uint8_t* p = s_sq + ((size_t)(ux + 1) * G::CY + uy + 1) * G::CZP + uz + 1;
            t = *p != kFar;
            const bool interior = x >= 1 && x <= b.nx - 2 && y >= 1 && y <= b.ny - 2 && z >= 1 && z <= b.nz - 2;
            if (interior && t)
                t = p[-G::CY * G::CZP] != kFar && p[G::CY * G::CZP] != kFar && p[-G::CZP] != kFar && p[G::CZP] != kFar &&
                    p[-1] != kFar && p[1] != kFar;
        }
        s_tg[i] = t ? 1 : 0;
    }
    __syncthreads();
    // flags: bit0 band = dilate6(tight) (narrowBand.cpp:46-79), bit1 tight; sq for the brick builder
    int nb = 0, nt = 0;
    {
        const int lz = tid & 31, lx = tid >> 5;
        const int x = x0 + lx, z = z0 + lz;
        if (x < b.nx && z < b.nz)
            for (int ly = 0; ly < kTY && y0 + ly < b.ny; ly++) {
                const int y = y0 + ly;
                const uint8_t* p = s_tg + ((size_t)(lx + 1) * G::UY + ly + 1) * G::UZP + lz + 1;
                const bool tg = *p != 0;
                const bool interior = x >= 1 && x <= b.nx - 2 && y >= 1 && y <= b.ny - 2 && z >= 1 && z <= b.nz - 2;
                bool bd = tg;
                if (interior && !bd) bd = p[-G::UY * G::UZP] || p[G::UY * G::UZP] || p[-G::UZP] || p[G::UZP] || p[-1] || p[1];
                if (tg && !interior) counts[2] = 1;
                const size_t idx = ((size_t)x * b.ny + y) * b.nz + z;
                flags_out[idx] = (uint8_t)((bd ? 1 : 0) | (tg ? 2 : 0));
                sq_out[idx] = s_sq[((size_t)(lx + 2) * G::CY + ly + 2) * G::CZP + lz + 2];
                nb += bd; nt += tg;
            }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { nb += __shfl_xor_sync(0xffffffffu, nb, o); nt += __shfl_xor_sync(0xffffffffu, nt, o); }
    if ((threadIdx.x & 31) == 0) { if (nb) atomicAdd(&counts[0], nb); if (nt) atomicAdd(&counts[1], nt); }
}

// ---- bit-sliced version for band_size <= 4 ---------------------------------------------------------------------------------
// k_tile_stage12 still spends ~600 instructions per output voxel on byte arithmetic over the halo.  Here a column of
// 32 + 2H voxels along z is ONE 64-bit mask and every step is word-parallel:
//   X:  MX[r]  = "a site within |dx| <= r"               (cumulative in r, OR of 2r+1 site words)
//   Y:  G[v]   = "d1^2 + dy^2 <= v"                      (OR of MX[r](y+dy) over r^2 + dy^2 <= v)
//   Z:  S[u]   = "sq <= u" = OR_dz shift(G[u - dz^2], dz);  band0 = S[W^2]
//   erode / dilate: AND / OR of the four neighbour columns and the two z shifts (narrowBand.cpp:10-79)
// The squared distance itself is only needed where the final band bit is set (the brick builder reads it there):
// it is read back from five bit planes built from the exact levels S[u] & ~S[u-1].
template <int W>
__global__ void __launch_bounds__(256) k_tile_bits(const uint8_t* __restrict__ tile_site, int tile_r, const uint32_t* __restrict__ site, BitDims b,
                                                   uint8_t* __restrict__ sq_out, uint8_t* __restrict__ flags_out,
                                                   int* __restrict__ counts, uint8_t* __restrict__ brick_bits) {
    typedef unsigned long long u64;
    constexpr int H = W + 2, W2 = W * W;
    constexpr int RX = kTX + 2 * H, RY = kTY + 2 * H;
    constexpr int CX = kTX + 4, CY = kTY + 4, UX = kTX + 2, UY = kTY + 2;
    __shared__ u64 s_site[RX * RY];
    __shared__ u64 s_mx[(W + 1) * CX * RY];
    __shared__ u64 s_b0[CX * CY];
    __shared__ u64 s_tg[UX * UY];
    __shared__ unsigned s_colz[kTX * kTY];       // per own column: which of its 8 z-groups of 4 voxels hold a band voxel
    // one block per tile of the grid; a tile with no perimeter site within reach is left at the memset values
    const int tile = blockIdx.x;
    if (!ActiveTilePred{tile_site, b, tile_r}((size_t)tile)) return;
    const int tz = tile % b.tz, ty = (tile / b.tz) % b.ty, tx = tile / (b.tz * b.ty);
    const int x0 = tx * kTX, y0 = ty * kTY, z0 = tz * 32;
    const int tid = threadIdx.x;
    // bit i <-> z = z0 - H + i.  valid: inside the grid; zint: 1 <= z <= nz - 2; tilebits: the 32 voxels of this tile
    u64 valid = 0, zint = 0;
    {
        const int lo = max(0, H - z0), hi = min(32 + 2 * H, b.nz - z0 + H);            // [lo, hi) valid bit range
        if (hi > lo) valid = ((hi >= 64 ? ~0ull : ((1ull << hi) - 1ull)) >> lo) << lo;
        const int lo1 = max(0, H - z0 + 1), hi1 = min(32 + 2 * H, b.nz - 1 - z0 + H);
        if (hi1 > lo1) zint = ((hi1 >= 64 ? ~0ull : ((1ull << hi1) - 1ull)) >> lo1) << lo1;
    }
    const u64 tilebits = 0xffffffffull << H;
    if (tid < kTX * kTY) s_colz[tid] = 0;
    for (int i = tid; i < RX * RY; i += 256) {
        const int rx = i / RY, ry = i % RY;
        const int x = x0 - H + rx, y = y0 - H + ry;
        u64 v = 0;
        if (x >= 0 && x < b.nx && y >= 0 && y < b.ny) {
            const uint32_t* row = site + ((size_t)x * b.ny + y) * b.nzw;
            const u64 w1 = __ldg(row + tz);
            const u64 w0 = tz > 0 ? __ldg(row + tz - 1) : 0u;
            const u64 w2 = tz < b.nzw - 1 ? __ldg(row + tz + 1) : 0u;
            v = (w0 >> (32 - H)) | (w1 << H) | (w2 << (32 + H));
        }
        s_site[i] = v;
    }
    __syncthreads();
    // X: column (cx, ry) <-> x = x0 - 2 + cx, site row rx = cx + W
    for (int i = tid; i < CX * RY; i += 256) {
        const int cx = i / RY, ry = i % RY, rx = cx + W;
        u64 m = s_site[rx * RY + ry];
        s_mx[i] = m;
#pragma unroll
        for (int r = 1; r <= W; r++) {
            m |= s_site[(rx - r) * RY + ry] | s_site[(rx + r) * RY + ry];
            s_mx[r * CX * RY + i] = m;
        }
    }
    __syncthreads();
    // Y + Z: thread <-> column (cx, cy), y = y0 - 2 + cy (MX row ry = cy + W).  G stays in registers for the own columns.
    const int cx = tid / CY, cy = tid % CY;
    const bool col = tid < CX * CY;
    const int gx = x0 - 2 + cx, gy = y0 - 2 + cy;
    const bool col_in = col && gx >= 0 && gx < b.nx && gy >= 0 && gy < b.ny;
    u64 G[W2 + 1];
#pragma unroll
    for (int v = 0; v <= W2; v++) G[v] = 0;
    if (col_in) {
#pragma unroll
        for (int r = 0; r <= W; r++)
#pragma unroll
            for (int dy = -W; dy <= W; dy++)
                if (r * r + dy * dy <= W2) G[r * r + dy * dy] |= s_mx[r * CX * RY + cx * RY + cy + W + dy];
#pragma unroll
        for (int v = 1; v <= W2; v++) G[v] |= G[v - 1];
    }
    auto level = [&](int u) -> u64 {          // S[u] = "sq <= u" (u is a compile-time constant after unrolling)
        u64 acc = 0;
#pragma unroll
        for (int dz = -W; dz <= W; dz++)
            if (dz * dz <= u) {
                const u64 g = G[u - dz * dz];
                acc |= dz >= 0 ? (g >> dz) : (g << -dz);
            }
        return acc;
    };
    if (col) s_b0[tid] = col_in ? (level(W2) & valid) : 0ull;
    __syncthreads();
    // erosion on the tile + 1 columns: thread <-> (ux, uy), x = x0 - 1 + ux
    if (tid < UX * UY) {
        const int ux = tid / UY, uy = tid % UY;
        const int x = x0 - 1 + ux, y = y0 - 1 + uy;
        const u64* p = s_b0 + (ux + 1) * CY + uy + 1;
        const u64 B = *p;
        u64 t = B;                                                  // border voxels are copied (narrowBand.cpp:20-30)
        if (x >= 1 && x <= b.nx - 2 && y >= 1 && y <= b.ny - 2) {
            const u64 er = B & p[-CY] & p[CY] & p[-1] & p[1] & (B << 1) & (B >> 1);
            t = (er & zint) | (B & ~zint);
        }
        s_tg[tid] = t;
    }
    __syncthreads();
    // dilation + outputs for the 8 x 8 own columns (the threads that hold their G)
    int nb = 0, nt = 0;
    const bool own = col_in && cx >= 2 && cx < 2 + kTX && cy >= 2 && cy < 2 + kTY;
    if (own) {
        const u64* p = s_tg + (cx - 1) * UY + (cy - 1);
        const u64 T = *p;
        const bool interior_col = gx >= 1 && gx <= b.nx - 2 && gy >= 1 && gy <= b.ny - 2;
        u64 band = T;
        if (interior_col) {
            const u64 dil = T | p[-UY] | p[UY] | p[-1] | p[1] | (T << 1) | (T >> 1);
            band = (dil & zint) | (T & ~zint);
        }
        const u64 tg = T & tilebits & valid;
        band &= tilebits & valid;
        nb = __popcll(band);
        nt = __popcll(tg);
        if (tg & ~(interior_col ? zint : 0ull)) counts[2] = 1;     // tight voxel on the grid border
        const uint32_t bw = (uint32_t)(band >> H), tw = (uint32_t)(tg >> H);
        {
            uint32_t g4 = bw | (bw >> 1) | (bw >> 2) | (bw >> 3);        // bit 4k = any of bits 4k..4k+3
            uint32_t m8 = 0;
#pragma unroll
            for (int k = 0; k < 8; k++) m8 |= ((g4 >> (4 * k)) & 1u) << k;
            s_colz[(cx - 2) * kTY + (cy - 2)] = m8;
        }
        const size_t base = ((size_t)gx * b.ny + gy) * b.nz + z0;
        if (bw | tw) {
            if (z0 + 32 <= b.nz && (b.nz & 3) == 0) {
                uint32_t* out = (uint32_t*)(flags_out + base);
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const uint32_t bb = (bw >> (4 * k)) & 15u, tt = (tw >> (4 * k)) & 15u;
                    // spread 4 bits to 4 bytes: bit j -> byte j
                    const uint32_t be = (bb & 1u) | ((bb & 2u) << 7) | ((bb & 4u) << 14) | ((bb & 8u) << 21);
                    const uint32_t te = (tt & 1u) | ((tt & 2u) << 7) | ((tt & 4u) << 14) | ((tt & 8u) << 21);
                    out[k] = be | (te << 1);
                }
            } else {
                for (int k = 0; k < 32 && z0 + k < b.nz; k++) flags_out[base + k] = (uint8_t)(((bw >> k) & 1u) | (((tw >> k) & 1u) << 1));
            }
        }
        if (bw) {
            // squared distance of the band voxels from the bit planes of the exact levels (values up to W^2)
            constexpr int NPL = W2 < 32 ? 5 : 7;
            u64 pl[NPL];
#pragma unroll
            for (int k = 0; k < NPL; k++) pl[k] = 0;
            u64 prev = 0;
#pragma unroll
            for (int u = 0; u <= W2; u++) {
                const u64 S = level(u);
                const u64 exact = S & ~prev;
                prev = S;
#pragma unroll
                for (int k = 0; k < NPL; k++)
                    if ((u >> k) & 1) pl[k] |= exact;
            }
            uint32_t m = bw;
            while (m) {
                const int k = __ffs(m) - 1;
                m &= m - 1;
                const int i = k + H;
                uint32_t v = 0;
#pragma unroll
                for (int q = 0; q < NPL; q++) v |= (uint32_t)((pl[q] >> i) & 1ull) << q;
                sq_out[base + k] = (uint8_t)v;
            }
        }
    }
    // bricks (2 x 2 columns x 4 z) with a band voxel: one byte per (brick column, tile z) = its 8 bricks along z
    if (brick_bits) {
        __syncthreads();
        if (tid < (kTX / 2) * (kTY / 2)) {
            const int bxl = tid / (kTY / 2), byl = tid % (kTY / 2);
            const unsigned m = s_colz[(2 * bxl) * kTY + 2 * byl] | s_colz[(2 * bxl) * kTY + 2 * byl + 1] |
                               s_colz[(2 * bxl + 1) * kTY + 2 * byl] | s_colz[(2 * bxl + 1) * kTY + 2 * byl + 1];
            const int X = tx * (kTX / 2) + bxl, Y = ty * (kTY / 2) + byl;
            const int bx = (b.nx + 1) / 2, by = (b.ny + 1) / 2;
            if (X < bx && Y < by && m) brick_bits[((size_t)X * by + Y) * b.tz + tz] = (uint8_t)m;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { nb += __shfl_xor_sync(0xffffffffu, nb, o); nt += __shfl_xor_sync(0xffffffffu, nt, o); }
    if ((threadIdx.x & 31) == 0) { if (nb) atomicAdd(&counts[0], nb); if (nt) atomicAdd(&counts[1], nt); }
}

template <int W>
static int launch_tile_stage12(mrb_ctx* ctx, const uint8_t* tile_site, int tile_r, int ntiles, const uint32_t* site, BitDims b,
                               uint8_t* sq, uint8_t* flags, int* counts) {
    const size_t smem = TileGeom<W>::bytes;
    if (smem > 48 * 1024) MRB_CUDA(cudaFuncSetAttribute(k_tile_stage12<W>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    MRB_LAUNCH(ctx, k_tile_stage12<W>, ntiles, 256, smem, tile_site, tile_r, site, b, sq, flags, counts);
    return MRB_OK;
}

// Stages 1+2 of the fused path for a binary volume.  sq / flags: N^3 bytes each (outputs).  counts: device int[4]
// (band, tight, border error), zeroed by the caller.  *binary_out = 0 when the volume holds another value than
// 0 / 1 (nothing usable was produced: the caller runs the general kernels).
int stage12_binary(mrb_ctx* ctx, const float* vol, Dims d, int W, uint8_t* sq, uint8_t* flags, int* counts, int* binary_out,
                   uint8_t** brick_bits_out) {
    if (brick_bits_out) *brick_bits_out = nullptr;
    BitDims b{d.nx, d.ny, d.nz, (d.nz + 31) / 32, (d.nx + kTX - 1) / kTX, (d.ny + kTY - 1) / kTY, (d.nz + 31) / 32};
    const size_t n = d.count(), nw = b.words(), nt = b.tiles();
    uint32_t *occ = nullptr, *site = nullptr;
    uint8_t *tile_site = nullptr, *d1 = nullptr, *g2 = nullptr;
    const bool fused = W >= 1 && W <= 8 && !getenv("MRB_STAGE12_UNFUSED");
    MRB_TRY(arena_alloc(ctx, nw, &occ));
    MRB_TRY(arena_alloc(ctx, nw, &site));
    MRB_TRY(arena_alloc(ctx, nt, &tile_site));
    if (!fused) {
        MRB_TRY(arena_alloc(ctx, n, &d1));
        MRB_TRY(arena_alloc(ctx, n, &g2));
    }
    int* nonbinary = ctx->d_mail + 52;
    MRB_CUDA(cudaMemsetAsync(nonbinary, 0, sizeof(int), ctx->stream));
    MRB_CUDA(cudaMemsetAsync(tile_site, 0, nt, ctx->stream));
    const size_t rows = (size_t)d.nx * d.ny;
    size_t want = (rows + 7) / 8, cap = (size_t)ctx->sm_count * 32;
    if (d.nz % 128 == 0 && ((uintptr_t)vol & 15) == 0 && !getenv("MRB_PACK_SCALAR"))
        MRB_LAUNCH(ctx, k_pack_occ_v4, (unsigned)ctx->sm_count * 8, 256, 0, (const float4*)vol, b, occ, nonbinary);
    else
        MRB_LAUNCH(ctx, k_pack_occ, (unsigned)(want < cap ? want : cap), 256, 0, vol, b, occ, nonbinary);
    want = (nw + 255) / 256;
    MRB_LAUNCH(ctx, k_site_bits, (unsigned)(want < cap ? want : cap), 256, 0, occ, b, site, tile_site);
    MRB_CUDA(cudaMemcpyAsync(ctx->h_mail + 52, nonbinary, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    // far / empty everywhere; the active tiles overwrite their voxels
    if (!fused) {
        MRB_CUDA(cudaMemsetAsync(d1, kFar, n, ctx->stream));
        MRB_CUDA(cudaMemsetAsync(g2, kFar, n, ctx->stream));
    }
    MRB_CUDA(cudaMemsetAsync(sq, kFar, n, ctx->stream));
    MRB_CUDA(cudaMemsetAsync(flags, 0, n, ctx->stream));
    const int r = (W + 1 + kTX - 1) / kTX;          // tiles within band_size + 1 voxels in x / y (z: one 32-voxel tile)
    if (fused) {
        // one block per tile of the grid, inactive tiles return at once: no tile list, no count to wait for
        MRB_CUDA(cudaStreamSynchronize(ctx->stream));                   // the "other value than 0 / 1" flag
        *binary_out = ctx->h_mail[52] ? 0 : 1;
        if (!*binary_out) return MRB_OK;
        const int ntl = (int)nt;
        const int bits_max = getenv("MRB_STAGE12_BITS_MAX") ? atoi(getenv("MRB_STAGE12_BITS_MAX")) : 8;
        if (W <= bits_max && !getenv("MRB_STAGE12_BYTES")) {
            uint8_t* bits = nullptr;
            const size_t nbits = (size_t)((d.nx + 1) / 2) * ((d.ny + 1) / 2) * b.tz;
            MRB_TRY(arena_alloc(ctx, nbits, &bits));
            MRB_CUDA(cudaMemsetAsync(bits, 0, nbits, ctx->stream));
            switch (W) {
                case 1: MRB_LAUNCH(ctx, k_tile_bits<1>, ntl, 256, 0, tile_site, r, site, b, sq, flags, counts, bits); break;
                case 2: MRB_LAUNCH(ctx, k_tile_bits<2>, ntl, 256, 0, tile_site, r, site, b, sq, flags, counts, bits); break;
                case 3: MRB_LAUNCH(ctx, k_tile_bits<3>, ntl, 256, 0, tile_site, r, site, b, sq, flags, counts, bits); break;
                case 4: MRB_LAUNCH(ctx, k_tile_bits<4>, ntl, 256, 0, tile_site, r, site, b, sq, flags, counts, bits); break;
                case 5: MRB_LAUNCH(ctx, k_tile_bits<5>, ntl, 256, 0, tile_site, r, site, b, sq, flags, counts, bits); break;
                case 6: MRB_LAUNCH(ctx, k_tile_bits<6>, ntl, 256, 0, tile_site, r, site, b, sq, flags, counts, bits); break;
                case 7: MRB_LAUNCH(ctx, k_tile_bits<7>, ntl, 256, 0, tile_site, r, site, b, sq, flags, counts, bits); break;
                default: MRB_LAUNCH(ctx, k_tile_bits<8>, ntl, 256, 0, tile_site, r, site, b, sq, flags, counts, bits); break;
            }
            if (brick_bits_out) *brick_bits_out = bits;
            return MRB_OK;
        }
        switch (W) {
            case 1: return launch_tile_stage12<1>(ctx, tile_site, r, ntl, site, b, sq, flags, counts);
            case 2: return launch_tile_stage12<2>(ctx, tile_site, r, ntl, site, b, sq, flags, counts);
            case 3: return launch_tile_stage12<3>(ctx, tile_site, r, ntl, site, b, sq, flags, counts);
            case 4: return launch_tile_stage12<4>(ctx, tile_site, r, ntl, site, b, sq, flags, counts);
            case 5: return launch_tile_stage12<5>(ctx, tile_site, r, ntl, site, b, sq, flags, counts);
            case 6: return launch_tile_stage12<6>(ctx, tile_site, r, ntl, site, b, sq, flags, counts);
            case 7: return launch_tile_stage12<7>(ctx, tile_site, r, ntl, site, b, sq, flags, counts);
            default: return launch_tile_stage12<8>(ctx, tile_site, r, ntl, site, b, sq, flags, counts);
        }
    }
    int* tiles = nullptr;
    int ntiles = 0;
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

// mrb_strips.cu -- the fused, device-resident stages 1-3 of mrb_optimize / mrb_reconstruct.
//
// Same results as fastPerim -> getsqdist -> getsqrt -> createBands -> primeQP -> runQP(500) -> F
// (distance_fields/dfields.cpp:7-99,182-192, narrow_band/narrowBand.cpp:10-102,
// optimize/primeqp.cpp:7-223, optimize/quadprog.cpp:45-147), organised for the B200:
//
//  * Stage 1+2 work on byte grids.  Only voxels with squared distance <= band_size^2 can enter the
//    band and only their margins are ever read, so the three EDT passes are windowed to +-band_size
//    (exact wherever the true squared distance is <= band_size^2; everything else becomes "far").
//    A voxel whose true nearest perimeter voxel is within the window in every axis gets the exact
//    value; a windowed minimum is always the distance to some real perimeter voxel, so it can never
//    undershoot.  Result: band / tight / margins identical to the full EDT.
//  * Stage 3 works on z-strips: 4 consecutive voxels (z = 4k..4k+3) of one (x,y) column that contain
//    at least one band voxel.  A strip's iterate is one float4 (128-bit loads/stores), its 10
//    neighbour strips (x-2,x-1,y-2,y-1,z-,z+,y+1,y+2,x+1,x+2) are found through 10 strip ids, and one
//    32-bit code word carries the tight-mask bits of the 4 voxels and their neighbours.  Non-band
//    voxels inside a strip have zero coefficients, lb=-1000, ub=1000 and stay exactly 0.
//  * One persistent cooperative kernel runs all sweeps (grid barrier between sweeps).
//
// Per voxel and sweep (quadprog.cpp:49-60):
//     res = sum c*x over x-2 x-1 y-2 y-1 z-2 z-1 z+1 z+2 y+1 y+2 x+1 x+2   (ascending band id)
//     c(v,v+-e) = -2(T(v)+T(v+-e)),  c(v,v+-2e) = T(v+-e),  d = 12 T(v) + sum of the 6 neighbour T
//     t = (x - res*invdg[d]) * 0.5, clamped to [lb, ub]
// Every product is exact (coefficients 0, 1, -2, -4), zero-coefficient terms are skipped or add +-0,
// so the fp32 result is bit-identical to the reference's CSC loop.
#include "mrb_internal.cuh"
#include "mrb_compact.cuh"

#include <algorithm>
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

namespace mrb {

static const int kFar = 255;      // u8 "beyond the window"

__constant__ float c_invdg_s[19];   // (float)(1.0/(double)d), d = 0..18; entry 0 is 0 (primeqp.cpp:189-197)

static unsigned grid_for(mrb_ctx* ctx, size_t n, int block) {
    size_t want = (n + block - 1) / block;
    size_t cap = (size_t)ctx->sm_count * 16;
    return (unsigned)(want < cap ? want : cap);
}

#define MRB_VOXEL_LOOP(idx, n) \
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < (n); idx += (size_t)gridDim.x * blockDim.x)

// ---- stage 1: perimeter sites + windowed exact EDT on bytes ---------------------------------------
// site = 1 where fastPerim would write 0 (a voxel of the clamped 3x3x3 window differs), dfields.cpp:13-30
__global__ void __launch_bounds__(256) k_site(const float* __restrict__ vol, Dims d, uint8_t* __restrict__ site) {
    MRB_VOXEL_LOOP(idx, d.count()) {
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
        site[idx] = same ? 0 : 1;
    }
}

// pass X: |dx| to the nearest site within the window, else kFar
__global__ void __launch_bounds__(256) k_tedt_x(const uint8_t* __restrict__ site, Dims d, int W, uint8_t* __restrict__ d1) {
    const size_t sx = (size_t)d.ny * d.nz;
    MRB_VOXEL_LOOP(idx, d.count()) {
        int x = (int)(idx / sx);
        int best = kFar;
        if (__ldg(site + idx)) best = 0;
        else
            for (int r = 1; r <= W; r++) {
                bool hit = (x - r >= 0 && __ldg(site + idx - r * sx)) || (x + r < d.nx && __ldg(site + idx + r * sx));
                if (hit) { best = r; break; }
            }
        d1[idx] = (uint8_t)best;
    }
}

// pass Y: min d1(y')^2 + (y'-y)^2 over the window; values above W^2 are "far"
__global__ void __launch_bounds__(256) k_tedt_y(const uint8_t* __restrict__ d1, Dims d, int W, uint8_t* __restrict__ g2) {
    const int W2 = W * W;
    MRB_VOXEL_LOOP(idx, d.count()) {
        int y = (int)((idx / d.nz) % d.ny);
        int c = __ldg(d1 + idx);
        int best = c == kFar ? 0x7fff : c * c;
        for (int r = 1; r <= W; r++) {
            int r2 = r * r;
            if (r2 >= best) break;
            if (y - r >= 0) { int v = __ldg(d1 + idx - (size_t)r * d.nz); if (v != kFar) best = min(best, v * v + r2); }
            if (y + r < d.ny) { int v = __ldg(d1 + idx + (size_t)r * d.nz); if (v != kFar) best = min(best, v * v + r2); }
        }
        g2[idx] = (uint8_t)(best <= W2 ? best : kFar);
    }
}

// pass Z: min g2(z') + (z'-z)^2 over the window -> squared distance (<= W^2) or kFar
__global__ void __launch_bounds__(256) k_tedt_z(const uint8_t* __restrict__ g2, Dims d, int W, uint8_t* __restrict__ sq) {
    const int W2 = W * W;
    MRB_VOXEL_LOOP(idx, d.count()) {
        int z = (int)(idx % d.nz);
        int c = __ldg(g2 + idx);
        int best = c == kFar ? 0x7fff : c;
        for (int r = 1; r <= W; r++) {
            int r2 = r * r;
            if (r2 >= best) break;
            if (z - r >= 0) { int v = __ldg(g2 + idx - r); if (v != kFar) best = min(best, v + r2); }
            if (z + r < d.nz) { int v = __ldg(g2 + idx + r); if (v != kFar) best = min(best, v + r2); }
        }
        sq[idx] = (uint8_t)(best <= W2 ? best : kFar);
    }
}

// ---- stage 2: band0 -> tight (erode6) -> band (dilate6), border voxels copied (narrowBand.cpp:10-102) ----
// band0(v) = sq(v) <= W^2  <=>  margin <= band_size   (sq is an integer, band_size an integer)
__global__ void __launch_bounds__(256) k_tight(const uint8_t* __restrict__ sq, Dims d, uint8_t* __restrict__ tight) {
    const size_t sx = (size_t)d.ny * d.nz, sy = d.nz;
    MRB_VOXEL_LOOP(idx, d.count()) {
        int z = (int)(idx % d.nz);
        size_t t = idx / d.nz;
        int y = (int)(t % d.ny);
        int x = (int)(t / d.ny);
        bool b = __ldg(sq + idx) != kFar;
        bool interior = x >= 1 && x <= d.nx - 2 && y >= 1 && y <= d.ny - 2 && z >= 1 && z <= d.nz - 2;
        if (interior && b)
            b = __ldg(sq + idx - sx) != kFar && __ldg(sq + idx + sx) != kFar && __ldg(sq + idx - sy) != kFar &&
                __ldg(sq + idx + sy) != kFar && __ldg(sq + idx - 1) != kFar && __ldg(sq + idx + 1) != kFar;
        tight[idx] = b ? 1 : 0;
    }
}

// flags: bit0 = band, bit1 = tight.  counts[0] += band voxels, counts[1] += tight voxels,
// counts[2] = 1 if a tight voxel lies on the border (reference exits, primeqp.cpp:40-75 -> grid.cpp:20-37)
__global__ void __launch_bounds__(256) k_band_flags(const uint8_t* __restrict__ tight, Dims d, uint8_t* __restrict__ flags,
                                                    int* __restrict__ counts) {
    const size_t sx = (size_t)d.ny * d.nz, sy = d.nz;
    int nb = 0, nt = 0;
    MRB_VOXEL_LOOP(idx, d.count()) {
        int z = (int)(idx % d.nz);
        size_t t = idx / d.nz;
        int y = (int)(t % d.ny);
        int x = (int)(t / d.ny);
        bool tg = __ldg(tight + idx) != 0;
        bool interior = x >= 1 && x <= d.nx - 2 && y >= 1 && y <= d.ny - 2 && z >= 1 && z <= d.nz - 2;
        bool b = tg;
        if (interior && !b)
            b = __ldg(tight + idx - sx) || __ldg(tight + idx + sx) || __ldg(tight + idx - sy) || __ldg(tight + idx + sy) ||
                __ldg(tight + idx - 1) || __ldg(tight + idx + 1);
        if (tg && !interior) counts[2] = 1;
        flags[idx] = (uint8_t)((b ? 1 : 0) | (tg ? 2 : 0));
        nb += b; nt += tg;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { nb += __shfl_xor_sync(0xffffffffu, nb, o); nt += __shfl_xor_sync(0xffffffffu, nt, o); }
    if ((threadIdx.x & 31) == 0) { if (nb) atomicAdd(&counts[0], nb); if (nt) atomicAdd(&counts[1], nt); }
}

// general path (band_size > 15 or caller-provided float grids): flags from float band / tight grids
__global__ void __launch_bounds__(256) k_flags_from_float(const float* __restrict__ band, const float* __restrict__ tight, Dims d,
                                                          uint8_t* __restrict__ flags, int* __restrict__ counts) {
    int nb = 0, nt = 0;
    MRB_VOXEL_LOOP(idx, d.count()) {
        int z = (int)(idx % d.nz);
        size_t t = idx / d.nz;
        int y = (int)(t % d.ny);
        int x = (int)(t / d.ny);
        bool b = band[idx] != 0.0f, tg = tight[idx] != 0.0f;
        bool interior = x >= 1 && x <= d.nx - 2 && y >= 1 && y <= d.ny - 2 && z >= 1 && z <= d.nz - 2;
        if (tg && !interior) counts[2] = 1;
        flags[idx] = (uint8_t)((b ? 1 : 0) | (tg ? 2 : 0));
        nb += b; nt += tg;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { nb += __shfl_xor_sync(0xffffffffu, nb, o); nt += __shfl_xor_sync(0xffffffffu, nt, o); }
    if ((threadIdx.x & 31) == 0) { if (nb) atomicAdd(&counts[0], nb); if (nt) atomicAdd(&counts[1], nt); }
}

// ---- strips -------------------------------------------------------------------------------------------
struct StripDims {
    int nx, ny, nz, nz4;
    __host__ __device__ size_t count() const { return (size_t)nx * ny * nz4; }
};

struct ActiveStripPred {
    const uint8_t* flags;
    StripDims sd;
    __device__ bool operator()(size_t s) const {
        int z4 = (int)(s % sd.nz4);
        size_t col = s / sd.nz4;
        const uint8_t* p = flags + col * sd.nz + 4 * z4;
        int nzl = min(4, sd.nz - 4 * z4);
        bool any = false;
        for (int j = 0; j < nzl; j++) any = any || (p[j] & 1);
        return any;
    }
};

__global__ void k_scatter_strip_ids(const int* __restrict__ spos, int S, int* __restrict__ sid) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < S) sid[spos[i]] = i;
}

// code word: bits 0..3 T(voxel j); bit 4 T(z-1 of voxel 0); bit 5 T(z+1 of voxel 3);
//            bits 6..9 T(x-1) j=0..3; 10..13 T(x+1); 14..17 T(y-1); 18..21 T(y+1)
struct MarginSq {       // margin = correctly rounded sqrtf(sq), as getsqrt (dfields.cpp:187)
    const uint8_t* sq;
    __device__ float operator()(size_t idx) const { return __fsqrt_rn((float)sq[idx]); }
};
struct MarginF32 {
    const float* m;
    __device__ float operator()(size_t idx) const { return m[idx]; }
};

template <typename MarginFn>
__global__ void __launch_bounds__(128) k_strip_build(const float* __restrict__ vol, MarginFn margin, const uint8_t* __restrict__ flags,
                                                     const int* __restrict__ sid, const int* __restrict__ spos, int S, StripDims sd,
                                                     uint32_t* __restrict__ code, float4* __restrict__ lb4, float4* __restrict__ ub4,
                                                     float4* __restrict__ inv4, float4* __restrict__ x0, int4* __restrict__ nbA, int4* __restrict__ nbB,
                                                     int2* __restrict__ nbC, int* __restrict__ err) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S) return;
    int p = spos[s];
    int z4 = p % sd.nz4;
    int col = p / sd.nz4;
    int y = col % sd.ny;
    int x = col / sd.ny;
    const size_t sx = (size_t)sd.ny * sd.nz, sy = sd.nz;
    const size_t base = (size_t)col * sd.nz + 4 * z4;
    auto tight_at = [&](long long off, int z) -> uint32_t {     // off relative to base, z absolute
        if (z < 0 || z >= sd.nz) return 0u;
        return (flags[base + off] >> 1) & 1u;
    };
    uint32_t c = 0;
    float lbv[4], ubv[4], xv[4];
    bool bad = false;
    for (int j = 0; j < 4; j++) {
        int z = 4 * z4 + j;
        lbv[j] = -1000.0f; ubv[j] = 1000.0f; xv[j] = 0.0f;
        if (z >= sd.nz) continue;
        uint32_t f = flags[base + j];
        uint32_t t0 = (f >> 1) & 1u;
        c |= t0 << j;
        if (x >= 1) c |= ((flags[base + j - sx] >> 1) & 1u) << (6 + j);
        if (x <= sd.nx - 2) c |= ((flags[base + j + sx] >> 1) & 1u) << (10 + j);
        if (y >= 1) c |= ((flags[base + j - sy] >> 1) & 1u) << (14 + j);
        if (y <= sd.ny - 2) c |= ((flags[base + j + sy] >> 1) & 1u) << (18 + j);
        if (t0) {
            // a tight voxel needs its six neighbours in the band (reference: indexMap -1 -> column -1)
            bool in = x >= 1 && x <= sd.nx - 2 && y >= 1 && y <= sd.ny - 2 && z >= 1 && z <= sd.nz - 2;
            if (!in) bad = true;
            else bad = bad || !(flags[base + j - sx] & 1) || !(flags[base + j + sx] & 1) || !(flags[base + j - sy] & 1) ||
                       !(flags[base + j + sy] & 1) || !(flags[base + j - 1] & 1) || !(flags[base + j + 1] & 1);
        }
        if (f & 1u) {
            float v = vol[base + j], mg = margin(base + j);
            float l = (v == 0.0f) ? -1000.0f : mg;      // getlb, primeqp.cpp:122-127
            float u = (v == 1.0f) ? 1000.0f : -mg;      // getub, primeqp.cpp:148-153
            float xx = 0.0f;                            // x0 = clamp(0, lb, ub), primeqp.cpp:178-187
            if (xx < l) xx = l;
            if (xx > u) xx = u;
            lbv[j] = l; ubv[j] = u; xv[j] = xx;
        }
    }
    c |= tight_at(-1, 4 * z4 - 1) << 4;
    c |= tight_at(4, 4 * z4 + 4) << 5;
    if (bad) atomicExch(err, 1);
    code[s] = c;
    lb4[s] = make_float4(lbv[0], lbv[1], lbv[2], lbv[3]);
    ub4[s] = make_float4(ubv[0], ubv[1], ubv[2], ubv[3]);
    x0[s] = make_float4(xv[0], xv[1], xv[2], xv[3]);
    {   // d = 12 T(v) + the six neighbour T bits; invdg from the table (0 where d == 0)
        float iv[4];
        for (int j = 0; j < 4; j++) {
            uint32_t tzm = j > 0 ? (c >> (j - 1)) & 1u : (c >> 4) & 1u;
            uint32_t tzp = j < 3 ? (c >> (j + 1)) & 1u : (c >> 5) & 1u;
            int dg = 12 * (int)((c >> j) & 1u) + (int)((c >> (6 + j)) & 1u) + (int)((c >> (10 + j)) & 1u) +
                     (int)((c >> (14 + j)) & 1u) + (int)((c >> (18 + j)) & 1u) + (int)tzm + (int)tzp;
            iv[j] = c_invdg_s[dg];
        }
        inv4[s] = make_float4(iv[0], iv[1], iv[2], iv[3]);
    }
    auto nid = [&](bool ok, long long dstrip) {
        int id = ok ? sid[p + dstrip] : -1;
        return id >= 0 ? id : S;           // slot S = the all-zero dummy strip
    };
    const long long cx = (long long)sd.ny * sd.nz4, cy = sd.nz4;
    nbA[s] = make_int4(nid(x >= 2, -2 * cx), nid(x >= 1, -cx), nid(y >= 2, -2 * cy), nid(y >= 1, -cy));
    nbB[s] = make_int4(nid(z4 >= 1, -1), nid(z4 <= sd.nz4 - 2, 1), nid(y <= sd.ny - 2, cy), nid(y <= sd.ny - 3, 2 * cy));
    nbC[s] = make_int2(nid(x <= sd.nx - 2, cx), nid(x <= sd.nx - 3, 2 * cx));
}

struct StripBand {
    int S = 0;
    int n_band = 0, n_tight = 0;
    StripDims sd{};
    uint8_t* flags = nullptr;   // N^3
    int* sid = nullptr;         // nx*ny*nz4 strip id or -1
    int* spos = nullptr;        // S
    uint32_t* code = nullptr;   // S
    float4 *lb4 = nullptr, *ub4 = nullptr, *inv4 = nullptr;   // S
    float4 *xa = nullptr, *xb = nullptr;     // S+1 (slot S = zeros)
    int4 *nbA = nullptr, *nbB = nullptr;     // S
    int2* nbC = nullptr;                     // S
};

static int upload_invdg_s(int device) {
    static bool ready[64] = {false};
    if (ready[device & 63]) return MRB_OK;
    float h[19];
    h[0] = 0.0f;
    for (int d = 1; d <= 18; d++) h[d] = (float)(1.0 / (double)(float)d);    // primeqp.cpp:196
    MRB_CUDA(cudaMemcpyToSymbol(c_invdg_s, h, sizeof(h)));
    ready[device & 63] = true;
    return MRB_OK;
}

__global__ void k_fill_i32(int* __restrict__ p, size_t n, int v) {
    MRB_VOXEL_LOOP(idx, n) p[idx] = v;
}

template <typename MarginFn>
static int build_strips(mrb_ctx* ctx, const float* vol, MarginFn margin, uint8_t* flags, Dims d, StripBand* sb) {
    MRB_TRY(upload_invdg_s(ctx->device));
    sb->flags = flags;
    sb->sd = StripDims{d.nx, d.ny, d.nz, (d.nz + 3) / 4};
    size_t ns = sb->sd.count();
    MRB_TRY(compact(ctx, ActiveStripPred{flags, sb->sd}, ns, &sb->spos, &sb->S));
    const int S = sb->S;
    MRB_TRY(arena_alloc(ctx, ns, &sb->sid));
    MRB_CUDA(cudaMemsetAsync(sb->sid, 0xFF, (size_t)ns * sizeof(int), ctx->stream));     // every int = -1
    size_t Sn = (size_t)(S > 0 ? S : 1);
    MRB_TRY(arena_alloc(ctx, Sn, &sb->code));
    MRB_TRY(arena_alloc(ctx, Sn, &sb->lb4));
    MRB_TRY(arena_alloc(ctx, Sn, &sb->ub4));
    MRB_TRY(arena_alloc(ctx, Sn, &sb->inv4));
    MRB_TRY(arena_alloc(ctx, Sn + 1, &sb->xa));
    MRB_TRY(arena_alloc(ctx, Sn + 1, &sb->xb));
    MRB_TRY(arena_alloc(ctx, Sn, &sb->nbA));
    MRB_TRY(arena_alloc(ctx, Sn, &sb->nbB));
    MRB_TRY(arena_alloc(ctx, Sn, &sb->nbC));
    if (S == 0) return MRB_OK;
    MRB_LAUNCH(ctx, k_scatter_strip_ids, cdiv(S, 256), 256, 0, sb->spos, S, sb->sid);
    int* err = ctx->d_mail + 24;
    MRB_CUDA(cudaMemsetAsync(err, 0, sizeof(int), ctx->stream));
    MRB_CUDA(cudaMemsetAsync(sb->xa + S, 0, sizeof(float4), ctx->stream));
    MRB_CUDA(cudaMemsetAsync(sb->xb, 0, (Sn + 1) * sizeof(float4), ctx->stream));
    MRB_LAUNCH(ctx, (k_strip_build<MarginFn>), cdiv(S, 128), 128, 0, vol, margin, flags, sb->sid, sb->spos, S, sb->sd, sb->code,
               sb->lb4, sb->ub4, sb->inv4, sb->xa, sb->nbA, sb->nbB, sb->nbC, err);
    return MRB_OK;
}

// ---- the persistent sweep kernel --------------------------------------------------------------------------
// One voxel update.  Bits of the strip code word are tested in place (one LOP3 -> predicate each):
//   MA: T(v); MXM/MXP/MYM/MYP/MZM/MZP: T of the six neighbours.
//   c(+-e)  = T(v) ? (t ? -4 : -2) : (t ? -2 : 0)      c(+-2e) = t ? 1 : 0
// inv = (float)(1.0/(double)d) precomputed per voxel (0 where d == 0), primeqp.cpp:189-197.
template <uint32_t MA, uint32_t MXM, uint32_t MXP, uint32_t MYM, uint32_t MYP, uint32_t MZM, uint32_t MZP>
__device__ __forceinline__ float qp_voxel(uint32_t c, float xc, float xm2, float xm1, float ym2, float ym1, float zm2,
                                          float zm1, float zp1, float zp2, float yp1, float yp2, float xp1, float xp2,
                                          float inv, float lb, float ub) {
    const bool a = c & MA;
    const float cA = a ? -4.0f : -2.0f;      // neighbour tight
    const float cB = a ? -2.0f : 0.0f;       // neighbour not tight
    float res = (c & MXM) ? xm2 : 0.0f;
    res = __fmaf_rn((c & MXM) ? cA : cB, xm1, res);
    if (c & MYM) res = __fadd_rn(res, ym2);
    res = __fmaf_rn((c & MYM) ? cA : cB, ym1, res);
    if (c & MZM) res = __fadd_rn(res, zm2);
    res = __fmaf_rn((c & MZM) ? cA : cB, zm1, res);
    res = __fmaf_rn((c & MZP) ? cA : cB, zp1, res);
    if (c & MZP) res = __fadd_rn(res, zp2);
    res = __fmaf_rn((c & MYP) ? cA : cB, yp1, res);
    if (c & MYP) res = __fadd_rn(res, yp2);
    res = __fmaf_rn((c & MXP) ? cA : cB, xp1, res);
    if (c & MXP) res = __fadd_rn(res, xp2);
    float t = __fsub_rn(xc, __fmul_rn(res, inv));
    t = t * 0.5f;
    if (t < lb) t = lb;
    if (t > ub) t = ub;
    return t;
}

__device__ __forceinline__ float4 ld4(const float4* p) { return *p; }

// Lightweight grid barrier: one arrival counter that only grows during a launch; every block adds 1
// (release) and polls it (acquire) until it reaches epoch * nblocks.  Together with the two bar.sync
// this makes every block's sweep-k stores visible to every block's sweep-k+1 loads.
// (Measured on B200, 148..296 blocks: ~1.5 us per barrier; a separate release flag written by the
// last arriver was slower, ~1.9 us.)
__device__ __forceinline__ void grid_barrier(unsigned* bar, unsigned epoch, unsigned nblocks) {
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned target = epoch * nblocks;
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(bar) : "memory");
        unsigned v;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(bar) : "memory");
        } while (v < target);
    }
    __syncthreads();
}

// Streaming (evict-first, no L1 allocation) loads for the per-strip constants, so that the L1 keeps
// the iterate lines that neighbouring strips of the same SM re-read.
__device__ __forceinline__ float4 ldcs4(const float4* p) { return __ldcs(p); }
__device__ __forceinline__ int4 ldcs4i(const int4* p) { return __ldcs(p); }
__device__ __forceinline__ int2 ldcs2i(const int2* p) { return __ldcs(p); }

struct StripIds { int4 a, b; int2 c; };

// Mode 0: neighbour ids of the block's strips resident in shared memory (copied once).
// Mode 1: ids from L2 every sweep.
// Mode 2: ids from L2, software-pipelined one strip ahead (the ids -> x dependent round trip of the
//         next strip overlaps the arithmetic of the current one); constants with streaming loads.
template <int Mode, int Threads>
__global__ void __launch_bounds__(Threads) k_qp_strips(const uint32_t* __restrict__ code, const float4* __restrict__ lb4,
                                                       const float4* __restrict__ ub4, const float4* __restrict__ inv4,
                                                       const int4* __restrict__ nbA, const int4* __restrict__ nbB,
                                                       const int2* __restrict__ nbC, float4* xa, float4* xb, int S, int iters,
                                                       int cap, unsigned* __restrict__ maxdelta_bits,
                                                       unsigned* barrier_counter, int dbg) {
    extern __shared__ int4 s_ids[];      // Mode 0: [cap] A, [cap] B, then [cap] int2 C
    // each block owns a contiguous range of strips (balanced to +-1 strip across blocks; neighbouring
    // y rows stay on the same SM, which keeps most y/z neighbour loads in L1)
    const int s_begin = (int)((long long)S * blockIdx.x / gridDim.x);
    int s_end = (int)((long long)S * (blockIdx.x + 1) / gridDim.x);
    if (dbg & 2) s_end = s_begin;     // tuning aid: barrier cost only
    int4* sA = s_ids;
    int4* sB = s_ids + cap;
    int2* sC = (int2*)(s_ids + 2 * cap);
    if (Mode == 0) {
        for (int s = s_begin + threadIdx.x; s < s_end; s += Threads) {
            sA[s - s_begin] = __ldg(nbA + s);
            sB[s - s_begin] = __ldg(nbB + s);
            sC[s - s_begin] = __ldg(nbC + s);
        }
        __syncthreads();
    }
    auto load_ids = [&](int s) {
        StripIds r;
        if (Mode == 0) { r.a = sA[s - s_begin]; r.b = sB[s - s_begin]; r.c = sC[s - s_begin]; }
        else if (Mode == 1) { r.a = __ldg(nbA + s); r.b = __ldg(nbB + s); r.c = __ldg(nbC + s); }
        else { r.a = ldcs4i(nbA + s); r.b = ldcs4i(nbB + s); r.c = ldcs2i(nbC + s); }
        return r;
    };
    const float4* in = xa;
    float4* out = xb;
    const int s0 = s_begin + threadIdx.x;
    StripIds nxt;
    if (Mode == 2 && s0 < s_end) nxt = load_ids(s0);
    for (int it = 0; it < iters; it++) {
        float dmax = 0.0f;
        for (int s = s0; s < s_end; s += Threads) {
            StripIds id;
            if (Mode == 2) {
                id = nxt;
                int sn = s + Threads;
                if (sn >= s_end) sn = s0;              // wrap: first strip of the next sweep
                nxt = load_ids(sn);
            } else {
                id = load_ids(s);
            }
            uint32_t c;
            float4 lb, ub, inv;
            if (Mode == 2) { c = __ldcs(code + s); lb = ldcs4(lb4 + s); ub = ldcs4(ub4 + s); inv = ldcs4(inv4 + s); }
            else if (dbg & 4) {     // tuning aid: no lb/ub/inv traffic (wrong results)
                c = __ldg(code + s);
                lb = make_float4(-1000.f, -1000.f, -1000.f, -1000.f); ub = make_float4(1000.f, 1000.f, 1000.f, 1000.f);
                inv = make_float4(0.055f, 0.055f, 0.055f, 0.055f);
            }
            else { c = __ldg(code + s); lb = __ldg(lb4 + s); ub = __ldg(ub4 + s); inv = __ldg(inv4 + s); }
            if (dbg & 8) { id.a.x = s; id.a.z = s; id.b.w = s; id.c.y = s; }   // tuning aid: +-2 rows alias the own row
            if (dbg & 16) { id.a.y = s; id.a.w = s; id.b.z = s; id.c.x = s; }  // tuning aid: +-1 rows alias the own row
            // the iterate is rewritten by other SMs every sweep: plain loads, ordered by the grid barrier
            const float4 xc = ld4(in + s);
            const float4 xm2 = ld4(in + id.a.x), xm1 = ld4(in + id.a.y), ym2 = ld4(in + id.a.z), ym1 = ld4(in + id.a.w);
            const float4 zlo = ld4(in + id.b.x), zhi = ld4(in + id.b.y), yp1 = ld4(in + id.b.z), yp2 = ld4(in + id.b.w);
            const float4 xp1 = ld4(in + id.c.x), xp2 = ld4(in + id.c.y);
            float4 r;
            r.x = qp_voxel<1u << 0, 1u << 6, 1u << 10, 1u << 14, 1u << 18, 1u << 4, 1u << 1>(
                c, xc.x, xm2.x, xm1.x, ym2.x, ym1.x, zlo.z, zlo.w, xc.y, xc.z, yp1.x, yp2.x, xp1.x, xp2.x, inv.x, lb.x, ub.x);
            r.y = qp_voxel<1u << 1, 1u << 7, 1u << 11, 1u << 15, 1u << 19, 1u << 0, 1u << 2>(
                c, xc.y, xm2.y, xm1.y, ym2.y, ym1.y, zlo.w, xc.x, xc.z, xc.w, yp1.y, yp2.y, xp1.y, xp2.y, inv.y, lb.y, ub.y);
            r.z = qp_voxel<1u << 2, 1u << 8, 1u << 12, 1u << 16, 1u << 20, 1u << 1, 1u << 3>(
                c, xc.z, xm2.z, xm1.z, ym2.z, ym1.z, xc.x, xc.y, xc.w, zhi.x, yp1.z, yp2.z, xp1.z, xp2.z, inv.z, lb.z, ub.z);
            r.w = qp_voxel<1u << 3, 1u << 9, 1u << 13, 1u << 17, 1u << 21, 1u << 2, 1u << 5>(
                c, xc.w, xm2.w, xm1.w, ym2.w, ym1.w, xc.y, xc.z, zhi.x, zhi.y, yp1.w, yp2.w, xp1.w, xp2.w, inv.w, lb.w, ub.w);
            out[s] = r;
            if (it == iters - 1)
                dmax = fmaxf(fmaxf(dmax, fmaxf(fabsf(r.x - xc.x), fabsf(r.y - xc.y))), fmaxf(fabsf(r.z - xc.z), fabsf(r.w - xc.w)));
        }
        if (it == iters - 1) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) dmax = fmaxf(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
            if ((threadIdx.x & 31) == 0) atomicMax(maxdelta_bits, __float_as_uint(dmax));
        }
        const float4* t = in; in = out; out = (float4*)t;
        if (it + 1 < iters && !(dbg & 1)) grid_barrier(barrier_counter, (unsigned)(it + 1), gridDim.x);
    }
}

// Kernel variant (tuning knob): MRB_QP_VARIANT = "<mode>:<threads>", e.g. 0:1024, 1:512, 2:512, 2:1024
static void qp_variant(int* mode, int* threads) {
    *mode = 1; *threads = 512;       // measured best on B200 at 256^3 (see profiles/)
    const char* e = getenv("MRB_QP_VARIANT");
    if (e && e[0] >= '0' && e[0] <= '2' && e[1] == ':') { *mode = e[0] - '0'; *threads = atoi(e + 2) == 1024 ? 1024 : 512; }
}

static int solve_strips(mrb_ctx* ctx, const StripBand& sb, int iters, float4** result, float* last_delta_host) {
    const int S = sb.S;
    *result = sb.xa;
    if (last_delta_host) *last_delta_host = 0.0f;
    if (S == 0 || iters == 0) return MRB_OK;
    unsigned* md = (unsigned*)(ctx->d_mail + 8);
    MRB_CUDA(cudaMemsetAsync(md, 0, sizeof(unsigned), ctx->stream));
    int mode, block;
    qp_variant(&mode, &block);
    const size_t kSmemBudget = 200 * 1024;
    if (mode == 0 && (size_t)((S + ctx->sm_count - 1) / ctx->sm_count) * 40 > kSmemBudget) mode = 2;
    void* kern = nullptr;
    if (mode == 0) kern = block == 1024 ? (void*)k_qp_strips<0, 1024> : (void*)k_qp_strips<0, 512>;
    else if (mode == 1) kern = block == 1024 ? (void*)k_qp_strips<1, 1024> : (void*)k_qp_strips<1, 512>;
    else kern = block == 1024 ? (void*)k_qp_strips<2, 1024> : (void*)k_qp_strips<2, 512>;
    int per_sm = 2048 / block;                       // blocks per SM we want resident
    if (mode == 0) per_sm = 1;
    int grid = per_sm * ctx->sm_count;
    int need = (S + block - 1) / block;
    if (grid > need) grid = need;
    int cap = (S + grid - 1) / grid;
    size_t smem = mode == 0 ? (size_t)cap * 40 : 0;
    MRB_CUDA(allow_dyn_smem((const void*)kern, smem));
    int fit = 0;
    MRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessorWithFlags(&fit, kern, block, smem, 0));
    if (fit < 1) { set_error("k_qp_strips does not fit on an SM"); return MRB_ECUDA; }
    if (grid > fit * ctx->sm_count) {                // registers allow fewer blocks than wanted
        grid = fit * ctx->sm_count;
        cap = (S + grid - 1) / grid;
        if (mode == 0) { smem = (size_t)cap * 40; if (smem > kSmemBudget) { set_error("k_qp_strips: ids do not fit in shared memory"); return MRB_ECUDA; } }
        MRB_CUDA(allow_dyn_smem((const void*)kern, smem));
    }
    const uint32_t* code = sb.code;
    const float4 *lb4 = sb.lb4, *ub4 = sb.ub4, *inv4 = sb.inv4;
    unsigned* bar = (unsigned*)ctx->d_bar;
    MRB_CUDA(cudaMemsetAsync(bar, 0, 64 * sizeof(unsigned), ctx->stream));
    const int4 *nbA = sb.nbA, *nbB = sb.nbB;
    const int2* nbC = sb.nbC;
    float4 *xa = sb.xa, *xb = sb.xb;
    int SS = S, its = iters;
    int dbg = getenv("MRB_QP_DEBUG") ? atoi(getenv("MRB_QP_DEBUG")) : 0;   // 1: no barrier, 2: no work (timing aids; wrong results)
    void* args[] = {(void*)&code, (void*)&lb4, (void*)&ub4, (void*)&inv4, (void*)&nbA, (void*)&nbB, (void*)&nbC,
                    (void*)&xa, (void*)&xb, (void*)&SS, (void*)&its, (void*)&cap, (void*)&md, (void*)&bar, (void*)&dbg};
    MRB_CUDA(cudaEventRecord(ctx->ev[6], ctx->stream));
    MRB_CUDA(cudaLaunchCooperativeKernel(kern, dim3(grid), dim3(block), args, smem, ctx->stream));
    MRB_CUDA(cudaEventRecord(ctx->ev[7], ctx->stream));
    ctx->launches++;
    *result = (iters % 2) ? sb.xb : sb.xa;
    if (last_delta_host) {
        MRB_CUDA(cudaMemcpyAsync(ctx->h_mail + 8, md, sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
        MRB_CUDA(cudaStreamSynchronize(ctx->stream));
        unsigned bits = (unsigned)ctx->h_mail[8];
        memcpy(last_delta_host, &bits, sizeof(float));
    }
    return MRB_OK;
}

// ---- small bands: the whole solve inside ONE thread-block cluster ---------------------------------------------------
// Grids of the size the reference was written for (40^3 completions: 40 .. 64 voxels per edge, up to ~30 000 strips) are
// pure synchronisation latency in k_qp_strips: 2.7 us per sweep whatever the size (flag or grid-barrier round trips through
// L2), 1.35 ms for 500 sweeps.  Here the band lives in the shared memory of a cluster of up to 16 CTAs for the whole solve:
//   * CTA r owns strips [r*cap, (r+1)*cap): both iterate buffers, the constants and, per strip, the ten neighbour strips as
//     32-bit shared::cluster addresses (mapa once at start-up), so a neighbour in another CTA is one ld.shared::cluster
//     (distributed shared memory, ~215 cycles) and costs no index arithmetic in the sweep;
//   * one hardware cluster barrier per sweep (barrier.cluster arrive.release / wait.acquire, ~380 cycles) instead of flags
//     in global memory; nothing touches L2 or HBM between the initial load and the final store.
// The per-voxel arithmetic is qp_voxel, the same instruction sequence as k_qp_strips: bit-identical results.
static const int kClThreads = 512;
static const int kClBytesPerStrip = 2 * 16 + 3 * 16 + 4 + 10 * 4;          // xa, xb, lb, ub, inv, code, 10 neighbour addresses
static const int kClCapMax = 2 * kClThreads;                               // strips per CTA: at most two per thread
__device__ __forceinline__ float4 ld_cluster4(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t mapa_cluster(uint32_t local_addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// REGS: one strip per thread at most -- the thread keeps it for the whole solve, so its ten neighbour addresses, tight-mask word
// and 1/d live in registers and shared memory holds the two iterate buffers and the bounds only.  !REGS: up to kClCapMax strips
// per CTA, those per-strip words in shared memory as well.
template <bool REGS>
__global__ void __launch_bounds__(kClThreads, 1) k_qp_cluster(const uint32_t* __restrict__ code, const float4* __restrict__ lb4,
                                                              const float4* __restrict__ ub4, const float4* __restrict__ inv4,
                                                              const int4* __restrict__ nbA, const int4* __restrict__ nbB,
                                                              const int2* __restrict__ nbC, float4* xa, float4* xb, int S, int iters, int cap,
                                                              unsigned* __restrict__ maxdelta_bits) {
    extern __shared__ float4 s_cl[];
    // layout (the same in every CTA of the cluster): xA[cap+1] | xB[cap+1] | lb[cap] | ub[cap] | !REGS: inv[cap] | code[cap] | nb[cap][10]
    float4* s_xa = s_cl;
    float4* s_xb = s_xa + (cap + 1);
    float4* s_lb = s_xb + (cap + 1);
    float4* s_ub = s_lb + cap;
    float4* s_inv = s_ub + cap;
    uint32_t* s_code = (uint32_t*)(s_inv + cap);
    uint32_t* s_nb = s_code + cap;
    uint32_t rank;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    const int s_begin = (int)rank * cap;
    const int n_own = max(0, min(S - s_begin, cap));
    const uint32_t xa_local = (uint32_t)__cvta_generic_to_shared(s_xa);
    const uint32_t buf_off = (uint32_t)(cap + 1) * 16u;                     // xB = xA + buf_off in every CTA
    // a neighbour strip id -> shared::cluster address of its slot in buffer A (slot cap of the own CTA = the all-zero dummy)
    auto slot_addr = [&](int id) -> uint32_t {
        if (id >= S) return xa_local + (uint32_t)cap * 16u;
        const uint32_t r = (uint32_t)(id / cap), l = (uint32_t)(id - (int)r * cap);
        return mapa_cluster(xa_local + l * 16u, r);
    };
    uint32_t rnb[10], rcw = 0;
    float4 rinv = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll
    for (int q = 0; q < 10; q++) rnb[q] = xa_local + (uint32_t)cap * 16u;
    for (int l = threadIdx.x; l < n_own; l += kClThreads) {
        const int s = s_begin + l;
        s_xa[l] = xa[s];
        s_xb[l] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        s_lb[l] = __ldg(lb4 + s); s_ub[l] = __ldg(ub4 + s);
        const int4 a = __ldg(nbA + s), b = __ldg(nbB + s);
        const int2 c = __ldg(nbC + s);
        const uint32_t n[10] = {slot_addr(a.x), slot_addr(a.y), slot_addr(a.z), slot_addr(a.w), slot_addr(b.x),
                                slot_addr(b.y), slot_addr(b.z), slot_addr(b.w), slot_addr(c.x), slot_addr(c.y)};
        if (REGS) {
            rinv = __ldg(inv4 + s); rcw = __ldg(code + s);
#pragma unroll
            for (int q = 0; q < 10; q++) rnb[q] = n[q];
        } else {
            s_inv[l] = __ldg(inv4 + s); s_code[l] = __ldg(code + s);
#pragma unroll
            for (int q = 0; q < 10; q++) s_nb[(size_t)l * 10 + q] = n[q];
        }
    }
    if (threadIdx.x == 0) { s_xa[cap] = make_float4(0.0f, 0.0f, 0.0f, 0.0f); s_xb[cap] = make_float4(0.0f, 0.0f, 0.0f, 0.0f); }
    cluster_sync_all();
    uint32_t off = 0;                                                          // buffer the sweep reads: 0 = A, buf_off = B
    float dmax = 0.0f;
    auto update = [&](int l, uint32_t c, const float4 iv, const uint32_t (&n)[10], float4* s_out, bool last) {
        const float4 xc = ld_cluster4(xa_local + (uint32_t)l * 16u + off);
        const float4 xm2 = ld_cluster4(n[0] + off), xm1 = ld_cluster4(n[1] + off), ym2 = ld_cluster4(n[2] + off), ym1 = ld_cluster4(n[3] + off);
        const float4 zlo = ld_cluster4(n[4] + off), zhi = ld_cluster4(n[5] + off), yp1 = ld_cluster4(n[6] + off), yp2 = ld_cluster4(n[7] + off);
        const float4 xp1 = ld_cluster4(n[8] + off), xp2 = ld_cluster4(n[9] + off);
        const float4 lb = s_lb[l], ub = s_ub[l];
        float4 r;
        r.x = qp_voxel<1u << 0, 1u << 6, 1u << 10, 1u << 14, 1u << 18, 1u << 4, 1u << 1>(
            c, xc.x, xm2.x, xm1.x, ym2.x, ym1.x, zlo.z, zlo.w, xc.y, xc.z, yp1.x, yp2.x, xp1.x, xp2.x, iv.x, lb.x, ub.x);
        r.y = qp_voxel<1u << 1, 1u << 7, 1u << 11, 1u << 15, 1u << 19, 1u << 0, 1u << 2>(
            c, xc.y, xm2.y, xm1.y, ym2.y, ym1.y, zlo.w, xc.x, xc.z, xc.w, yp1.y, yp2.y, xp1.y, xp2.y, iv.y, lb.y, ub.y);
        r.z = qp_voxel<1u << 2, 1u << 8, 1u << 12, 1u << 16, 1u << 20, 1u << 1, 1u << 3>(
            c, xc.z, xm2.z, xm1.z, ym2.z, ym1.z, xc.x, xc.y, xc.w, zhi.x, yp1.z, yp2.z, xp1.z, xp2.z, iv.z, lb.z, ub.z);
        r.w = qp_voxel<1u << 3, 1u << 9, 1u << 13, 1u << 17, 1u << 21, 1u << 2, 1u << 5>(
            c, xc.w, xm2.w, xm1.w, ym2.w, ym1.w, xc.y, xc.z, zhi.x, zhi.y, yp1.w, yp2.w, xp1.w, xp2.w, iv.w, lb.w, ub.w);
        s_out[l] = r;
        if (last) dmax = fmaxf(fmaxf(dmax, fmaxf(fabsf(r.x - xc.x), fabsf(r.y - xc.y))), fmaxf(fabsf(r.z - xc.z), fabsf(r.w - xc.w)));
    };
    for (int it = 0; it < iters; it++) {
        float4* s_out = off ? s_xa : s_xb;
        const bool last = it == iters - 1;
        if (REGS) {
            if ((int)threadIdx.x < n_own) update((int)threadIdx.x, rcw, rinv, rnb, s_out, last);
        } else {
            for (int l = threadIdx.x; l < n_own; l += kClThreads) {
                uint32_t n[10];
#pragma unroll
                for (int q = 0; q < 10; q++) n[q] = s_nb[(size_t)l * 10 + q];
                update(l, s_code[l], s_inv[l], n, s_out, last);
            }
        }
        off ^= buf_off;
        // every CTA's stores of this sweep before anybody's loads of the next (and their loads of this sweep before the next
        // sweep's stores into the same buffer); after the last sweep nobody reads remote memory any more
        if (it + 1 < iters) cluster_sync_all();
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dmax = fmaxf(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
    if ((threadIdx.x & 31) == 0 && iters > 0) atomicMax(maxdelta_bits, __float_as_uint(dmax));
    // the result goes where k_qp_strips leaves it: buffer B after an odd number of sweeps, A after an even one
    float4* gout = (iters & 1) ? xb : xa;
    const float4* s_res = (iters & 1) ? s_xb : s_xa;
    for (int l = threadIdx.x; l < n_own; l += kClThreads) gout[s_begin + l] = s_res[l];
    // a CTA must not exit while another may still read its shared memory (the last sweep's loads): one more barrier
    cluster_sync_all();
}

// Returns MRB_OK with *used = true when the cluster kernel ran.
static int solve_strips_cluster(mrb_ctx* ctx, const StripBand& sb, int iters, float4** result, float* last_delta_host, bool* used) {
    *used = false;
    const int S = sb.S;
    if (S == 0 || iters == 0) return MRB_OK;
    const char* e = getenv("MRB_QP_CLUSTER");
    if (e && atoi(e) == 0) return MRB_OK;
    // Sixteen SMs do all the arithmetic: measured on B200 1.4 us per sweep with one strip per thread (42^3), 1.9 us with two
    // (52^3), 3.0 us with three (64^3) against 2.7 us for k_qp_strips on all SMs -- so up to two strips per thread.
    if (S > 16 * kClCapMax) return MRB_OK;
    int cs = 1;
    while (cs < 16 && (cs * kClThreads < S || (S + cs - 1) / cs > kClCapMax)) cs *= 2;
    if (e && atoi(e) > 1) cs = atoi(e);                       // tuning knob: the cluster size
    if (cs != 1 && cs != 2 && cs != 4 && cs != 8 && cs != 16) return MRB_OK;
    const int cap = (S + cs - 1) / cs;
    if (cap > kClCapMax) return MRB_OK;                       // does not fit: the L2-resident kernel
    const size_t smem = (size_t)cap * kClBytesPerStrip + 2 * 16 + 16;
    {
        // function attributes are process-wide: set once to the largest value any call can need (contexts on several host
        // threads launch this kernel with different sizes)
        static bool attr_set[64] = {false};
        if (!attr_set[ctx->device & 63]) {
            MRB_CUDA(cudaFuncSetAttribute(k_qp_cluster<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
            MRB_CUDA(cudaFuncSetAttribute(k_qp_cluster<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
            MRB_CUDA(cudaFuncSetAttribute(k_qp_cluster<true>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
            MRB_CUDA(cudaFuncSetAttribute(k_qp_cluster<false>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
            attr_set[ctx->device & 63] = true;
        }
    }
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(cs);
    cfg.blockDim = dim3(kClThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = ctx->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    int nclusters = 0;
    auto kern = cap <= kClThreads ? k_qp_cluster<true> : k_qp_cluster<false>;
    if (cudaOccupancyMaxActiveClusters(&nclusters, kern, &cfg) != cudaSuccess || nclusters < 1) {
        cudaGetLastError();
        return MRB_OK;                                        // this GPU cannot co-schedule such a cluster
    }
    unsigned* md = (unsigned*)(ctx->d_mail + 8);
    MRB_CUDA(cudaMemsetAsync(md, 0, sizeof(unsigned), ctx->stream));
    MRB_CUDA(cudaEventRecord(ctx->ev[6], ctx->stream));
    MRB_CUDA(cudaLaunchKernelEx(&cfg, kern, (const uint32_t*)sb.code, (const float4*)sb.lb4, (const float4*)sb.ub4,
                                (const float4*)sb.inv4, (const int4*)sb.nbA, (const int4*)sb.nbB, (const int2*)sb.nbC, sb.xa, sb.xb, S,
                                iters, cap, md));
    MRB_CUDA(cudaEventRecord(ctx->ev[7], ctx->stream));
    ctx->launches++;
    *result = (iters % 2) ? sb.xb : sb.xa;
    *used = true;
    if (last_delta_host) {
        MRB_CUDA(cudaMemcpyAsync(ctx->h_mail + 8, md, sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
        MRB_CUDA(cudaStreamSynchronize(ctx->stream));
        unsigned bits = (unsigned)ctx->h_mail[8];
        memcpy(last_delta_host, &bits, sizeof(float));
    }
    return MRB_OK;
}

// ---- feature reset (quadprog.cpp:92-101 with getFeatureIndexes, primeqp.cpp:231-244) ---------------------
__global__ void k_feat_check_s(const float* __restrict__ fmap, const uint8_t* __restrict__ flags, size_t n, int* __restrict__ any_off) {
    MRB_VOXEL_LOOP(idx, n) if (fmap[idx] == 1.0f && !(flags[idx] & 1)) *any_off = 1;
}
__global__ void k_feat_apply_s(const float* __restrict__ fmap, const uint8_t* __restrict__ flags, const int* __restrict__ sid,
                               StripDims sd, const int* __restrict__ any_off, float* __restrict__ x /* float4 strips as floats */) {
    if (*any_off) return;
    MRB_VOXEL_LOOP(idx, (size_t)sd.nx * sd.ny * sd.nz) {
        if (fmap[idx] == 1.0f && (flags[idx] & 1)) {
            int z = (int)(idx % sd.nz);
            size_t col = idx / sd.nz;
            int s = sid[col * sd.nz4 + (z >> 2)];
            x[(size_t)s * 4 + (z & 3)] = 0.2f;
        }
    }
}

// ---- F assembly (quadprog.cpp:131-145) -----------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_assemble_F_s(const float* __restrict__ vol, const uint8_t* __restrict__ flags,
                                                      const int* __restrict__ sid, const float* __restrict__ x, StripDims sd,
                                                      double scale, float* __restrict__ F) {
    MRB_VOXEL_LOOP(idx, (size_t)sd.nx * sd.ny * sd.nz) {
        float f;
        if (flags[idx] & 1) {
            int z = (int)(idx % sd.nz);
            size_t col = idx / sd.nz;
            int s = sid[col * sd.nz4 + (z >> 2)];
            f = x[(size_t)s * 4 + (z & 3)];
        } else {
            f = (float)__dmul_rn(__dsub_rn(__dmul_rn(2.0, (double)vol[idx]), 1.0), scale);
        }
        F[idx] = f;
    }
}

int stage12_binary(mrb_ctx* ctx, const float* vol, Dims d, int W, uint8_t* sq, uint8_t* flags, int* counts, int* binary_out,
                   uint8_t** brick_bits_out);
int bricks_solve(mrb_ctx* ctx, const float* vol, const float* fmap, const uint8_t* flags, const uint8_t* sq, Dims d,
                 int band_size, int iters, int using_features, float* F_dev, float* last_delta_host, int* n_bricks,
                 int* closure_err_host, int binary_volume, const uint8_t* brick_bits);

// Stages 1-3 on device pointers.  F_dev: N^3 floats.  Records events ev[0..3] (+ ev[6..7] around the sweeps).
int optimize_strips(mrb_ctx* ctx, const float* vol, const float* fmap, Dims d, int using_features, int band_size,
                    int iters, float* F_dev, mrb_qp_stats* stats) {
    const size_t n = d.count();
    MRB_CUDA(cudaEventRecord(ctx->ev[0], ctx->stream));
    uint8_t *flags = nullptr;
    int* counts = ctx->d_mail + 32;
    MRB_CUDA(cudaMemsetAsync(counts, 0, 4 * sizeof(int), ctx->stream));
    StripBand sb;
    MRB_TRY(arena_alloc(ctx, n, &flags));
    if (band_size <= 15) {
        uint8_t *b0 = nullptr, *b1 = nullptr, *sq = nullptr;
        MRB_TRY(arena_alloc(ctx, n, &b0));
        MRB_TRY(arena_alloc(ctx, n, &b1));
        MRB_TRY(arena_alloc(ctx, n, &sq));
        const unsigned g = grid_for(ctx, n, 256);
        // binary occupancy grids (the normal case): bit-packed perimeter + tile-sparse passes (mrb_stage12.cu);
        // any other value in the grid -> the general byte-grid kernels below
        int binary = 0;
        uint8_t* brick_bits = nullptr;          // by-product of the bit-sliced tile kernel: which bricks hold a band voxel
        if (!getenv("MRB_STAGE12_GENERAL")) MRB_TRY(stage12_binary(ctx, vol, d, band_size, sq, flags, counts, &binary, &brick_bits));
        if (binary) {
            MRB_CUDA(cudaEventRecord(ctx->ev[1], ctx->stream));
        } else {
            MRB_CUDA(cudaMemsetAsync(counts, 0, 4 * sizeof(int), ctx->stream));
            MRB_LAUNCH(ctx, k_site, g, 256, 0, vol, d, b0);
            MRB_LAUNCH(ctx, k_tedt_x, g, 256, 0, b0, d, band_size, b1);
            MRB_LAUNCH(ctx, k_tedt_y, g, 256, 0, b1, d, band_size, b0);
            MRB_LAUNCH(ctx, k_tedt_z, g, 256, 0, b0, d, band_size, sq);
            MRB_CUDA(cudaEventRecord(ctx->ev[1], ctx->stream));
            MRB_LAUNCH(ctx, k_tight, g, 256, 0, sq, d, b1);
            MRB_LAUNCH(ctx, k_band_flags, g, 256, 0, b1, d, flags, counts);
        }
        // Layout of stage 3: 2x2x4 bricks (mrb_bricks.cu) for grids that fill the GPU; small grids are latency-bound (one
        // brick update per thread and sweep is a 2 us dependency chain), there the z-strip kernel with four times as many,
        // four times lighter threads finishes a sweep sooner (500 sweeps, strips / bricks: 92^3 1.4 / 2.1 ms, 128^3 1.8 / 2.0,
        // 144^3 1.9 / 2.0, 160^3 2.4 / 2.0, 172^3 2.5 / 2.0).  MRB_QP_LAYOUT=strips|bricks overrides.
        const char* layout = getenv("MRB_QP_LAYOUT");
        const bool use_strips = layout ? strcmp(layout, "strips") == 0 : n <= (size_t)150 * 150 * 150;
        if (!use_strips) {
            // default: 2x2x4 bricks (mrb_bricks.cu)
            MRB_CUDA(cudaMemcpyAsync(ctx->h_mail + 32, counts, 4 * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
            int Q = 0, closure_err = 0;
            float last_delta = 0.0f;
            ctx->h_mail[8] = 0;
            MRB_TRY(bricks_solve(ctx, vol, fmap, flags, sq, d, band_size, iters, using_features, F_dev,
                                 stats ? &last_delta : nullptr, &Q, &closure_err, binary, brick_bits));
            MRB_CUDA(cudaEventRecord(ctx->ev[3], ctx->stream));
            MRB_CUDA(cudaStreamSynchronize(ctx->stream));
            if (stats && last_delta == 0.0f) memcpy(&last_delta, &ctx->h_mail[8], sizeof(float));     // tiled kernel: copied with the stream
            if (ctx->h_mail[34] || closure_err) {
                set_error("narrow band touches the grid border (tight-band voxel on the border or with a neighbour outside "
                          "the band); the reference exits here (grid.cpp:20-37 via primeqp.cpp:40-75)");
                return MRB_EBORDER;
            }
            if (stats) {
                stats->n_band = ctx->h_mail[32];
                stats->n_tight = ctx->h_mail[33];
                stats->iters = iters;
                stats->last_max_delta = last_delta;
                cudaEventElapsedTime(&stats->ms_dfield, ctx->ev[0], ctx->ev[1]);
                stats->ms_mc = 0.0f;
                stats->ms_qp_kernel = 0.0f;
                stats->ms_band = 0.0f;
                if (iters > 0 && Q > 0) {
                    cudaEventElapsedTime(&stats->ms_qp_kernel, ctx->ev[6], ctx->ev[7]);
                    cudaEventElapsedTime(&stats->ms_band, ctx->ev[1], ctx->ev[6]);
                    cudaEventElapsedTime(&stats->ms_qp, ctx->ev[6], ctx->ev[3]);
                } else {
                    cudaEventElapsedTime(&stats->ms_qp, ctx->ev[1], ctx->ev[3]);
                }
            }
            return MRB_OK;
        }
        MRB_TRY(build_strips(ctx, vol, MarginSq{sq}, flags, d, &sb));
    } else {
        // wide bands: full exact EDT + float morphology, then the same strip machinery
        float *margin = nullptr, *b0 = nullptr, *tight = nullptr;
        MRB_TRY(arena_alloc(ctx, n, &margin));
        MRB_TRY(arena_alloc(ctx, n, &b0));
        MRB_TRY(arena_alloc(ctx, n, &tight));
        MRB_TRY(dev_dfield(ctx, vol, d, margin));
        MRB_CUDA(cudaEventRecord(ctx->ev[1], ctx->stream));
        MRB_TRY(dev_threshold(ctx, margin, n, (float)band_size, b0));
        MRB_TRY(dev_erode(ctx, b0, d, tight));
        MRB_TRY(dev_dilate(ctx, tight, d, b0));       // b0 now holds the band
        MRB_LAUNCH(ctx, k_flags_from_float, grid_for(ctx, n, 256), 256, 0, b0, tight, d, flags, counts);
        MRB_TRY(build_strips(ctx, vol, MarginF32{margin}, flags, d, &sb));
    }
    // counts: band, tight, border error; d_mail[24]: neighbour-closure error
    MRB_CUDA(cudaMemcpyAsync(ctx->h_mail + 32, counts, 4 * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    MRB_CUDA(cudaMemcpyAsync(ctx->h_mail + 24, ctx->d_mail + 24, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    MRB_CUDA(cudaEventRecord(ctx->ev[2], ctx->stream));
    MRB_CUDA(cudaStreamSynchronize(ctx->stream));
    sb.n_band = ctx->h_mail[32];
    sb.n_tight = ctx->h_mail[33];
    if (ctx->h_mail[34] || (sb.S > 0 && ctx->h_mail[24])) {
        set_error("narrow band touches the grid border (tight-band voxel on the border or with a neighbour outside the "
                  "band); the reference exits here (grid.cpp:20-37 via primeqp.cpp:40-75)");
        return MRB_EBORDER;
    }
    float4* xres = nullptr;
    float last_delta = 0.0f;
    bool in_cluster = false;
    xres = sb.xa;
    MRB_TRY(solve_strips_cluster(ctx, sb, iters, &xres, stats ? &last_delta : nullptr, &in_cluster));
    if (!in_cluster) MRB_TRY(solve_strips(ctx, sb, iters, &xres, stats ? &last_delta : nullptr));
    if (using_features && fmap && sb.S > 0) {
        int* flag = ctx->d_mail + 16;
        MRB_CUDA(cudaMemsetAsync(flag, 0, sizeof(int), ctx->stream));
        const unsigned g = grid_for(ctx, n, 256);
        MRB_LAUNCH(ctx, k_feat_check_s, g, 256, 0, fmap, flags, n, flag);
        MRB_LAUNCH(ctx, k_feat_apply_s, g, 256, 0, fmap, flags, sb.sid, sb.sd, flag, (float*)xres);
    }
    MRB_LAUNCH(ctx, k_assemble_F_s, grid_for(ctx, n, 256), 256, 0, vol, flags, sb.sid, (const float*)xres, sb.sd,
               (double)band_size + 1.0, F_dev);
    MRB_CUDA(cudaEventRecord(ctx->ev[3], ctx->stream));
    if (stats) {
        MRB_CUDA(cudaEventSynchronize(ctx->ev[3]));
        stats->n_band = sb.n_band;
        stats->n_tight = sb.n_tight;
        stats->iters = iters;
        stats->last_max_delta = last_delta;
        cudaEventElapsedTime(&stats->ms_dfield, ctx->ev[0], ctx->ev[1]);
        cudaEventElapsedTime(&stats->ms_band, ctx->ev[1], ctx->ev[2]);
        cudaEventElapsedTime(&stats->ms_qp, ctx->ev[2], ctx->ev[3]);
        stats->ms_mc = 0.0f;
        stats->ms_qp_kernel = 0.0f;
        if (iters > 0 && sb.S > 0) cudaEventElapsedTime(&stats->ms_qp_kernel, ctx->ev[6], ctx->ev[7]);
    }
    return MRB_OK;
}

}  // namespace mrb

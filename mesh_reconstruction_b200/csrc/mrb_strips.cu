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

#include <cooperative_groups.h>
namespace cg = cooperative_groups;

namespace mrb {

static const int kFar = 255;      // u8 "beyond the window"

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
                                                     float4* __restrict__ x0, int4* __restrict__ nbA, int4* __restrict__ nbB,
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
    float4 *lb4 = nullptr, *ub4 = nullptr;   // S
    float4 *xa = nullptr, *xb = nullptr;     // S+1 (slot S = zeros)
    int4 *nbA = nullptr, *nbB = nullptr;     // S
    int2* nbC = nullptr;                     // S
};

__constant__ float c_invdg_s[19];
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
    MRB_LAUNCH(ctx, k_fill_i32, grid_for(ctx, ns, 256), 256, 0, sb->sid, ns, -1);
    size_t Sn = (size_t)(S > 0 ? S : 1);
    MRB_TRY(arena_alloc(ctx, Sn, &sb->code));
    MRB_TRY(arena_alloc(ctx, Sn, &sb->lb4));
    MRB_TRY(arena_alloc(ctx, Sn, &sb->ub4));
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
               sb->lb4, sb->ub4, sb->xa, sb->nbA, sb->nbB, sb->nbC, err);
    return MRB_OK;
}

// ---- the persistent sweep kernel --------------------------------------------------------------------------
// One voxel update.  T bits: a = T(v); the six neighbour bits select the coefficients.
//   c(+-e)  = a ? (t ? -4 : -2) : (t ? -2 : 0)      c(+-2e) = t ? 1 : 0
__device__ __forceinline__ float qp_voxel(float xc, float xm2, float xm1, float ym2, float ym1, float zm2, float zm1,
                                          float zp1, float zp2, float yp1, float yp2, float xp1, float xp2, bool a, bool txm,
                                          bool txp, bool tym, bool typ, bool tzm, bool tzp, float lb, float ub) {
    const float cA = a ? -4.0f : -2.0f;      // neighbour tight
    const float cB = a ? -2.0f : 0.0f;       // neighbour not tight
    float res = 0.0f;
    if (txm) res = __fadd_rn(res, xm2);
    res = __fmaf_rn(txm ? cA : cB, xm1, res);
    if (tym) res = __fadd_rn(res, ym2);
    res = __fmaf_rn(tym ? cA : cB, ym1, res);
    if (tzm) res = __fadd_rn(res, zm2);
    res = __fmaf_rn(tzm ? cA : cB, zm1, res);
    res = __fmaf_rn(tzp ? cA : cB, zp1, res);
    if (tzp) res = __fadd_rn(res, zp2);
    res = __fmaf_rn(typ ? cA : cB, yp1, res);
    if (typ) res = __fadd_rn(res, yp2);
    res = __fmaf_rn(txp ? cA : cB, xp1, res);
    if (txp) res = __fadd_rn(res, xp2);
    const int dg = (a ? 12 : 0) + (int)txm + (int)txp + (int)tym + (int)typ + (int)tzm + (int)tzp;
    float t = __fsub_rn(xc, __fmul_rn(res, c_invdg_s[dg]));
    t = t * 0.5f;
    if (t < lb) t = lb;
    if (t > ub) t = ub;
    return t;
}

__device__ __forceinline__ float4 ld4(const float4* p) { return *p; }

__global__ void __launch_bounds__(512) k_qp_strips(const uint32_t* __restrict__ code, const float4* __restrict__ lb4,
                                                   const float4* __restrict__ ub4, const int4* __restrict__ nbA,
                                                   const int4* __restrict__ nbB, const int2* __restrict__ nbC, float4* xa,
                                                   float4* xb, int S, int iters, unsigned* __restrict__ maxdelta_bits) {
    cg::grid_group grid = cg::this_grid();
    const int gsize = gridDim.x * blockDim.x;
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
    const float4* in = xa;
    float4* out = xb;
    for (int it = 0; it < iters; it++) {
        float dmax = 0.0f;
        for (int s = gtid; s < S; s += gsize) {
            const int4 a4 = __ldg(nbA + s);
            const int4 b4 = __ldg(nbB + s);
            const int2 c2 = __ldg(nbC + s);
            const uint32_t c = __ldg(code + s);
            const float4 lb = __ldg(lb4 + s), ub = __ldg(ub4 + s);
            // the iterate is rewritten by other SMs every sweep: plain (coherent after grid.sync) loads
            const float4 xc = ld4(in + s);
            const float4 xm2 = ld4(in + a4.x), xm1 = ld4(in + a4.y), ym2 = ld4(in + a4.z), ym1 = ld4(in + a4.w);
            const float4 zlo = ld4(in + b4.x), zhi = ld4(in + b4.y), yp1 = ld4(in + b4.z), yp2 = ld4(in + b4.w);
            const float4 xp1 = ld4(in + c2.x), xp2 = ld4(in + c2.y);
            const bool t0 = c & 1u, t1 = c & 2u, t2 = c & 4u, t3 = c & 8u, tlo = c & 16u, thi = c & 32u;
            float4 r;
            r.x = qp_voxel(xc.x, xm2.x, xm1.x, ym2.x, ym1.x, zlo.z, zlo.w, xc.y, xc.z, yp1.x, yp2.x, xp1.x, xp2.x, t0,
                           c & (1u << 6), c & (1u << 10), c & (1u << 14), c & (1u << 18), tlo, t1, lb.x, ub.x);
            r.y = qp_voxel(xc.y, xm2.y, xm1.y, ym2.y, ym1.y, zlo.w, xc.x, xc.z, xc.w, yp1.y, yp2.y, xp1.y, xp2.y, t1,
                           c & (1u << 7), c & (1u << 11), c & (1u << 15), c & (1u << 19), t0, t2, lb.y, ub.y);
            r.z = qp_voxel(xc.z, xm2.z, xm1.z, ym2.z, ym1.z, xc.x, xc.y, xc.w, zhi.x, yp1.z, yp2.z, xp1.z, xp2.z, t2,
                           c & (1u << 8), c & (1u << 12), c & (1u << 16), c & (1u << 20), t1, t3, lb.z, ub.z);
            r.w = qp_voxel(xc.w, xm2.w, xm1.w, ym2.w, ym1.w, xc.y, xc.z, zhi.x, zhi.y, yp1.w, yp2.w, xp1.w, xp2.w, t3,
                           c & (1u << 9), c & (1u << 13), c & (1u << 17), c & (1u << 21), t2, thi, lb.w, ub.w);
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
        if (it + 1 < iters) grid.sync();
    }
}

static int solve_strips(mrb_ctx* ctx, const StripBand& sb, int iters, float4** result, float* last_delta_host) {
    const int S = sb.S;
    *result = sb.xa;
    if (last_delta_host) *last_delta_host = 0.0f;
    if (S == 0 || iters == 0) return MRB_OK;
    unsigned* md = (unsigned*)(ctx->d_mail + 8);
    MRB_CUDA(cudaMemsetAsync(md, 0, sizeof(unsigned), ctx->stream));
    int per_sm = 0;
    const int block = 512;
    MRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_qp_strips, block, 0));
    if (per_sm < 1) { set_error("k_qp_strips does not fit on an SM"); return MRB_ECUDA; }
    int grid = per_sm * ctx->sm_count;
    int need = (S + block - 1) / block;
    if (grid > need) grid = need;
    const uint32_t* code = sb.code;
    const float4 *lb4 = sb.lb4, *ub4 = sb.ub4;
    const int4 *nbA = sb.nbA, *nbB = sb.nbB;
    const int2* nbC = sb.nbC;
    float4 *xa = sb.xa, *xb = sb.xb;
    int SS = S, its = iters;
    void* args[] = {(void*)&code, (void*)&lb4, (void*)&ub4, (void*)&nbA, (void*)&nbB, (void*)&nbC,
                    (void*)&xa, (void*)&xb, (void*)&SS, (void*)&its, (void*)&md};
    MRB_CUDA(cudaEventRecord(ctx->ev[6], ctx->stream));
    MRB_CUDA(cudaLaunchCooperativeKernel((void*)k_qp_strips, dim3(grid), dim3(block), args, 0, ctx->stream));
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
        MRB_LAUNCH(ctx, k_site, g, 256, 0, vol, d, b0);
        MRB_LAUNCH(ctx, k_tedt_x, g, 256, 0, b0, d, band_size, b1);
        MRB_LAUNCH(ctx, k_tedt_y, g, 256, 0, b1, d, band_size, b0);
        MRB_LAUNCH(ctx, k_tedt_z, g, 256, 0, b0, d, band_size, sq);
        MRB_CUDA(cudaEventRecord(ctx->ev[1], ctx->stream));
        MRB_LAUNCH(ctx, k_tight, g, 256, 0, sq, d, b1);
        MRB_LAUNCH(ctx, k_band_flags, g, 256, 0, b1, d, flags, counts);
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
    MRB_TRY(solve_strips(ctx, sb, iters, &xres, stats ? &last_delta : nullptr));
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

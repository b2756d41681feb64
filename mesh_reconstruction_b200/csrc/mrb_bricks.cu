// mrb_bricks.cu -- stage 3 on 2x2x4 voxel bricks: the sweep kernel of mrb_optimize / mrb_reconstruct.
//
// Same arithmetic as the z-strip kernel in mrb_strips.cu (doStep, optimize/quadprog.cpp:45-61, on the
// matrix-free operator of primeqp.cpp:7-112; see the header of mrb_strips.cu for the bit-exactness
// argument), different data layout.  Profiling the strip kernel on B200 (profiles/r01_summary.md)
// showed it bound by L1 data-pipe wavefronts, not by HBM or L2: every 4-voxel strip pulled 11 float4
// of the iterate, 10 neighbour ids and 52 B of constants through the LSU.  A brick is 2(x) x 2(y) x
// 4(z) voxels aligned to even x, even y and z % 4 == 0, stored as 4 consecutive float4 (64 B):
//   * the radius-2 star of all 16 voxels reaches exactly ONE brick in each of the six directions, so
//     a brick needs 6 neighbour ids (24 B) instead of 40 B per strip, and
//   * the in-brick x/y neighbours are register values: 20 float4 + 8 float2 loads per 16 voxels
//     instead of 44 float4 for four strips;
//   * the per-voxel constants shrink to one 16-bit field (squared distance, diagonal d, volume==0,
//     volume==1) decoded through two small constant-memory tables: margin = sqrtf(sq) and
//     invdg = (float)(1.0/(double)d) -- the same values getsqrt (dfields.cpp:187) and primeQP
//     (primeqp.cpp:196) compute.
// Non-band voxels inside a brick have all tight bits 0, d = 0 (invdg 0), lb = -1000, ub = 1000 and
// therefore stay exactly 0; they are never read with a non-zero coefficient.
#include "mrb_internal.cuh"
#include "mrb_compact.cuh"

#include <vector>
#include <algorithm>

namespace mrb {

#define MRB_VOXEL_LOOP(idx, n) \
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < (n); idx += (size_t)gridDim.x * blockDim.x)

static unsigned grid_for(mrb_ctx* ctx, size_t n, int block) {
    size_t want = (n + block - 1) / block;
    size_t cap = (size_t)ctx->sm_count * 16;
    return (unsigned)(want < cap ? want : cap);
}

__constant__ float c_brick_sqrt[256];    // (float)sqrt((double)i): correctly rounded sqrtf(i)
__constant__ float c_brick_invdg[32];    // (float)(1.0/(double)d), d = 1..18; 0 elsewhere

static int upload_tables(int device) {
    static bool ready[64] = {false};
    if (ready[device & 63]) return MRB_OK;
    float s[256], v[32];
    for (int i = 0; i < 256; i++) s[i] = (float)sqrt((double)i);
    for (int d = 0; d < 32; d++) v[d] = (d >= 1 && d <= 18) ? (float)(1.0 / (double)(float)d) : 0.0f;
    MRB_CUDA(cudaMemcpyToSymbol(c_brick_sqrt, s, sizeof(s)));
    MRB_CUDA(cudaMemcpyToSymbol(c_brick_invdg, v, sizeof(v)));
    ready[device & 63] = true;
    return MRB_OK;
}

struct BrickDims {
    int nx, ny, nz;     // voxel grid
    int bx, by, bz;     // brick grid = ceil(nx/2), ceil(ny/2), ceil(nz/4)
    __host__ __device__ size_t count() const { return (size_t)bx * by * bz; }
};

struct ActiveBrickPred {
    const uint8_t* flags;
    BrickDims bd;
    __device__ bool operator()(size_t p) const {
        int Z = (int)(p % bd.bz);
        size_t t = p / bd.bz;
        int Y = (int)(t % bd.by);
        int X = (int)(t / bd.by);
        bool any = false;
        for (int dx = 0; dx < 2; dx++)
            for (int dy = 0; dy < 2; dy++) {
                int x = 2 * X + dx, y = 2 * Y + dy;
                if (x >= bd.nx || y >= bd.ny) continue;
                const uint8_t* row = flags + ((size_t)x * bd.ny + y) * bd.nz;
                for (int j = 0; j < 4; j++) {
                    int z = 4 * Z + j;
                    if (z < bd.nz) any = any || (row[z] & 1);
                }
            }
        return any;
    }
};

__global__ void k_fill_i32b(int* __restrict__ p, size_t n, int v) {
    MRB_VOXEL_LOOP(idx, n) p[idx] = v;
}
__global__ void k_scatter_brick_ids(const int* __restrict__ bpos, int Q, int* __restrict__ bid) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < Q) bid[bpos[i]] = i;
}

// Voxel v = (dx*2+dy)*4 + j of brick (X,Y,Z) is (2X+dx, 2Y+dy, 4Z+j).
// code.x: bits 0..15 T(v);  16..23 T(2X-1, y, z) [dy*4+j];  24..31 T(2X+2, y, z) [dy*4+j]
// code.y: bits 0..7  T(x, 2Y-1, z) [dx*4+j];  8..15 T(x, 2Y+2, z) [dx*4+j];
//         16..19 T(x, y, 4Z-1) [dx*2+dy];     20..23 T(x, y, 4Z+4) [dx*2+dy]
// field (u16 per voxel): bits 0..7 squared distance, 8..12 d = 12 T(v) + six neighbour T,
//         13 volume==0 (lb=-1000), 14 volume==1 (ub=1000), 15 in band
__global__ void __launch_bounds__(128) k_brick_build(const float* __restrict__ vol, const uint8_t* __restrict__ sq,
                                                     const uint8_t* __restrict__ flags, const int* __restrict__ bid,
                                                     const int* __restrict__ bpos, int Q, BrickDims bd,
                                                     uint2* __restrict__ code, uint4* __restrict__ fields,
                                                     float4* __restrict__ x0, int4* __restrict__ nbA, int2* __restrict__ nbB,
                                                     int* __restrict__ err, size_t P) {
    // P = plane stride of the SoA iterate: x[k][q], k = dx*2+dy (>= Q + 1; slot Q is the all-zero dummy brick)
    int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= Q) return;
    int p = bpos[q];
    int Z = p % bd.bz;
    int t = p / bd.bz;
    int Y = t % bd.by;
    int X = t / bd.by;
    auto flag_at = [&](int x, int y, int z) -> uint32_t {
        if (x < 0 || y < 0 || z < 0 || x >= bd.nx || y >= bd.ny || z >= bd.nz) return 0u;
        return flags[((size_t)x * bd.ny + y) * bd.nz + z];
    };
    auto T = [&](int x, int y, int z) -> uint32_t { return (flag_at(x, y, z) >> 1) & 1u; };
    uint32_t lo = 0, hi = 0;
    uint32_t f16[16];
    float xv[16];
    bool bad = false;
    for (int dx = 0; dx < 2; dx++)
        for (int dy = 0; dy < 2; dy++)
            for (int j = 0; j < 4; j++) {
                const int k = dx * 2 + dy, v = k * 4 + j;
                const int x = 2 * X + dx, y = 2 * Y + dy, z = 4 * Z + j;
                const uint32_t fl = flag_at(x, y, z);
                const uint32_t t0 = (fl >> 1) & 1u;
                const uint32_t txm = T(x - 1, y, z), txp = T(x + 1, y, z), tym = T(x, y - 1, z), typ = T(x, y + 1, z);
                const uint32_t tzm = T(x, y, z - 1), tzp = T(x, y, z + 1);
                lo |= t0 << v;
                if (dx == 0) lo |= txm << (16 + dy * 4 + j); else lo |= txp << (24 + dy * 4 + j);
                if (dy == 0) hi |= tym << (dx * 4 + j); else hi |= typ << (8 + dx * 4 + j);
                if (j == 0) hi |= tzm << (16 + k);
                if (j == 3) hi |= tzp << (20 + k);
                if (t0) {
                    // a tight voxel needs its six neighbours inside the grid and in the band
                    // (reference: grid.cpp:20-37 exit / indexMap -1 column, primeqp.cpp:40-75)
                    bool in = x >= 1 && y >= 1 && z >= 1 && x <= bd.nx - 2 && y <= bd.ny - 2 && z <= bd.nz - 2;
                    if (!in) bad = true;
                    else bad = bad || !(flag_at(x - 1, y, z) & 1) || !(flag_at(x + 1, y, z) & 1) || !(flag_at(x, y - 1, z) & 1) ||
                               !(flag_at(x, y + 1, z) & 1) || !(flag_at(x, y, z - 1) & 1) || !(flag_at(x, y, z + 1) & 1);
                }
                uint32_t f = (1u << 13) | (1u << 14);      // not in band: lb=-1000, ub=1000, d=0, margin 0
                float xx = 0.0f;
                if (fl & 1u) {
                    const size_t a = ((size_t)x * bd.ny + y) * bd.nz + z;
                    const uint32_t s = sq[a];
                    const uint32_t dg = 12u * t0 + txm + txp + tym + typ + tzm + tzp;
                    const float vv = vol[a];
                    const float mg = c_brick_sqrt[s];
                    f = s | (dg << 8) | (1u << 15);
                    if (vv == 0.0f) f |= 1u << 13;          // getlb, primeqp.cpp:122-127
                    if (vv == 1.0f) f |= 1u << 14;          // getub, primeqp.cpp:148-153
                    const float l = (vv == 0.0f) ? -1000.0f : mg;
                    const float u = (vv == 1.0f) ? 1000.0f : -mg;
                    if (xx < l) xx = l;                     // x0 = clamp(0, lb, ub), primeqp.cpp:178-187
                    if (xx > u) xx = u;
                }
                f16[v] = f;
                xv[v] = xx;
            }
    if (bad) atomicExch(err, 1);
    code[q] = make_uint2(lo, hi);
    uint32_t w[8];
    for (int i = 0; i < 8; i++) w[i] = f16[2 * i] | (f16[2 * i + 1] << 16);
    fields[q] = make_uint4(w[0], w[1], w[2], w[3]);
    fields[(size_t)Q + q] = make_uint4(w[4], w[5], w[6], w[7]);
    for (int k = 0; k < 4; k++) x0[k * P + q] = make_float4(xv[4 * k], xv[4 * k + 1], xv[4 * k + 2], xv[4 * k + 3]);
    auto nid = [&](bool ok, long long dp) {
        int id = ok ? bid[p + dp] : -1;
        return id >= 0 ? id : Q;             // slot Q = the all-zero dummy brick
    };
    const long long cx = (long long)bd.by * bd.bz, cy = bd.bz;
    nbA[q] = make_int4(nid(X >= 1, -cx), nid(X <= bd.bx - 2, cx), nid(Y >= 1, -cy), nid(Y <= bd.by - 2, cy));
    nbB[q] = make_int2(nid(Z >= 1, -1), nid(Z <= bd.bz - 2, 1));
}

// ---- sweep ---------------------------------------------------------------------------------------------------
struct BrickRegs {
    float own[4][4];      // [k = dx*2+dy][j]
    float xm[4][4], xp[4][4], ym[4][4], yp[4][4];   // the four full neighbour bricks
    float zl[4][2], zh[4][2];                       // last two z of brick Z-1, first two z of brick Z+1
};

__device__ __forceinline__ void put4(float (&d)[4], const float4 v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w; }

template <int DX, int DY, int J>
__device__ __forceinline__ float brick_voxel(const BrickRegs& r, uint32_t lo, uint32_t hi, uint32_t field,
                                             const float* __restrict__ s_sqrt, const float* __restrict__ s_inv) {
    constexpr int K = DX * 2 + DY, V = K * 4 + J;
    const bool a = lo & (1u << V);
    const bool txm = DX == 0 ? (lo & (1u << (16 + DY * 4 + J))) : (lo & (1u << ((0 * 2 + DY) * 4 + J)));
    const bool txp = DX == 0 ? (lo & (1u << ((1 * 2 + DY) * 4 + J))) : (lo & (1u << (24 + DY * 4 + J)));
    const bool tym = DY == 0 ? (hi & (1u << (DX * 4 + J))) : (lo & (1u << ((DX * 2 + 0) * 4 + J)));
    const bool typ = DY == 0 ? (lo & (1u << ((DX * 2 + 1) * 4 + J))) : (hi & (1u << (8 + DX * 4 + J)));
    const bool tzm = J > 0 ? (lo & (1u << (V > 0 ? V - 1 : 0))) : (hi & (1u << (16 + K)));
    const bool tzp = J < 3 ? (lo & (1u << (V + 1))) : (hi & (1u << (20 + K)));
    const float xm2 = DX == 0 ? r.xm[0 * 2 + DY][J] : r.xm[1 * 2 + DY][J];
    const float xm1 = DX == 0 ? r.xm[1 * 2 + DY][J] : r.own[0 * 2 + DY][J];
    const float xp1 = DX == 0 ? r.own[1 * 2 + DY][J] : r.xp[0 * 2 + DY][J];
    const float xp2 = DX == 0 ? r.xp[0 * 2 + DY][J] : r.xp[1 * 2 + DY][J];
    const float ym2 = DY == 0 ? r.ym[DX * 2 + 0][J] : r.ym[DX * 2 + 1][J];
    const float ym1 = DY == 0 ? r.ym[DX * 2 + 1][J] : r.own[DX * 2 + 0][J];
    const float yp1 = DY == 0 ? r.own[DX * 2 + 1][J] : r.yp[DX * 2 + 0][J];
    const float yp2 = DY == 0 ? r.yp[DX * 2 + 0][J] : r.yp[DX * 2 + 1][J];
    const float zm2 = J >= 2 ? r.own[K][J >= 2 ? J - 2 : 0] : r.zl[K][J < 2 ? J : 0];
    const float zm1 = J >= 1 ? r.own[K][J >= 1 ? J - 1 : 0] : r.zl[K][1];
    const float zp1 = J <= 2 ? r.own[K][J <= 2 ? J + 1 : 0] : r.zh[K][0];
    const float zp2 = J <= 1 ? r.own[K][J <= 1 ? J + 2 : 0] : r.zh[K][J >= 2 ? J - 2 : 0];
    const float xc = r.own[K][J];
    // c(+-e) = T(v) ? (t ? -4 : -2) : (t ? -2 : 0),  c(+-2e) = t ? 1 : 0; ascending band id order
    const float cA = a ? -4.0f : -2.0f;
    const float cB = a ? -2.0f : 0.0f;
    float res = txm ? xm2 : 0.0f;
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
    const float m = s_sqrt[field & 0xffu];
    const float inv = s_inv[(field >> 8) & 31u];
    const float lb = (field & (1u << 13)) ? -1000.0f : m;
    const float ub = (field & (1u << 14)) ? 1000.0f : -m;
    float t = __fsub_rn(xc, __fmul_rn(res, inv));
    t = t * 0.5f;
    // "if (t < lb) t = lb; if (t > ub) t = ub;" (quadprog.cpp:58-59) == min(max(t, lb), ub) for lb <= ub and no
    // NaN; the two forms can only differ in the sign of a zero result, which no later stage can observe
    // (see mrb_strips.cu header: res never becomes -0, val < iso and the interpolation treat +-0 alike).
    return fminf(fmaxf(t, lb), ub);
}

template <int K>
__device__ __forceinline__ float4 brick_strip(const BrickRegs& r, uint32_t lo, uint32_t hi, uint32_t w0, uint32_t w1,
                                              const float* __restrict__ s_sqrt, const float* __restrict__ s_inv) {
    constexpr int DX = K >> 1, DY = K & 1;
    float4 o;
    o.x = brick_voxel<DX, DY, 0>(r, lo, hi, w0 & 0xffffu, s_sqrt, s_inv);
    o.y = brick_voxel<DX, DY, 1>(r, lo, hi, w0 >> 16, s_sqrt, s_inv);
    o.z = brick_voxel<DX, DY, 2>(r, lo, hi, w1 & 0xffffu, s_sqrt, s_inv);
    o.w = brick_voxel<DX, DY, 3>(r, lo, hi, w1 >> 16, s_sqrt, s_inv);
    return o;
}

// Grid barrier: one arrival counter that only grows during a launch (see mrb_strips.cu).
__device__ __forceinline__ void brick_grid_barrier(unsigned* bar, unsigned epoch, unsigned nblocks) {
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

// One Jacobi update of brick q: reads the iterate planes `in` (stride P), writes `out`.  Returns max |delta|.
__device__ __forceinline__ float brick_update(int q, size_t P, const int4 na, const int2 nz, const uint2 c, const uint4 f0,
                                              const uint4 f1, const float4* in, float4* out, const float* __restrict__ s_sqrt,
                                              const float* __restrict__ s_inv, bool want_delta) {
    BrickRegs r;
    // the iterate is rewritten by other SMs every sweep: plain loads, ordered by the sweep synchronisation
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const float4* pl = in + k * P;
        put4(r.own[k], pl[q]);
        put4(r.xm[k], pl[na.x]);
        put4(r.xp[k], pl[na.y]);
        put4(r.ym[k], pl[na.z]);
        put4(r.yp[k], pl[na.w]);
        const float2 l = *reinterpret_cast<const float2*>(reinterpret_cast<const float*>(pl + nz.x) + 2);
        const float2 h = *reinterpret_cast<const float2*>(pl + nz.y);
        r.zl[k][0] = l.x; r.zl[k][1] = l.y;
        r.zh[k][0] = h.x; r.zh[k][1] = h.y;
    }
    const float4 o0 = brick_strip<0>(r, c.x, c.y, f0.x, f0.y, s_sqrt, s_inv);
    const float4 o1 = brick_strip<1>(r, c.x, c.y, f0.z, f0.w, s_sqrt, s_inv);
    const float4 o2 = brick_strip<2>(r, c.x, c.y, f1.x, f1.y, s_sqrt, s_inv);
    const float4 o3 = brick_strip<3>(r, c.x, c.y, f1.z, f1.w, s_sqrt, s_inv);
    out[q] = o0; out[P + q] = o1; out[2 * P + q] = o2; out[3 * P + q] = o3;
    float dmax = 0.0f;
    if (want_delta) {
        const float4 os[4] = {o0, o1, o2, o3};
#pragma unroll
        for (int k = 0; k < 4; k++)
            dmax = fmaxf(fmaxf(dmax, fmaxf(fabsf(os[k].x - r.own[k][0]), fabsf(os[k].y - r.own[k][1]))),
                         fmaxf(fabsf(os[k].z - r.own[k][2]), fabsf(os[k].w - r.own[k][3])));
    }
    return dmax;
}

// ---- expanded per-brick constants (shared-memory resident) --------------------------------------------------
// When a block's bricks fit (and the volume is binary), the sweep kernel decodes their constants ONCE into shared
// memory: the signed margin and invdg as floats and one byte of seven tight bits per voxel (self, x-, y-, z-, z+, y+,
// x+ : the order of use).  The sweep then spends no instructions on field decoding, table look-ups or bit gathering
// (about 700 -> 500 instructions per brick update); the global layout (64 B per brick) is unchanged.
// Signed margin s: +margin where volume == 1 (lb = margin, ub = 1000), -margin where volume == 0 (lb = -1000,
// ub = -margin; -0.0f on the perimeter itself), so one float carries both bounds (getlb/getub, primeqp.cpp:115-165):
//   volume == 1:  min(max(t,  s), 1000)          volume == 0:  min(max(t, -1000), s) == max(min(t, s), -1000)
// Voxels of a brick outside the band get s = +0 (invdg = 0 keeps them at exactly 0).  A voxel whose volume is neither
// 0 nor 1 has lb = margin > ub = -margin, which this encoding cannot express: such grids use the 64-byte form.
struct BrickConstsX {
    float4 sm[4], inv[4];            // [k = dx*2+dy], components j = 0..3
    uint4 pred;                      // word k, byte j
};

__device__ __forceinline__ uint32_t brick_pred_byte(int DX, int DY, int J, uint32_t lo, uint32_t hi) {
    const int K = DX * 2 + DY, V = K * 4 + J;
    auto L = [&](int b) { return (lo >> b) & 1u; };
    auto H = [&](int b) { return (hi >> b) & 1u; };
    const uint32_t a = L(V);
    const uint32_t txm = DX == 0 ? L(16 + DY * 4 + J) : L((0 * 2 + DY) * 4 + J);
    const uint32_t txp = DX == 0 ? L((1 * 2 + DY) * 4 + J) : L(24 + DY * 4 + J);
    const uint32_t tym = DY == 0 ? H(DX * 4 + J) : L((DX * 2 + 0) * 4 + J);
    const uint32_t typ = DY == 0 ? L((DX * 2 + 1) * 4 + J) : H(8 + DX * 4 + J);
    const uint32_t tzm = J > 0 ? L(V - 1) : H(16 + K);
    const uint32_t tzp = J < 3 ? L(V + 1) : H(20 + K);
    return a | (txm << 1) | (tym << 2) | (tzm << 3) | (tzp << 4) | (typ << 5) | (txp << 6);
}

__device__ __forceinline__ BrickConstsX brick_expand(const uint2 c, const uint4 f0, const uint4 f1, const float* __restrict__ s_sqrt,
                                                     const float* __restrict__ s_inv) {
    BrickConstsX x;
    const uint32_t w[8] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
    uint32_t pw[4];
    float sm[16], iv[16];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        pw[k] = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int v = k * 4 + j;
            const uint32_t field = (w[v >> 1] >> ((v & 1) * 16)) & 0xffffu;
            const float m = s_sqrt[field & 0xffu];
            iv[v] = s_inv[(field >> 8) & 31u];
            // band voxel of a binary volume: exactly one of bit 13 (volume == 0) / bit 14 (volume == 1)
            sm[v] = (field & (1u << 15)) ? ((field & (1u << 13)) ? -m : m) : 0.0f;
            // bit 7: band voxel of the outside (volume == 0): its bounds are [-1000, sm] instead of [sm, 1000]
            const uint32_t outside = ((field >> 15) & (field >> 13) & 1u) << 7;
            pw[k] |= (brick_pred_byte(k >> 1, k & 1, j, c.x, c.y) | outside) << (8 * j);
        }
        x.sm[k] = make_float4(sm[k * 4], sm[k * 4 + 1], sm[k * 4 + 2], sm[k * 4 + 3]);
        x.inv[k] = make_float4(iv[k * 4], iv[k * 4 + 1], iv[k * 4 + 2], iv[k * 4 + 3]);
    }
    x.pred = make_uint4(pw[0], pw[1], pw[2], pw[3]);
    return x;
}

template <int DX, int DY, int J>
__device__ __forceinline__ float brick_voxel_x(const BrickRegs& r, uint32_t pw, float sm, float inv) {
    constexpr int K = DX * 2 + DY;
    const uint32_t pb = pw >> (8 * J);
    const bool a = pb & 1u, txm = pb & 2u, tym = pb & 4u, tzm = pb & 8u, tzp = pb & 16u, typ = pb & 32u, txp = pb & 64u;
    const float xm2 = DX == 0 ? r.xm[0 * 2 + DY][J] : r.xm[1 * 2 + DY][J];
    const float xm1 = DX == 0 ? r.xm[1 * 2 + DY][J] : r.own[0 * 2 + DY][J];
    const float xp1 = DX == 0 ? r.own[1 * 2 + DY][J] : r.xp[0 * 2 + DY][J];
    const float xp2 = DX == 0 ? r.xp[0 * 2 + DY][J] : r.xp[1 * 2 + DY][J];
    const float ym2 = DY == 0 ? r.ym[DX * 2 + 0][J] : r.ym[DX * 2 + 1][J];
    const float ym1 = DY == 0 ? r.ym[DX * 2 + 1][J] : r.own[DX * 2 + 0][J];
    const float yp1 = DY == 0 ? r.own[DX * 2 + 1][J] : r.yp[DX * 2 + 0][J];
    const float yp2 = DY == 0 ? r.yp[DX * 2 + 0][J] : r.yp[DX * 2 + 1][J];
    const float zm2 = J >= 2 ? r.own[K][J >= 2 ? J - 2 : 0] : r.zl[K][J < 2 ? J : 0];
    const float zm1 = J >= 1 ? r.own[K][J >= 1 ? J - 1 : 0] : r.zl[K][1];
    const float zp1 = J <= 2 ? r.own[K][J <= 2 ? J + 1 : 0] : r.zh[K][0];
    const float zp2 = J <= 1 ? r.own[K][J <= 1 ? J + 2 : 0] : r.zh[K][J >= 2 ? J - 2 : 0];
    const float xc = r.own[K][J];
    const float cA = a ? -4.0f : -2.0f;       // same arithmetic, same order as brick_voxel
    const float cB = a ? -2.0f : 0.0f;
    float res = txm ? xm2 : 0.0f;
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
    float t = __fsub_rn(xc, __fmul_rn(res, inv));
    t = t * 0.5f;
    // clamp to [lb, ub] from the signed margin (see BrickConstsX); differs from "if (t < lb) t = lb; if (t > ub) t = ub"
    // at most in the sign of a zero, as in brick_voxel
    // [lb, ub] = volume == 0 ? [-1000, sm] : [sm, 1000]; min(max(t, lb), ub) is one three-input FMNMX3 on sm_100
    const bool outside = pb & 128u;
    const float lb = outside ? -1000.0f : sm, ub = outside ? sm : 1000.0f;
    return fminf(fmaxf(t, lb), ub);
}

template <int K>
__device__ __forceinline__ float4 brick_strip_x(const BrickRegs& r, uint32_t pw, const float4 sm, const float4 inv) {
    constexpr int DX = K >> 1, DY = K & 1;
    float4 o;
    o.x = brick_voxel_x<DX, DY, 0>(r, pw, sm.x, inv.x);
    o.y = brick_voxel_x<DX, DY, 1>(r, pw, sm.y, inv.y);
    o.z = brick_voxel_x<DX, DY, 2>(r, pw, sm.z, inv.z);
    o.w = brick_voxel_x<DX, DY, 3>(r, pw, sm.w, inv.w);
    return o;
}

__device__ __forceinline__ void brick_load(BrickRegs& r, int q, size_t P, const int4 na, const int2 nz, const float4* in) {
    // the iterate is rewritten by other SMs every sweep: plain loads, ordered by the sweep synchronisation
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const float4* pl = in + k * P;
        put4(r.own[k], pl[q]);
        put4(r.xm[k], pl[na.x]);
        put4(r.xp[k], pl[na.y]);
        put4(r.ym[k], pl[na.z]);
        put4(r.yp[k], pl[na.w]);
        const float2 l = *reinterpret_cast<const float2*>(reinterpret_cast<const float*>(pl + nz.x) + 2);
        const float2 h = *reinterpret_cast<const float2*>(pl + nz.y);
        r.zl[k][0] = l.x; r.zl[k][1] = l.y;
        r.zh[k][0] = h.x; r.zh[k][1] = h.y;
    }
}

// Same as brick_update with the constants of brick q read from the expanded shared-memory arrays (slot l, stride cap).
__device__ __forceinline__ float brick_update_x(int q, size_t P, int l, int cap, const float4* __restrict__ s_sm,
                                                const float4* __restrict__ s_iv, const uint4* __restrict__ s_pred, const int4 na,
                                                const int2 nz, const float4* in, float4* out, bool want_delta) {
    BrickRegs r;
    brick_load(r, q, P, na, nz, in);
    const uint4 pw = s_pred[l];
    const float4 o0 = brick_strip_x<0>(r, pw.x, s_sm[l], s_iv[l]);
    const float4 o1 = brick_strip_x<1>(r, pw.y, s_sm[cap + l], s_iv[cap + l]);
    const float4 o2 = brick_strip_x<2>(r, pw.z, s_sm[2 * cap + l], s_iv[2 * cap + l]);
    const float4 o3 = brick_strip_x<3>(r, pw.w, s_sm[3 * cap + l], s_iv[3 * cap + l]);
    out[q] = o0; out[P + q] = o1; out[2 * P + q] = o2; out[3 * P + q] = o3;
    float dmax = 0.0f;
    if (want_delta) {
        const float4 os[4] = {o0, o1, o2, o3};
#pragma unroll
        for (int k = 0; k < 4; k++)
            dmax = fmaxf(fmaxf(dmax, fmaxf(fabsf(os[k].x - r.own[k][0]), fabsf(os[k].y - r.own[k][1]))),
                         fmaxf(fabsf(os[k].z - r.own[k][2]), fabsf(os[k].w - r.own[k][3])));
    }
    return dmax;
}

// ---- neighbour-only synchronisation --------------------------------------------------------------------------
// Block b owns bricks [Q*b/G, Q*(b+1)/G).  Its sweep k+1 reads bricks owned by a few other blocks only
// (bricks are ordered (X,Y,Z): the X+-1 neighbours sit about one x-plane of bricks away), so instead of
// a grid-wide barrier every block publishes "sweep k done" in its own progress word and waits for the
// progress words of exactly the blocks [lo_b, hi_b] that own one of its neighbour bricks.  The same
// condition covers the write-after-read hazard of the ping-pong buffers (a neighbour that finished
// sweep k no longer reads the buffer sweep k+1 overwrites).  All blocks are co-resident (cooperative
// launch), so the waits cannot deadlock.
__host__ __device__ __forceinline__ int brick_owner(int q, int Q, int G) {
    int b = (int)(((long long)q * G) / Q);
    while (b + 1 < G && (int)(((long long)Q * (b + 1)) / G) <= q) b++;
    while (b > 0 && (int)(((long long)Q * b) / G) > q) b--;
    return b;
}

__global__ void k_block_ranges_init(int* __restrict__ lo, int* __restrict__ hi, int G) {
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < G) { lo[b] = b; hi[b] = b; }
}
__global__ void k_block_ranges(const int4* __restrict__ nbA, const int2* __restrict__ nbB, int Q, int G,
                               int* __restrict__ lo, int* __restrict__ hi) {
    int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= Q) return;
    const int b = brick_owner(q, Q, G);
    const int4 a = nbA[q];
    const int2 z = nbB[q];
    const int n[6] = {a.x, a.y, a.z, a.w, z.x, z.y};
    int mn = b, mx = b;
    for (int i = 0; i < 6; i++)
        if (n[i] < Q) { int o = brick_owner(n[i], Q, G); mn = min(mn, o); mx = max(mx, o); }
    if (mn < b) atomicMin(&lo[b], mn);
    if (mx > b) atomicMax(&hi[b], mx);
}

static const int kProgressStride = 32;     // one 128-byte line per block / per chunk

// ---- warp-level dataflow synchronisation ------------------------------------------------------------------------
// Finer than the block-level progress words: every warp owns a contiguous chunk of its block's bricks and
// publishes its own "sweeps completed" word; before sweep k it waits only for the chunks that own one of
// its neighbour bricks (typically 6-12 warps).  There is no block-wide barrier inside the sweep loop, so
// while one warp pays the release/poll latency the other warps of the SM keep computing.  Producer and
// consumer sets coincide (the neighbour relation is symmetric), so the one wait also covers the
// write-after-read hazard of the ping-pong buffers.
static const int kChunkListCap = 64;

__host__ __device__ __forceinline__ void chunk_range(int b, int w, int Q, int G, int W, int* c0, int* c1) {
    const int qb = (int)(((long long)Q * b) / G), qe = (int)(((long long)Q * (b + 1)) / G);
    const int cw = (qe - qb + W - 1) / W;
    int a = qb + w * cw, e = a + cw;
    if (a > qe) a = qe;
    if (e > qe) e = qe;
    *c0 = a; *c1 = e;
}
__host__ __device__ __forceinline__ int chunk_of(int q, int Q, int G, int W) {
    const int b = brick_owner(q, Q, G);
    const int qb = (int)(((long long)Q * b) / G), qe = (int)(((long long)Q * (b + 1)) / G);
    const int cw = (qe - qb + W - 1) / W;
    return b * W + (q - qb) / (cw > 0 ? cw : 1);
}

__global__ void k_chunk_bitmap(const int4* __restrict__ nbA, const int2* __restrict__ nbB, int Q, int G, int W, int NW,
                               unsigned* __restrict__ bitmap) {
    int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= Q) return;
    const int g = chunk_of(q, Q, G, W);
    const int4 a = nbA[q];
    const int2 z = nbB[q];
    const int n[6] = {a.x, a.y, a.z, a.w, z.x, z.y};
    for (int i = 0; i < 6; i++)
        if (n[i] < Q) {
            int o = chunk_of(n[i], Q, G, W);
            if (o != g) atomicOr(&bitmap[(size_t)g * NW + (o >> 5)], 1u << (o & 31));
        }
}
// one warp per chunk: bitmap row -> producer list (order irrelevant)
__global__ void k_chunk_lists(const unsigned* __restrict__ bitmap, int NC, int NW, int* __restrict__ plist,
                              int* __restrict__ pcount, int* __restrict__ overflow) {
    const int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (g >= NC) return;
    for (int wi = lane; wi < NW; wi += 32) {
        unsigned m = bitmap[(size_t)g * NW + wi];
        while (m) {
            int bit = __ffs(m) - 1;
            m &= m - 1;
            int slot = atomicAdd(&pcount[g], 1);
            if (slot < kChunkListCap) plist[(size_t)g * kChunkListCap + slot] = wi * 32 + bit;
            else *overflow = 1;
        }
    }
}


__device__ __forceinline__ void neighbour_sync(unsigned* progress, unsigned epoch, int lo, int hi) {
    __syncthreads();                                   // the block's stores of this sweep are issued
    if (threadIdx.x == 0)
        asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(progress + (size_t)blockIdx.x * kProgressStride), "r"(epoch) : "memory");
    for (int c = lo + (int)threadIdx.x; c <= hi; c += blockDim.x) {
        if (c == (int)blockIdx.x) continue;
        unsigned v;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(progress + (size_t)c * kProgressStride) : "memory");
        } while (v < epoch);
    }
    __syncthreads();
}

template <int Threads, int MinBlocks, bool WarpFlow>
__global__ void __launch_bounds__(Threads, MinBlocks) k_qp_bricks(const uint2* __restrict__ code, const uint4* __restrict__ fields,
                                                                  const int4* __restrict__ nbA, const int2* __restrict__ nbB,
                                                                  float4* xa, float4* xb, int Q, int iters,
                                                                  unsigned* __restrict__ maxdelta_bits, unsigned* bar, int dbg,
                                                                  const int* __restrict__ blk_lo, const int* __restrict__ blk_hi,
                                                                  unsigned* progress, int cap, const int* __restrict__ plist,
                                                                  const int* __restrict__ pcount) {
    // contiguous range of bricks per block (bricks are ordered (X,Y,Z), Z fastest)
    // margin / invdg tables in shared memory: conflict-free LDS (distinct entries sit in distinct banks)
    __shared__ float s_sqrt[256];
    __shared__ float s_inv[32];
    for (int i = threadIdx.x; i < 256; i += Threads) s_sqrt[i] = c_brick_sqrt[i];
    if (threadIdx.x < 32) s_inv[threadIdx.x] = c_brick_invdg[threadIdx.x];
    __syncthreads();
    const size_t P = (size_t)Q + 1;      // plane stride of the SoA iterate
    const int q_begin = (int)((long long)Q * blockIdx.x / gridDim.x);
    int q_end = (int)((long long)Q * (blockIdx.x + 1) / gridDim.x);
    if (dbg & 2) q_end = q_begin;
    const int nlo = blk_lo[blockIdx.x], nhi = blk_hi[blockIdx.x];
    // The per-brick constants (6 neighbour ids, code, fields: 64 B) never change: when a block's share
    // fits they are copied to shared memory once, which takes the ids -> iterate dependent L2 round
    // trip out of every sweep.
    extern __shared__ uint4 s_dyn[];
    uint4* s_f0 = s_dyn;
    uint4* s_f1 = s_dyn + cap;
    int4* s_na = (int4*)(s_dyn + 2 * cap);
    int2* s_nz = (int2*)(s_dyn + 3 * cap);
    uint2* s_code = (uint2*)(s_nz + cap);
    const bool smem_consts = cap > 0;
    if (smem_consts) {
        for (int q = q_begin + threadIdx.x; q < q_end; q += Threads) {
            const int l = q - q_begin;
            s_na[l] = __ldg(nbA + q); s_nz[l] = __ldg(nbB + q); s_code[l] = __ldg(code + q);
            s_f0[l] = __ldg(fields + q); s_f1[l] = __ldg(fields + (size_t)Q + q);
        }
        __syncthreads();
    }
    const float4* in = xa;
    float4* out = xb;
    // work assignment: block-strided (block-level sync) or one contiguous chunk per warp (warp dataflow)
    constexpr int W = Threads / 32;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int q_first = q_begin + threadIdx.x, q_last = q_end, q_step = Threads;
    int np = 0, p0 = 0, p1 = 0;
    const int chunk = blockIdx.x * W + wid;
    if (WarpFlow) {
        int c0, c1;
        chunk_range(blockIdx.x, wid, Q, gridDim.x, W, &c0, &c1);
        if (dbg & 2) c1 = c0;
        q_first = c0 + lane; q_last = c1; q_step = 32;
        np = min(pcount[chunk], kChunkListCap);
        if (lane < np) p0 = plist[(size_t)chunk * kChunkListCap + lane];
        if (lane + 32 < np) p1 = plist[(size_t)chunk * kChunkListCap + lane + 32];
    }
    for (int it = 0; it < iters; it++) {
        float dmax = 0.0f;
        if (WarpFlow && it > 0 && !(dbg & 1)) {
            // wait until every producer chunk has completed sweep it-1
            unsigned spins = 0;
            if (lane < np) {
                unsigned v;
                do {
                    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(progress + (size_t)p0 * kProgressStride) : "memory");
                    if (++spins == (1u << 28)) __trap();       // a lost producer must not hang the GPU
                } while (v < (unsigned)it);
            }
            if (lane + 32 < np) {
                unsigned v;
                do {
                    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(progress + (size_t)p1 * kProgressStride) : "memory");
                    if (++spins == (1u << 28)) __trap();
                } while (v < (unsigned)it);
            }
            asm volatile("fence.acq_rel.gpu;" ::: "memory");
            __syncwarp();
        }
        for (int q = q_first; q < q_last; q += q_step) {
            int4 na; int2 nz; uint2 c; uint4 f0, f1;
            if (smem_consts) {            // brick constants resident in shared memory (copied once)
                const int l = q - q_begin;
                na = s_na[l]; nz = s_nz[l]; c = s_code[l]; f0 = s_f0[l]; f1 = s_f1[l];
            } else {
                na = __ldg(nbA + q); nz = __ldg(nbB + q); c = __ldg(code + q);
                f0 = __ldg(fields + q); f1 = __ldg(fields + (size_t)Q + q);
            }
            dmax = fmaxf(dmax, brick_update(q, P, na, nz, c, f0, f1, in, out, s_sqrt, s_inv, it == iters - 1));
        }
        if (it == iters - 1) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) dmax = fmaxf(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
            if ((threadIdx.x & 31) == 0) atomicMax(maxdelta_bits, __float_as_uint(dmax));
        }
        const float4* t = in; in = out; out = (float4*)t;
        if (WarpFlow) {
            __syncwarp();
            if (lane == 0)
                asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(progress + (size_t)chunk * kProgressStride), "r"((unsigned)(it + 1)) : "memory");
        } else if (it + 1 < iters && !(dbg & 1)) {
            if (dbg & 4) brick_grid_barrier(bar, (unsigned)(it + 1), gridDim.x);     // tuning aid: global barrier
            else neighbour_sync(progress, (unsigned)(it + 1), nlo, nhi);
        }
    }
}

// ---- tiled brick order, rim bricks first ------------------------------------------------------------------------
// In (X,Y,Z) order the X+-1 neighbours of a brick sit a whole x-plane of bricks away.  Here bricks are numbered along
// x-slabs of T brick columns, (slab, Y, x in slab, Z): a block's contiguous id range is a compact T x L patch of
// columns and most neighbour bricks belong to the same block or to a handful of adjacent ones.  Inside each block's
// range the rim bricks (those with a neighbour in another block) are numbered first and updated first, the interior
// bricks (which touch only this block's own bricks) second, dealt from the last warp down.  Results do not depend
// on the numbering: every brick update reads the previous iterate only.  Measured on B200 at 256^3, 512 threads x 1
// block per SM: 3.49 ms (X,Y,Z order, 256 x 2) -> 3.33 ms.
// Tried on top of this and measured slower (profiles/r01_summary.md): publishing the sweep as soon as the rim bricks
// are stored, so that the flag travels while the interior is updated -- with a barrier between the two phases
// 3.85 ms, with per-warp fences and an arrive counter instead of the barrier 3.57-3.75 ms, polling during the
// interior phase 4.4-5.5 ms.
// Block b of the tiled kernel owns bricks [tiled_begin(b), tiled_begin(b+1)): whole groups of 8 bricks, and the plane
// stride is a multiple of 8 float4, so no 128-byte line of the iterate holds bricks of two blocks (see the kernel).
__host__ __device__ __forceinline__ int tiled_begin(int b, int Q, int G) {
    const long long U = ((long long)Q + 7) / 8;
    const long long q = (U * b / G) * 8;
    return (int)(q < Q ? q : Q);
}
__host__ __device__ __forceinline__ int tiled_owner(int q, int Q, int G) {
    const long long U = ((long long)Q + 7) / 8, u = q / 8;
    int b = (int)((u * G) / U);
    while (b + 1 < G && U * (b + 1) / G <= u) b++;
    while (b > 0 && U * b / G > u) b--;
    return b;
}
__host__ __device__ __forceinline__ size_t tiled_stride(int Q) { return ((size_t)Q + 1 + 7) & ~(size_t)7; }

struct TiledOrder {
    BrickDims bd;
    int T;
    // Morton mode (mx > 0): the brick grid is cut into tiles of mx x my x mz bricks; the tiles are visited in Morton
    // (Z-curve) order of their coordinates (nbx / nby / nbz bits per axis, interleaved from the least significant bit while
    // an axis has bits left), the bricks of a tile in (x, y, z) order with z fastest.  A contiguous range of this order
    // is a compact 3-D patch whatever the orientation of the surface, so the blocks of the sweep kernel have about the
    // same share of rim bricks everywhere (x-slab order: 33 % in the median, 61 % where the surface is tangent to z).
    int mx, my, mz, nbx, nby, nbz;
    __host__ __device__ size_t count() const {
        if (mx > 0) return ((size_t)1 << (nbx + nby + nbz)) * (size_t)(mx * my * mz);
        return (size_t)((bd.bx + T - 1) / T) * T * bd.by * bd.bz;
    }
    // traversal index -> linear brick-grid position (X*by + Y)*bz + Z, or -1 in the padding
    __host__ __device__ long long position(size_t i) const {
        if (mx > 0) {
            const int per = mx * my * mz;
            const int w = (int)(i % per);
            size_t m = i / per;
            int tx = 0, ty = 0, tz = 0;
            for (int b = 0; b < 16; b++) {
                if (b < nbx) { tx |= (int)(m & 1) << b; m >>= 1; }
                if (b < nby) { ty |= (int)(m & 1) << b; m >>= 1; }
                if (b < nbz) { tz |= (int)(m & 1) << b; m >>= 1; }
            }
            const int X = tx * mx + w / (my * mz), Y = ty * my + (w / mz) % my, Z = tz * mz + w % mz;
            if (X >= bd.bx || Y >= bd.by || Z >= bd.bz) return -1;
            return ((long long)X * bd.by + Y) * bd.bz + Z;
        }
        const int Z = (int)(i % bd.bz);
        size_t t = i / bd.bz;
        const int x = (int)(t % T);
        t /= T;
        const int Y = (int)(t % bd.by);
        const int X = (int)(t / bd.by) * T + x;
        if (X >= bd.bx) return -1;
        return ((long long)X * bd.by + Y) * bd.bz + Z;
    }
};
static int bits_for(int n) { int b = 0; while ((1 << b) < n) b++; return b; }
static TiledOrder make_order(const BrickDims& bd, int tile_T) {
    TiledOrder o{bd, tile_T, 0, 0, 0, 0, 0, 0};
    const char* e = getenv("MRB_QP_ORDER");
    if (e && strncmp(e, "morton", 6) == 0) {
        int mx = 4, my = 4, mz = 8;
        if (e[6] == ':') sscanf(e + 7, "%d,%d,%d", &mx, &my, &mz);
        if (mx < 1) mx = 1; if (my < 1) my = 1; if (mz < 1) mz = 1;
        o.mx = mx; o.my = my; o.mz = mz;
        o.nbx = bits_for((bd.bx + mx - 1) / mx); o.nby = bits_for((bd.by + my - 1) / my); o.nbz = bits_for((bd.bz + mz - 1) / mz);
    }
    return o;
}
// brick activity from the stage-1+2 tile kernel's bitmap: one byte per (brick column, 8 bricks along z)
struct BrickBitPred {
    const uint8_t* bits;
    BrickDims bd;
    int tz;
    __device__ bool operator()(size_t p) const {
        const int Z = (int)(p % bd.bz);
        const size_t col = p / bd.bz;
        return (bits[col * tz + (Z >> 3)] >> (Z & 7)) & 1;
    }
};
template <typename Base>
struct TiledBrickPred {
    Base base;
    TiledOrder ord;
    __device__ bool operator()(size_t i) const {
        const long long p = ord.position(i);
        return p >= 0 && base((size_t)p);
    }
};
// How many bricks of a block run in the second ("rim") phase of a sweep.  The rim bricks must; interior bricks may, and a
// few do when that saves the interior phase a pass: a pass of a brick update is a ~1.5 us dependency chain however few
// threads take part, so 505 interior bricks on 480 threads cost two passes (measured: 3.2 us against 1.5 us) and the block
// falls behind its neighbours.  T threads; warp 0 collects the neighbours' flags first, so a single interior pass should
// not reach it (T - 32 bricks).  threads <= 0: off (rim bricks only).
__host__ __device__ __forceinline__ int late_count(int nrim, int nblock, int T) {
    if (T <= 0) return nrim;
    const int I = nblock - nrim;
    // interior passes deal T bricks each, warp 0's 32 last: keep the last pass off warp 0 (at most T - 32 bricks in it) and
    // drop a last pass that is less than half full
    const int full = I / T, rem = I % T;
    int keep = full * T + (rem < T - 32 ? rem : T - 32);
    if (full > 0 && rem < T / 2) keep = full * T - 32;
    int r = I - keep;
    const int room = ((nrim > 0 ? nrim : 1) + T - 1) / T * T;        // the rim phase keeps its number of passes
    if (r > 0 && nrim + r <= room) return nrim + r;
    return nrim;
}
__global__ void k_tiled_scatter(const int* __restrict__ tpos, int Q, TiledOrder ord, int* __restrict__ bid1) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < Q) bid1[ord.position((size_t)tpos[i])] = i;
}
// isb[i] = 1 if a neighbour brick of i (tile-order id) belongs to another block
__global__ void k_tiled_boundary(const int* __restrict__ tpos, int Q, int G, TiledOrder ord, const int* __restrict__ bid1,
                                 int* __restrict__ isb) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= Q) return;
    const BrickDims bd = ord.bd;
    const long long p = ord.position((size_t)tpos[i]);
    const int Z = (int)(p % bd.bz);
    const long long t = p / bd.bz;
    const int Y = (int)(t % bd.by), X = (int)(t / bd.by);
    const int own = tiled_owner(i, Q, G);
    const long long cx = (long long)bd.by * bd.bz, cy = bd.bz;
    bool rim = false;
    auto look = [&](bool ok, long long dp) {
        if (!ok) return;
        const int id = bid1[p + dp];
        if (id >= 0 && tiled_owner(id, Q, G) != own) rim = true;
    };
    look(X >= 1, -cx); look(X <= bd.bx - 2, cx); look(Y >= 1, -cy); look(Y <= bd.by - 2, cy);
    look(Z >= 1, -1); look(Z <= bd.bz - 2, 1);
    isb[i] = rim ? 1 : 0;
}
// final id = block begin + (boundary ? rank among the block's boundary bricks : #boundary + rank among its interior bricks)
__global__ void k_tiled_final(const int* __restrict__ tpos, const int* __restrict__ sb, int Q, int G, TiledOrder ord,
                              int* __restrict__ bpos, int* __restrict__ bid, int* __restrict__ blk_nb, int late_T) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= Q) return;
    const int b = tiled_owner(i, Q, G);
    const int qb = tiled_begin(b, Q, G), qe = tiled_begin(b + 1, Q, G);
    const int nbc = sb[qe] - sb[qb];
    const int rank_b = sb[i] - sb[qb];
    const bool rim = sb[i + 1] != sb[i];
    const int id = rim ? qb + rank_b : qb + nbc + (i - qb) - rank_b;
    const int p = (int)ord.position((size_t)tpos[i]);
    bpos[id] = p;
    bid[p] = id;
    if (i == qb) blk_nb[b] = late_count(nbc, qe - qb, late_T);
}

__global__ void k_tiled_bitmap(const int4* __restrict__ nbA, const int2* __restrict__ nbB, int Q, int G, int NW,
                               unsigned* __restrict__ bitmap) {
    int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= Q) return;
    const int g = tiled_owner(q, Q, G);
    const int4 a = nbA[q];
    const int2 z = nbB[q];
    const int n[6] = {a.x, a.y, a.z, a.w, z.x, z.y};
    for (int i = 0; i < 6; i++)
        if (n[i] < Q) {
            int o = tiled_owner(n[i], Q, G);
            if (o != g) atomicOr(&bitmap[(size_t)g * NW + (o >> 5)], 1u << (o & 31));
        }
}

// Slab-sharded runs (mrb_qp_plan_solve_sharded): the same kernel on every GPU over its own x-slab.  The bricks of the
// neighbour GPUs' adjacent planes are halo slots behind the owned bricks (never updated here); a rim brick of this
// GPU's first / last plane is also stored into the neighbour's halo slot over NVLink peer memory, so the halo travels
// with the sweep's own stores.  Blocks that own such bricks publish the sweep in a slot of the neighbour's flag array
// (system scope) and wait for the neighbour's slots the same way as for local blocks: behind the interior phase.
static const int kShardSlots = 128;
// Timeline instrumentation (MRB_QP_TRACE=1): per block and sweep eight %globaltimer stamps for sweeps [kTraceFirst,
// kTraceFirst + kTraceSweeps); tools/qp_trace.py prints the phase lengths.  trace == nullptr: off.
static const int kTraceFirst = 200, kTraceSweeps = 32;
__device__ __forceinline__ unsigned long long gtimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
struct TShard {
    float4* peer_lo[2];          // lower / upper neighbour's iterate buffers (A, B); nullptr = no neighbour
    float4* peer_hi[2];
    size_t P_lo, P_hi;           // their plane strides
    const int2* peer_ids;        // per owned brick: slot in the lower (.x) / upper (.y) neighbour's layout, -1 = none
    const int* slot_lo;          // per block: its slot among this GPU's blocks that push down / up, -1 = none
    const int* slot_hi;
    unsigned* flag_lo_remote;    // lower neighbour's "from upper" slots / upper neighbour's "from lower" slots
    unsigned* flag_hi_remote;
    unsigned* flags_local;       // [0,128) from lower, [128,256) from upper, [256] / [257]: pusher blocks of the lower / upper neighbour + 1
};

// The constants of one brick for the tiled kernel, from wherever they live (expanded / compact shared memory, global).
struct TiledConsts {
    bool expanded, smem_consts;
    int q_begin, cap, Q;
    const float4 *s_sm, *s_iv;
    const uint4* s_pred;
    const int4* s_na;
    const int2* s_nz;
    const uint2* s_code;
    const uint4 *s_f0, *s_f1;
    const uint2* code;
    const uint4* fields;
    const int4* nbA;
    const int2* nbB;
    const float *s_sqrt, *s_inv;
};
__device__ __forceinline__ float tiled_update(const TiledConsts& tc, int q, size_t P, const float4* in, float4* out, bool last) {
    const int l = q - tc.q_begin;
    if (tc.expanded) return brick_update_x(q, P, l, tc.cap, tc.s_sm, tc.s_iv, tc.s_pred, tc.s_na[l], tc.s_nz[l], in, out, last);
    int4 na; int2 nz; uint2 c; uint4 f0, f1;
    if (tc.smem_consts) {
        na = tc.s_na[l]; nz = tc.s_nz[l]; c = tc.s_code[l]; f0 = tc.s_f0[l]; f1 = tc.s_f1[l];
    } else {
        na = __ldg(tc.nbA + q); nz = __ldg(tc.nbB + q); c = __ldg(tc.code + q);
        f0 = __ldg(tc.fields + q); f1 = __ldg(tc.fields + (size_t)tc.Q + q);
    }
    return brick_update(q, P, na, nz, c, f0, f1, in, out, tc.s_sqrt, tc.s_inv, last);
}

template <int Threads, int MinBlocks, bool Sharded, bool Trace = false>
__global__ void __launch_bounds__(Threads, MinBlocks) k_qp_bricks_tiled(const uint2* __restrict__ code, const uint4* __restrict__ fields,
                                                                        const int4* __restrict__ nbA, const int2* __restrict__ nbB,
                                                                        float4* xa, float4* xb, int Q, int iters,
                                                                        unsigned* __restrict__ maxdelta_bits, unsigned* bar, int dbg,
                                                                        const int* __restrict__ blk_nb, unsigned* progress, int cap,
                                                                        const int* __restrict__ plist, const int* __restrict__ pcount,
                                                                        int expanded, size_t P, TShard sh,
                                                                        unsigned long long* trace) {
    __shared__ float s_sqrt[256];
    __shared__ float s_inv[32];
    // Threads is the launch bound; the block may be launched with fewer threads (grids that do not fill the GPU)
    const int nthreads = (int)blockDim.x;
    for (int i = threadIdx.x; i < 256; i += nthreads) s_sqrt[i] = c_brick_sqrt[i];
    if (threadIdx.x < 32) s_inv[threadIdx.x] = c_brick_invdg[threadIdx.x];
    const int q_begin = tiled_begin(blockIdx.x, Q, gridDim.x);
    const int q_end = tiled_begin(blockIdx.x + 1, Q, gridDim.x);
    const int q_int = q_begin + blk_nb[blockIdx.x];          // first interior brick
    __syncthreads();
    // Per-brick constants in shared memory, copied once: expanded (168 B per brick, see BrickConstsX) when they fit,
    // else in the global 64-byte form, else not at all (cap == 0: read through L1/L2 every sweep).
    extern __shared__ uint4 s_dyn[];
    uint4* s_f0 = s_dyn;
    uint4* s_f1 = s_dyn + cap;
    int4* s_na = (int4*)(s_dyn + 2 * cap);
    int2* s_nz = (int2*)(s_dyn + 3 * cap);
    uint2* s_code = (uint2*)(s_nz + cap);
    float4* s_sm = (float4*)s_dyn;
    float4* s_iv = s_sm + 4 * cap;
    uint4* s_pred = (uint4*)(s_iv + 4 * cap);
    if (expanded) { s_na = (int4*)(s_pred + cap); s_nz = (int2*)(s_na + cap); }
    const bool smem_consts = cap > 0;
    if (smem_consts)
        for (int q = q_begin + threadIdx.x; q < q_end; q += nthreads) {
            const int l = q - q_begin;
            s_na[l] = __ldg(nbA + q); s_nz[l] = __ldg(nbB + q);
            if (expanded) {
                const BrickConstsX x = brick_expand(__ldg(code + q), __ldg(fields + q), __ldg(fields + (size_t)Q + q), s_sqrt, s_inv);
#pragma unroll
                for (int k = 0; k < 4; k++) { s_sm[k * cap + l] = x.sm[k]; s_iv[k * cap + l] = x.inv[k]; }
                s_pred[l] = x.pred;
            } else {
                s_code[l] = __ldg(code + q); s_f0[l] = __ldg(fields + q); s_f1[l] = __ldg(fields + (size_t)Q + q);
            }
        }
    __syncthreads();
    // the blocks whose sweep this block has to wait for (at most kChunkListCap, checked by the host)
    const int np = pcount[blockIdx.x];
    const unsigned *watch = nullptr, *watch2 = nullptr;      // warp 0 polls: up to two producers per lane
    if (threadIdx.x < 32) {
        if ((int)threadIdx.x < np) watch = progress + (size_t)plist[(size_t)blockIdx.x * kChunkListCap + threadIdx.x] * kProgressStride;
        if ((int)threadIdx.x + 32 < np) watch2 = progress + (size_t)plist[(size_t)blockIdx.x * kChunkListCap + threadIdx.x + 32] * kProgressStride;
    }
    unsigned* mine = progress + (size_t)blockIdx.x * kProgressStride;
    // Work order inside a sweep: interior bricks first, dealt from warp 0 up; rim bricks second, dealt from the last warp
    // down, so that warp 0 (which publishes and polls) has the least to do.
    const int W = nthreads / 32;
    const int rtid = (W - 1 - (int)(threadIdx.x >> 5)) * 32 + (int)(threadIdx.x & 31);
    __shared__ unsigned s_seen;              // number of sweeps whose neighbour flags have all been observed
    __shared__ int s_nrem[2];                // slots to wait for in the lower / upper neighbour's flag rows
    if (threadIdx.x == 0) s_seen = 0;
    const unsigned s_seen_addr = (unsigned)__cvta_generic_to_shared(&s_seen);
    int my_lo = -1, my_hi = -1;
    if (Sharded) {
        my_lo = sh.slot_lo[blockIdx.x];
        my_hi = sh.slot_hi[blockIdx.x];
        if (threadIdx.x < 2) {
            // the neighbour wrote "number of its pusher blocks + 1" before launching its kernel
            unsigned v = 1;
            if (threadIdx.x == 0 ? my_lo >= 0 : my_hi >= 0) {
                unsigned spins = 0;
                do {
                    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(sh.flags_local + 2 * kShardSlots + threadIdx.x) : "memory");
                    if (++spins == (1u << 27)) __trap();
                } while (v == 0);
            }
            s_nrem[threadIdx.x] = (int)v - 1;
        }
    }
    __syncthreads();
    const float4* in = xa;
    float4* out = xb;
    const TiledConsts tc{expanded != 0, smem_consts, q_begin, cap, Q, s_sm, s_iv, s_pred, s_na, s_nz, s_code, s_f0, s_f1,
                         code, fields, nbA, nbB, s_sqrt, s_inv};
    // Sweep it: (1) interior bricks -- they read this block's own bricks only, so they start right after the barrier
    // that ended sweep it-1, while that sweep's flag is still travelling; (2) rim bricks -- they wait until warp 0 has seen
    // "sweep it-1 done" from every neighbour block (which also means the neighbours no longer read the buffer this sweep
    // overwrites); (3) barrier, publish.  One block-wide barrier per sweep; the release -> poll round trip (about 1.3 us)
    // is hidden behind the interior phase instead of being paid by every warp.
    // L1 coherence: warp 0's acquire loads invalidate this SM's L1 after the neighbours' stores are visible.  Loads that
    // are in flight at that moment belong to interior bricks and touch this block's own lines only (block ranges and the
    // plane stride are multiples of 8 bricks = 128 bytes), so no line with another block's bricks can be cached stale.
    for (int it = 0; it < iters; it++) {
        const bool last = it == iters - 1;
        float dmax = 0.0f;
        const bool flags_on = !(dbg & 5);
        const bool tr = Trace && it >= kTraceFirst && it < kTraceFirst + kTraceSweeps;
        unsigned long long* trow = Trace ? trace + ((size_t)blockIdx.x * kTraceSweeps + (tr ? it - kTraceFirst : 0)) * 8 : nullptr;
        if (tr && threadIdx.x == 0) trow[0] = gtimer();
        if (it > 0 && (dbg & 4) && !(dbg & 1)) brick_grid_barrier(bar, (unsigned)it, gridDim.x);
        if (it > 0 && flags_on && threadIdx.x < 32) {
            // warp 0: collect the neighbours' flags of sweep it-1, then tell the block
            unsigned v, spins = 0;
            if (watch) do {
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(watch) : "memory");
                if (++spins == (1u << 26)) __trap();       // a lost producer must not hang the GPU
            } while (v < (unsigned)it);
            if (watch2) do {
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(watch2) : "memory");
                if (++spins == (1u << 26)) __trap();
            } while (v < (unsigned)it);
            if (Sharded) {
                // the neighbour GPUs' blocks that push into this GPU's halo: one slot each, system scope
                if (my_lo >= 0)
                    for (int sl = threadIdx.x; sl < s_nrem[0]; sl += 32) do {
                        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(sh.flags_local + sl) : "memory");
                        if (++spins == (1u << 26)) __trap();
                    } while (v < (unsigned)it);
                if (my_hi >= 0)
                    for (int sl = threadIdx.x; sl < s_nrem[1]; sl += 32) do {
                        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(sh.flags_local + kShardSlots + sl) : "memory");
                        if (++spins == (1u << 26)) __trap();
                    } while (v < (unsigned)it);
            }
            __syncwarp();
            if (threadIdx.x == 0) asm volatile("st.release.cta.shared.u32 [%0], %1;" ::"r"(s_seen_addr), "r"((unsigned)it) : "memory");
            if (tr && threadIdx.x == 0) trow[1] = gtimer();
        }
        // interior bricks are dealt from warp 1 up and wrap around to warp 0 last: warp 0 starts late (it collects the flags
        // first), so the stray second brick of a 481-brick interior must not be its lane 0's (measured: the block barrier
        // was waiting for exactly that brick)
        for (int q = q_int + ((int)threadIdx.x >= 32 ? (int)threadIdx.x - 32 : nthreads - 32 + (int)threadIdx.x); q < q_end; q += nthreads)
            dmax = fmaxf(dmax, tiled_update(tc, q, P, in, out, last));
        if (tr && threadIdx.x == 0) trow[2] = gtimer();
        if (tr && (int)threadIdx.x == nthreads - 32) trow[3] = gtimer();
        if (it > 0 && flags_on && q_begin + rtid < q_int) {
            unsigned v;
            do {
                asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(v) : "r"(s_seen_addr) : "memory");
            } while (v < (unsigned)it);
        }
        if (tr && (int)threadIdx.x == nthreads - 32) trow[4] = gtimer();
        for (int q = q_begin + rtid; q < q_int; q += nthreads) {
            dmax = fmaxf(dmax, tiled_update(tc, q, P, in, out, last));
            if (Sharded && !last && (my_lo >= 0 || my_hi >= 0)) {
                // first / last plane of the slab: the same values go into the neighbour's halo slot over NVLink
                const int2 pid = sh.peer_ids[q];
                if (pid.x >= 0) {
                    float4* dst = (it & 1) ? sh.peer_lo[0] : sh.peer_lo[1];       // sweep `it` writes buffer B when it is even
#pragma unroll
                    for (int k = 0; k < 4; k++) dst[k * sh.P_lo + pid.x] = out[k * P + q];
                }
                if (pid.y >= 0) {
                    float4* dst = (it & 1) ? sh.peer_hi[0] : sh.peer_hi[1];
#pragma unroll
                    for (int k = 0; k < 4; k++) dst[k * sh.P_hi + pid.y] = out[k * P + q];
                }
            }
        }
        if (last) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) dmax = fmaxf(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
            if ((threadIdx.x & 31) == 0) atomicMax(maxdelta_bits, __float_as_uint(dmax));
        } else if (!(dbg & 1)) {
            if (tr && (int)threadIdx.x == nthreads - 32) trow[5] = gtimer();
            __syncthreads();        // the block's stores of this sweep are issued; its reads of the neighbours' bricks are done
            if (tr && (int)threadIdx.x == nthreads - 32) trow[6] = gtimer();
            // Published by the LAST warp's first lane (MRB_QP_DEBUG=32: by thread 0): the release waits for the block's
            // stores to be visible, and warp 0 should already be polling the neighbours' flags during that time.
            if (flags_on && (int)threadIdx.x == ((dbg & 32) ? 0 : nthreads - 32)) {
                if (Sharded && (my_lo >= 0 || my_hi >= 0)) {
                    // one system-scope fence covers the stores into the neighbour GPUs (made cumulative by the barrier)
                    asm volatile("fence.acq_rel.sys;" ::: "memory");
                    if (my_lo >= 0) asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(sh.flag_lo_remote + my_lo), "r"((unsigned)(it + 1)) : "memory");
                    if (my_hi >= 0) asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(sh.flag_hi_remote + my_hi), "r"((unsigned)(it + 1)) : "memory");
                }
                asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(mine), "r"((unsigned)(it + 1)) : "memory");
                if (tr) trow[7] = gtimer();
            }
        }
        const float4* t = in; in = out; out = (float4*)t;
    }
}

// ---- on-chip sweep kernel: the block's iterate lives in shared memory for all sweeps ----------------------------------
// k_qp_bricks_tiled above reads the iterate through L1/L2 every sweep (20 LDG.128 + 8 LDG.64 per brick update, L1 hit
// rate 56 %) and pays, per sweep, a block barrier -> release fence (0.9 us) -> flag hop -> acquire poll chain in front of
// its rim bricks (measured timeline: profiles/r02_summary.md).  Here
//   * a block keeps BOTH ping-pong copies of its own bricks in shared memory (2 x 64 B per brick), a copy of the foreign
//     bricks its rim reads (the HALO, 64 B each) and the constants (signed margins as floats, one predicate byte per
//     voxel, six 16-bit neighbour slots: 92 B per brick): every load of a sweep is an LDS;
//   * the only thing other blocks need, the RIM bricks (a neighbour in another block), travels through an exchange buffer
//     in global memory (L2) in which every 64-bit word carries one fp32 value and the number of the sweep that produced
//     it.  A 64-bit store is single-copy atomic, so a consumer that reads the expected sweep number has the value too:
//     no fence, no flag, no release/acquire pair, and the dependency is per brick instead of per block;
//   * a sweep updates the rim bricks FIRST and sends them at once; while a thread updates its second (interior) brick, its
//     loads of ONE halo brick for the next sweep are in flight (the producers sent it one interior phase ago); the halo is
//     put into shared memory between the two block barriers that end the sweep.
// invdg is not stored: d = 12 T(v) + (number of tight neighbours) is recomputed from the predicate byte (popc) and looked
// up in the 32-entry table (a conflict-free LDS: distinct entries sit in distinct banks).
// Write-after-read on the two exchange parities is safe because a producer can only reach sweep s+2 after its neighbours
// have produced sweep s+1, i.e. after they have consumed sweep s.
// Shared-memory planes (k = dx*2+dy), S = 2*(CAP+1) + CAPH float4 each:  [buffer 0: own slots 0..CAP][halo][buffer 1,
// own slots in REVERSE order].  Neighbour slot r (own brick, CAP' = the all-zero dummy brick, CAP+1+h = halo brick h)
// is element r of a plane in even sweeps and element S-1-r in odd sweeps (the halo is written in that order too), so a
// sweep addresses everything as base + step * r with step = +-16 bytes and the plane offsets are immediates.
static const int kOcCap = 848, kOcHalo = 608, kOcThreadsMax = 480;
__device__ __forceinline__ unsigned long long ll_pack(float v, unsigned tag) {
    return (unsigned long long)__float_as_uint(v) | ((unsigned long long)tag << 32);
}
__device__ __forceinline__ void ll_store2(ulonglong2* dst, float a, float b, unsigned tag) {
    asm volatile("st.global.v2.u64 [%0], {%1, %2};" ::"l"(dst), "l"(ll_pack(a, tag)), "l"(ll_pack(b, tag)) : "memory");
}
__device__ __forceinline__ ulonglong2 ll_load2(const ulonglong2* src) {
    ulonglong2 v;
    asm volatile("ld.global.cg.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(src) : "memory");
    return v;
}
__device__ __forceinline__ bool ll_valid(const ulonglong2 (&v)[8], unsigned tag) {
    bool ok = true;
#pragma unroll
    for (int i = 0; i < 8; i++) ok = ok && (unsigned)(v[i].x >> 32) == tag && (unsigned)(v[i].y >> 32) == tag;
    return ok;
}
// Waiting for a late producer: poll ONE word with a pause between the attempts (a waiting warp leaves the SM's issue slots
// and the L2 to the warps that have work), then load the brick again.
__device__ __forceinline__ void ll_wait_reload(const ulonglong2* src, unsigned tag, ulonglong2 (&v)[8]) {
    unsigned spins = 0;
    for (;;) {
        unsigned long long w;
        asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(w) : "l"(src) : "memory");
        if ((unsigned)(w >> 32) == tag) {
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = ll_load2(src + i);
            if (ll_valid(v, tag)) return;
        }
        if (++spins == (1u << 22)) __trap();               // a lost producer must not hang the GPU
        __nanosleep(200);
    }
}

// For every brick g and direction d (0 x-, 1 x+, 2 y-, 3 y+, 4 z-, 5 z+) whose neighbour lives in another block o: g is
// halo brick number hslot[g*6+d] of block o (one number per distinct o), listed in hgid[o*caph + number].
__global__ void k_onchip_halo(const int4* __restrict__ nbA, const int2* __restrict__ nbB, int Q, int G, int caph,
                              int* __restrict__ hcount, int* __restrict__ hgid, unsigned short* __restrict__ hslot,
                              int* __restrict__ hmax) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= Q) return;
    const int own = tiled_owner(g, Q, G);
    const int4 a = nbA[g];
    const int2 z = nbB[g];
    const int n[6] = {a.x, a.y, a.z, a.w, z.x, z.y};
    int o[6], sl[6];
#pragma unroll
    for (int d = 0; d < 6; d++) {
        o[d] = n[d] < Q ? tiled_owner(n[d], Q, G) : own;
        sl[d] = -1;
        if (o[d] != own) {
            for (int e = 0; e < d; e++) if (o[e] == o[d]) sl[d] = sl[e];
            if (sl[d] < 0) {
                sl[d] = atomicAdd(&hcount[o[d]], 1);
                if (sl[d] < caph) hgid[(size_t)o[d] * caph + sl[d]] = g;
                atomicMax(hmax, sl[d] + 1);
            }
        }
        hslot[(size_t)g * 6 + d] = (unsigned short)(sl[d] < 0 ? 0 : sl[d]);
    }
}

template <int CAP, int CAPH>
struct OnchipLayout {
    static constexpr int CAPX = CAP + 1;                    // own slots + the dummy brick
    static constexpr int S = 2 * CAPX + CAPH;               // float4 per plane
    static constexpr size_t kIterBytes = (size_t)4 * S * 16;
    static constexpr size_t kBytes = kIterBytes + (size_t)CAP * (64 + 16 + 8 + 4) + (size_t)CAPH * 4;
};

// One brick update out of shared memory: slot l, neighbour slots from s_na / s_nz; `cb` = byte address (shared window) of
// element 0 of plane 0 as the current sweep sees it, `cs` = +-16; `nb` / `ns` the same for the buffer being written.
template <int CAP, int CAPH>
__device__ __forceinline__ float onchip_update(int l, float4* __restrict__ s_x, int c0, int cs, int n0, const float4* __restrict__ s_sm,
                                               const uint4* __restrict__ s_pred, const uint2* __restrict__ s_na,
                                               const unsigned* __restrict__ s_nz, const char* __restrict__ s_inv_bytes,
                                               float4* gout, size_t P, int q, bool to_global, bool want_delta) {
    // element r of the buffer this sweep reads: s_x[k*S + c0 + cs*r] (cs = +-1); of the buffer it writes: s_x[k*S + n0 - cs*r]
    using L = OnchipLayout<CAP, CAPH>;
    constexpr int S = L::S;
    BrickRegs r;
    const uint2 na = s_na[l];
    const unsigned nz = s_nz[l];
    const float4* p_own = s_x + c0 + cs * l;
    const float4* p_xm = s_x + c0 + cs * (int)(na.x & 0xffffu);
    const float4* p_xp = s_x + c0 + cs * (int)(na.x >> 16);
    const float4* p_ym = s_x + c0 + cs * (int)(na.y & 0xffffu);
    const float4* p_yp = s_x + c0 + cs * (int)(na.y >> 16);
    const float4* p_zl = s_x + c0 + cs * (int)(nz & 0xffffu);
    const float4* p_zh = s_x + c0 + cs * (int)(nz >> 16);
#pragma unroll
    for (int k = 0; k < 4; k++) {
        put4(r.own[k], p_own[k * S]);
        put4(r.xm[k], p_xm[k * S]);
        put4(r.xp[k], p_xp[k * S]);
        put4(r.ym[k], p_ym[k * S]);
        put4(r.yp[k], p_yp[k * S]);
        const float2 lo = *reinterpret_cast<const float2*>(reinterpret_cast<const float*>(p_zl + k * S) + 2);
        const float2 hi = *reinterpret_cast<const float2*>(p_zh + k * S);
        r.zl[k][0] = lo.x; r.zl[k][1] = lo.y;
        r.zh[k][0] = hi.x; r.zh[k][1] = hi.y;
    }
    const uint4 pw = s_pred[l];
    // invdg of the four voxels of a strip: byte j of w = predicate byte (bit 0 T(v), bits 1..6 the six neighbours' T)
    auto inv4 = [&](uint32_t w) {
        auto one = [&](uint32_t pb) {
            const uint32_t d4 = __popc(pb & 0x7eu) * 4u + (pb & 1u) * 48u;       // 4 * (12 T(v) + tight neighbours)
            return *reinterpret_cast<const float*>(s_inv_bytes + d4);
        };
        return make_float4(one(w & 0xffu), one((w >> 8) & 0xffu), one((w >> 16) & 0xffu), one(w >> 24));
    };
    const float4 o0 = brick_strip_x<0>(r, pw.x, s_sm[l], inv4(pw.x));
    const float4 o1 = brick_strip_x<1>(r, pw.y, s_sm[CAP + l], inv4(pw.y));
    const float4 o2 = brick_strip_x<2>(r, pw.z, s_sm[2 * CAP + l], inv4(pw.z));
    const float4 o3 = brick_strip_x<3>(r, pw.w, s_sm[3 * CAP + l], inv4(pw.w));
    float4* p_out = s_x + n0 - cs * l;
    p_out[0] = o0; p_out[S] = o1; p_out[2 * S] = o2; p_out[3 * S] = o3;
    if (to_global) { gout[q] = o0; gout[P + q] = o1; gout[2 * P + q] = o2; gout[3 * P + q] = o3; }
    float dmax = 0.0f;
    if (want_delta) {
        const float4 os[4] = {o0, o1, o2, o3};
#pragma unroll
        for (int k = 0; k < 4; k++)
            dmax = fmaxf(fmaxf(dmax, fmaxf(fabsf(os[k].x - r.own[k][0]), fabsf(os[k].y - r.own[k][1]))),
                         fmaxf(fabsf(os[k].z - r.own[k][2]), fabsf(os[k].w - r.own[k][3])));
    }
    return dmax;
}

// ex: exchange buffer, 2 parities x Q bricks x 16 words of {fp32 value, sweep number}; zeroed by the host before the launch
// (sweep numbers start at 1, so nothing stale can match).  Sweep `it` consumes number it + 1 (staged into the halo before
// the sweep starts) and produces number it + 2 into parity it & 1; the set-up publishes the initial iterate as number 1.
template <int Threads, int CAP, int CAPH, bool Trace>
__global__ void __launch_bounds__(Threads, 1) k_qp_bricks_onchip(const uint2* __restrict__ code, const uint4* __restrict__ fields,
                                                                 const int4* __restrict__ nbA, const int2* __restrict__ nbB,
                                                                 float4* xa, float4* xb, int Q, int iters,
                                                                 unsigned* __restrict__ maxdelta_bits, const int* __restrict__ blk_nb,
                                                                 size_t P, ulonglong2* ex, const int* __restrict__ hcount,
                                                                 const int* __restrict__ hgid, const unsigned short* __restrict__ hslot,
                                                                 int mode, unsigned long long* trace) {
    using L = OnchipLayout<CAP, CAPH>;
    constexpr int CAPX = L::CAPX, S = L::S;
    constexpr unsigned PS = (unsigned)S * 16u;
    __shared__ float s_sqrt[256];
    __shared__ float s_inv[32];
    const int nthreads = (int)blockDim.x;
    for (int i = threadIdx.x; i < 256; i += nthreads) s_sqrt[i] = c_brick_sqrt[i];
    if (threadIdx.x < 32) s_inv[threadIdx.x] = c_brick_invdg[threadIdx.x];
    const int q_begin = tiled_begin(blockIdx.x, Q, gridDim.x);
    const int q_end = tiled_begin(blockIdx.x + 1, Q, gridDim.x);
    const int q_int = q_begin + blk_nb[blockIdx.x];          // first interior brick; [q_begin, q_int) are the rim bricks
    const int nloc = q_end - q_begin;                        // slot nloc = dummy brick (all zero in both buffers)
    extern __shared__ uint4 s_dyn[];
    float4* s_x = (float4*)s_dyn;                            // [4][S]
    float4* s_sm = s_x + 4 * S;                              // [4][CAP]
    uint4* s_pred = (uint4*)(s_sm + 4 * CAP);
    uint2* s_na = (uint2*)(s_pred + CAP);
    unsigned* s_nz = (unsigned*)(s_na + CAP);
    ulonglong2* ex0 = ex;                                    // parity 0 / 1
    ulonglong2* ex1 = ex + (size_t)Q * 8;
    __syncthreads();
    for (int q = q_begin + threadIdx.x; q < q_end; q += nthreads) {
        const int l = q - q_begin;
        const int4 a = __ldg(nbA + q);
        const int2 z = __ldg(nbB + q);
        const int n[6] = {a.x, a.y, a.z, a.w, z.x, z.y};
        unsigned sl[6];
#pragma unroll
        for (int d = 0; d < 6; d++)
            sl[d] = (n[d] >= q_begin && n[d] < q_end) ? (unsigned)(n[d] - q_begin)
                    : (n[d] >= Q ? (unsigned)nloc : (unsigned)(CAPX + hslot[(size_t)n[d] * 6 + (d ^ 1)]));
        s_na[l] = make_uint2(sl[0] | (sl[1] << 16), sl[2] | (sl[3] << 16));
        s_nz[l] = sl[4] | (sl[5] << 16);
        const uint4 f0 = __ldg(fields + q), f1 = __ldg(fields + (size_t)Q + q);
        const BrickConstsX x = brick_expand(__ldg(code + q), f0, f1, s_sqrt, s_inv);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            s_sm[k * CAP + l] = x.sm[k];
            const float4 xv = xa[k * P + q];
            s_x[k * S + l] = xv;                                        // buffer 0: forward order
            s_x[k * S + S - 1 - l] = make_float4(0.f, 0.f, 0.f, 0.f);   // buffer 1: reverse order
        }
        s_pred[l] = x.pred;
    }
    if (threadIdx.x < 8) {
        const int k = threadIdx.x & 3;
        s_x[k * S + ((threadIdx.x < 4) ? nloc : S - 1 - nloc)] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    // The exchange is WARP-COOPERATIVE so that every global access is a full 128-byte line per brick: lane L of a warp
    // handles piece L % 8 (16 bytes: two {value, number} words) of brick L / 8 of a group of four bricks.  (One thread
    // reading or writing "its" brick in eight 16-byte pieces touches 32 different lines per warp instruction: measured,
    // that alone took 6 us per sweep.)
    //   send: after a warp has updated its 32 bricks of a pass, it copies the rim bricks among them from the freshly
    //         written shared-memory planes into the exchange buffer (8 coalesced STG.128);
    //   fetch: groups of four halo bricks, 133 groups for the largest halo: LDG.128, number check with a warp vote,
    //         STS.64 into the halo slots as the next sweep will address them.
    // Passes: a thread updates bricks q_begin + threadIdx.x + p * nthreads, p = 0 .. npass-1 (rim bricks first).
    // Fast path (two passes, every rim brick in the first one, at least one warp without a brick in the second): those
    // FREE warps fetch the halo of the next sweep while the others update their second brick; they know that the rim
    // bricks are done (nobody reads the halo any more) from a split barrier: the busy warps arrive on barrier 1 after the
    // first pass without waiting, the free warps wait on it.  Otherwise: all warps fetch after the last pass, between
    // two block barriers.
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nwarps = nthreads >> 5;
    const int npass = (nloc + nthreads - 1) / nthreads;
    const int nh = hcount[blockIdx.x];
    int* s_hgid = (int*)(s_nz + CAP);
    for (int h = threadIdx.x; h < nh; h += nthreads) s_hgid[h] = hgid[(size_t)blockIdx.x * CAPH + h];
    const int nrim = q_int - q_begin;
    const int first_free_warp = (nloc - (npass - 1) * nthreads + 31) >> 5;      // warps >= this have no brick in the last pass
    const int nfree = nwarps - first_free_warp;
    const bool fast = npass == 2 && nrim <= nthreads && nfree >= 1 && !(mode & 2);
    const bool free_warp = fast && wid >= first_free_warp;
    const int sync_count = nthreads;
    auto send = [&](int q0, int n0, int cs, ulonglong2* ex_out, unsigned tag) {
        // bricks q0 .. q0+31 of this warp's pass; rim bricks are those below q_int
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int qb = q0 + 4 * i + (lane >> 3);
            if (qb < q_int) {
                const int piece = lane & 7;
                const float2 v = *reinterpret_cast<const float2*>(
                    reinterpret_cast<const float*>(s_x + (piece >> 1) * S + n0 - cs * (qb - q_begin)) + 2 * (piece & 1));
                ll_store2(ex_out + (size_t)qb * 8 + piece, v.x, v.y, tag);
            }
        }
    };
    // fetch groups g = w0, w0 + nw, ... of the halo for the sweep of parity `parity`, numbers `tag`
    auto fetch = [&](int w0, int nw, const ulonglong2* ex_in, unsigned tag, int parity) {
        const int ngroups = (nh + 3) >> 2;
        for (int g0 = w0; g0 < ngroups; g0 += nw * 4) {
            ulonglong2 v[4];
            int hb[4];
            bool ok[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                hb[u] = 4 * (g0 + u * nw) + (lane >> 3);
                ok[u] = true;
                if (g0 + u * nw < ngroups && hb[u] < nh) v[u] = ll_load2(ex_in + (size_t)s_hgid[hb[u]] * 8 + (lane & 7));
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                if (g0 + u * nw >= ngroups) continue;                    // warp-uniform
                const bool mine = hb[u] < nh;
                unsigned spins = 0;
                for (;;) {
                    const bool good = !mine || ((unsigned)(v[u].x >> 32) == tag && (unsigned)(v[u].y >> 32) == tag);
                    if (__all_sync(0xffffffffu, good)) break;
                    if (++spins == (1u << 22)) __trap();                 // a lost producer must not hang the GPU
                    __nanosleep(100);
                    if (mine) v[u] = ll_load2(ex_in + (size_t)s_hgid[hb[u]] * 8 + (lane & 7));
                }
                if (mine) {
                    const int piece = lane & 7;
                    const int e = parity ? S - 1 - (CAPX + hb[u]) : CAPX + hb[u];
                    *reinterpret_cast<float2*>(reinterpret_cast<float*>(s_x + (piece >> 1) * S + e) + 2 * (piece & 1)) =
                        make_float2(__uint_as_float((unsigned)v[u].x), __uint_as_float((unsigned)v[u].y));
                }
            }
        }
    };
    __syncthreads();
    // the initial iterate of the rim bricks is "number 1"; fetch the halo of sweep 0
    for (int q0 = q_begin + wid * 32; q0 < q_int; q0 += nthreads) send(q0, 0, -1, ex1, 1u);      // buffer 0, forward order
    fetch(wid, nwarps, ex1, 1u, 0);
    const char* s_inv_bytes = reinterpret_cast<const char*>(s_inv);
    __syncthreads();
    for (int it = 0; it < iters; it++) {
        const bool last = it == iters - 1;
        const int par = it & 1;
        // element r of the buffer read by this sweep: c0 + cs * r; of the buffer it writes: n0 - cs * r
        const int c0 = par ? S - 1 : 0, cs = par ? -1 : 1, n0 = par ? 0 : S - 1;
        float4* gout = par ? xa : xb;
        ulonglong2* ex_out = par ? ex1 : ex0;                   // this sweep's rim results: number it + 2
        const unsigned tag_out = (unsigned)it + 2u;
        float dmax = 0.0f;
        const bool tr = Trace && it >= kTraceFirst && it < kTraceFirst + kTraceSweeps;
        unsigned long long* trow = Trace ? trace + ((size_t)blockIdx.x * kTraceSweeps + (it - kTraceFirst)) * 8 : nullptr;
        if (tr && threadIdx.x == 0) trow[0] = gtimer();
        int q = q_begin + threadIdx.x;
        for (int p = 0; p < npass; p++, q += nthreads) {
            if (fast && p == 1 && free_warp) break;
            if (q < q_end)
                dmax = fmaxf(dmax, onchip_update<CAP, CAPH>(q - q_begin, s_x, c0, cs, n0, s_sm, s_pred, s_na, s_nz, s_inv_bytes,
                                                            gout, P, q, last, last));
            const int q0 = q - lane;                                     // first brick of this warp in this pass
            if (!last && q0 < q_int) {
                __syncwarp();                                            // the warp's results are in shared memory
                send(q0, n0, cs, ex_out, tag_out);
            }
            if (tr && threadIdx.x == 0 && p == 0) trow[1] = gtimer();
            if (tr && (int)threadIdx.x == nthreads - 32 && p == 0) trow[3] = gtimer();
            if (fast && p == 0 && !last) {
                // every rim brick of the block is in this pass: after it nobody reads the halo any more
                if (free_warp) asm volatile("barrier.cta.sync 1, %0;" ::"r"(sync_count) : "memory");
                else asm volatile("barrier.cta.arrive 1, %0;" ::"r"(sync_count) : "memory");
            }
        }
        if (tr && threadIdx.x == 0) trow[2] = gtimer();
        if (last) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) dmax = fmaxf(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
            if ((threadIdx.x & 31) == 0) atomicMax(maxdelta_bits, __float_as_uint(dmax));
        } else if (fast) {
            if (free_warp) fetch(wid - first_free_warp, nfree, ex_out, tag_out, par ^ 1);
            if (tr && (int)threadIdx.x == nthreads - 32) trow[4] = gtimer();
            __syncthreads();        // the next sweep's inputs (own planes, halo) are complete
            if (tr && threadIdx.x == 0) trow[5] = gtimer();
        } else {
            __syncthreads();        // every read of this sweep (own slots, halo) is done
            fetch(wid, nwarps, ex_out, tag_out, par ^ 1);
            if (tr && (int)threadIdx.x == nthreads - 32) trow[4] = gtimer();
            __syncthreads();        // the next sweep's inputs are complete
            if (tr && threadIdx.x == 0) trow[5] = gtimer();
        }
    }
}

// ---- feature reset + F assembly ---------------------------------------------------------------------------------
__device__ __forceinline__ size_t brick_slot(const int* __restrict__ bid, const BrickDims& bd, size_t P, size_t idx) {
    int z = (int)(idx % bd.nz);
    size_t t = idx / bd.nz;
    int y = (int)(t % bd.ny);
    int x = (int)(t / bd.ny);
    int q = bid[((size_t)(x >> 1) * bd.by + (y >> 1)) * bd.bz + (z >> 2)];
    return ((size_t)((x & 1) * 2 + (y & 1)) * P + (size_t)q) * 4 + (size_t)(z & 3);
}

// getFeatureIndexes (primeqp.cpp:231-244) + the reset loop of runQP (quadprog.cpp:92-101): a single
// feature voxel outside the band (-1 in the sorted id list) blocks every reset.
__global__ void k_feat_check_b(const float* __restrict__ fmap, const uint8_t* __restrict__ flags, size_t n, int* __restrict__ any_off) {
    MRB_VOXEL_LOOP(idx, n) if (fmap[idx] == 1.0f && !(flags[idx] & 1)) *any_off = 1;
}
__global__ void k_feat_apply_b(const float* __restrict__ fmap, const uint8_t* __restrict__ flags, const int* __restrict__ bid,
                               BrickDims bd, size_t P, const int* __restrict__ any_off, float* __restrict__ x) {
    if (*any_off) return;
    MRB_VOXEL_LOOP(idx, (size_t)bd.nx * bd.ny * bd.nz)
        if (fmap[idx] == 1.0f && (flags[idx] & 1)) x[brick_slot(bid, bd, P, idx)] = 0.2f;
}

// F = (2 volume - 1)(band_size + 1) off the band, the QP solution on it (quadprog.cpp:131-145)
__global__ void __launch_bounds__(256) k_assemble_F_b(const float* __restrict__ vol, const uint8_t* __restrict__ flags,
                                                      const int* __restrict__ bid, const float* __restrict__ x, BrickDims bd,
                                                      size_t P, double scale, float* __restrict__ F) {
    MRB_VOXEL_LOOP(idx, (size_t)bd.nx * bd.ny * bd.nz) {
        float f;
        if (flags[idx] & 1) f = x[brick_slot(bid, bd, P, idx)];
        else f = (float)__dmul_rn(__dsub_rn(__dmul_rn(2.0, (double)vol[idx]), 1.0), scale);
        F[idx] = f;
    }
}

// k_assemble_F_b + the sign bits of F for marching cubes (layout of k_sign_bits: one word per 32 z, nz % 32 == 0):
// one warp per word, lanes = 32 consecutive z.
__global__ void __launch_bounds__(256) k_assemble_F_sign(const float* __restrict__ vol, const uint8_t* __restrict__ flags,
                                                         const int* __restrict__ bid, const float* __restrict__ x, BrickDims bd,
                                                         size_t P, double scale, float iso, float* __restrict__ F,
                                                         uint32_t* __restrict__ sb) {
    const size_t n = (size_t)bd.nx * bd.ny * bd.nz;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (size_t)gridDim.x * blockDim.x) {
        float f;
        if (flags[idx] & 1) f = x[brick_slot(bid, bd, P, idx)];
        else f = (float)__dmul_rn(__dsub_rn(__dmul_rn(2.0, (double)vol[idx]), 1.0), scale);
        F[idx] = f;
        const unsigned m = __ballot_sync(0xffffffffu, f < iso);
        if ((threadIdx.x & 31) == 0) sb[idx >> 5] = m;
    }
}

// The same, four consecutive z per thread (nz % 128 == 0, 16-byte aligned grids): LDG.128 of the volume, one 32-bit load of
// four flag bytes, one LDG.128 of the brick's strip (the four z of a thread are one strip of one brick), STG.128 of F; the four
// sign-bit words of a 128-voxel segment are OR-reductions over groups of eight lanes.
__global__ void __launch_bounds__(256) k_assemble_F_sign_v4(const float4* __restrict__ vol, const uint32_t* __restrict__ flags4,
                                                            const int* __restrict__ bid, const float4* __restrict__ x, BrickDims bd,
                                                            size_t P, double scale, float iso, float4* __restrict__ F,
                                                            uint32_t* __restrict__ sb) {
    const size_t n4 = (size_t)bd.nx * bd.ny * bd.nz / 4;
    const int nz4 = bd.nz / 4;
    const int lane = threadIdx.x & 31;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 v = __ldg(vol + i);
        const uint32_t fl = __ldg(flags4 + i);
        float4 f;
        f.x = (float)__dmul_rn(__dsub_rn(__dmul_rn(2.0, (double)v.x), 1.0), scale);
        f.y = (float)__dmul_rn(__dsub_rn(__dmul_rn(2.0, (double)v.y), 1.0), scale);
        f.z = (float)__dmul_rn(__dsub_rn(__dmul_rn(2.0, (double)v.z), 1.0), scale);
        f.w = (float)__dmul_rn(__dsub_rn(__dmul_rn(2.0, (double)v.w), 1.0), scale);
        if (fl & 0x01010101u) {
            const int Z = (int)(i % nz4);
            const size_t t = i / nz4;
            const int y = (int)(t % bd.ny), xx = (int)(t / bd.ny);
            const int q = bid[((size_t)(xx >> 1) * bd.by + (y >> 1)) * bd.bz + Z];
            const float4 s = x[(size_t)((xx & 1) * 2 + (y & 1)) * P + (size_t)q];
            if (fl & 0x00000001u) f.x = s.x;
            if (fl & 0x00000100u) f.y = s.y;
            if (fl & 0x00010000u) f.z = s.z;
            if (fl & 0x01000000u) f.w = s.w;
        }
        F[i] = f;
        uint32_t m = ((f.x < iso) ? 1u : 0u) | ((f.y < iso) ? 2u : 0u) | ((f.z < iso) ? 4u : 0u) | ((f.w < iso) ? 8u : 0u);
        m <<= 4 * (lane & 7);
        m |= __shfl_xor_sync(0xffffffffu, m, 1);
        m |= __shfl_xor_sync(0xffffffffu, m, 2);
        m |= __shfl_xor_sync(0xffffffffu, m, 4);
        if ((lane & 7) == 0) sb[i >> 3] = m;
    }
}

// F for the voxels [v0, v0 + count) of the grid only (an x-slab: whole planes), written to F[0 .. count): scalar and
// four-voxels-per-thread versions (count, v0 multiples of 4 and 16-byte aligned pointers for the latter).
__global__ void __launch_bounds__(256) k_assemble_F_range(const float* __restrict__ vol, const uint8_t* __restrict__ flags,
                                                          const int* __restrict__ bid, const float* __restrict__ x, BrickDims bd,
                                                          size_t P, double scale, size_t v0, size_t count, float* __restrict__ F) {
    MRB_VOXEL_LOOP(i, count) {
        const size_t idx = v0 + i;
        float f;
        if (flags[idx] & 1) f = x[brick_slot(bid, bd, P, idx)];
        else f = (float)__dmul_rn(__dsub_rn(__dmul_rn(2.0, (double)vol[idx]), 1.0), scale);
        F[i] = f;
    }
}
__global__ void __launch_bounds__(256) k_assemble_F_range_v4(const float4* __restrict__ vol, const uint32_t* __restrict__ flags4,
                                                             const int* __restrict__ bid, const float4* __restrict__ x, BrickDims bd,
                                                             size_t P, double scale, size_t v04, size_t count4, float4* __restrict__ F) {
    const int nz4 = bd.nz / 4;
    for (size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x; j < count4; j += (size_t)gridDim.x * blockDim.x) {
        const size_t i = v04 + j;
        const float4 v = __ldg(vol + i);
        const uint32_t fl = __ldg(flags4 + i);
        float4 f;
        f.x = (float)__dmul_rn(__dsub_rn(__dmul_rn(2.0, (double)v.x), 1.0), scale);
        f.y = (float)__dmul_rn(__dsub_rn(__dmul_rn(2.0, (double)v.y), 1.0), scale);
        f.z = (float)__dmul_rn(__dsub_rn(__dmul_rn(2.0, (double)v.z), 1.0), scale);
        f.w = (float)__dmul_rn(__dsub_rn(__dmul_rn(2.0, (double)v.w), 1.0), scale);
        if (fl & 0x01010101u) {
            const int Z = (int)(i % nz4);
            const size_t t = i / nz4;
            const int y = (int)(t % bd.ny), xx = (int)(t / bd.ny);
            const int q = bid[((size_t)(xx >> 1) * bd.by + (y >> 1)) * bd.bz + Z];
            const float4 s = x[(size_t)((xx & 1) * 2 + (y & 1)) * P + (size_t)q];
            if (fl & 0x00000001u) f.x = s.x;
            if (fl & 0x00000100u) f.y = s.y;
            if (fl & 0x00010000u) f.z = s.z;
            if (fl & 0x01000000u) f.w = s.w;
        }
        F[j] = f;
    }
}

// Builds the bricks from the stage-2 byte grids, runs `iters` sweeps, applies the feature reset and writes F.
// *n_bricks / *closure_err_host are outputs; ev[6], ev[7] bracket the sweep kernel.
int bricks_solve(mrb_ctx* ctx, const float* vol, const float* fmap, const uint8_t* flags, const uint8_t* sq, Dims d,
                 int band_size, int iters, int using_features, float* F_dev, float* last_delta_host, int* n_bricks,
                 int* closure_err_host, int binary_volume, const uint8_t* brick_bits) {
    MRB_TRY(upload_tables(ctx->device));
    BrickDims bd{d.nx, d.ny, d.nz, (d.nx + 1) / 2, (d.ny + 1) / 2, (d.nz + 3) / 4};
    const size_t nb = bd.count();
    int* bpos = nullptr;
    int Q = 0;
    // brick numbering: x-slab tiles, rim bricks first (default), or plain (X,Y,Z) (MRB_QP_ORDER=xyz)
    const char* eo = getenv("MRB_QP_ORDER");
    const bool tiled = !(eo && strcmp(eo, "xyz") == 0) && iters > 0;
    int tile_T = 8;
    if (getenv("MRB_QP_TILE")) tile_T = atoi(getenv("MRB_QP_TILE"));
    if (tile_T < 1) tile_T = 1;
    // One block of 15 warps per SM: the patches are larger than with two 256-thread blocks, so fewer of their bricks
    // are rim bricks, and 480 threads leave ptxas 136 registers per thread (the update needs 127; at 512 threads the
    // limit is 128 and the allocation is fragile: one build put part of the brick on the stack, 3.0 -> 4.3 ms).
    int t_threads = 480;
    if (getenv("MRB_QP_BRICK_THREADS")) t_threads = atoi(getenv("MRB_QP_BRICK_THREADS"));
    void* t_kern = nullptr;
    switch (t_threads) {
        case 256: t_kern = (void*)k_qp_bricks_tiled<256, 2, false>; break;
        case 448: t_kern = (void*)k_qp_bricks_tiled<448, 1, false>; break;
        case 512: t_kern = (void*)k_qp_bricks_tiled<512, 1, false>; break;
        default: t_threads = 480; t_kern = (void*)k_qp_bricks_tiled<480, 1, false>; break;
    }
    int t_grid = 0, t_fit = 0;
    int* blk_nb = nullptr;
    int* tpos = nullptr;
    const TiledOrder ord = make_order(bd, tile_T);
    const BrickBitPred bitpred{brick_bits, bd, (d.nz + 31) / 32};
    if (tiled && brick_bits) MRB_TRY(compact(ctx, TiledBrickPred<BrickBitPred>{bitpred, ord}, ord.count(), &tpos, &Q));
    else if (tiled) MRB_TRY(compact(ctx, TiledBrickPred<ActiveBrickPred>{ActiveBrickPred{flags, bd}, ord}, ord.count(), &tpos, &Q));
    else MRB_TRY(compact(ctx, ActiveBrickPred{flags, bd}, nb, &bpos, &Q));
    *n_bricks = Q;
    *closure_err_host = 0;
    if (last_delta_host) *last_delta_host = 0.0f;
    int* bid = nullptr;
    MRB_TRY(arena_alloc(ctx, nb, &bid));
    MRB_CUDA(cudaMemsetAsync(bid, 0xFF, (size_t)nb * sizeof(int), ctx->stream));     // every int = -1
    if (tiled && Q > 0) {
        MRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&t_fit, t_kern, t_threads, 0));
        if (t_fit < 1) { set_error("k_qp_bricks_tiled does not fit on an SM"); return MRB_ECUDA; }
        t_grid = t_fit * ctx->sm_count;
        const int need = (Q + t_threads - 1) / t_threads;
        if (t_grid > need) {
            // Fewer bricks than threads: spread them over all SMs with smaller blocks instead of leaving SMs idle; a
            // sweep's length is then set by fewer warps per SM queueing on the L1 data pipe.
            t_grid = need;
            if (t_threads == 480 && !getenv("MRB_QP_BRICK_THREADS")) {
                const int per_sm = (Q + t_fit * ctx->sm_count - 1) / (t_fit * ctx->sm_count);
                int th = (per_sm + 31) / 32 * 32;
                const int th_min = getenv("MRB_QP_MIN_THREADS") ? atoi(getenv("MRB_QP_MIN_THREADS")) : 352;   // measured 120^3..200^3
                if (th < th_min) th = th_min;
                if (th < t_threads) { t_threads = th; t_grid = std::min(t_fit * ctx->sm_count, (Q + th - 1) / th); }
            }
        }
        int *bid1 = nullptr, *sb = nullptr;
        MRB_TRY(arena_alloc(ctx, nb, &bid1));
        MRB_TRY(arena_alloc(ctx, (size_t)Q + 1, &sb));
        MRB_TRY(arena_alloc(ctx, (size_t)Q, &bpos));
        MRB_TRY(arena_alloc(ctx, (size_t)t_grid, &blk_nb));
        MRB_CUDA(cudaMemsetAsync(blk_nb, 0, (size_t)t_grid * sizeof(int), ctx->stream));      // blocks with an empty range
        MRB_CUDA(cudaMemsetAsync(bid1, 0xFF, (size_t)nb * sizeof(int), ctx->stream));     // every int = -1
        MRB_LAUNCH(ctx, k_tiled_scatter, cdiv(Q, 256), 256, 0, tpos, Q, ord, bid1);
        MRB_LAUNCH(ctx, k_tiled_boundary, cdiv(Q, 256), 256, 0, tpos, Q, t_grid, ord, bid1, sb);
        MRB_TRY(scan_exclusive(ctx, sb, sb, (size_t)Q, sb + Q));
        MRB_LAUNCH(ctx, k_tiled_final, cdiv(Q, 256), 256, 0, tpos, sb, Q, t_grid, ord, bpos, bid, blk_nb,
                   ((getenv("MRB_QP_LATE") && atoi(getenv("MRB_QP_LATE")) == 0) || (getenv("MRB_QP_ONCHIP") && atoi(getenv("MRB_QP_ONCHIP")))) ? 0 : t_threads);
    }
    const size_t Qn = (size_t)(Q > 0 ? Q : 1);
    const size_t P = tiled ? tiled_stride(Q) : (size_t)Q + 1;       // plane stride of the SoA iterate
    uint2* code = nullptr;
    uint4* fields = nullptr;
    float4 *xa = nullptr, *xb = nullptr;
    int4* nbA = nullptr;
    int2* nbB = nullptr;
    MRB_TRY(arena_alloc(ctx, Qn, &code));
    MRB_TRY(arena_alloc(ctx, 2 * Qn, &fields));
    MRB_TRY(arena_alloc(ctx, 4 * P, &xa));
    MRB_TRY(arena_alloc(ctx, 4 * P, &xb));
    MRB_TRY(arena_alloc(ctx, Qn, &nbA));
    MRB_TRY(arena_alloc(ctx, Qn, &nbB));
    const float* xres = (const float*)xa;
    if (Q > 0) {
        if (!tiled) MRB_LAUNCH(ctx, k_scatter_brick_ids, cdiv(Q, 256), 256, 0, bpos, Q, bid);
        int* err = ctx->d_mail + 24;
        MRB_CUDA(cudaMemsetAsync(err, 0, sizeof(int), ctx->stream));
        for (int k = 0; k < 4; k++) MRB_CUDA(cudaMemsetAsync(xa + (size_t)k * P + Q, 0, (P - Q) * sizeof(float4), ctx->stream));
        MRB_CUDA(cudaMemsetAsync(xb, 0, 4 * P * sizeof(float4), ctx->stream));
        MRB_LAUNCH(ctx, k_brick_build, cdiv(Q, 128), 128, 0, vol, sq, flags, bid, bpos, Q, bd, code, fields, xa, nbA, nbB, err, P);
        MRB_CUDA(cudaMemcpyAsync(ctx->h_mail + 24, err, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        if (!(iters > 0 && tiled)) {
            MRB_CUDA(cudaStreamSynchronize(ctx->stream));
            *closure_err_host = ctx->h_mail[24];
            if (*closure_err_host) return MRB_OK;      // caller reports MRB_EBORDER
        }
        if (iters > 0 && tiled) {
            unsigned* md = (unsigned*)(ctx->d_mail + 8);
            unsigned* bar = (unsigned*)ctx->d_bar;
            MRB_CUDA(cudaMemsetAsync(md, 0, sizeof(unsigned), ctx->stream));
            MRB_CUDA(cudaMemsetAsync(bar, 0, 64 * sizeof(unsigned), ctx->stream));
            const int grid = t_grid, NW = (grid + 31) / 32;
            unsigned *progress = nullptr, *bitmap = nullptr;
            int *plist = nullptr, *pcount = nullptr;
            MRB_TRY(arena_alloc(ctx, (size_t)grid * kProgressStride, &progress));
            MRB_TRY(arena_alloc(ctx, (size_t)grid * kChunkListCap, &plist));
            MRB_TRY(arena_alloc(ctx, (size_t)grid, &pcount));
            MRB_TRY(arena_alloc(ctx, (size_t)grid * NW, &bitmap));
            int* ovf = ctx->d_mail + 50;
            MRB_CUDA(cudaMemsetAsync(progress, 0, (size_t)grid * kProgressStride * sizeof(unsigned), ctx->stream));
            MRB_CUDA(cudaMemsetAsync(bitmap, 0, (size_t)grid * NW * sizeof(unsigned), ctx->stream));
            MRB_CUDA(cudaMemsetAsync(pcount, 0, (size_t)grid * sizeof(int), ctx->stream));
            MRB_CUDA(cudaMemsetAsync(ovf, 0, sizeof(int), ctx->stream));
            MRB_LAUNCH(ctx, k_tiled_bitmap, cdiv(Q, 256), 256, 0, nbA, nbB, Q, grid, NW, bitmap);
            MRB_LAUNCH(ctx, k_chunk_lists, cdiv((size_t)grid * 32, 256), 256, 0, bitmap, grid, NW, plist, pcount, ovf);
            int overflow = 0;
            MRB_TRY(read_int(ctx, ovf, &overflow));            // one host round trip for this and the closure check above
            *closure_err_host = ctx->h_mail[24];
            if (*closure_err_host) return MRB_OK;              // caller reports MRB_EBORDER
            int dbg = getenv("MRB_QP_DEBUG") ? atoi(getenv("MRB_QP_DEBUG")) : 0;
            // who publishes a block's sweep: the last warp when the grid fills the GPU (warp 0 polls meanwhile: 2.69 -> 2.66 ms at
            // 256^3), thread 0 when the blocks were shrunk (there the last warp is first in line for rim bricks: 172^3 2.19 vs 1.99 ms)
            if (t_threads < 480) dbg |= 32;
            if (overflow) dbg |= 4;        // a block with more than kChunkListCap neighbour blocks: grid barrier per sweep
            int cap = (Q + grid - 1) / grid + 9;       // block ranges are whole groups of 8 bricks
            // expanded constants (168 B per brick) for binary volumes if one block's share fits, else the 64-byte form
            int expanded = binary_volume ? 1 : 0;
            size_t smem = (size_t)cap * 168;
            const size_t smem_max = (size_t)220 * 1024 / (size_t)t_fit;
            if (getenv("MRB_QP_BRICK_COMPACT") || !expanded || smem > smem_max) { expanded = 0; smem = (size_t)cap * 64; }
            if (getenv("MRB_QP_BRICK_NOSMEM") || smem > 96 * 1024 / (size_t)t_fit) { if (!expanded) { cap = 0; smem = 0; } }
            MRB_CUDA(allow_dyn_smem((const void*)t_kern, smem));
            int its = iters, QQ = Q;
            // On-chip kernel (the block's iterate in shared memory for all sweeps): binary volumes whose per-block share fits
            // (up to kOcCap bricks and kOcHalo foreign neighbour bricks per SM: band 4 up to about 256^3).  OPT-IN with
            // MRB_QP_ONCHIP=1: measured on B200 it is bit-identical but slower than the tiled kernel (3.7 ms against 2.7 ms
            // at 256^3; profiles/r02_summary.md says why), so the tiled kernel stays the default.
            const int need_cap = (Q + grid - 1) / grid + 9;
            const char* eoc = getenv("MRB_QP_ONCHIP");
            bool onchip = binary_volume && t_fit == 1 && !(dbg & 5) && (eoc && atoi(eoc) != 0) && need_cap <= kOcCap &&
                          (t_threads == 480 || t_threads < 480);
            int oc_threads = t_threads;
            int *hcount = nullptr, *hgid = nullptr;
            unsigned short* hslot = nullptr;
            if (onchip) {
                int* hmax = ctx->d_mail + 51;
                MRB_TRY(arena_alloc(ctx, (size_t)grid, &hcount));
                MRB_TRY(arena_alloc(ctx, (size_t)grid * kOcHalo, &hgid));
                MRB_TRY(arena_alloc(ctx, (size_t)Q * 6, &hslot));
                MRB_CUDA(cudaMemsetAsync(hcount, 0, (size_t)grid * sizeof(int), ctx->stream));
                MRB_CUDA(cudaMemsetAsync(hmax, 0, sizeof(int), ctx->stream));
                MRB_LAUNCH(ctx, k_onchip_halo, cdiv(Q, 256), 256, 0, nbA, nbB, Q, grid, kOcHalo, hcount, hgid, hslot, hmax);
                int hm = 0;
                MRB_TRY(read_int(ctx, hmax, &hm));
                if (getenv("MRB_QP_VERBOSE")) {
                    std::vector<int> hb(grid), hh(grid);
                    MRB_CUDA(cudaMemcpyAsync(hb.data(), blk_nb, grid * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
                    MRB_CUDA(cudaMemcpyAsync(hh.data(), hcount, grid * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
                    MRB_CUDA(cudaStreamSynchronize(ctx->stream));
                    std::sort(hb.begin(), hb.end()); std::sort(hh.begin(), hh.end());
                    fprintf(stderr, "mrb: rim bricks per block min %d p25 %d median %d p75 %d p90 %d max %d; halo min %d median %d p90 %d max %d\n",
                            hb[0], hb[grid / 4], hb[grid / 2], hb[3 * grid / 4], hb[9 * grid / 10], hb[grid - 1], hh[0], hh[grid / 2], hh[9 * grid / 10], hh[grid - 1]);
                }
                if (getenv("MRB_QP_VERBOSE")) fprintf(stderr, "mrb: on-chip kernel: %d bricks per block (cap %d), largest halo %d (cap %d), %d threads\n", need_cap, kOcCap, hm, kOcHalo, oc_threads > t_threads ? oc_threads : (hm > oc_threads ? (hm + 31) / 32 * 32 : oc_threads));
                // Threads per block: MRB_QP_OC_THREADS (default: the tiled kernel's choice, 480 for grids that fill the GPU = two
                // passes over ~830 bricks with four warps free in the second one to fetch the halo).
                if (getenv("MRB_QP_OC_THREADS")) oc_threads = atoi(getenv("MRB_QP_OC_THREADS"));
                oc_threads = std::max(32, std::min(kOcThreadsMax, oc_threads / 32 * 32));
                if (hm > kOcHalo) onchip = false;
            }
            if (onchip) {
                typedef OnchipLayout<kOcCap, kOcHalo> OL;
                void* oc_kern = oc_threads <= 352 ? (void*)k_qp_bricks_onchip<352, kOcCap, kOcHalo, false>
                                                  : (void*)k_qp_bricks_onchip<kOcThreadsMax, kOcCap, kOcHalo, false>;
                if (getenv("MRB_QP_OC_BOUND")) {       // tuning aid: register budget of the kernel (launch bound), not its block size
                    const int bnd = atoi(getenv("MRB_QP_OC_BOUND"));
                    if (bnd == 480) oc_kern = (void*)k_qp_bricks_onchip<480, kOcCap, kOcHalo, false>;
                    if (bnd == 640) oc_kern = (void*)k_qp_bricks_onchip<640, kOcCap, kOcHalo, false>;
                    if (bnd == 544) oc_kern = (void*)k_qp_bricks_onchip<544, kOcCap, kOcHalo, false>;
                }
                int prefetch = getenv("MRB_QP_OC_MODE") ? atoi(getenv("MRB_QP_OC_MODE")) : 0;       // 2: no free-warp fetch (timing aid)
                unsigned long long* trace = nullptr;
                const bool want_trace = getenv("MRB_QP_TRACE") && atoi(getenv("MRB_QP_TRACE")) && iters >= kTraceFirst + kTraceSweeps;
                if (want_trace) {
                    oc_kern = oc_threads <= 352 ? (void*)k_qp_bricks_onchip<352, kOcCap, kOcHalo, true>
                                                : (void*)k_qp_bricks_onchip<kOcThreadsMax, kOcCap, kOcHalo, true>;
                    MRB_TRY(arena_alloc(ctx, (size_t)grid * kTraceSweeps * 8, &trace));
                    MRB_CUDA(cudaMemsetAsync(trace, 0, (size_t)grid * kTraceSweeps * 8 * sizeof(unsigned long long), ctx->stream));
                }
                // exchange buffer of the rim bricks: 2 parities x Q bricks x 16 words {value, sweep number}, numbers start at 1
                ulonglong2* ex = nullptr;
                MRB_TRY(arena_alloc(ctx, (size_t)Q * 16, &ex));
                MRB_CUDA(cudaMemsetAsync(ex, 0, (size_t)Q * 16 * sizeof(ulonglong2), ctx->stream));
                MRB_CUDA(cudaFuncSetAttribute(oc_kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)OL::kBytes));
                void* args[] = {(void*)&code, (void*)&fields, (void*)&nbA, (void*)&nbB, (void*)&xa, (void*)&xb, (void*)&QQ, (void*)&its,
                                (void*)&md, (void*)&blk_nb, (void*)&P, (void*)&ex, (void*)&hcount, (void*)&hgid, (void*)&hslot, (void*)&prefetch,
                                (void*)&trace};
                MRB_CUDA(cudaEventRecord(ctx->ev[6], ctx->stream));
                MRB_CUDA(cudaLaunchCooperativeKernel(oc_kern, dim3(grid), dim3(oc_threads), args, OL::kBytes, ctx->stream));
                MRB_CUDA(cudaEventRecord(ctx->ev[7], ctx->stream));
                if (want_trace) {
                    // tuning aid: dump the stamps (ns) for tools/qp_trace.py
                    std::vector<unsigned long long> h((size_t)grid * kTraceSweeps * 8);
                    MRB_CUDA(cudaMemcpyAsync(h.data(), trace, h.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
                    MRB_CUDA(cudaStreamSynchronize(ctx->stream));
                    const char* path = getenv("MRB_QP_TRACE_FILE") ? getenv("MRB_QP_TRACE_FILE") : "qp_trace.bin";
                    if (FILE* f = fopen(path, "wb")) {
                        int hdr[4] = {grid, kTraceSweeps, 8, 0};
                        fwrite(hdr, sizeof(int), 4, f);
                        fwrite(h.data(), sizeof(unsigned long long), h.size(), f);
                        fclose(f);
                    }
                }
            } else {
                TShard no_shard;
                memset(&no_shard, 0, sizeof(no_shard));
                unsigned long long* trace = nullptr;
                const bool want_trace = getenv("MRB_QP_TRACE") && atoi(getenv("MRB_QP_TRACE")) && iters >= kTraceFirst + kTraceSweeps && t_threads <= 480;
                if (want_trace) {
                    if (t_threads <= 480) t_kern = (void*)k_qp_bricks_tiled<480, 1, false, true>;     // launch bound >= the block size
                    MRB_CUDA(allow_dyn_smem((const void*)t_kern, smem));
                    MRB_TRY(arena_alloc(ctx, (size_t)grid * kTraceSweeps * 8, &trace));
                    MRB_CUDA(cudaMemsetAsync(trace, 0, (size_t)grid * kTraceSweeps * 8 * sizeof(unsigned long long), ctx->stream));
                }
                void* args[] = {(void*)&code, (void*)&fields, (void*)&nbA, (void*)&nbB, (void*)&xa, (void*)&xb, (void*)&QQ, (void*)&its,
                                (void*)&md, (void*)&bar, (void*)&dbg, (void*)&blk_nb, (void*)&progress, (void*)&cap, (void*)&plist, (void*)&pcount,
                                (void*)&expanded, (void*)&P, (void*)&no_shard, (void*)&trace};
                MRB_CUDA(cudaEventRecord(ctx->ev[6], ctx->stream));
                MRB_CUDA(cudaLaunchCooperativeKernel(t_kern, dim3(grid), dim3(t_threads), args, smem, ctx->stream));
                MRB_CUDA(cudaEventRecord(ctx->ev[7], ctx->stream));
                if (want_trace) {
                    std::vector<unsigned long long> h((size_t)grid * kTraceSweeps * 8);
                    MRB_CUDA(cudaMemcpyAsync(h.data(), trace, h.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
                    MRB_CUDA(cudaStreamSynchronize(ctx->stream));
                    const char* path = getenv("MRB_QP_TRACE_FILE") ? getenv("MRB_QP_TRACE_FILE") : "qp_trace.bin";
                    std::vector<int> hnb(grid), hsz(grid);
                    MRB_CUDA(cudaMemcpy(hnb.data(), blk_nb, grid * sizeof(int), cudaMemcpyDeviceToHost));
                    for (int b = 0; b < grid; b++) hsz[b] = tiled_begin(b + 1, Q, grid) - tiled_begin(b, Q, grid);
                    if (FILE* f = fopen(path, "wb")) {
                        int hdr[4] = {grid, kTraceSweeps, 8, 1};
                        fwrite(hdr, sizeof(int), 4, f);
                        fwrite(h.data(), sizeof(unsigned long long), h.size(), f);
                        fwrite(hnb.data(), sizeof(int), hnb.size(), f);       // rim bricks per block, bricks per block
                        fwrite(hsz.data(), sizeof(int), hsz.size(), f);
                        fclose(f);
                    }
                }
            }
            ctx->launches++;
            xres = (iters % 2) ? (const float*)xb : (const float*)xa;
            // the residual travels to h_mail[8] with the stream; the caller reads it after its own synchronisation
            // (no host round trip of its own between the sweeps and the F assembly)
            if (last_delta_host) MRB_CUDA(cudaMemcpyAsync(ctx->h_mail + 8, md, sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
        } else if (iters > 0) {
            unsigned* md = (unsigned*)(ctx->d_mail + 8);
            unsigned* bar = (unsigned*)ctx->d_bar;
            MRB_CUDA(cudaMemsetAsync(md, 0, sizeof(unsigned), ctx->stream));
            MRB_CUDA(cudaMemsetAsync(bar, 0, 64 * sizeof(unsigned), ctx->stream));
            int threads = 256;
            const char* e = getenv("MRB_QP_BRICK_THREADS");
            if (e) threads = atoi(e);
            void* kern = nullptr;
            // tuning knob: MRB_QP_SYNC=flow selects the warp-dataflow synchronisation (measured slower on B200:
            // 3.9-4.4 ms vs 3.5 ms -- one fence per warp and sweep costs more than the overlap wins)
            bool flow = getenv("MRB_QP_SYNC") && strcmp(getenv("MRB_QP_SYNC"), "flow") == 0;
            void* kern_block = nullptr;
            if (threads == 128) { kern = (void*)k_qp_bricks<128, 4, true>; kern_block = (void*)k_qp_bricks<128, 4, false>; }
            else if (threads == 512) { kern = (void*)k_qp_bricks<512, 1, true>; kern_block = (void*)k_qp_bricks<512, 1, false>; }
            else { threads = 256; kern = (void*)k_qp_bricks<256, 2, true>; kern_block = (void*)k_qp_bricks<256, 2, false>; }
            if (!flow) kern = kern_block;
            int fit = 0;
            MRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&fit, kern, threads, 0));
            if (fit < 1) { set_error("k_qp_bricks does not fit on an SM"); return MRB_ECUDA; }
            int grid = fit * ctx->sm_count;
            int need = (Q + threads - 1) / threads;
            if (grid > need) grid = need;
            // neighbour block ranges + progress words for the point-to-point synchronisation
            int *blk_lo = nullptr, *blk_hi = nullptr;
            unsigned* progress = nullptr;
            MRB_TRY(arena_alloc(ctx, (size_t)grid, &blk_lo));
            MRB_TRY(arena_alloc(ctx, (size_t)grid, &blk_hi));
            MRB_TRY(arena_alloc(ctx, (size_t)grid * kProgressStride, &progress));
            MRB_CUDA(cudaMemsetAsync(progress, 0, (size_t)grid * kProgressStride * sizeof(unsigned), ctx->stream));
            MRB_LAUNCH(ctx, k_block_ranges_init, cdiv(grid, 256), 256, 0, blk_lo, blk_hi, grid);
            MRB_LAUNCH(ctx, k_block_ranges, cdiv(Q, 256), 256, 0, nbA, nbB, Q, grid, blk_lo, blk_hi);
            int cap = (Q + grid - 1) / grid + 1;
            size_t smem = (size_t)cap * 64;
            if (getenv("MRB_QP_BRICK_NOSMEM") || smem > 96 * 1024 / (size_t)fit) { cap = 0; smem = 0; }
            if (smem > 48 * 1024) {
                MRB_CUDA(allow_dyn_smem((const void*)kern, smem));
                MRB_CUDA(allow_dyn_smem((const void*)kern_block, smem));
            }
            // producer lists of the warp chunks (warp-dataflow synchronisation)
            int *plist = nullptr, *pcount = nullptr;
            const int W = threads / 32, NC = grid * W, NW = (NC + 31) / 32;
            MRB_TRY(arena_alloc(ctx, (size_t)NC * kChunkListCap, &plist));
            MRB_TRY(arena_alloc(ctx, (size_t)NC, &pcount));
            if (flow) {
                unsigned* bitmap = nullptr;
                MRB_TRY(arena_alloc(ctx, (size_t)NC * NW, &bitmap));
                int* ovf = ctx->d_mail + 50;
                MRB_CUDA(cudaMemsetAsync(bitmap, 0, (size_t)NC * NW * sizeof(unsigned), ctx->stream));
                MRB_CUDA(cudaMemsetAsync(pcount, 0, (size_t)NC * sizeof(int), ctx->stream));
                MRB_CUDA(cudaMemsetAsync(ovf, 0, sizeof(int), ctx->stream));
                MRB_LAUNCH(ctx, k_chunk_bitmap, cdiv(Q, 256), 256, 0, nbA, nbB, Q, grid, W, NW, bitmap);
                MRB_LAUNCH(ctx, k_chunk_lists, cdiv((size_t)NC * 32, 256), 256, 0, bitmap, NC, NW, plist, pcount, ovf);
                int overflow = 0;
                MRB_TRY(read_int(ctx, ovf, &overflow));
                if (overflow) { flow = false; kern = kern_block; }      // a chunk with > 64 producers: block-level sync
                // progress words: one line per chunk
                MRB_TRY(arena_alloc(ctx, (size_t)NC * kProgressStride, &progress));
                MRB_CUDA(cudaMemsetAsync(progress, 0, (size_t)NC * kProgressStride * sizeof(unsigned), ctx->stream));
            }
            int its = iters, QQ = Q;
            int dbg = getenv("MRB_QP_DEBUG") ? atoi(getenv("MRB_QP_DEBUG")) : 0;   // timing aids (wrong results): 1 no barrier, 2 no work
            void* args[] = {(void*)&code, (void*)&fields, (void*)&nbA, (void*)&nbB, (void*)&xa, (void*)&xb, (void*)&QQ,
                            (void*)&its, (void*)&md, (void*)&bar, (void*)&dbg, (void*)&blk_lo, (void*)&blk_hi, (void*)&progress, (void*)&cap, (void*)&plist, (void*)&pcount};
            MRB_CUDA(cudaEventRecord(ctx->ev[6], ctx->stream));
            MRB_CUDA(cudaLaunchCooperativeKernel(kern, dim3(grid), dim3(threads), args, smem, ctx->stream));
            MRB_CUDA(cudaEventRecord(ctx->ev[7], ctx->stream));
            ctx->launches++;
            xres = (iters % 2) ? (const float*)xb : (const float*)xa;
            if (last_delta_host) {
                MRB_CUDA(cudaMemcpyAsync(ctx->h_mail + 8, md, sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
                MRB_CUDA(cudaStreamSynchronize(ctx->stream));
                unsigned bits = (unsigned)ctx->h_mail[8];
                memcpy(last_delta_host, &bits, sizeof(float));
            }
        }
    }
    const size_t n = d.count();
    if (using_features && fmap && Q > 0) {
        int* flag = ctx->d_mail + 16;
        MRB_CUDA(cudaMemsetAsync(flag, 0, sizeof(int), ctx->stream));
        const unsigned g = grid_for(ctx, n, 256);
        MRB_LAUNCH(ctx, k_feat_check_b, g, 256, 0, fmap, flags, n, flag);
        MRB_LAUNCH(ctx, k_feat_apply_b, g, 256, 0, fmap, flags, bid, bd, P, flag, (float*)xres);
    }
    ctx->sign_bits = nullptr;
    if (ctx->want_sign_bits && d.nz % 32 == 0 && n % 256 == 0) {
        // full warps over whole words: the grid-stride loop keeps every warp on 32 consecutive z of one row
        uint32_t* sb = nullptr;
        MRB_TRY(arena_alloc(ctx, n / 32, &sb));
        if (d.nz % 128 == 0 && (((uintptr_t)vol | (uintptr_t)F_dev | (uintptr_t)flags) & 15) == 0 && !getenv("MRB_PACK_SCALAR"))
            MRB_LAUNCH(ctx, k_assemble_F_sign_v4, (unsigned)ctx->sm_count * 16, 256, 0, (const float4*)vol, (const uint32_t*)flags, bid,
                       (const float4*)xres, bd, P, (double)band_size + 1.0, ctx->sign_iso, (float4*)F_dev, sb);
        else
            MRB_LAUNCH(ctx, k_assemble_F_sign, grid_for(ctx, n, 256), 256, 0, vol, flags, bid, xres, bd, P, (double)band_size + 1.0,
                       ctx->sign_iso, F_dev, sb);
        ctx->sign_bits = sb;
        ctx->sign_F = F_dev;
        return MRB_OK;
    }
    MRB_LAUNCH(ctx, k_assemble_F_b, grid_for(ctx, n, 256), 256, 0, vol, flags, bid, xres, bd, P, (double)band_size + 1.0,
               F_dev);
    return MRB_OK;
}

// ---- stepwise solver ("plan") for slab-sharded multi-GPU runs -----------------------------------------------
// The same bricks and the same per-brick update, but driven one sweep at a time over a caller-chosen brick
// range, with the iterate in caller-owned device buffers.  Bricks are numbered (X,Y,Z) with X slowest, so an
// x-slab of the grid is a contiguous brick range and its one-brick-plane halo is the contiguous range of the
// neighbouring X plane: ranks that built the plan from the same volume share the numbering and exchange
// plain sub-ranges of the four SoA planes between sweeps (mesh_reconstruction_b200/slab.py).
int stage12_binary(mrb_ctx* ctx, const float* vol, Dims d, int W, uint8_t* sq, uint8_t* flags, int* counts, int* binary_out,
                   uint8_t** brick_bits_out);

__global__ void __launch_bounds__(256) k_qp_bricks_range(const uint2* __restrict__ code, const uint4* __restrict__ fields,
                                                         const int4* __restrict__ nbA, const int2* __restrict__ nbB,
                                                         const float4* in, float4* out, int Q, int q0, int q1) {
    __shared__ float s_sqrt[256];
    __shared__ float s_inv[32];
    s_sqrt[threadIdx.x] = c_brick_sqrt[threadIdx.x];
    if (threadIdx.x < 32) s_inv[threadIdx.x] = c_brick_invdg[threadIdx.x];
    __syncthreads();
    const int q = q0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= q1) return;
    brick_update(q, (size_t)Q + 1, __ldg(nbA + q), __ldg(nbB + q), __ldg(code + q), __ldg(fields + q),
                 __ldg(fields + (size_t)Q + q), in, out, s_sqrt, s_inv, false);
}

// ---- slab-sharded persistent kernel: sweep + halo exchange over NVLink peer memory in ONE kernel ---------------
// Rank r owns the brick range [qa, qb) (its x-slab).  The iterate buffers of all ranks are peer-mapped (symmetric
// memory); the bricks of the first / last owned brick plane are stored into the neighbour rank's buffer as well,
// so the halo travels with the sweep's own stores, tile by tile, instead of in a separate exchange step.
// Synchronisation per sweep: the local neighbour-only progress words (as in k_qp_bricks) plus, for the blocks
// that own boundary-plane bricks, one flag per direction in the neighbour's memory: each such block fences at
// system scope and bumps a local counter; the last one stores the sweep number into the neighbour's flag; the
// same blocks wait for the neighbour's flag before the next sweep (which also protects the ping-pong buffers:
// a neighbour only starts writing sweep k+1 halos after it has seen our sweep-k flag).
struct ShardArgs {
    int qa, qb;               // owned brick range
    int lo0, lo1, hi0, hi1;   // first / last owned plane: [lo0,lo1) is sent down, [hi0,hi1) is sent up (empty if no neighbour)
    float4* peer_lo[2];       // lower neighbour's iterate buffers (A, B), nullptr if none
    float4* peer_hi[2];
    unsigned* flag_lo_remote; // lower neighbour's "from_upper" slots (one per boundary block of this rank)
    unsigned* flag_hi_remote; // upper neighbour's "from_lower" slots
    unsigned* flags_local;    // [0, 128) from_lower slots, [128, 256) from_upper slots
    int first_lo, first_hi;   // first block of this rank that owns lo / hi plane bricks (slot = block - first)
    int nrem_lo, nrem_hi;     // boundary blocks of the lower / upper neighbour (slots to wait for)
};

__device__ __forceinline__ void spin_until(const unsigned* p, unsigned target) {
    unsigned v, spins = 0;
    do {
        asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
        if (++spins == (1u << 27)) __trap();          // a lost neighbour must not hang the GPU
    } while (v < target);
}

template <int Threads, int MinBlocks>
__global__ void __launch_bounds__(Threads, MinBlocks) k_qp_bricks_sharded(const uint2* __restrict__ code, const uint4* __restrict__ fields,
                                                                          const int4* __restrict__ nbA, const int2* __restrict__ nbB,
                                                                          float4* xa, float4* xb, int Q, int iters, ShardArgs sh,
                                                                          const int* __restrict__ blk_lo, const int* __restrict__ blk_hi,
                                                                          unsigned* progress) {
    __shared__ float s_sqrt[256];
    __shared__ float s_inv[32];
    for (int i = threadIdx.x; i < 256; i += Threads) s_sqrt[i] = c_brick_sqrt[i];
    if (threadIdx.x < 32) s_inv[threadIdx.x] = c_brick_invdg[threadIdx.x];
    __syncthreads();
    const size_t P = (size_t)Q + 1;
    const int Ql = sh.qb - sh.qa;
    const int q_begin = sh.qa + (int)((long long)Ql * blockIdx.x / gridDim.x);
    const int q_end = sh.qa + (int)((long long)Ql * (blockIdx.x + 1) / gridDim.x);
    const int nlo = blk_lo[blockIdx.x], nhi = blk_hi[blockIdx.x];
    const bool has_lo = sh.peer_lo[0] != nullptr && q_begin < sh.lo1 && q_end > sh.lo0;
    const bool has_hi = sh.peer_hi[0] != nullptr && q_begin < sh.hi1 && q_end > sh.hi0;
    float4* in = xa;
    float4* out = xb;
    for (int it = 0; it < iters; it++) {
        float4* plo = (it & 1) ? sh.peer_lo[0] : sh.peer_lo[1];       // sweep `it` writes buffer B when it is even
        float4* phi = (it & 1) ? sh.peer_hi[0] : sh.peer_hi[1];
        for (int q = q_begin + threadIdx.x; q < q_end; q += Threads) {
            const int4 na = __ldg(nbA + q);
            const int2 nz = __ldg(nbB + q);
            brick_update(q, P, na, nz, __ldg(code + q), __ldg(fields + q), __ldg(fields + (size_t)Q + q), in, out, s_sqrt, s_inv, false);
            // boundary planes: the same values go into the neighbour's halo over NVLink
            if (plo && q >= sh.lo0 && q < sh.lo1) {
#pragma unroll
                for (int k = 0; k < 4; k++) plo[k * P + q] = out[k * P + q];
            }
            if (phi && q >= sh.hi0 && q < sh.hi1) {
#pragma unroll
                for (int k = 0; k < 4; k++) phi[k * P + q] = out[k * P + q];
            }
        }
        float4* t = in; in = out; out = t;
        if (it + 1 == iters) break;
        const unsigned epoch = (unsigned)(it + 1);
        __syncthreads();
        if (threadIdx.x == 0) {
            // ONE fence (system scope for boundary blocks: it also covers the stores into the neighbour GPU, which
            // the bar.sync above made cumulative), then relaxed stores of the sweep number: the block's own progress
            // word and, for boundary blocks, its slot in the neighbour's memory
            if (has_lo || has_hi) asm volatile("fence.acq_rel.sys;" ::: "memory");
            else asm volatile("fence.acq_rel.gpu;" ::: "memory");
            asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(progress + (size_t)blockIdx.x * kProgressStride), "r"(epoch) : "memory");
            if (has_lo)
                asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(sh.flag_lo_remote + (blockIdx.x - sh.first_lo)), "r"(epoch) : "memory");
            if (has_hi)
                asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(sh.flag_hi_remote + (blockIdx.x - sh.first_hi)), "r"(epoch) : "memory");
        }
        // local neighbours (block-level progress words)
        for (int c = nlo + (int)threadIdx.x; c <= nhi; c += Threads) {
            if (c == (int)blockIdx.x) continue;
            unsigned v;
            do {
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(progress + (size_t)c * kProgressStride) : "memory");
            } while (v < epoch);
        }
        // remote neighbours (halo planes written by the other GPU)
        if (has_lo && (int)threadIdx.x < sh.nrem_lo) { spin_until(sh.flags_local + threadIdx.x, epoch); __threadfence_system(); }
        if (has_hi && (int)threadIdx.x < sh.nrem_hi) { spin_until(sh.flags_local + kShardSlots + threadIdx.x, epoch); __threadfence_system(); }
        __syncthreads();
    }
}

// block -> [lo, hi] neighbour blocks for a local range [qa, qb) split over G blocks
__global__ void k_block_ranges_local(const int4* __restrict__ nbA, const int2* __restrict__ nbB, int qa, int qb, int G,
                                     int* __restrict__ lo, int* __restrict__ hi) {
    int q = qa + blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= qb) return;
    const int Ql = qb - qa;
    const int b = brick_owner(q - qa, Ql, G);
    const int4 a = nbA[q];
    const int2 z = nbB[q];
    const int n[6] = {a.x, a.y, a.z, a.w, z.x, z.y};
    int mn = b, mx = b;
    for (int i = 0; i < 6; i++)
        if (n[i] >= qa && n[i] < qb) { int o = brick_owner(n[i] - qa, Ql, G); mn = min(mn, o); mx = max(mx, o); }
    if (mn < b) atomicMin(&lo[b], mn);
    if (mx > b) atomicMax(&hi[b], mx);
}

}  // namespace mrb

namespace mrb {

// ---- slab-local brick numbering for the sharded tiled kernel -------------------------------------------------------
// A rank owns the brick planes [X0, X0 + nX).  Its bricks are renumbered like the single-GPU path (x-slab tiles, rim
// bricks first inside each block's range); the bricks of the planes X0 - 1 and X0 + nX (owned by the neighbour ranks)
// follow as halo slots in the plan's global (Y,Z) order, which both sides know without talking to each other.
struct ShardSub {
    BrickDims bd;
    int T, X0, nX;
    __host__ __device__ size_t count() const { return (size_t)((nX + T - 1) / T) * T * bd.by * bd.bz; }
    __host__ __device__ long long position(size_t i) const {        // global brick-grid position, -1 in the padding
        const int Z = (int)(i % bd.bz);
        size_t t = i / bd.bz;
        const int x = (int)(t % T);
        t /= T;
        const int Y = (int)(t % bd.by);
        const int Xl = (int)(t / bd.by) * T + x;
        if (Xl >= nX) return -1;
        return ((long long)(X0 + Xl) * bd.by + Y) * bd.bz + Z;
    }
    // index into the local id map, which covers the planes X0 - 1 .. X0 + nX
    __host__ __device__ size_t map_index(long long p) const { return (size_t)(p - (long long)(X0 - 1) * bd.by * bd.bz); }
    __host__ __device__ size_t map_size() const { return (size_t)(nX + 2) * bd.by * bd.bz; }
};
struct ShardSubPred {
    const int* gbid;
    ShardSub sub;
    __device__ bool operator()(size_t i) const {
        const long long p = sub.position(i);
        return p >= 0 && gbid[p] >= 0;
    }
};
__global__ void k_sh_scatter1(const int* __restrict__ tpos, int Qo, ShardSub sub, int* __restrict__ lmap1) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < Qo) lmap1[sub.map_index(sub.position((size_t)tpos[i]))] = i;
}
// halo plane X: local id = base + rank of the brick inside its plane in the global numbering
__global__ void k_sh_halo_ids(const int* __restrict__ gbid, ShardSub sub, int X, int base, int plane_first, int* __restrict__ lmap1,
                              int* __restrict__ lmap) {
    const int n = sub.bd.by * sub.bd.bz;
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const long long p = (long long)X * n + i;
    const int g = gbid[p];
    if (g < 0) return;
    const int id = base + (g - plane_first);
    lmap1[sub.map_index(p)] = id;
    lmap[sub.map_index(p)] = id;
}
__global__ void k_sh_boundary(const int* __restrict__ tpos, int Qo, int G, ShardSub sub, const int* __restrict__ lmap1,
                              int* __restrict__ isb) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= Qo) return;
    const BrickDims bd = sub.bd;
    const long long p = sub.position((size_t)tpos[i]);
    const int Z = (int)(p % bd.bz);
    const long long t = p / bd.bz;
    const int Y = (int)(t % bd.by), X = (int)(t / bd.by);
    const int own = tiled_owner(i, Qo, G);
    const long long cx = (long long)bd.by * bd.bz, cy = bd.bz;
    bool rim = false;
    auto look = [&](bool ok, long long dp) {
        if (!ok) return;
        const int id = lmap1[sub.map_index(p + dp)];
        if (id >= Qo || (id >= 0 && tiled_owner(id, Qo, G) != own)) rim = true;     // halo slot or another block's brick
    };
    look(X >= 1, -cx); look(X <= bd.bx - 2, cx); look(Y >= 1, -cy); look(Y <= bd.by - 2, cy);
    look(Z >= 1, -1); look(Z <= bd.bz - 2, 1);
    isb[i] = rim ? 1 : 0;
}
__global__ void k_sh_final(const int* __restrict__ tpos, const int* __restrict__ sb, int Qo, int G, ShardSub sub,
                           const int* __restrict__ gbid, int* __restrict__ lmap, int* __restrict__ gid_of, int* __restrict__ pos_of,
                           int* __restrict__ blk_nb, int late_T) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= Qo) return;
    const int b = tiled_owner(i, Qo, G);
    const int qb = tiled_begin(b, Qo, G), qe = tiled_begin(b + 1, Qo, G);
    const int nbc = sb[qe] - sb[qb];
    const int rank_b = sb[i] - sb[qb];
    const bool rim = sb[i + 1] != sb[i];
    const int id = rim ? qb + rank_b : qb + nbc + (i - qb) - rank_b;
    const long long p = sub.position((size_t)tpos[i]);
    lmap[sub.map_index(p)] = id;
    gid_of[id] = gbid[p];
    pos_of[id] = (int)p;
    if (i == qb) blk_nb[b] = late_count(nbc, qe - qb, late_T);
}
struct ShardBuild {
    int Qo, D, Qg, G;             // owned bricks, dummy slot, bricks of the whole plan, blocks
    int has_lo, has_hi, base_lo, base_hi, first_lo_plane_id, first_hi_plane_id;   // peer slot = base + (global id - first id of the plane)
};
__global__ void k_sh_build(ShardBuild sbd, ShardSub sub, size_t P, const int* __restrict__ lmap, const int* __restrict__ gid_of,
                           const int* __restrict__ pos_of, const uint2* __restrict__ gcode, const uint4* __restrict__ gfields,
                           const float4* __restrict__ gx0, uint2* __restrict__ code, uint4* __restrict__ fields,
                           int4* __restrict__ nbA, int2* __restrict__ nbB, float4* __restrict__ xa, int2* __restrict__ peer_ids,
                           int* __restrict__ blk_has) {
    int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= sbd.Qo) return;
    const BrickDims bd = sub.bd;
    const int g = gid_of[q];
    const long long p = pos_of[q];
    const int Z = (int)(p % bd.bz);
    const long long t = p / bd.bz;
    const int Y = (int)(t % bd.by), X = (int)(t / bd.by);
    code[q] = gcode[g];
    fields[q] = gfields[g];
    fields[(size_t)sbd.Qo + q] = gfields[(size_t)sbd.Qg + g];
    for (int k = 0; k < 4; k++) xa[k * P + q] = gx0[(size_t)k * ((size_t)sbd.Qg + 1) + g];
    const long long cx = (long long)bd.by * bd.bz, cy = bd.bz;
    auto nid = [&](bool ok, long long dp) {
        int id = ok ? lmap[sub.map_index(p + dp)] : -1;
        return id >= 0 ? id : sbd.D;
    };
    nbA[q] = make_int4(nid(X >= 1, -cx), nid(X <= bd.bx - 2, cx), nid(Y >= 1, -cy), nid(Y <= bd.by - 2, cy));
    nbB[q] = make_int2(nid(Z >= 1, -1), nid(Z <= bd.bz - 2, 1));
    int2 pid = make_int2(-1, -1);
    if (sbd.has_lo && X == sub.X0) pid.x = sbd.base_lo + (g - sbd.first_lo_plane_id);
    if (sbd.has_hi && X == sub.X0 + sub.nX - 1) pid.y = sbd.base_hi + (g - sbd.first_hi_plane_id);
    peer_ids[q] = pid;
    const int b = tiled_owner(q, sbd.Qo, sbd.G);
    if (pid.x >= 0) blk_has[2 * b] = 1;
    if (pid.y >= 0) blk_has[2 * b + 1] = 1;
}
// initial values of the halo slots of plane X
__global__ void k_sh_halo_x(const int* __restrict__ gbid, ShardSub sub, int X, const int* __restrict__ lmap, int Qg, size_t P,
                            const float4* __restrict__ gx0, float4* __restrict__ xa) {
    const int n = sub.bd.by * sub.bd.bz;
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const long long p = (long long)X * n + i;
    const int g = gbid[p];
    if (g < 0) return;
    const int id = lmap[sub.map_index(p)];
    for (int k = 0; k < 4; k++) xa[k * P + id] = gx0[(size_t)k * ((size_t)Qg + 1) + g];
}
__global__ void k_sh_result(const float4* __restrict__ tmp, const int* __restrict__ gid_of, int Qo, int Qg, float4* __restrict__ out) {
    int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= Qo) return;
    const int g = gid_of[q];
    for (int k = 0; k < 4; k++) out[(size_t)k * ((size_t)Qg + 1) + g] = tmp[(size_t)k * Qo + q];
}
__global__ void k_sh_gather_owned(const float4* __restrict__ x, size_t P, int Qo, float4* __restrict__ tmp) {
    int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= Qo) return;
    for (int k = 0; k < 4; k++) tmp[(size_t)k * Qo + q] = x[k * P + q];
}
__global__ void k_sh_ctrl(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

}  // namespace mrb

namespace mrb {
// bricks sorted by grid position: thread q (0 .. Q) writes plane_off[X] = q for every plane X in (plane(q-1), plane(q)]
// (plane(-1) = -1, plane(Q) = bx)
__global__ void k_plane_offsets(const int* __restrict__ bpos, int Q, int per_plane, int bx, int* __restrict__ plane_off) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q > Q) return;
    const int hi = q < Q ? bpos[q] / per_plane : bx;
    const int lo = q > 0 ? bpos[q - 1] / per_plane : -1;
    for (int X = lo + 1; X <= hi; X++) plane_off[X] = q;
}
}  // namespace mrb

struct mrb_qp_plan {
    mrb_ctx* ctx = nullptr;
    mrb::Arena arena;                 // owns every device buffer of the plan
    mrb::Dims d{};
    mrb::BrickDims bd{};
    int band_size = 0, Q = 0, n_band = 0, n_tight = 0;
    const float* vol = nullptr;
    uint8_t *flags = nullptr, *sq = nullptr;
    int* bid = nullptr;
    uint2* code = nullptr;
    uint4* fields = nullptr;
    int4* nbA = nullptr;
    int2* nbB = nullptr;
    float4* x0 = nullptr;
    std::vector<int> plane_off;       // bx + 1 entries: first brick id of every X plane
};

using namespace mrb;

extern "C" {

void mrb_qp_plan_destroy(mrb_qp_plan* plan) {
    if (!plan) return;
    if (plan->ctx) { cudaSetDevice(plan->ctx->device); cudaStreamSynchronize(plan->ctx->stream); }
    if (plan->ctx && plan->ctx->plan_arenas.size() < 2) plan->ctx->plan_arenas.push_back(std::move(plan->arena));   // kept for the next plan
    else plan->arena.release();
    delete plan;
}

static int plan_create_impl(mrb_ctx* ctx, const float* volume, const int dims[3], int band_size, bool allow_border, mrb_qp_plan** out);

int mrb_qp_plan_create(mrb_ctx* ctx, const float* volume, const int dims[3], int band_size, mrb_qp_plan** out) {
    return plan_create_impl(ctx, volume, dims, band_size, false, out);
}
int mrb_qp_plan_create_slab(mrb_ctx* ctx, const float* volume_sub, const int dims_sub[3], int band_size, mrb_qp_plan** out) {
    return plan_create_impl(ctx, volume_sub, dims_sub, band_size, true, out);
}

static int plan_create_impl(mrb_ctx* ctx, const float* volume, const int dims[3], int band_size, bool allow_border, mrb_qp_plan** out) {
    MRB_REQUIRE(ctx && volume && dims && out, "NULL argument");
    MRB_REQUIRE(band_size >= 1 && band_size <= 15, "plans support band_size 1..15");
    *out = nullptr;
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(upload_tables(ctx->device));
    mrb_qp_plan* p = new mrb_qp_plan();
    p->ctx = ctx;
    p->d = mkdims(dims);
    p->band_size = band_size;
    const Dims d = p->d;
    const size_t n = d.count();
    // every allocation below (scratch included) lands in the plan's own arena and lives as long as the plan
    if (!ctx->plan_arenas.empty()) {
        p->arena = std::move(ctx->plan_arenas.back());
        ctx->plan_arenas.pop_back();
        if (p->arena.reset() != MRB_OK) p->arena = Arena();
    }
    std::swap(ctx->arena, p->arena);
    auto build = [&]() -> int {
        const float* dv = nullptr;
        if (is_device_ptr(volume)) {
            float* copy = nullptr;
            MRB_TRY(arena_alloc(ctx, n, &copy));
            MRB_CUDA(cudaMemcpyAsync(copy, volume, n * sizeof(float), cudaMemcpyDeviceToDevice, ctx->stream));
            dv = copy;
        } else {
            MRB_TRY(stage_in(ctx, volume, n, &dv));
        }
        p->vol = dv;
        MRB_TRY(arena_alloc(ctx, n, &p->flags));
        MRB_TRY(arena_alloc(ctx, n, &p->sq));
        int* counts = ctx->d_mail + 32;
        MRB_CUDA(cudaMemsetAsync(counts, 0, 4 * sizeof(int), ctx->stream));
        int binary = 0;
        uint8_t* brick_bits = nullptr;     // by-product of the tile kernel: which bricks hold a band voxel
        MRB_TRY(stage12_binary(ctx, dv, d, band_size, p->sq, p->flags, counts, &binary, &brick_bits));
        if (!binary) { set_error("mrb_qp_plan_create: the volume must contain only 0 and 1"); return MRB_EINVAL; }
        MRB_CUDA(cudaMemcpyAsync(ctx->h_mail + 32, counts, 4 * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        p->bd = BrickDims{d.nx, d.ny, d.nz, (d.nx + 1) / 2, (d.ny + 1) / 2, (d.nz + 3) / 4};
        const size_t nb = p->bd.count();
        int* bpos = nullptr;
        if (brick_bits) MRB_TRY(compact(ctx, BrickBitPred{brick_bits, p->bd, (d.nz + 31) / 32}, nb, &bpos, &p->Q));
        else MRB_TRY(compact(ctx, ActiveBrickPred{p->flags, p->bd}, nb, &bpos, &p->Q));
        p->n_band = ctx->h_mail[32];
        p->n_tight = ctx->h_mail[33];
        if (ctx->h_mail[34] && !allow_border) { set_error("narrow band touches the grid border (reference exits, grid.cpp:20-37)"); return MRB_EBORDER; }
        const int Q = p->Q;
        const size_t Qn = (size_t)(Q > 0 ? Q : 1);
        MRB_TRY(arena_alloc(ctx, nb, &p->bid));
        MRB_CUDA(cudaMemsetAsync(p->bid, 0xFF, (size_t)nb * sizeof(int), ctx->stream));     // every int = -1
        MRB_TRY(arena_alloc(ctx, Qn, &p->code));
        MRB_TRY(arena_alloc(ctx, 2 * Qn, &p->fields));
        MRB_TRY(arena_alloc(ctx, 4 * (Qn + 1), &p->x0));
        MRB_TRY(arena_alloc(ctx, Qn, &p->nbA));
        MRB_TRY(arena_alloc(ctx, Qn, &p->nbB));
        p->plane_off.assign((size_t)p->bd.bx + 1, Q);
        if (Q == 0) return MRB_OK;
        MRB_LAUNCH(ctx, k_scatter_brick_ids, cdiv(Q, 256), 256, 0, bpos, Q, p->bid);
        int* err = ctx->d_mail + 24;
        MRB_CUDA(cudaMemsetAsync(err, 0, sizeof(int), ctx->stream));
        MRB_CUDA(cudaMemsetAsync(p->x0, 0, 4 * (Qn + 1) * sizeof(float4), ctx->stream));
        MRB_LAUNCH(ctx, k_brick_build, cdiv(Q, 128), 128, 0, dv, p->sq, p->flags, p->bid, bpos, Q, p->bd, p->code, p->fields,
                   p->x0, p->nbA, p->nbB, err, (size_t)Q + 1);
        // plane_off[X] = first brick id whose plane is >= X (bricks are sorted by position): computed on the device, bx + 1 ints come back
        int* poff = nullptr;
        MRB_TRY(arena_alloc(ctx, (size_t)p->bd.bx + 1, &poff));
        MRB_LAUNCH(ctx, k_plane_offsets, cdiv(Q + 1, 256), 256, 0, bpos, Q, p->bd.by * p->bd.bz, p->bd.bx, poff);
        MRB_CUDA(cudaMemcpyAsync(p->plane_off.data(), poff, ((size_t)p->bd.bx + 1) * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        MRB_CUDA(cudaMemcpyAsync(ctx->h_mail + 24, err, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        MRB_CUDA(cudaStreamSynchronize(ctx->stream));
        if (ctx->h_mail[24] && !allow_border) { set_error("narrow band touches the grid border (reference exits, grid.cpp:20-37)"); return MRB_EBORDER; }
        return MRB_OK;
    };
    int rc = build();
    std::swap(ctx->arena, p->arena);
    if (rc != MRB_OK) { p->arena.release(); delete p; return rc; }
    *out = p;
    return MRB_OK;
}

int mrb_qp_plan_info(mrb_qp_plan* plan, int* n_bricks, int* n_band, int* n_tight, int* n_brick_planes) {
    MRB_REQUIRE(plan, "NULL plan");
    if (n_bricks) *n_bricks = plan->Q;
    if (n_band) *n_band = plan->n_band;
    if (n_tight) *n_tight = plan->n_tight;
    if (n_brick_planes) *n_brick_planes = plan->bd.bx;
    return MRB_OK;
}

int mrb_qp_plan_plane_offsets(mrb_qp_plan* plan, int* off) {
    MRB_REQUIRE(plan && off, "NULL argument");
    for (size_t i = 0; i < plan->plane_off.size(); i++) off[i] = plan->plane_off[i];
    return MRB_OK;
}

int mrb_qp_plan_init_x(mrb_qp_plan* plan, float* x_dev) {
    MRB_REQUIRE(plan && x_dev && is_device_ptr(x_dev), "x must be a device buffer of 16*(Q+1) floats");
    mrb_ctx* ctx = plan->ctx;
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_CUDA(cudaMemcpyAsync(x_dev, plan->x0, 4 * ((size_t)plan->Q + 1) * sizeof(float4), cudaMemcpyDeviceToDevice, ctx->stream));
    return sync(ctx);
}

int mrb_qp_plan_sweep(mrb_qp_plan* plan, const float* x_in_dev, float* x_out_dev, int q_begin, int q_end) {
    MRB_REQUIRE(plan && x_in_dev && x_out_dev && x_in_dev != x_out_dev, "bad buffers");
    MRB_REQUIRE(q_begin >= 0 && q_end <= plan->Q && q_begin <= q_end, "bad brick range");
    mrb_ctx* ctx = plan->ctx;
    MRB_CUDA(cudaSetDevice(ctx->device));
    if (q_end > q_begin)
        MRB_LAUNCH(ctx, k_qp_bricks_range, cdiv((size_t)(q_end - q_begin), 256), 256, 0, plan->code, plan->fields, plan->nbA,
                   plan->nbB, (const float4*)x_in_dev, (float4*)x_out_dev, plan->Q, q_begin, q_end);
    return sync(ctx);
}

// Blocks of a G-block split of [qa, qb) whose range intersects [p0, p1): count and first block.
static int blocks_touching(int qa, int qb, int G, int p0, int p1, int* first) {
    *first = 0;
    if (p1 <= p0) return 0;
    int n = 0;
    const long long Ql = qb - qa;
    for (int b = 0; b < G; b++) {
        int s = qa + (int)(Ql * b / G), e = qa + (int)(Ql * (b + 1) / G);
        if (s < p1 && e > p0) { if (n == 0) *first = b; n++; }
    }
    return n;
}
// Grid size every rank uses for a range of Ql bricks (all ranks run the same GPU model and kernel).
static int shard_grid(int fit, int sm_count, int Ql, int threads) {
    int grid = fit * sm_count, need = (Ql + threads - 1) / threads;
    return grid > need ? need : grid;
}

// Slab-sharded solve with the tiled kernel (default; MRB_SHARD_KERNEL=xyz selects the older range kernel below).
// Buffers: 16 * (Q + 9) floats each (the slab-local layout needs a plane stride that is a multiple of 8 bricks);
// flags: 512 ints, zeroed by the caller before a barrier across the ranks.  On return the owned bricks' values sit in
// xa (even iters) / xb (odd iters) in the plan's GLOBAL layout (four planes of Q + 1 float4), like the older kernel.
// cuts: brick-plane boundaries of all ranks in this plan's plane indices, or nullptr when the plan covers only this rank's
// neighbourhood (mrb_qp_plan_solve_sharded_sub): then [Xa_sub, Xb_sub) is the owned range and nbl = {owned bricks, lower halo
// bricks, upper halo bricks} of the lower and of the upper neighbour (what layout() computes from a whole-grid plan).
static int solve_sharded_tiled(mrb_qp_plan* plan, int rank, int world, void* const* xa_ptrs, void* const* xb_ptrs,
                               void* const* flag_ptrs, const int* cuts, int iters, int Xa_sub = 0, int Xb_sub = 0,
                               const int* nbl = nullptr) {
    mrb_ctx* ctx = plan->ctx;
    const std::vector<int>& po = plan->plane_off;
    const BrickDims bd = plan->bd;
    const int Qg = plan->Q;
    const int Xa = cuts ? cuts[rank] : Xa_sub, Xb = cuts ? cuts[rank + 1] : Xb_sub, nX = Xb - Xa;
    const int Qo = po[Xb] - po[Xa];
    const bool has_lo = rank > 0, has_hi = rank < world - 1;
    const int nh_lo = has_lo ? po[Xa] - po[Xa - 1] : 0, nh_hi = has_hi ? po[Xb + 1] - po[Xb] : 0;
    const int D = Qo + nh_lo + nh_hi;
    const size_t P = tiled_stride(D);
    auto layout = [&](int r, int* Qo_r, int* nlo_r, int* nhi_r) {
        if (!cuts) {
            const int* t = nbl + (r < rank ? 0 : 3);
            *Qo_r = t[0]; *nlo_r = t[1]; *nhi_r = t[2];
            return;
        }
        *Qo_r = po[cuts[r + 1]] - po[cuts[r]];
        *nlo_r = r > 0 ? po[cuts[r]] - po[cuts[r] - 1] : 0;
        *nhi_r = r < world - 1 ? po[cuts[r + 1] + 1] - po[cuts[r + 1]] : 0;
    };
    float4* xa = (float4*)xa_ptrs[rank];
    float4* xb = (float4*)xb_ptrs[rank];
    int tile_T = 8;
    if (getenv("MRB_QP_TILE")) tile_T = atoi(getenv("MRB_QP_TILE"));
    if (tile_T < 1) tile_T = 1;
    const int threads = 480;
    void* kern = world > 1 ? (void*)k_qp_bricks_tiled<480, 1, true> : (void*)k_qp_bricks_tiled<480, 1, false>;
    int fit = 0;
    MRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&fit, kern, threads, 0));
    if (fit < 1) { set_error("k_qp_bricks_tiled does not fit on an SM"); return MRB_ECUDA; }
    int grid = fit * ctx->sm_count;
    const int need = (Qo + threads - 1) / threads;
    if (grid > need) grid = need;
    if (grid < 1) grid = 1;

    // ---- slab-local numbering
    const ShardSub sub{bd, tile_T, Xa, nX};
    int* tpos = nullptr;
    int Qc = 0;
    MRB_TRY(compact(ctx, ShardSubPred{plan->bid, sub}, sub.count(), &tpos, &Qc));
    if (Qc != Qo) { set_error("mrb_qp_plan_solve_sharded: slab brick count mismatch (%d vs %d)", Qc, Qo); return MRB_ECUDA; }
    int *lmap1 = nullptr, *lmap = nullptr, *sbuf = nullptr, *gid_of = nullptr, *pos_of = nullptr, *blk_nb = nullptr, *blk_has = nullptr;
    const size_t msz = sub.map_size();
    MRB_TRY(arena_alloc(ctx, msz, &lmap1));
    MRB_TRY(arena_alloc(ctx, msz, &lmap));
    MRB_TRY(arena_alloc(ctx, (size_t)Qo + 1, &sbuf));
    MRB_TRY(arena_alloc(ctx, (size_t)Qo, &gid_of));
    MRB_TRY(arena_alloc(ctx, (size_t)Qo, &pos_of));
    MRB_TRY(arena_alloc(ctx, (size_t)grid, &blk_nb));
    MRB_TRY(arena_alloc(ctx, (size_t)grid * 2, &blk_has));
    MRB_CUDA(cudaMemsetAsync(lmap1, 0xFF, (size_t)msz * sizeof(int), ctx->stream));     // every int = -1
    MRB_CUDA(cudaMemsetAsync(lmap, 0xFF, (size_t)msz * sizeof(int), ctx->stream));     // every int = -1
    MRB_CUDA(cudaMemsetAsync(blk_nb, 0, (size_t)grid * sizeof(int), ctx->stream));
    MRB_CUDA(cudaMemsetAsync(blk_has, 0, (size_t)grid * 2 * sizeof(int), ctx->stream));
    const int plane_n = bd.by * bd.bz;
    MRB_LAUNCH(ctx, k_sh_scatter1, cdiv(Qo, 256), 256, 0, tpos, Qo, sub, lmap1);
    if (has_lo) MRB_LAUNCH(ctx, k_sh_halo_ids, cdiv(plane_n, 256), 256, 0, plan->bid, sub, Xa - 1, Qo, po[Xa - 1], lmap1, lmap);
    if (has_hi) MRB_LAUNCH(ctx, k_sh_halo_ids, cdiv(plane_n, 256), 256, 0, plan->bid, sub, Xb, Qo + nh_lo, po[Xb], lmap1, lmap);
    MRB_LAUNCH(ctx, k_sh_boundary, cdiv(Qo, 256), 256, 0, tpos, Qo, grid, sub, lmap1, sbuf);
    MRB_TRY(scan_exclusive(ctx, sbuf, sbuf, (size_t)Qo, sbuf + Qo));
    MRB_LAUNCH(ctx, k_sh_final, cdiv(Qo, 256), 256, 0, tpos, sbuf, Qo, grid, sub, plan->bid, lmap, gid_of, pos_of, blk_nb,
               (getenv("MRB_QP_LATE") && atoi(getenv("MRB_QP_LATE")) == 0) ? 0 : threads);

    // ---- slab-local constants, neighbour ids, initial iterate, peer slots
    uint2* code = nullptr;
    uint4* fields = nullptr;
    int4* nbA = nullptr;
    int2 *nbB = nullptr, *peer_ids = nullptr;
    MRB_TRY(arena_alloc(ctx, (size_t)Qo, &code));
    MRB_TRY(arena_alloc(ctx, (size_t)Qo * 2, &fields));
    MRB_TRY(arena_alloc(ctx, (size_t)Qo, &nbA));
    MRB_TRY(arena_alloc(ctx, (size_t)Qo, &nbB));
    MRB_TRY(arena_alloc(ctx, (size_t)Qo, &peer_ids));
    ShardBuild sbd;
    memset(&sbd, 0, sizeof(sbd));
    sbd.Qo = Qo; sbd.D = D; sbd.Qg = Qg; sbd.G = grid;
    sbd.has_lo = has_lo; sbd.has_hi = has_hi;
    TShard sh;
    memset(&sh, 0, sizeof(sh));
    if (has_lo) {
        int q, nl, nh;
        layout(rank - 1, &q, &nl, &nh);
        sbd.base_lo = q + nl;                       // my first plane is its upper halo
        sbd.first_lo_plane_id = po[Xa];
        sh.P_lo = tiled_stride(q + nl + nh);
        sh.peer_lo[0] = (float4*)xa_ptrs[rank - 1]; sh.peer_lo[1] = (float4*)xb_ptrs[rank - 1];
        sh.flag_lo_remote = (unsigned*)flag_ptrs[rank - 1] + kShardSlots;       // its "from upper" slots
    }
    if (has_hi) {
        int q, nl, nh;
        layout(rank + 1, &q, &nl, &nh);
        sbd.base_hi = q;                            // my last plane is its lower halo
        sbd.first_hi_plane_id = po[Xb - 1];
        sh.P_hi = tiled_stride(q + nl + nh);
        sh.peer_hi[0] = (float4*)xa_ptrs[rank + 1]; sh.peer_hi[1] = (float4*)xb_ptrs[rank + 1];
        sh.flag_hi_remote = (unsigned*)flag_ptrs[rank + 1];                      // its "from lower" slots
    }
    sh.flags_local = (unsigned*)flag_ptrs[rank];
    // xa: owned + halo values + zero dummy / padding; xb: only the dummy / padding (the neighbours may already be
    // pushing their first sweep into xb's halo slots)
    for (int k = 0; k < 4; k++) {
        MRB_CUDA(cudaMemsetAsync(xa + (size_t)k * P + D, 0, (P - D) * sizeof(float4), ctx->stream));
        MRB_CUDA(cudaMemsetAsync(xb + (size_t)k * P + D, 0, (P - D) * sizeof(float4), ctx->stream));
    }
    MRB_LAUNCH(ctx, k_sh_build, cdiv(Qo, 128), 128, 0, sbd, sub, P, lmap, gid_of, pos_of, plan->code, plan->fields, plan->x0, code, fields,
               nbA, nbB, xa, peer_ids, blk_has);
    if (has_lo) MRB_LAUNCH(ctx, k_sh_halo_x, cdiv(plane_n, 256), 256, 0, plan->bid, sub, Xa - 1, lmap, Qg, P, plan->x0, xa);
    if (has_hi) MRB_LAUNCH(ctx, k_sh_halo_x, cdiv(plane_n, 256), 256, 0, plan->bid, sub, Xb, lmap, Qg, P, plan->x0, xa);
    // pusher blocks -> slots; tell the neighbours how many slots to wait for
    std::vector<int> h_has((size_t)grid * 2), h_slot_lo((size_t)grid, -1), h_slot_hi((size_t)grid, -1);
    MRB_CUDA(cudaMemcpyAsync(h_has.data(), blk_has, h_has.size() * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    MRB_CUDA(cudaStreamSynchronize(ctx->stream));
    int n_lo = 0, n_hi = 0;
    for (int b = 0; b < grid; b++) {
        if (h_has[2 * (size_t)b]) h_slot_lo[b] = n_lo++;
        if (h_has[2 * (size_t)b + 1]) h_slot_hi[b] = n_hi++;
    }
    if (n_lo > kShardSlots || n_hi > kShardSlots) { set_error("mrb_qp_plan_solve_sharded: more than %d boundary blocks per side", kShardSlots); return MRB_EINVAL; }
    int *slot_lo = nullptr, *slot_hi = nullptr;
    MRB_TRY(arena_alloc(ctx, (size_t)grid, &slot_lo));
    MRB_TRY(arena_alloc(ctx, (size_t)grid, &slot_hi));
    MRB_CUDA(cudaMemcpyAsync(slot_lo, h_slot_lo.data(), (size_t)grid * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    MRB_CUDA(cudaMemcpyAsync(slot_hi, h_slot_hi.data(), (size_t)grid * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    sh.peer_ids = peer_ids; sh.slot_lo = slot_lo; sh.slot_hi = slot_hi;
    if (has_lo) MRB_LAUNCH(ctx, k_sh_ctrl, 1, 1, 0, (unsigned*)flag_ptrs[rank - 1] + 2 * kShardSlots + 1, (unsigned)n_lo + 1u);
    if (has_hi) MRB_LAUNCH(ctx, k_sh_ctrl, 1, 1, 0, (unsigned*)flag_ptrs[rank + 1] + 2 * kShardSlots + 0, (unsigned)n_hi + 1u);

    if (iters > 0 && Qo > 0) {
        unsigned* md = (unsigned*)(ctx->d_mail + 8);
        unsigned* bar = (unsigned*)ctx->d_bar;
        MRB_CUDA(cudaMemsetAsync(md, 0, sizeof(unsigned), ctx->stream));
        MRB_CUDA(cudaMemsetAsync(bar, 0, 64 * sizeof(unsigned), ctx->stream));
        const int NW = (grid + 31) / 32;
        unsigned *progress = nullptr, *bitmap = nullptr;
        int *plist = nullptr, *pcount = nullptr;
        MRB_TRY(arena_alloc(ctx, (size_t)grid * kProgressStride, &progress));
        MRB_TRY(arena_alloc(ctx, (size_t)grid * kChunkListCap, &plist));
        MRB_TRY(arena_alloc(ctx, (size_t)grid, &pcount));
        MRB_TRY(arena_alloc(ctx, (size_t)grid * NW, &bitmap));
        int* ovf = ctx->d_mail + 50;
        MRB_CUDA(cudaMemsetAsync(progress, 0, (size_t)grid * kProgressStride * sizeof(unsigned), ctx->stream));
        MRB_CUDA(cudaMemsetAsync(bitmap, 0, (size_t)grid * NW * sizeof(unsigned), ctx->stream));
        MRB_CUDA(cudaMemsetAsync(pcount, 0, (size_t)grid * sizeof(int), ctx->stream));
        MRB_CUDA(cudaMemsetAsync(ovf, 0, sizeof(int), ctx->stream));
        MRB_LAUNCH(ctx, k_tiled_bitmap, cdiv(Qo, 256), 256, 0, nbA, nbB, Qo, grid, NW, bitmap);
        MRB_LAUNCH(ctx, k_chunk_lists, cdiv((size_t)grid * 32, 256), 256, 0, bitmap, grid, NW, plist, pcount, ovf);
        int overflow = 0;
        MRB_TRY(read_int(ctx, ovf, &overflow));
        if (overflow) { set_error("mrb_qp_plan_solve_sharded: a block has more than %d neighbour blocks", kChunkListCap); return MRB_EINVAL; }
        int cap = (Qo + grid - 1) / grid + 9;
        int expanded = 1;
        size_t smem = (size_t)cap * 168;
        if (getenv("MRB_QP_BRICK_COMPACT") || smem > (size_t)220 * 1024 / (size_t)fit) { expanded = 0; smem = (size_t)cap * 64; }
        if (!expanded && smem > 96 * 1024 / (size_t)fit) { cap = 0; smem = 0; }
        MRB_CUDA(allow_dyn_smem((const void*)kern, smem));
        int its = iters, QQ = Qo, dbg = 32;      // thread 0 publishes (the configuration the multi-GPU runs were validated with)
        size_t PP = P;
        unsigned long long* no_trace = nullptr;
        void* args[] = {(void*)&code, (void*)&fields, (void*)&nbA, (void*)&nbB, (void*)&xa, (void*)&xb, (void*)&QQ, (void*)&its,
                        (void*)&md, (void*)&bar, (void*)&dbg, (void*)&blk_nb, (void*)&progress, (void*)&cap, (void*)&plist, (void*)&pcount,
                        (void*)&expanded, (void*)&PP, (void*)&sh, (void*)&no_trace};
        MRB_CUDA(cudaEventRecord(ctx->ev[6], ctx->stream));
        MRB_CUDA(cudaLaunchCooperativeKernel(kern, dim3(grid), dim3(threads), args, smem, ctx->stream));
        MRB_CUDA(cudaEventRecord(ctx->ev[7], ctx->stream));
        ctx->launches++;
    }
    // owned values back into the plan's global layout, in the buffer that holds the result
    float4* res = (iters % 2) ? xb : xa;
    float4* tmp = nullptr;
    MRB_TRY(arena_alloc(ctx, (size_t)(Qo > 0 ? Qo : 1) * 4, &tmp));
    if (Qo > 0) {
        MRB_LAUNCH(ctx, k_sh_gather_owned, cdiv(Qo, 256), 256, 0, res, P, Qo, tmp);
        MRB_LAUNCH(ctx, k_sh_result, cdiv(Qo, 256), 256, 0, tmp, gid_of, Qo, Qg, res);
    }
    return sync(ctx);
}

int mrb_qp_plan_solve_sharded(mrb_qp_plan* plan, int rank, int world, void* const* xa_ptrs, void* const* xb_ptrs,
                              void* const* flag_ptrs, const int* plane_cuts, int iters) {
    MRB_REQUIRE(plan && xa_ptrs && xb_ptrs && flag_ptrs && plane_cuts && world >= 1 && rank >= 0 && rank < world && iters >= 0,
                "bad argument");
    mrb_ctx* ctx = plan->ctx;
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    const std::vector<int>& po = plan->plane_off;
    for (int r = 0; r < world; r++)
        MRB_REQUIRE(plane_cuts[r] < plane_cuts[r + 1] && po[plane_cuts[r]] < po[plane_cuts[r + 1]], "every rank must own at least one brick");
    MRB_REQUIRE(plane_cuts[0] == 0 && plane_cuts[world] == plan->bd.bx, "plane cuts must cover the grid");
    {
        const char* ek = getenv("MRB_SHARD_KERNEL");
        if (!(ek && strcmp(ek, "xyz") == 0)) return solve_sharded_tiled(plan, rank, world, xa_ptrs, xb_ptrs, flag_ptrs, plane_cuts, iters);
    }
    const int Xa = plane_cuts[rank], Xb = plane_cuts[rank + 1];
    ShardArgs sh;
    memset(&sh, 0, sizeof(sh));
    sh.qa = po[Xa]; sh.qb = po[Xb];
    if (rank > 0) {
        sh.lo0 = po[Xa]; sh.lo1 = po[Xa + 1];
        sh.peer_lo[0] = (float4*)xa_ptrs[rank - 1]; sh.peer_lo[1] = (float4*)xb_ptrs[rank - 1];
        sh.flag_lo_remote = (unsigned*)flag_ptrs[rank - 1] + kShardSlots;   // the lower rank's "from_upper" slots
    }
    if (rank < world - 1) {
        sh.hi0 = po[Xb - 1]; sh.hi1 = po[Xb];
        sh.peer_hi[0] = (float4*)xa_ptrs[rank + 1]; sh.peer_hi[1] = (float4*)xb_ptrs[rank + 1];
        sh.flag_hi_remote = (unsigned*)flag_ptrs[rank + 1] + 0;             // the upper rank's "from_lower" slots
    }
    sh.flags_local = (unsigned*)flag_ptrs[rank];
    if (iters == 0 || sh.qb == sh.qa) return sync(ctx);
    const int threads = 256;
    void* kern = (void*)k_qp_bricks_sharded<256, 2>;
    int fit = 0;
    MRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&fit, kern, threads, 0));
    if (fit < 1) { set_error("k_qp_bricks_sharded does not fit on an SM"); return MRB_ECUDA; }
    const int Ql = sh.qb - sh.qa;
    const int grid = shard_grid(fit, ctx->sm_count, Ql, threads);
    int dummy = 0;
    const int nlo_mine = blocks_touching(sh.qa, sh.qb, grid, sh.lo0, sh.lo1, &sh.first_lo);
    const int nhi_mine = blocks_touching(sh.qa, sh.qb, grid, sh.hi0, sh.hi1, &sh.first_hi);
    if (rank > 0) {            // the lower neighbour's blocks that own its LAST plane (they fill my from_lower slots)
        const int a = po[plane_cuts[rank - 1]], b = po[plane_cuts[rank]];
        sh.nrem_lo = blocks_touching(a, b, shard_grid(fit, ctx->sm_count, b - a, threads), po[plane_cuts[rank] - 1], b, &dummy);
    }
    if (rank < world - 1) {    // the upper neighbour's blocks that own its FIRST plane
        const int a = po[plane_cuts[rank + 1]], b = po[plane_cuts[rank + 2]];
        sh.nrem_hi = blocks_touching(a, b, shard_grid(fit, ctx->sm_count, b - a, threads), a, po[plane_cuts[rank + 1] + 1], &dummy);
    }
    if (nlo_mine > kShardSlots || nhi_mine > kShardSlots || sh.nrem_lo > kShardSlots || sh.nrem_hi > kShardSlots) {
        set_error("mrb_qp_plan_solve_sharded: more than %d boundary blocks per side", kShardSlots);
        return MRB_EINVAL;
    }
    int *blk_lo = nullptr, *blk_hi = nullptr;
    unsigned* progress = nullptr;
    MRB_TRY(arena_alloc(ctx, (size_t)grid, &blk_lo));
    MRB_TRY(arena_alloc(ctx, (size_t)grid, &blk_hi));
    MRB_TRY(arena_alloc(ctx, (size_t)grid * kProgressStride, &progress));
    MRB_CUDA(cudaMemsetAsync(progress, 0, (size_t)grid * kProgressStride * sizeof(unsigned), ctx->stream));
    MRB_LAUNCH(ctx, k_block_ranges_init, cdiv(grid, 256), 256, 0, blk_lo, blk_hi, grid);
    MRB_LAUNCH(ctx, k_block_ranges_local, cdiv((size_t)Ql, 256), 256, 0, plan->nbA, plan->nbB, sh.qa, sh.qb, grid, blk_lo, blk_hi);
    float4* xa = (float4*)xa_ptrs[rank];
    float4* xb = (float4*)xb_ptrs[rank];
    int Q = plan->Q, its = iters;
    void* args[] = {(void*)&plan->code, (void*)&plan->fields, (void*)&plan->nbA, (void*)&plan->nbB, (void*)&xa, (void*)&xb,
                    (void*)&Q, (void*)&its, (void*)&sh, (void*)&blk_lo, (void*)&blk_hi, (void*)&progress};
    MRB_CUDA(cudaEventRecord(ctx->ev[6], ctx->stream));
    MRB_CUDA(cudaLaunchCooperativeKernel(kern, dim3(grid), dim3(threads), args, 0, ctx->stream));
    MRB_CUDA(cudaEventRecord(ctx->ev[7], ctx->stream));
    ctx->launches++;
    return sync(ctx);
}

/* Device time (ms) of the last persistent sweep kernel launched on this plan's context. */
float mrb_qp_plan_last_kernel_ms(mrb_qp_plan* plan) {
    float ms = 0.0f;
    if (plan) cudaEventElapsedTime(&ms, plan->ctx->ev[6], plan->ctx->ev[7]);
    return ms;
}

int mrb_qp_plan_assemble(mrb_qp_plan* plan, const float* x_dev, float* F_out) {
    MRB_REQUIRE(plan && x_dev && F_out, "NULL argument");
    mrb_ctx* ctx = plan->ctx;
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    const size_t n = plan->d.count();
    Out<float> o;
    MRB_TRY(o.init(ctx, F_out, n));
    const Dims d = plan->d;
    if (d.nz % 128 == 0 && n % 256 == 0 && (((uintptr_t)plan->vol | (uintptr_t)o.dev | (uintptr_t)plan->flags | (uintptr_t)x_dev) & 15) == 0) {
        // four voxels per thread; the sign bits it packs on the way are not needed here
        uint32_t* sb = nullptr;
        MRB_TRY(arena_alloc(ctx, n / 32, &sb));
        MRB_LAUNCH(ctx, k_assemble_F_sign_v4, (unsigned)ctx->sm_count * 16, 256, 0, (const float4*)plan->vol, (const uint32_t*)plan->flags,
                   plan->bid, (const float4*)x_dev, plan->bd, (size_t)plan->Q + 1, (double)plan->band_size + 1.0, 0.0f, (float4*)o.dev, sb);
    } else {
        MRB_LAUNCH(ctx, k_assemble_F_b, grid_for(ctx, n, 256), 256, 0, plan->vol, plan->flags, plan->bid, x_dev, plan->bd,
                   (size_t)plan->Q + 1, (double)plan->band_size + 1.0, o.dev);
    }
    MRB_TRY(o.commit(ctx));
    return sync(ctx);
}

int mrb_qp_plan_assemble_slab(mrb_qp_plan* plan, const float* x_dev, int x_begin, int x_end, float* F_slab_out) {
    MRB_REQUIRE(plan && x_dev && F_slab_out, "NULL argument");
    MRB_REQUIRE(x_begin >= 0 && x_begin < x_end && x_end <= plan->d.nx, "bad plane range");
    mrb_ctx* ctx = plan->ctx;
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    const Dims d = plan->d;
    const size_t plane = (size_t)d.ny * d.nz, v0 = (size_t)x_begin * plane, count = (size_t)(x_end - x_begin) * plane;
    Out<float> o;
    MRB_TRY(o.init(ctx, F_slab_out, count));
    if (d.nz % 4 == 0 && (((uintptr_t)plan->vol | (uintptr_t)o.dev | (uintptr_t)plan->flags | (uintptr_t)x_dev) & 15) == 0) {
        MRB_LAUNCH(ctx, k_assemble_F_range_v4, (unsigned)ctx->sm_count * 16, 256, 0, (const float4*)plan->vol, (const uint32_t*)plan->flags,
                   plan->bid, (const float4*)x_dev, plan->bd, (size_t)plan->Q + 1, (double)plan->band_size + 1.0, v0 / 4, count / 4,
                   (float4*)o.dev);
    } else {
        MRB_LAUNCH(ctx, k_assemble_F_range, grid_for(ctx, count, 256), 256, 0, plan->vol, plan->flags, plan->bid, x_dev, plan->bd,
                   (size_t)plan->Q + 1, (double)plan->band_size + 1.0, v0, count, o.dev);
    }
    MRB_TRY(o.commit(ctx));
    return sync(ctx);
}

int mrb_qp_plan_solve_sharded_sub(mrb_qp_plan* plan, int rank, int world, void* const* xa_ptrs, void* const* xb_ptrs,
                                  void* const* flag_ptrs, int plane_lo, int plane_hi, const int* nb_layout, int iters) {
    MRB_REQUIRE(plan && xa_ptrs && xb_ptrs && flag_ptrs && nb_layout && world >= 1 && rank >= 0 && rank < world && iters >= 0,
                "bad argument");
    const std::vector<int>& po = plan->plane_off;
    MRB_REQUIRE(plane_lo >= 0 && plane_lo < plane_hi && plane_hi <= plan->bd.bx && po[plane_lo] < po[plane_hi], "bad owned plane range");
    MRB_REQUIRE((rank == 0 || plane_lo >= 1) && (rank == world - 1 || plane_hi + 1 <= plan->bd.bx), "the plan must hold the neighbours' planes");
    mrb_ctx* ctx = plan->ctx;
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    MRB_TRY(upload_tables(ctx->device));
    return solve_sharded_tiled(plan, rank, world, xa_ptrs, xb_ptrs, flag_ptrs, nullptr, iters, plane_lo, plane_hi, nb_layout);
}

}  // extern "C"

// mrb_qp.cu -- stage 3: the bound-constrained higher-order-smoothness QP.
//
// Replaces primeQP / getHMat / getlb / getub (optimize/primeqp.cpp:7-223), the CSC export
// (quadprog.cpp:8-42), doStep / runQP (quadprog.cpp:45-105) and doStepDevice / runQPGPU
// (quadprog.cu:58-186).
//
// The reference assembles H = D^T D with Eigen, exports R = H - diag(H) as CSC and runs 500
// launches of a gather kernel.  Here the operator is never assembled: for band voxel v with tight
// mask T and unit step e along an axis (verified against the assembled H by
// tests/test_oracle_qp.py):
//     d(v)        = sum_axes 4 T(v) + T(v-e) + T(v+e)           (1..18)
//     c(v, v+-e)  = -2 (T(v) + T(v+-e))                         (0, -2, -4)
//     c(v, v+-2e) = T(v+-e)                                     (0, 1)
// One sweep (quadprog.cpp:49-60):  res = sum_n c(v,n) x(n)  in ascending band id order, i.e.
//     x-2 x-1 y-2 y-1 z-2 z-1 z+1 z+2 y+1 y+2 x+1 x+2,
//     t = (x(v) - res * invdg(v)) * 0.5, clamped to [lb, ub].
// Bit-exactness: every coefficient is 0 or a power of two times +-1, so each product is exact and
// fma(c, x, res) == res + c*x rounded once, exactly the reference's fp32 "res += pr*in".  Terms the
// reference skips (pruned zeros) contribute +-0 here, which never changes res (res starts at +0
// and can therefore never be -0).  res*invdg and the subtraction are rounded separately
// (__fmul_rn / __fsub_rn: no contraction).  invdg = (float)(1.0/(double)d) comes from a table.
//
// Kernel: ONE persistent cooperative launch runs all sweeps; band voxels are processed in band-id
// order (= ascending address in the z-fastest grid), so the +-1/+-2 z neighbours are the adjacent
// band ids and the x/y neighbour ids of consecutive lanes are consecutive too (coalesced gathers).
// The iterate ping-pongs between two B-float buffers that stay L2 resident; a grid barrier
// separates sweeps.
#include "mrb_internal.cuh"

#include <cooperative_groups.h>
namespace cg = cooperative_groups;

namespace mrb {

// per-band-voxel code word
//   bit 0      T(v)
//   bit 1,2    T(x-1), T(x+1)
//   bit 3,4    T(y-1), T(y+1)
//   bit 5,6    T(z-1), T(z+1)
//   bit 7      volume == 0   (lb = -1000, primeqp.cpp:122-124)
//   bit 8      volume == 1   (ub = +1000, primeqp.cpp:148-150)
enum : uint32_t { kT0 = 1u, kTxm = 2u, kTxp = 4u, kTym = 8u, kTyp = 16u, kTzm = 32u, kTzp = 64u, kZero = 128u, kOne = 256u };

__constant__ float c_invdg[19];   // (float)(1.0/(double)d), d = 0..18 (entry 0 is 0: primeqp.cpp:191-197)

static int upload_invdg() {
    float h[19];
    h[0] = 0.0f;
    for (int d = 1; d <= 18; d++) h[d] = (float)(1.0 / (double)(float)d);
    MRB_CUDA(cudaMemcpyToSymbol(c_invdg, h, sizeof(h)));
    return MRB_OK;
}

// ---- build ------------------------------------------------------------------------------------
__global__ void k_fill_int(int* __restrict__ p, size_t n, int v) {
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (size_t)gridDim.x * blockDim.x) p[idx] = v;
}
__global__ void k_scatter_ids(const int* __restrict__ pos, int n, int* __restrict__ band_id) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) band_id[pos[i]] = i;
}

// err[0]: tight voxel on the border / tight voxel with a neighbour outside the band;  err[1]: tight count
__global__ void __launch_bounds__(256) k_qp_code(const float* __restrict__ volume, const float* __restrict__ margin,
                                                 const float* __restrict__ tight, const int* __restrict__ band_id,
                                                 const int* __restrict__ pos, int n, Dims d,
                                                 uint32_t* __restrict__ code, float* __restrict__ m,
                                                 int* __restrict__ nbr, int* __restrict__ err) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    bool t0 = false;
    if (i < n) {
        int p = pos[i];
        int z = p % d.nz;
        int t = p / d.nz;
        int y = t % d.ny;
        int x = t / d.ny;
        const int sx = d.ny * d.nz, sy = d.nz;
        t0 = tight[p] != 0.0f;
        bool border = x < 1 || y < 1 || z < 1 || x > d.nx - 2 || y > d.ny - 2 || z > d.nz - 2;
        uint32_t c = 0;
        if (t0) {
            c |= kT0;
            if (border) atomicExch(&err[0], 1);
        }
        bool bad = false;
        auto tbit = [&](bool ok, int off, uint32_t bit) {
            if (!ok) return;
            bool tn = tight[p + off] != 0.0f;
            if (tn) c |= bit;
            // a tight voxel needs all six neighbours in the band (reference: indexMap == -1 -> column -1)
            if (t0 && band_id[p + off] < 0) bad = true;
        };
        tbit(x >= 1, -sx, kTxm); tbit(x <= d.nx - 2, sx, kTxp);
        tbit(y >= 1, -sy, kTym); tbit(y <= d.ny - 2, sy, kTyp);
        tbit(z >= 1, -1, kTzm);  tbit(z <= d.nz - 2, 1, kTzp);
        if (bad) atomicExch(&err[0], 1);
        float v = volume[p];
        if (v == 0.0f) c |= kZero;
        if (v == 1.0f) c |= kOne;
        code[i] = c;
        m[i] = margin[p];
        // x/y neighbour ids; absent or out-of-range neighbours carry coefficient 0 -> self
        auto nid = [&](bool ok, int off) {
            int id = ok ? band_id[p + off] : -1;
            return id >= 0 ? id : i;
        };
        nbr[0 * (size_t)n + i] = nid(x >= 2, -2 * sx);
        nbr[1 * (size_t)n + i] = nid(x >= 1, -sx);
        nbr[2 * (size_t)n + i] = nid(y >= 2, -2 * sy);
        nbr[3 * (size_t)n + i] = nid(y >= 1, -sy);
        nbr[4 * (size_t)n + i] = nid(y <= d.ny - 2, sy);
        nbr[5 * (size_t)n + i] = nid(y <= d.ny - 3, 2 * sy);
        nbr[6 * (size_t)n + i] = nid(x <= d.nx - 2, sx);
        nbr[7 * (size_t)n + i] = nid(x <= d.nx - 3, 2 * sx);
    }
    unsigned b = __ballot_sync(0xffffffffu, t0);
    if ((threadIdx.x & 31) == 0 && b) atomicAdd(&err[1], __popc(b));
}

int dev_qp_build(mrb_ctx* ctx, const float* volume, const float* margin, const float* band, const float* tight,
                 Dims d, BandDev* bd) {
    static bool lut_ready[64] = {false};
    if (!lut_ready[ctx->device & 63]) { MRB_TRY(upload_invdg()); lut_ready[ctx->device & 63] = true; }
    int n = 0;
    MRB_TRY(dev_compact_nonzero(ctx, band, d, &bd->band_pos, &n));
    bd->n_band = n;
    MRB_TRY(arena_alloc(ctx, d.count(), &bd->band_id));
    MRB_LAUNCH(ctx, k_fill_int, cdiv(d.count(), 256 * 8), 256, 0, bd->band_id, d.count(), -1);
    size_t nn = (size_t)(n > 0 ? n : 1);
    MRB_TRY(arena_alloc(ctx, nn, &bd->code));
    MRB_TRY(arena_alloc(ctx, nn, &bd->margin));
    MRB_TRY(arena_alloc(ctx, 8 * nn, &bd->nbr));
    bd->n_tight = 0;
    if (n == 0) return MRB_OK;
    MRB_LAUNCH(ctx, k_scatter_ids, cdiv(n, 256), 256, 0, bd->band_pos, n, bd->band_id);
    MRB_CUDA(cudaMemsetAsync(ctx->d_mail, 0, 2 * sizeof(int), ctx->stream));
    MRB_LAUNCH(ctx, k_qp_code, cdiv(n, 256), 256, 0, volume, margin, tight, bd->band_id, bd->band_pos, n, d, bd->code,
               bd->margin, bd->nbr, ctx->d_mail);
    MRB_CUDA(cudaMemcpyAsync(ctx->h_mail, ctx->d_mail, 2 * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    MRB_CUDA(cudaStreamSynchronize(ctx->stream));
    if (ctx->h_mail[0]) {
        set_error("narrow band touches the grid border (tight-band voxel on the border or with a neighbour outside the "
                  "band); the reference exits here (grid.cpp:20-37 via primeqp.cpp:40-75)");
        return MRB_EBORDER;
    }
    bd->n_tight = ctx->h_mail[1];
    return MRB_OK;
}

// ---- one Jacobi update ----------------------------------------------------------------------------
// xin points at element 0 of a buffer padded by 2 finite floats on both sides.
__device__ __forceinline__ float qp_update(const float* xin, int i, int n, uint32_t c, float mi,
                                           const int* __restrict__ nbr, float* delta) {
    const float t0 = (c & kT0) ? 1.0f : 0.0f;
    const float txm = (c & kTxm) ? 1.0f : 0.0f, txp = (c & kTxp) ? 1.0f : 0.0f;
    const float tym = (c & kTym) ? 1.0f : 0.0f, typ = (c & kTyp) ? 1.0f : 0.0f;
    const float tzm = (c & kTzm) ? 1.0f : 0.0f, tzp = (c & kTzp) ? 1.0f : 0.0f;
    const size_t N = (size_t)n;
    const float xc = xin[i];
    float res = 0.0f;
    res = __fmaf_rn(txm, xin[nbr[0 * N + i]], res);
    res = __fmaf_rn(-2.0f * (t0 + txm), xin[nbr[1 * N + i]], res);
    res = __fmaf_rn(tym, xin[nbr[2 * N + i]], res);
    res = __fmaf_rn(-2.0f * (t0 + tym), xin[nbr[3 * N + i]], res);
    res = __fmaf_rn(tzm, xin[i - 2], res);
    res = __fmaf_rn(-2.0f * (t0 + tzm), xin[i - 1], res);
    res = __fmaf_rn(-2.0f * (t0 + tzp), xin[i + 1], res);
    res = __fmaf_rn(tzp, xin[i + 2], res);
    res = __fmaf_rn(-2.0f * (t0 + typ), xin[nbr[4 * N + i]], res);
    res = __fmaf_rn(typ, xin[nbr[5 * N + i]], res);
    res = __fmaf_rn(-2.0f * (t0 + txp), xin[nbr[6 * N + i]], res);
    res = __fmaf_rn(txp, xin[nbr[7 * N + i]], res);
    const int dg = 12 * (int)(c & 1u) + __popc(c & 0x7eu);
    float t = __fsub_rn(xc, __fmul_rn(res, c_invdg[dg]));
    t = t * 0.5f;
    const float lb = (c & kZero) ? -1000.0f : mi;
    const float ub = (c & kOne) ? 1000.0f : -mi;
    if (t < lb) t = lb;
    if (t > ub) t = ub;
    *delta = fabsf(t - xc);
    return t;
}

__global__ void __launch_bounds__(512) k_qp_persistent(const uint32_t* __restrict__ code, const float* __restrict__ m,
                                                       const int* __restrict__ nbr, float* xa, float* xb, int n,
                                                       int iters, unsigned* __restrict__ maxdelta_bits) {
    cg::grid_group grid = cg::this_grid();
    const int gsize = gridDim.x * blockDim.x;
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
    float* in = xa;
    float* out = xb;
    for (int it = 0; it < iters; it++) {
        float dmax = 0.0f;
        for (int i = gtid; i < n; i += gsize) {
            float dl;
            // plain loads: the iterate is rewritten by other SMs every sweep (grid.sync orders them)
            out[i] = qp_update(in, i, n, __ldg(code + i), __ldg(m + i), nbr, &dl);
            dmax = fmaxf(dmax, dl);
        }
        if (it == iters - 1) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) dmax = fmaxf(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
            if ((threadIdx.x & 31) == 0) atomicMax(maxdelta_bits, __float_as_uint(dmax));
        }
        float* t = in; in = out; out = t;
        if (it + 1 < iters) grid.sync();
    }
}

// x0 = clamp(0, lb, ub) (primeqp.cpp:178-187) into the padded buffer; also clears the pads of both buffers.
__global__ void k_qp_init(const uint32_t* __restrict__ code, const float* __restrict__ m, int n, float* __restrict__ xa_base,
                          float* __restrict__ xb_base) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 2) { xa_base[i] = 0.0f; xb_base[i] = 0.0f; xa_base[n + 2 + i] = 0.0f; xb_base[n + 2 + i] = 0.0f; }
    if (i >= n) return;
    uint32_t c = code[i];
    float mi = m[i];
    float lb = (c & kZero) ? -1000.0f : mi;
    float ub = (c & kOne) ? 1000.0f : -mi;
    float x = 0.0f;
    if (x < lb) x = lb;
    if (x > ub) x = ub;
    xa_base[2 + i] = x;
    xb_base[2 + i] = 0.0f;
}

int dev_qp_solve(mrb_ctx* ctx, const BandDev& bd, Dims d, int band_size, int iters, float* x_out, float* last_delta_host) {
    (void)d; (void)band_size;
    const int n = bd.n_band;
    if (last_delta_host) *last_delta_host = 0.0f;
    if (n == 0) return MRB_OK;
    float *xa = nullptr, *xb = nullptr;
    MRB_TRY(arena_alloc(ctx, (size_t)n + 4, &xa));
    MRB_TRY(arena_alloc(ctx, (size_t)n + 4, &xb));
    MRB_LAUNCH(ctx, k_qp_init, cdiv((size_t)n, 256), 256, 0, bd.code, bd.margin, n, xa, xb);
    unsigned* md = (unsigned*)(ctx->d_mail + 8);
    MRB_CUDA(cudaMemsetAsync(md, 0, sizeof(unsigned), ctx->stream));
    if (iters > 0) {
        int per_sm = 0;
        const int block = 512;
        MRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_qp_persistent, block, 0));
        if (per_sm < 1) { set_error("k_qp_persistent does not fit on an SM"); return MRB_ECUDA; }
        int grid = per_sm * ctx->sm_count;
        int need = (n + block - 1) / block;
        if (grid > need) grid = need;
        const uint32_t* code = bd.code;
        const float* m = bd.margin;
        const int* nbr = bd.nbr;
        float* pa = xa + 2;
        float* pb = xb + 2;
        int nn = n, its = iters;
        void* args[] = {(void*)&code, (void*)&m, (void*)&nbr, (void*)&pa, (void*)&pb, (void*)&nn, (void*)&its, (void*)&md};
        MRB_CUDA(cudaEventRecord(ctx->ev[6], ctx->stream));
        MRB_CUDA(cudaLaunchCooperativeKernel((void*)k_qp_persistent, dim3(grid), dim3(block), args, 0, ctx->stream));
        MRB_CUDA(cudaEventRecord(ctx->ev[7], ctx->stream));
        ctx->launches++;
    }
    // result = buffer written by the last sweep (iters odd -> xb, even -> xa; iters == 0 -> x0)
    const float* res = (iters % 2) ? xb + 2 : xa + 2;
    MRB_CUDA(cudaMemcpyAsync(x_out, res, (size_t)n * sizeof(float), cudaMemcpyDeviceToDevice, ctx->stream));
    if (last_delta_host) {
        MRB_CUDA(cudaMemcpyAsync(ctx->h_mail + 8, md, sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
        MRB_CUDA(cudaStreamSynchronize(ctx->stream));
        unsigned bits = (unsigned)ctx->h_mail[8];
        memcpy(last_delta_host, &bits, sizeof(float));
    }
    return MRB_OK;
}

// ---- F assembly (quadprog.cpp:131-145) --------------------------------------------------------------
__global__ void __launch_bounds__(256) k_assemble_F(const float* __restrict__ volume, const int* __restrict__ band_id,
                                                    const float* __restrict__ x, size_t n, double scale, float* __restrict__ F) {
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (size_t)gridDim.x * blockDim.x) {
        int id = band_id[idx];
        float f;
        if (id >= 0) f = x[id];
        else f = (float)__dmul_rn(__dsub_rn(__dmul_rn(2.0, (double)volume[idx]), 1.0), scale);
        F[idx] = f;
    }
}

int dev_assemble_F(mrb_ctx* ctx, const float* volume, const BandDev& bd, const float* x, Dims d, int band_size, float* F) {
    size_t n = d.count();
    size_t want = (n + 255) / 256, cap = (size_t)ctx->sm_count * 16;
    MRB_LAUNCH(ctx, k_assemble_F, (unsigned)(want < cap ? want : cap), 256, 0, volume, bd.band_id, x, n,
               (double)band_size + 1.0, F);
    return MRB_OK;
}

// ---- explicit problem data for the module-level API (primeQP + getIr/getJc/getPr) -------------------
__global__ void k_qp_vectors(const float* __restrict__ volume, const float* __restrict__ margin, const int* __restrict__ pos,
                             const uint32_t* __restrict__ code, int n, float* __restrict__ lb, float* __restrict__ ub,
                             float* __restrict__ x0, float* __restrict__ invdg, int* __restrict__ nnz_row) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int p = pos[i];
    float v = volume[p], mg = margin[p];
    float l = (v == 0.0f) ? -1000.0f : mg;
    float u = (v == 1.0f) ? 1000.0f : -mg;
    if (lb) lb[i] = l;
    if (ub) ub[i] = u;
    if (x0) { float x = 0.0f; if (x < l) x = l; if (x > u) x = u; x0[i] = x; }
    uint32_t c = code[i];
    int dg = 12 * (int)(c & 1u) + __popc(c & 0x7eu);
    if (invdg) invdg[i] = c_invdg[dg];
    // off-diagonal non-zeros: +-e present iff T(v)|T(v+-e); +-2e present iff T(v+-e)
    int t0 = c & 1u;
    int cnt = 0;
    for (int b = 1; b <= 6; b++) {
        int tn = (c >> b) & 1u;
        cnt += tn;               // +-2e
        cnt += (t0 | tn);        // +-e
    }
    nnz_row[i] = cnt;
}

__global__ void k_qp_csc(const uint32_t* __restrict__ code, const int* __restrict__ nbr, int n, const int* __restrict__ jc,
                         int* __restrict__ ir, float* __restrict__ pr) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t c = code[i];
    const size_t N = (size_t)n;
    const float t0 = (c & kT0) ? 1.0f : 0.0f;
    // ascending band id: x-2 x-1 y-2 y-1 z-2 z-1 z+1 z+2 y+1 y+2 x+1 x+2
    int ids[12] = {nbr[0 * N + i], nbr[1 * N + i], nbr[2 * N + i], nbr[3 * N + i], i - 2, i - 1,
                   i + 1, i + 2, nbr[4 * N + i], nbr[5 * N + i], nbr[6 * N + i], nbr[7 * N + i]};
    const uint32_t bit[12] = {kTxm, kTxm, kTym, kTym, kTzm, kTzm, kTzp, kTzp, kTyp, kTyp, kTxp, kTxp};
    const bool far[12] = {true, false, true, false, true, false, false, true, false, true, false, true};
    int w = jc[i];
    for (int k = 0; k < 12; k++) {
        float tn = (c & bit[k]) ? 1.0f : 0.0f;
        float coef = far[k] ? tn : -2.0f * (t0 + tn);
        if (coef != 0.0f) {
            if (ir) ir[w] = ids[k];
            if (pr) pr[w] = coef;
            w++;
        }
    }
}

// ---- generic CSC solver (module-level runQP on caller-provided matrices) --------------------------------
__global__ void __launch_bounds__(512) k_csc_persistent(const int* __restrict__ jc, const int* __restrict__ ir,
                                                        const float* __restrict__ pr, const float* __restrict__ invdg,
                                                        const float* __restrict__ lb, const float* __restrict__ ub,
                                                        float* xa, float* xb, int n, int iters) {
    cg::grid_group grid = cg::this_grid();
    const int gsize = gridDim.x * blockDim.x;
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
    float* in = xa;
    float* out = xb;
    for (int it = 0; it < iters; it++) {
        for (int row = gtid; row < n; row += gsize) {
            float res = 0.0f;
            int s = __ldg(jc + row), e = __ldg(jc + row + 1);
            for (int k = s; k < e; k++) res = __fadd_rn(res, __fmul_rn(__ldg(pr + k), in[__ldg(ir + k)]));
            float t = __fsub_rn(in[row], __fmul_rn(res, __ldg(invdg + row)));
            t = t * 0.5f;
            float l = __ldg(lb + row), u = __ldg(ub + row);
            if (t < l) t = l;
            if (t > u) t = u;
            out[row] = t;
        }
        float* t = in; in = out; out = t;
        if (it + 1 < iters) grid.sync();
    }
}

// quadprog.cpp:92-101: count=0; for i: if (i == featureIndexes[count]) { out[i]=0.2; count++; }
// With a sorted list this resets the strictly increasing prefix of feature ids and stalls at the
// first id that is not larger than the previous one (any -1 stalls everything).
__global__ void k_feature_reset(const int* __restrict__ feat, int nfeat, int n, float* __restrict__ x) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    int prev = -1;
    for (int k = 0; k < nfeat; k++) {
        int f = feat[k];
        if (f <= prev || f >= n) break;
        x[f] = 0.2f;
        prev = f;
    }
}

int dev_feature_reset(mrb_ctx* ctx, const int* feat_dev, int nfeat, int n, float* x) {
    if (nfeat <= 0) return MRB_OK;
    MRB_LAUNCH(ctx, k_feature_reset, 1, 32, 0, feat_dev, nfeat, n, x);
    return MRB_OK;
}

}  // namespace mrb

using namespace mrb;

extern "C" {

int mrb_prime_qp(mrb_ctx* ctx, const float* volume, const float* margin, const float* band, const float* tight,
                 const int dims[3], int n, float* lb, float* ub, float* x0, float* invdg, int* jc, int* ir, float* pr,
                 int nnz_capacity, int* nnz_out) {
    MRB_REQUIRE(ctx && volume && margin && band && tight && dims, "NULL argument");
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    Dims d = mkdims(dims);
    const float *dv, *dm, *db, *dt;
    MRB_TRY(stage_in(ctx, volume, d.count(), &dv));
    MRB_TRY(stage_in(ctx, margin, d.count(), &dm));
    MRB_TRY(stage_in(ctx, band, d.count(), &db));
    MRB_TRY(stage_in(ctx, tight, d.count(), &dt));
    BandDev bd;
    MRB_TRY(dev_qp_build(ctx, dv, dm, db, dt, d, &bd));
    if (bd.n_band != n) { set_error("mrb_prime_qp: n (%d) != band voxel count (%d)", n, bd.n_band); return MRB_EINVAL; }
    if (n == 0) { if (nnz_out) *nnz_out = 0; if (jc) { int z = 0; if (is_device_ptr(jc)) cudaMemcpy(jc, &z, 4, cudaMemcpyHostToDevice); else jc[0] = 0; } return MRB_OK; }
    Out<float> olb, oub, ox0, oinv, opr;
    Out<int> ojc, oir;
    MRB_TRY(olb.init(ctx, lb, n)); MRB_TRY(oub.init(ctx, ub, n)); MRB_TRY(ox0.init(ctx, x0, n)); MRB_TRY(oinv.init(ctx, invdg, n));
    int* jc_dev = nullptr;   // always needed internally
    MRB_TRY(ojc.init(ctx, jc, (size_t)n + 1));
    if (ojc.dev) jc_dev = ojc.dev; else MRB_TRY(arena_alloc(ctx, (size_t)n + 1, &jc_dev));
    MRB_LAUNCH(ctx, k_qp_vectors, cdiv(n, 256), 256, 0, dv, dm, bd.band_pos, bd.code, n, olb.dev, oub.dev, ox0.dev, oinv.dev, jc_dev);
    MRB_TRY(scan_exclusive(ctx, jc_dev, jc_dev, (size_t)n, jc_dev + n));
    int nnz = 0;
    MRB_TRY(read_int(ctx, jc_dev + n, &nnz));
    if (nnz_out) *nnz_out = nnz;
    if (ir || pr) {
        if (nnz > nnz_capacity) { set_error("mrb_prime_qp: nnz capacity %d < %d", nnz_capacity, nnz); return MRB_ECAPACITY; }
        MRB_TRY(oir.init(ctx, ir, (size_t)nnz)); MRB_TRY(opr.init(ctx, pr, (size_t)nnz));
        MRB_LAUNCH(ctx, k_qp_csc, cdiv(n, 256), 256, 0, bd.code, bd.nbr, n, jc_dev, oir.dev, opr.dev);
    }
    MRB_TRY(olb.commit(ctx)); MRB_TRY(oub.commit(ctx)); MRB_TRY(ox0.commit(ctx)); MRB_TRY(oinv.commit(ctx));
    MRB_TRY(ojc.commit(ctx)); MRB_TRY(oir.commit(ctx)); MRB_TRY(opr.commit(ctx));
    return sync(ctx);
}

int mrb_run_qp(mrb_ctx* ctx, int n, const int* jc, const int* ir, const float* pr, const float* invdg, const float* lb,
               const float* ub, const float* x0, int iters, const int* feature_idx, int nfeat, int using_features,
               float* x_out) {
    MRB_REQUIRE(ctx && jc && invdg && lb && ub && x0 && x_out && n >= 0 && iters >= 0, "bad argument");
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    if (n == 0) return MRB_OK;
    // nnz = jc[n]
    int nnz = 0;
    if (is_device_ptr(jc)) MRB_TRY(read_int(ctx, jc + n, &nnz)); else nnz = jc[n];
    MRB_REQUIRE(nnz >= 0 && (nnz == 0 || (ir && pr)), "ir/pr missing");
    const int *djc, *dir, *dfeat = nullptr;
    const float *dpr, *dinv, *dlb, *dub, *dx0;
    MRB_TRY(stage_in(ctx, jc, (size_t)n + 1, &djc));
    MRB_TRY(stage_in(ctx, ir, (size_t)(nnz > 0 ? nnz : 1), &dir));
    MRB_TRY(stage_in(ctx, pr, (size_t)(nnz > 0 ? nnz : 1), &dpr));
    MRB_TRY(stage_in(ctx, invdg, (size_t)n, &dinv));
    MRB_TRY(stage_in(ctx, lb, (size_t)n, &dlb));
    MRB_TRY(stage_in(ctx, ub, (size_t)n, &dub));
    MRB_TRY(stage_in(ctx, x0, (size_t)n, &dx0));
    if (using_features && nfeat > 0) MRB_TRY(stage_in(ctx, feature_idx, (size_t)nfeat, &dfeat));
    float *xa, *xb;
    MRB_TRY(arena_alloc(ctx, (size_t)n, &xa));
    MRB_TRY(arena_alloc(ctx, (size_t)n, &xb));
    MRB_CUDA(cudaMemcpyAsync(xa, dx0, (size_t)n * sizeof(float), cudaMemcpyDeviceToDevice, ctx->stream));
    MRB_CUDA(cudaMemsetAsync(xb, 0, (size_t)n * sizeof(float), ctx->stream));
    if (iters > 0) {
        int per_sm = 0;
        const int block = 512;
        MRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_csc_persistent, block, 0));
        if (per_sm < 1) { set_error("k_csc_persistent does not fit on an SM"); return MRB_ECUDA; }
        int grid = per_sm * ctx->sm_count;
        int need = (n + block - 1) / block;
        if (grid > need) grid = need;
        void* args[] = {(void*)&djc, (void*)&dir, (void*)&dpr, (void*)&dinv, (void*)&dlb, (void*)&dub,
                        (void*)&xa, (void*)&xb, (void*)&n, (void*)&iters};
        MRB_CUDA(cudaLaunchCooperativeKernel((void*)k_csc_persistent, dim3(grid), dim3(block), args, 0, ctx->stream));
        ctx->launches++;
    }
    float* res = (iters % 2) ? xb : xa;
    if (using_features) MRB_TRY(dev_feature_reset(ctx, dfeat, nfeat, n, res));
    if (is_device_ptr(x_out)) MRB_CUDA(cudaMemcpyAsync(x_out, res, (size_t)n * sizeof(float), cudaMemcpyDeviceToDevice, ctx->stream));
    else MRB_CUDA(cudaMemcpyAsync(x_out, res, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    return sync(ctx);
}

}  // extern "C"

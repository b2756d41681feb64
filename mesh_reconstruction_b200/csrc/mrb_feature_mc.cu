// mrb_feature_mc.cu -- marching cubes with feature points and edge flips (mcubes with USING_FEATURES).
//
// Replaces getGridCells + isFeature, the feature branch of Polygonise, flipEdges and the concatenation in mcubes
// (marching_cubes/mcubes.cpp:9-152, 157-312, 396-588, 592-616).  Per-cell arithmetic = include/mrb_feature_mc.h (shared
// with the oracle; parity unpinned against the reference's Eigen::JacobiSVD, bit-exact against our oracle).
//
//   1. classify   sign-bit planes -> ordered list of crossing cells (mrb_mcubes.cu)
//   2. polygonise one thread per crossing cell: feature class, feature point, triangle fan or table triangles
//   3. flip       flipEdges is order dependent (a cell's triangle list is edited by every earlier neighbour), so
//                 the feature cells are visited in cell order by ONE warp: for each (cell, neighbour) pair the
//                 lanes test the triangle pairs for a shared edge in parallel, lane 0 applies the flips in the
//                 reference's order.  Only feature cells take part (a few thousand on sharp-edged shapes).
//   4. emit       remaining per-cell triangles in cell order, then the flipped triangles; hashXYZ keys; de-dup.
#include "mrb_internal.cuh"
#include <algorithm>
#include <vector>
#include "mrb_compact.cuh"
#include "mc_tables.h"
#include "../../include/mrb_feature_mc.h"

#include <string.h>

namespace mrb {

__constant__ uint64_t c_tri_f[256] = MRB_MC_TRI_WORDS;
__constant__ int c_ea_f[12] = {0, 1, 2, 3, 4, 5, 6, 7, 0, 1, 2, 3};
__constant__ int c_eb_f[12] = {1, 2, 3, 0, 5, 6, 7, 4, 4, 5, 6, 7};

__global__ void k_fill_i32f(int* __restrict__ p, size_t n, int v) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}
__global__ void k_scatter_slots(const int* __restrict__ cells, int n, int* __restrict__ cellmap) {
    int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a < n) cellmap[cells[a]] = a;
}

__global__ void __launch_bounds__(64) k_feat_polygonise(const float* __restrict__ F, const float* __restrict__ smap,
                                                        const float* __restrict__ normals, int nnormals, Dims d,
                                                        const int* __restrict__ cells, int ncells, float iso, float ft,
                                                        float ct, MrbFeatCell* __restrict__ out, int* __restrict__ err) {
    int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= ncells) return;
    const int cy = d.ny - 1, cz = d.nz - 1;
    int cell = cells[a];
    int k = cell % cz, t = cell / cz, j = t % cy, i = t / cy;
    float val[8];
    int nidx[8], cubeindex = 0, ea[12], eb[12];
    for (int n = 0; n < 8; n++) {
        size_t q = d.at(i + ((((n % 4) == 1) || ((n % 4) == 2)) ? 1 : 0), j + (n % 4) / 2, k + n / 4);
        val[n] = F[q];
        nidx[n] = (int)smap[q];
        if (nidx[n] < -1 || nidx[n] >= nnormals) { *err = 1; nidx[n] = -1; }
        if (val[n] < iso) cubeindex |= (1 << n);
    }
    for (int e = 0; e < 12; e++) { ea[e] = c_ea_f[e]; eb[e] = c_eb_f[e]; }
    MrbFeatCell c;
    mrb_polygonise_feature(i, j, k, val, nidx, normals, iso, ft, ct, c_tri_f[cubeindex], cubeindex, ea, eb, &c);
    out[a] = c;
}

struct FeaturePred {
    const MrbFeatCell* c;
    __device__ bool operator()(size_t a) const { return c[a].feature != 0; }
};

// flipEdges (mcubes.cpp:434-588) by one warp.  fcells: slots of the feature cells in cell order.
__global__ void __launch_bounds__(32) k_flip_edges(MrbFeatCell* cellsv, const int* __restrict__ cells, const int* __restrict__ fcells,
                                                   int nf, const int* __restrict__ cellmap, Dims d, MrbTri* __restrict__ new_tris,
                                                   int cap, int* __restrict__ nnew_out, int* __restrict__ err) {
    const int lane = threadIdx.x;
    const int cx = d.nx - 1, cy = d.ny - 1, cz = d.nz - 1;
    __shared__ int s_si[MRB_FEAT_MAX_TRI * MRB_FEAT_MAX_TRI], s_sj[MRB_FEAT_MAX_TRI * MRB_FEAT_MAX_TRI];
    __shared__ unsigned char s_flag[MRB_FEAT_MAX_TRI * MRB_FEAT_MAX_TRI];
    int nnew = 0;
    for (int f = 0; f < nf; f++) {
        const int a = fcells[f];
        MrbFeatCell* c = cellsv + a;
        if (c->feature == 0) continue;
        const int id = cells[a];
        const int z = id % cz, y = (id / cz) % cy, x = id / (cz * cy);
        int nb[6], nn = 0;
        if (x != 0) nb[nn++] = id - cy * cz;
        if (x != cx - 1) nb[nn++] = id + cy * cz;
        if (y != 0) nb[nn++] = id - cz;
        if (y != cy - 1) nb[nn++] = id + cz;
        if (z != 0) nb[nn++] = id - 1;
        if (z != cz - 1) nb[nn++] = id + 1;
        for (int q = 0; q < nn; q++) {
            const int b = cellmap[nb[q]];
            if (b < 0) continue;                       // not a surface cell
            MrbFeatCell* n = cellsv + b;
            if (n->feature == 0) continue;
            const int nc = c->ntri, nnb = n->ntri, np = nc * nnb;
            for (int p = lane; p < np; p += 32) s_flag[p] = (unsigned char)mrb_flip_shared(c, n, p / nnb, p % nnb);
            __syncwarp();
            if (lane == 0) {
                int ns = 0;
                for (int p = 0; p < np; p++)
                    if (s_flag[p]) { s_si[ns] = p / nnb; s_sj[ns] = p % nnb; ns++; }
                if (ns > 0) {
                    if (nnew + 2 * ns > cap) { *err = 2; }
                    else nnew += mrb_flip_apply(c, n, s_si, s_sj, ns, new_tris + nnew, cap - nnew);
                }
            }
            nnew = __shfl_sync(0xffffffffu, nnew, 0);
            __threadfence_block();
            __syncwarp();
        }
    }
    if (lane == 0) *nnew_out = nnew;
}

// flipEdges in dependency order instead of list order.  A step (feature cell f against its feature neighbours) reads and
// rewrites only f's cell and those neighbours, so two steps whose cells are more than two cell-steps apart commute.  One
// warp per feature cell: it takes its place in the list from a ticket counter (so every earlier cell's warp is already
// running), waits until the earlier feature cells within Manhattan distance 2 are done, runs the reference's step, and
// marks itself done.  Triangles the reference would append to its global list go into a slot per (cell, neighbour)
// and are emitted in (cell, neighbour) order afterwards, which is the reference's append order.  The result is the
// sequential one bit for bit; the critical path is the longest run of adjacent feature cells (an edge of the object),
// not the number of feature cells (172^3 box + sphere: 14.9 ms -> see profiles).
// mrb_flip_shared without the square roots: sameEdge compares (double)sqrtf(d) < 0.00001 for two squared distances d;
// sqrtf is monotone, so that is d <= D with D the largest float whose root passes (found once on the host by bisection
// over the float bit patterns).  Same booleans as mrb_flip_shared, about a third of the instructions.
static float same_edge_threshold() {
    uint32_t lo = 0u, hi = 0x7f800000u;            // passes(lo), !passes(hi)
    while (hi - lo > 1u) {
        const uint32_t mid = lo + (hi - lo) / 2u;
        float dd;
        memcpy(&dd, &mid, 4);
        if ((double)sqrtf(dd) < 0.00001) lo = mid; else hi = mid;
    }
    float out;
    memcpy(&out, &lo, 4);
    return out;
}
__device__ __forceinline__ float p3_sqdist(const MrbP3& l, const MrbP3& r) {       // the sum inside mrb_p3_dist, same operation order
    float dd = __fmul_rn(__fsub_rn(r.x, l.x), __fsub_rn(r.x, l.x));
    dd = __fadd_rn(dd, __fmul_rn(__fsub_rn(r.y, l.y), __fsub_rn(r.y, l.y)));
    dd = __fadd_rn(dd, __fmul_rn(__fsub_rn(r.z, l.z), __fsub_rn(r.z, l.z)));
    return dd;
}
__device__ __forceinline__ bool flip_shared_fast(const MrbFeatCell* c, const MrbFeatCell* nb, int i, int j, float D) {
    for (int n = 0; n < 3; n++) {
        MrbP3 a0 = c->tri[i].p[n], a1 = c->tri[i].p[(n + 1) % 3];
        if (mrb_p3_less(a1, a0)) { MrbP3 t = a0; a0 = a1; a1 = t; }
        for (int m = 0; m < 3; m++) {
            MrbP3 b0 = nb->tri[j].p[m], b1 = nb->tri[j].p[(m + 1) % 3];
            if (mrb_p3_less(b1, b0)) { MrbP3 t = b0; b0 = b1; b1 = t; }
            if (p3_sqdist(a0, b0) <= D && p3_sqdist(a1, b1) <= D) return true;
        }
    }
    return false;
}

static const int kFlipSlot = 8;       // new triangles per (cell, neighbour) step; more -> *err = 3, the caller runs k_flip_edges
__global__ void k_findex(const int* __restrict__ fcells, int nf, int* __restrict__ findex) {
    int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f < nf) findex[fcells[f]] = f;
}
__global__ void __launch_bounds__(128) k_flip_edges_par(MrbFeatCell* cellsv, const int* __restrict__ cells, const int* __restrict__ fcells,
                                                        int nf, const int* __restrict__ cellmap, const int* __restrict__ findex, Dims d,
                                                        MrbTri* __restrict__ slots, int* __restrict__ slot_cnt, unsigned* done,
                                                        unsigned* ticket, int* __restrict__ err, int chunk, long long* __restrict__ prof, float sqD) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int cx = d.nx - 1, cy = d.ny - 1, cz = d.nz - 1;
    __shared__ int s_si_all[4][MRB_FEAT_MAX_TRI * MRB_FEAT_MAX_TRI], s_sj_all[4][MRB_FEAT_MAX_TRI * MRB_FEAT_MAX_TRI];
    int* s_si = s_si_all[w];
    int* s_sj = s_sj_all[w];
    // a warp takes `chunk` consecutive cells of the list: cells that follow each other in the list are usually adjacent in
    // z, and inside one warp their order costs no flag hand-over
    int t = 0;
    if (lane == 0) t = (int)atomicAdd(ticket, 1u);
    t = __shfl_sync(0xffffffffu, t, 0);
    const int f_begin = t * chunk;
    const long long t_start = prof ? clock64() : 0;
    for (int f = f_begin; f < min(nf, f_begin + chunk); f++) {
    const int a = fcells[f];
    unsigned depth = 0;          // done[f] = 1 + the longest chain of earlier steps this one waited for (a statistic, and the flag)
    MrbFeatCell* c = cellsv + a;
    const int id = cells[a];
    const int z = id % cz, y = (id / cz) % cy, x = id / (cz * cy);
    // Which neighbours take part does not change during the kernel (feature flags and the cell map are read-only here):
    // looked up BEFORE the wait, lane q for neighbour q (two dependent loads), shared by shuffles after it.
    const bool is_feature = c->feature != 0;
    int myb = -1;
    {
        int nb[6], nn = 0;
        if (x != 0) nb[nn++] = id - cy * cz;
        if (x != cx - 1) nb[nn++] = id + cy * cz;
        if (y != 0) nb[nn++] = id - cz;
        if (y != cy - 1) nb[nn++] = id + cz;
        if (z != 0) nb[nn++] = id - 1;
        if (z != cz - 1) nb[nn++] = id + 1;
        if (lane < nn) {
            const int b = cellmap[nb[lane]];
            if (b >= 0 && cellsv[b].feature != 0) myb = b;           // a surface cell that is a feature cell
        }
    }
    // earlier feature cells within Manhattan distance 2: 6 at distance 1, 6 straight + 12 diagonal at distance 2
    int fp = -1;                 // the earlier feature cell this lane has to wait for
    if (lane < 24) {
        int dx = 0, dy = 0, dz = 0;
        if (lane < 6) { const int ax = lane >> 1, sg = (lane & 1) ? 1 : -1; (ax == 0 ? dx : ax == 1 ? dy : dz) = sg; }
        else if (lane < 12) { const int ax = (lane - 6) >> 1, sg = (lane & 1) ? 2 : -2; (ax == 0 ? dx : ax == 1 ? dy : dz) = sg; }
        else {
            const int k = lane - 12, pair = k >> 2, s1 = (k & 1) ? 1 : -1, s2 = (k & 2) ? 1 : -1;      // pairs: xy, xz, yz
            if (pair == 0) { dx = s1; dy = s2; } else if (pair == 1) { dx = s1; dz = s2; } else { dy = s1; dz = s2; }
        }
        const int xx = x + dx, yy = y + dy, zz = z + dz;
        if (xx >= 0 && xx < cx && yy >= 0 && yy < cy && zz >= 0 && zz < cz) {
            const int b = cellmap[(xx * cy + yy) * cz + zz];
            const int fq = b >= 0 ? findex[b] : -1;
            if (fq >= 0 && fq < f_begin) fp = fq;          // earlier cells of this warp's own chunk are done by construction
        }
    }
    // Poll the latest predecessor first (the others have usually finished by then), with relaxed loads and a short sleep:
    // 1800 warps polling 24 flags each with acquire loads slow every other warp's L2 round trip down several times.
    int fmax = fp;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) fmax = max(fmax, __shfl_xor_sync(0xffffffffu, fmax, o));
    for (int pass = 0; pass < 2; pass++) {
        if (fp >= 0 && (pass == 1 || fp == fmax)) {
            unsigned v, spins = 0;
            for (;;) {
                asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(done + fp) : "memory");
                if (v != 0) break;
                __nanosleep(200);
                if (++spins == (1u << 26)) __trap();     // ~15 s: a lost predecessor must not hang the GPU
            }
            depth = max(depth, v);
        }
        __syncwarp();
    }
    asm volatile("fence.acq_rel.gpu;" ::: "memory");        // the flags were read relaxed: order the cell reads after them
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) depth = max(depth, __shfl_xor_sync(0xffffffffu, depth, o));
    __syncwarp();
    const long long t_waited = prof ? clock64() : 0;
    // The cell and its (up to six) feature neighbours are fetched into shared memory in ONE round of loads, worked on
    // there, and the changed ones written back: the step itself is a chain of dependent accesses by one lane, far too
    // slow against L2, and fetching the neighbours one after the other would pay the L2 latency six times.
    __shared__ MrbFeatCell s_cell[4][7];
    MrbFeatCell* sc = &s_cell[w][0];
    constexpr int kWords = (int)(sizeof(MrbFeatCell) / 4);
    if (is_feature) {
        MrbFeatCell* gsrc[7];
        gsrc[0] = c;
#pragma unroll
        for (int q = 0; q < 6; q++) {
            const int b = __shfl_sync(0xffffffffu, myb, q);
            gsrc[q + 1] = b >= 0 ? cellsv + b : nullptr;
        }
        // all loads first (independent, one L2 round trip), then the shared-memory stores: interleaved, every generic load
        // would have to wait for the store before it
        constexpr int kPer = (kWords + 31) / 32;
        uint32_t regs[7][kPer];
#pragma unroll
        for (int k2 = 0; k2 < 7; k2++)
#pragma unroll
            for (int i = 0; i < kPer; i++) {
                const int wi = lane + 32 * i;
                regs[k2][i] = (gsrc[k2] && wi < kWords) ? __ldcg((const uint32_t*)gsrc[k2] + wi) : 0u;
            }
#pragma unroll
        for (int k2 = 0; k2 < 7; k2++)
#pragma unroll
            for (int i = 0; i < kPer; i++) {
                const int wi = lane + 32 * i;
                if (gsrc[k2] && wi < kWords) ((uint32_t*)&s_cell[w][k2])[wi] = regs[k2][i];
            }
        __syncwarp();
        const long long t_loaded = prof ? clock64() : 0;
        if (prof && lane == 0) prof[4 * f + 1] = t_loaded - t_waited;
        unsigned dirty = 0;
#pragma unroll
        for (int q = 0; q < 6; q++) {
            if (!gsrc[q + 1]) continue;
            MrbFeatCell* sn = &s_cell[w][q + 1];
            const int nc = sc->ntri, nnb = sn->ntri, np = nc * nnb;
            int ns = 0;                                 // (i, j) pairs that share an edge, in (i, j) order: lane 0 collects them
            for (int base = 0; base < np; base += 32) {
                const int p = base + lane;
                const bool fl = p < np && flip_shared_fast(sc, sn, p / nnb, p % nnb, sqD);
                unsigned m = __ballot_sync(0xffffffffu, fl);
                if (lane == 0)
                    while (m) {
                        const int pp = base + __ffs(m) - 1;
                        m &= m - 1;
                        s_si[ns] = pp / nnb; s_sj[ns] = pp % nnb; ns++;
                    }
            }
            if (lane == 0 && ns > 0) {
                if (2 * ns > kFlipSlot) *err = 3;
                else slot_cnt[f * 6 + q] = mrb_flip_apply(sc, sn, s_si, s_sj, ns, slots + (size_t)(f * 6 + q) * kFlipSlot, kFlipSlot);
            }
            ns = __shfl_sync(0xffffffffu, ns, 0);       // (also orders lane 0's shared-memory writes before the next reads)
            if (ns > 0) dirty |= 1u | (2u << q);
        }
#pragma unroll
        for (int k2 = 0; k2 < 7; k2++)
            if (gsrc[k2] && ((dirty >> k2) & 1u)) {
                const uint32_t* s32 = (const uint32_t*)&s_cell[w][k2];
                uint32_t* d32 = (uint32_t*)gsrc[k2];
                for (int i = lane; i < kWords; i += 32) d32[i] = s32[i];
            }
    }
    __syncwarp();
    if (prof && lane == 0) { prof[4 * f + 0] = t_waited - t_start; prof[4 * f + 2] = clock64() - t_waited; }
    // one release per cell: it orders every lane's writes of this step (made cumulative by the __syncwarp) before the flag
    if (lane == 0) asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(done + f), "r"(depth + 1u) : "memory");
    if (prof && lane == 0) prof[4 * f + 3] = clock64() - t_start;
    }
}
__global__ void k_feat_emit_slots(const MrbTri* __restrict__ slots, const int* __restrict__ off, int nsteps, int first,
                                  float* __restrict__ soup, unsigned long long* __restrict__ keys, int ny, int nz);

__global__ void k_feat_counts(const MrbFeatCell* __restrict__ c, int n, int* __restrict__ cnt) {
    int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a < n) cnt[a] = c[a].ntri;
}

__device__ __forceinline__ unsigned long long hash_p3(MrbP3 s, int ny, int nz) {      // hashXYZ, hasher.h:99-104
    unsigned long long val = (unsigned long long)__dmul_rn(__dmul_rn((double)s.x, 1000000000.0), (double)(ny * nz));
    val += (unsigned long long)__dmul_rn(__dmul_rn((double)s.y, 1000000.0), (double)nz);
    val += (unsigned long long)__dmul_rn((double)s.z, 1000.0);
    return val;
}
__device__ __forceinline__ void put_tri(const MrbTri& t, size_t tri, float* __restrict__ soup, unsigned long long* __restrict__ keys,
                                        int ny, int nz) {
    for (int v = 0; v < 3; v++) {
        soup[tri * 9 + v * 3 + 0] = t.p[v].x; soup[tri * 9 + v * 3 + 1] = t.p[v].y; soup[tri * 9 + v * 3 + 2] = t.p[v].z;
        keys[tri * 3 + v] = hash_p3(t.p[v], ny, nz);
    }
}
__global__ void k_feat_emit_slots(const MrbTri* __restrict__ slots, const int* __restrict__ off, int nsteps, int first,
                                  float* __restrict__ soup, unsigned long long* __restrict__ keys, int ny, int nz) {
    int st = blockIdx.x * blockDim.x + threadIdx.x;
    if (st >= nsteps) return;
    const int n = off[st + 1] - off[st];
    for (int i = 0; i < n; i++) put_tri(slots[(size_t)st * kFlipSlot + i], (size_t)first + off[st] + i, soup, keys, ny, nz);
}
__global__ void k_feat_emit(const MrbFeatCell* __restrict__ c, int n, const int* __restrict__ off, float* __restrict__ soup,
                            unsigned long long* __restrict__ keys, int ny, int nz) {
    int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= n) return;
    for (int t = 0; t < c[a].ntri; t++) put_tri(c[a].tri[t], (size_t)off[a] + t, soup, keys, ny, nz);
}
__global__ void k_feat_emit_new(const MrbTri* __restrict__ tris, int n, int first, float* __restrict__ soup,
                                unsigned long long* __restrict__ keys, int ny, int nz) {
    int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a < n) put_tri(tris[a], (size_t)first + a, soup, keys, ny, nz);
}

}  // namespace mrb

using namespace mrb;

extern "C" int mrb_mcubes_features(mrb_ctx* ctx, const float* F, const float* surface_map, const float* normals, int nnormals,
                                   const int dims[3], float isolevel, float feat_thresh, float corner_thresh, int device_out,
                                   mrb_mesh* out) {
    MRB_REQUIRE(ctx && F && surface_map && dims && out && nnormals >= 0 && (nnormals == 0 || normals), "bad argument");
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    memset(out, 0, sizeof(*out));
    Dims d = mkdims(dims);
    if (d.nx < 2 || d.ny < 2 || d.nz < 2) return MRB_OK;
    const float *dF = nullptr, *dm = nullptr, *dn = nullptr;
    MRB_TRY(stage_in(ctx, F, d.count(), &dF));
    MRB_TRY(stage_in(ctx, surface_map, d.count(), &dm));
    MRB_TRY(stage_in(ctx, normals, (size_t)(nnormals > 0 ? nnormals : 1) * 3, &dn));
    int* cells = nullptr;
    int ncells = 0;
    MRB_TRY(classify_cells(ctx, dF, d, isolevel, &cells, &ncells));
    if (ncells == 0) return MRB_OK;
    const size_t ncell_total = (size_t)(d.nx - 1) * (d.ny - 1) * (d.nz - 1);
    int* cellmap = nullptr;
    MrbFeatCell* fc = nullptr;
    MRB_TRY(arena_alloc(ctx, ncell_total, &cellmap));
    MRB_TRY(arena_alloc(ctx, (size_t)ncells, &fc));
    size_t want = (ncell_total + 255) / 256, cap = (size_t)ctx->sm_count * 16;
    MRB_LAUNCH(ctx, k_fill_i32f, (unsigned)(want < cap ? want : cap), 256, 0, cellmap, ncell_total, -1);
    MRB_LAUNCH(ctx, k_scatter_slots, cdiv(ncells, 256), 256, 0, cells, ncells, cellmap);
    int* err = ctx->d_mail + 54;
    MRB_CUDA(cudaMemsetAsync(err, 0, 2 * sizeof(int), ctx->stream));
    MRB_LAUNCH(ctx, k_feat_polygonise, cdiv(ncells, 64), 64, 0, dF, dm, dn, nnormals, d, cells, ncells, isolevel, feat_thresh,
               corner_thresh, fc, err);
    int* fcells = nullptr;
    int nf = 0;
    MRB_TRY(compact(ctx, FeaturePred{fc}, (size_t)ncells, &fcells, &nf));
    // flipEdges: dependency-ordered on many warps (k_flip_edges_par); if a step needs more slots than kFlipSlot the cells
    // are rebuilt and the one-warp kernel in list order runs instead (MRB_FLIP_SEQUENTIAL=1 forces that).
    MrbTri *new_tris = nullptr, *slots = nullptr;
    int *slot_off = nullptr, *tri_off = nullptr;
    MRB_TRY(arena_alloc(ctx, (size_t)ncells, &tri_off));
    int* nnew_dev = ctx->d_mail + 55;
    bool par = nf > 0 && !getenv("MRB_FLIP_SEQUENTIAL");
    int ntri_cells = 0, nnew = 0;
    for (int attempt = 0; attempt < 2; attempt++) {
        const int nsteps = nf * 6;
        if (par) {
            int* findex = nullptr;
            unsigned* done = nullptr;
            MRB_TRY(arena_alloc(ctx, (size_t)ncells, &findex));
            MRB_TRY(arena_alloc(ctx, (size_t)nsteps * kFlipSlot, &slots));
            MRB_TRY(arena_alloc(ctx, (size_t)nsteps + 1, &slot_off));
            MRB_TRY(arena_alloc(ctx, (size_t)nf + 1, &done));
            MRB_LAUNCH(ctx, k_fill_i32f, cdiv(ncells, 256), 256, 0, findex, (size_t)ncells, -1);
            MRB_LAUNCH(ctx, k_findex, cdiv(nf, 256), 256, 0, fcells, nf, findex);
            MRB_CUDA(cudaMemsetAsync(slot_off, 0, ((size_t)nsteps + 1) * sizeof(int), ctx->stream));
            MRB_CUDA(cudaMemsetAsync(done, 0, ((size_t)nf + 1) * sizeof(unsigned), ctx->stream));
            const int chunk = 1;          // several consecutive cells per warp were measured slower (9.5 ms at 2, 30 ms at 16)
            long long* prof = nullptr;
            if (getenv("MRB_FLIP_STATS")) { MRB_TRY(arena_alloc(ctx, (size_t)nf * 4, &prof)); MRB_CUDA(cudaMemsetAsync(prof, 0, (size_t)nf * 32, ctx->stream)); }
            MRB_LAUNCH(ctx, k_flip_edges_par, cdiv(cdiv(nf, chunk), 4), 128, 0, fc, cells, fcells, nf, cellmap, findex, d, slots, slot_off, done,
                       done + nf, err, chunk, prof, same_edge_threshold());
            if (prof) {
                std::vector<long long> hp((size_t)nf * 4);
                MRB_CUDA(cudaMemcpyAsync(hp.data(), prof, hp.size() * 8, cudaMemcpyDeviceToHost, ctx->stream));
                MRB_CUDA(cudaStreamSynchronize(ctx->stream));
                double sum[4] = {0, 0, 0, 0}; long long mx[4] = {0, 0, 0, 0};
                for (int i = 0; i < nf; i++) for (int j = 0; j < 4; j++) { sum[j] += (double)hp[(size_t)i * 4 + j]; mx[j] = std::max(mx[j], hp[(size_t)i * 4 + j]); }
                fprintf(stderr, "[mrb] flip cycles avg (max): wait %.0f (%lld), load %.0f (%lld), step %.0f (%lld), total %.0f (%lld)\n",
                        sum[0] / nf, mx[0], sum[1] / nf, mx[1], sum[2] / nf, mx[2], sum[3] / nf, mx[3]);
            }
            if (getenv("MRB_FLIP_STATS")) {
                std::vector<unsigned> h((size_t)nf);
                MRB_CUDA(cudaMemcpyAsync(h.data(), done, (size_t)nf * sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
                MRB_CUDA(cudaStreamSynchronize(ctx->stream));
                unsigned mx = 0;
                for (unsigned v : h) mx = std::max(mx, v);
                fprintf(stderr, "[mrb] flipEdges: %d feature cells of %d surface cells, longest dependency chain %u\n", nf, ncells, mx);
            }
            MRB_TRY(scan_exclusive(ctx, slot_off, slot_off, (size_t)nsteps, slot_off + nsteps));
            MRB_CUDA(cudaMemcpyAsync(ctx->h_mail + 55, slot_off + nsteps, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        } else {
            const int ncap = nf > 0 ? nf * 24 : 1;             // every feature cell shares edges with a few neighbours
            MRB_TRY(arena_alloc(ctx, (size_t)ncap, &new_tris));
            if (nf > 0) {
                MRB_LAUNCH(ctx, k_flip_edges, 1, 32, 0, fc, cells, fcells, nf, cellmap, d, new_tris, ncap, nnew_dev, err);
                MRB_CUDA(cudaMemcpyAsync(ctx->h_mail + 55, nnew_dev, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
            }
        }
        MRB_LAUNCH(ctx, k_feat_counts, cdiv(ncells, 256), 256, 0, fc, ncells, tri_off);
        MRB_TRY(scan_exclusive(ctx, tri_off, tri_off, (size_t)ncells, ctx->d_mail));
        MRB_CUDA(cudaMemcpyAsync(ctx->h_mail, ctx->d_mail, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        MRB_CUDA(cudaMemcpyAsync(ctx->h_mail + 54, err, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        MRB_TRY(sync(ctx));
        if (ctx->h_mail[54] == 1) { set_error("mrb_mcubes_features: surface map refers to a normal index outside [0, nnormals)"); return MRB_EINVAL; }
        if (ctx->h_mail[54] == 2) { set_error("mrb_mcubes_features: flipped-triangle buffer overflow"); return MRB_ECAPACITY; }
        if (ctx->h_mail[54] == 3 && par) {
            // a step with more than kFlipSlot new triangles: rebuild the cells and flip in list order
            par = false;
            MRB_CUDA(cudaMemsetAsync(err, 0, 2 * sizeof(int), ctx->stream));
            MRB_LAUNCH(ctx, k_feat_polygonise, cdiv(ncells, 64), 64, 0, dF, dm, dn, nnormals, d, cells, ncells, isolevel, feat_thresh,
                       corner_thresh, fc, err);
            continue;
        }
        ntri_cells = ctx->h_mail[0];
        nnew = nf > 0 ? ctx->h_mail[55] : 0;
        break;
    }
    const int ntri = ntri_cells + nnew;
    if (ntri == 0) return MRB_OK;
    float* soup = nullptr;
    unsigned long long* keys = nullptr;
    int* faces = nullptr;
    MRB_TRY(arena_alloc(ctx, (size_t)ntri * 9, &soup));
    MRB_TRY(arena_alloc(ctx, (size_t)ntri * 3, &keys));
    MRB_TRY(arena_alloc(ctx, (size_t)ntri * 3, &faces));
    MRB_LAUNCH(ctx, k_feat_emit, cdiv(ncells, 128), 128, 0, fc, ncells, tri_off, soup, keys, d.ny, d.nz);
    if (nnew > 0 && par) MRB_LAUNCH(ctx, k_feat_emit_slots, cdiv(nf * 6, 128), 128, 0, slots, slot_off, nf * 6, ntri_cells, soup, keys, d.ny, d.nz);
    else if (nnew > 0) MRB_LAUNCH(ctx, k_feat_emit_new, cdiv(nnew, 128), 128, 0, new_tris, nnew, ntri_cells, soup, keys, d.ny, d.nz);
    float* verts = nullptr;
    int nv = 0;
    MRB_TRY(dev_dedup(ctx, soup, keys, ntri * 3, faces, &verts, &nv));
    return export_mesh(ctx, verts, faces, nv, ntri, device_out, out);
}

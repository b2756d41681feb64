// mrb_mcubes.cu -- stage 4: marching cubes with hasher-equivalent vertex de-duplication.
//
// Replaces getGridCells / Polygonise / getSamplePoints / VertexInterp (marching_cubes/mcubes.cpp:
// 157-392) and pointHashmap (hasher.h:84-106, hasher.cpp:10-32) for the default (non-feature) path.
//
// The reference materialises one ~200-byte GRIDCELL per cell and walks them sequentially.  Here:
//   1. classify   one thread per cell (lanes along z): cube index from the 8 corner values;
//                 surface-crossing cells are compacted in (i,j,k) order (ballot + scan);
//   2. count      per crossing cell: interpolated edge vertices, triangles of the case table,
//                 degenerate-triangle filter (any side < 1e-5)  -> triangle count; exclusive scan;
//   3. emit       triangle corners (xyz) and their 64-bit hashXYZ keys at the scanned offsets:
//                 the triangle soup is in exactly the reference's concatenation order;
//   4. de-dup     open-addressing hash table on the 64-bit key, value = atomicMin(first position);
//                 a corner is a "first appearance" iff it holds the minimum position of its key;
//                 exclusive scan of those flags = the reference's vertex ids (first-appearance
//                 order), coordinates of the first occurrence are kept (hasher.cpp:13-23).
// Everything order-defining is a scan or a min, so the result is deterministic and identical to the
// sequential reference regardless of thread scheduling.
#include "mrb_internal.cuh"
#include "mrb_compact.cuh"
#include "mc_tables.h"

#include <stdlib.h>
#include <string.h>

namespace mrb {

__constant__ uint64_t c_tri[256] = MRB_MC_TRI_WORDS;
// edge e joins corners kEdgeA[e], kEdgeB[e] (mcubes.cpp:317-364)
__constant__ int c_edge_a[12] = {0, 1, 2, 3, 4, 5, 6, 7, 0, 1, 2, 3};
__constant__ int c_edge_b[12] = {1, 2, 3, 0, 5, 6, 7, 4, 4, 5, 6, 7};

struct CellDims {
    int nx, ny, nz;      // grid dims
    int cx, cy, cz;      // cell dims = grid dims - 1
    // stacked batch (mrb_reconstruct_batch): the grid is nobj objects of xpitch x-planes each, one after the other along
    // x.  Vertex coordinates are local to the object and its hash keys are shifted by obj * kstride, so every object
    // gets exactly the vertices, keys and first-appearance order it would get alone.  0 = one object.
    int xpitch;
    unsigned long long kstride;
    // x-slab of a larger grid (mrb_mcubes_slab): plane 0 of this grid is plane xorg of the whole grid; vertex coordinates
    // (and through them the hash keys) are those of the whole grid.  i + xorg is added as an integer BEFORE the conversion to
    // float, so the interpolation rounds exactly as in an unsharded run.
    int xorg;
};

// corner n of cell (i,j,k): (i + [n%4 in {1,2}], j + (n%4)/2, k + n/4)  (mcubes.cpp:166-168)
__device__ __forceinline__ void cell_corner_values(const float* __restrict__ F, const CellDims& cd, int i, int j, int k,
                                                   float v[8]) {
    const size_t sx = (size_t)cd.ny * cd.nz, sy = cd.nz;
    const float* p = F + ((size_t)i * cd.ny + j) * cd.nz + k;
    v[0] = __ldg(p);
    v[1] = __ldg(p + sx);
    v[2] = __ldg(p + sx + sy);
    v[3] = __ldg(p + sy);
    v[4] = __ldg(p + 1);
    v[5] = __ldg(p + sx + 1);
    v[6] = __ldg(p + sx + sy + 1);
    v[7] = __ldg(p + sy + 1);
}

__device__ __forceinline__ int cube_index(const float v[8], float iso) {
    int c = 0;
#pragma unroll
    for (int n = 0; n < 8; n++) c |= (v[n] < iso) ? (1 << n) : 0;
    return c;
}

struct ActiveCellPred {
    const float* F;
    CellDims cd;
    float iso;
    __device__ bool operator()(size_t cell) const {
        int k = (int)(cell % cd.cz);
        size_t t = cell / cd.cz;
        int j = (int)(t % cd.cy);
        int i = (int)(t / cd.cy);
        float v[8];
        cell_corner_values(F, cd, i, j, k, v);
        int c = cube_index(v, iso);
        return c != 0 && c != 255;     // edgeTable[c] != 0
    }
};

struct F3 { float x, y, z; };

__device__ __forceinline__ bool lex_less(const F3& a, const F3& b) {   // XYZ::operator<, hasher.h:41-58
    if (a.x < b.x) return true; else if (a.x > b.x) return false;
    if (a.y < b.y) return true; else if (a.y > b.y) return false;
    if (a.z < b.z) return true;
    return false;
}

// VertexInterp, mcubes.cpp:369-392 (fp32, IEEE division, no contraction).
__device__ __forceinline__ F3 vertex_interp(float iso, F3 p1, F3 p2, float v1, float v2) {
    if (lex_less(p2, p1)) { F3 t = p1; p1 = p2; p2 = t; float tv = v1; v1 = v2; v2 = tv; }
    F3 p;
    if ((double)fabsf(__fsub_rn(v1, v2)) > 0.00001) {
        float den = __fsub_rn(v2, v1), num = __fsub_rn(iso, v1);
        p.x = __fadd_rn(p1.x, __fmul_rn(__fdiv_rn(__fsub_rn(p2.x, p1.x), den), num));
        p.y = __fadd_rn(p1.y, __fmul_rn(__fdiv_rn(__fsub_rn(p2.y, p1.y), den), num));
        p.z = __fadd_rn(p1.z, __fmul_rn(__fdiv_rn(__fsub_rn(p2.z, p1.z), den), num));
    } else {
        p = p1;
    }
    return p;
}

// XYZ::dist < 0.00001, hasher.h:28-33 / mcubes.cpp:227-229
__device__ __forceinline__ bool too_close(const F3& l, const F3& r) {
    float dx = __fsub_rn(r.x, l.x), dy = __fsub_rn(r.y, l.y), dz = __fsub_rn(r.z, l.z);
    float d = __fmul_rn(dx, dx);
    d = __fadd_rn(d, __fmul_rn(dy, dy));
    d = __fadd_rn(d, __fmul_rn(dz, dz));
    return (double)__fsqrt_rn(d) < 0.00001;
}

// pointHashmap::hashXYZ, hasher.h:99-104
__device__ __forceinline__ unsigned long long hash_xyz(const F3& s, int ny, int nz) {
    unsigned long long val = (unsigned long long)__dmul_rn(__dmul_rn((double)s.x, 1000000000.0), (double)(ny * nz));
    val += (unsigned long long)__dmul_rn(__dmul_rn((double)s.y, 1000000.0), (double)nz);
    val += (unsigned long long)__dmul_rn((double)s.z, 1000.0);
    return val;
}

// Polygonise (mcubes.cpp:194-232) for one crossing cell.  Emit == false: returns the triangle count only.
template <bool Emit>
__device__ __forceinline__ int polygonise(const float* __restrict__ F, const CellDims& cd, int cell, float iso,
                                          float* __restrict__ soup, unsigned long long* __restrict__ keys, int first_tri) {
    int k = cell % cd.cz;
    int t = cell / cd.cz;
    int j = t % cd.cy;
    int i = t / cd.cy;
    float v[8];
    cell_corner_values(F, cd, i, j, k, v);
    int ci = cube_index(v, iso);
    uint64_t w = c_tri[ci];
    int ntri = (int)(w & 15);
    const int ob = cd.xpitch ? i / cd.xpitch : 0;
    const unsigned long long kofs = (unsigned long long)ob * cd.kstride;
    i -= ob * cd.xpitch;             // coordinates local to the object (the corner values were read above)
    i += cd.xorg;                    // ... or global, for an x-slab of a larger grid
    F3 vert[12];
#pragma unroll
    for (int e = 0; e < 12; e++) {
        int a = c_edge_a[e], b = c_edge_b[e];
        if (((ci >> a) & 1) != ((ci >> b) & 1)) {
            F3 pa = {(float)(i + (((a & 3) == 1) || ((a & 3) == 2))), (float)(j + ((a & 3) >> 1)), (float)(k + (a >> 2))};
            F3 pb = {(float)(i + (((b & 3) == 1) || ((b & 3) == 2))), (float)(j + ((b & 3) >> 1)), (float)(k + (b >> 2))};
            vert[e] = vertex_interp(iso, pa, pb, v[a], v[b]);
        } else {
            vert[e] = F3{-1.0f, -1.0f, -1.0f};
        }
    }
    int out = 0;
    for (int q = 0; q < ntri; q++) {
        F3 a = vert[(w >> (4 + 12 * q)) & 15];
        F3 b = vert[(w >> (8 + 12 * q)) & 15];
        F3 c = vert[(w >> (12 + 12 * q)) & 15];
        if (too_close(a, b)) continue;
        if (too_close(b, c)) continue;
        if (too_close(c, a)) continue;
        if (Emit) {
            size_t o = (size_t)(first_tri + out) * 9;
            soup[o + 0] = a.x; soup[o + 1] = a.y; soup[o + 2] = a.z;
            soup[o + 3] = b.x; soup[o + 4] = b.y; soup[o + 5] = b.z;
            soup[o + 6] = c.x; soup[o + 7] = c.y; soup[o + 8] = c.z;
            size_t kq = (size_t)(first_tri + out) * 3;
            keys[kq + 0] = hash_xyz(a, cd.ny, cd.nz) + kofs;
            keys[kq + 1] = hash_xyz(b, cd.ny, cd.nz) + kofs;
            keys[kq + 2] = hash_xyz(c, cd.ny, cd.nz) + kofs;
        }
        out++;
    }
    return out;
}

__global__ void __launch_bounds__(128) k_mc_count(const float* __restrict__ F, CellDims cd, const int* __restrict__ cells,
                                                  int ncells, float iso, int* __restrict__ ntri, int* __restrict__ seam_err) {
    int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= ncells) return;
    if (cd.xpitch) {                 // a crossing cell that spans two objects of a stacked batch: the caller falls back
        const int i = cells[a] / (cd.cz * cd.cy);
        if (i % cd.xpitch == cd.xpitch - 1) *seam_err = 1;
    }
    ntri[a] = polygonise<false>(F, cd, cells[a], iso, nullptr, nullptr, 0);
}

// ---- per-object offsets of a stacked batch ------------------------------------------------------------------------
__global__ void k_stack_face_off(const int* __restrict__ cells, int ncells, const int* __restrict__ tri_off, int ntri, CellDims cd,
                                 int nobj, int* __restrict__ face_off) {
    int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a > ncells) return;
    const int plane = cd.cz * cd.cy;
    const int prev = a > 0 ? (cells[a - 1] / plane) / cd.xpitch : -1;
    const int cur = a < ncells ? (cells[a] / plane) / cd.xpitch : nobj;
    const int off = a < ncells ? tri_off[a] : ntri;
    for (int o = prev + 1; o <= cur; o++) face_off[o] = off;       // objects without a crossing cell get an empty range
}
__global__ void k_stack_vert_off(const int* __restrict__ face_off, const int* __restrict__ vid, int ntri, int nverts, int nobj,
                                 int* __restrict__ vert_off) {
    int o = blockIdx.x * blockDim.x + threadIdx.x;
    if (o > nobj) return;
    const int f = face_off[o];
    vert_off[o] = f < ntri ? vid[3 * (size_t)f] : nverts;          // first-appearance ids grow with the corner position
}
__global__ void k_stack_localise(int* __restrict__ faces, int ntri, const int* __restrict__ face_off, const int* __restrict__ vert_off,
                                 int nobj) {
    int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= ntri) return;
    int lo = 0, hi = nobj;           // largest o with face_off[o] <= f
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (face_off[mid] <= f) lo = mid; else hi = mid;
    }
    const int v0 = vert_off[lo];
    faces[3 * (size_t)f] -= v0; faces[3 * (size_t)f + 1] -= v0; faces[3 * (size_t)f + 2] -= v0;
}

__global__ void __launch_bounds__(128) k_mc_emit(const float* __restrict__ F, CellDims cd, const int* __restrict__ cells,
                                                 int ncells, float iso, const int* __restrict__ tri_off,
                                                 float* __restrict__ soup, unsigned long long* __restrict__ keys) {
    int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= ncells) return;
    polygonise<true>(F, cd, cells[a], iso, soup, keys, tri_off[a]);
}

// ---- hash de-duplication ------------------------------------------------------------------------------
static const unsigned long long kEmptyKey = 0xffffffffffffffffull;

__device__ __forceinline__ uint32_t mix64(unsigned long long k) {
    k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull; k ^= k >> 33;
    return (uint32_t)k;
}

__global__ void k_hash_insert(const unsigned long long* __restrict__ keys, int ncorner, unsigned long long* tkeys,
                              unsigned* tfirst, uint32_t mask, int* __restrict__ slot_of) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= ncorner) return;
    unsigned long long key = keys[p];
    uint32_t s = mix64(key) & mask;
    while (true) {
        unsigned long long prev = atomicCAS(&tkeys[s], kEmptyKey, key);
        if (prev == kEmptyKey || prev == key) break;
        s = (s + 1) & mask;
    }
    atomicMin(&tfirst[s], (unsigned)p);           // table entries start at 0xffffffff (one memset clears keys and positions)
    slot_of[p] = (int)s;
}

__global__ void k_mark_first(const int* __restrict__ slot_of, const int* __restrict__ tfirst, int ncorner,
                             int* __restrict__ is_first) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= ncorner) return;
    is_first[p] = (tfirst[slot_of[p]] == p) ? 1 : 0;
}

__global__ void k_assign_ids(const int* __restrict__ slot_of, const int* __restrict__ tfirst, const int* __restrict__ vid,
                             const float* __restrict__ soup, int ncorner, int* __restrict__ faces, float* __restrict__ verts) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= ncorner) return;
    int fp = tfirst[slot_of[p]];
    int id = vid[fp];
    faces[p] = id;
    if (fp == p) {
        verts[(size_t)id * 3 + 0] = soup[(size_t)p * 3 + 0];
        verts[(size_t)id * 3 + 1] = soup[(size_t)p * 3 + 1];
        verts[(size_t)id * 3 + 2] = soup[(size_t)p * 3 + 2];
    }
}

// keys -> (vertex ids per corner, unique vertices).  Scratch comes from the ctx arena.
int dev_dedup(mrb_ctx* ctx, const float* soup, const unsigned long long* keys, int ncorner, int* faces, float** verts_out,
              int* nverts_out, int** vid_out) {
    size_t cap = 1024;
    while (cap < (size_t)ncorner * 2) cap <<= 1;
    unsigned long long* tkeys = nullptr;
    int *tfirst = nullptr, *slot_of = nullptr, *is_first = nullptr;
    MRB_TRY(arena_alloc(ctx, cap, &tkeys));
    MRB_TRY(arena_alloc(ctx, cap, &tfirst));
    MRB_TRY(arena_alloc(ctx, (size_t)ncorner, &slot_of));
    MRB_TRY(arena_alloc(ctx, (size_t)ncorner, &is_first));
    // empty key = all ones, "no position yet" = 0xffffffff as unsigned: two memsets instead of a clearing kernel
    MRB_CUDA(cudaMemsetAsync(tkeys, 0xFF, cap * sizeof(unsigned long long), ctx->stream));
    MRB_CUDA(cudaMemsetAsync(tfirst, 0xFF, cap * sizeof(int), ctx->stream));
    MRB_LAUNCH(ctx, k_hash_insert, cdiv(ncorner, 256), 256, 0, keys, ncorner, tkeys, (unsigned*)tfirst, (uint32_t)(cap - 1), slot_of);
    MRB_LAUNCH(ctx, k_mark_first, cdiv(ncorner, 256), 256, 0, slot_of, tfirst, ncorner, is_first);
    MRB_TRY(scan_exclusive(ctx, is_first, is_first, (size_t)ncorner, ctx->d_mail));
    int nverts = 0;
    MRB_TRY(read_int(ctx, ctx->d_mail, &nverts));
    float* verts = nullptr;
    MRB_TRY(arena_alloc(ctx, (size_t)(nverts > 0 ? nverts : 1) * 3, &verts));
    MRB_LAUNCH(ctx, k_assign_ids, cdiv(ncorner, 256), 256, 0, slot_of, tfirst, is_first, soup, ncorner, faces, verts);
    *verts_out = verts;
    *nverts_out = nverts;
    if (vid_out) *vid_out = is_first;       // exclusive scan of the first-appearance marks = vertex id per first corner
    return MRB_OK;
}

// ---- bit-plane classification -------------------------------------------------------------------------------
// "val < isolevel" (mcubes.cpp:199-207) is one bit per voxel; a cell is surface-crossing iff its 8 corner bits
// are not all equal (edgeTable[cubeindex] != 0).  F is read once to pack the bits (32 z per word); the cell
// test then runs on words: OR and AND over the 2x2 (x,y) rows and the z / z+1 bit pair, 32 cells at a time.
// popc of the active-cell words + exclusive scan gives the (i,j,k)-ordered list of crossing cells.
__global__ void __launch_bounds__(256) k_sign_bits(const float* __restrict__ F, Dims d, int nzw, float iso,
                                                   uint32_t* __restrict__ sb) {
    const int lane = threadIdx.x & 31;
    const size_t rows = (size_t)d.nx * d.ny;
    for (size_t row = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < rows; row += (size_t)gridDim.x * (blockDim.x >> 5)) {
        const float* p = F + row * d.nz;
        for (int w = 0; w < nzw; w++) {
            int z = w * 32 + lane;
            bool below = z < d.nz && __ldg(p + z) < iso;
            unsigned m = __ballot_sync(0xffffffffu, below);
            if (lane == 0) sb[row * nzw + w] = m;
        }
    }
}

// one thread per (i, j, zw): active-cell word for cells k = 32 zw .. 32 zw + 31 and its popcount
__global__ void __launch_bounds__(256) k_cell_words(const uint32_t* __restrict__ sb, CellDims cd, int nzw,
                                                    uint32_t* __restrict__ aw, int* __restrict__ cnt) {
    const size_t n = (size_t)cd.cx * cd.cy * nzw;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (size_t)gridDim.x * blockDim.x) {
        const int w = (int)(idx % nzw);
        const size_t t = idx / nzw;
        const int j = (int)(t % cd.cy), i = (int)(t / cd.cy);
        uint32_t any = 0, all = 0xffffffffu;
        for (int a = 0; a < 2; a++)
            for (int b = 0; b < 2; b++) {
                const uint32_t* row = sb + ((size_t)(i + a) * cd.ny + (j + b)) * nzw;
                const uint32_t lo = __ldg(row + w);
                const uint32_t nx = w + 1 < nzw ? __ldg(row + w + 1) : 0u;
                const uint32_t hi = (lo >> 1) | (nx << 31);          // bit k+1
                any |= lo | hi;
                all &= lo & hi;
            }
        const int nvalid = min(32, cd.cz - 32 * w);                   // cells k <= nz - 2
        const uint32_t valid = nvalid >= 32 ? 0xffffffffu : (nvalid > 0 ? ((1u << nvalid) - 1u) : 0u);
        const uint32_t m = any & ~all & valid;
        aw[idx] = m;
        cnt[idx] = __popc(m);
    }
}

__global__ void __launch_bounds__(256) k_cell_list(const uint32_t* __restrict__ aw, const int* __restrict__ off, CellDims cd,
                                                   int nzw, int* __restrict__ cells) {
    const size_t n = (size_t)cd.cx * cd.cy * nzw;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (size_t)gridDim.x * blockDim.x) {
        uint32_t m = aw[idx];
        if (!m) continue;
        const int w = (int)(idx % nzw);
        const size_t col = idx / nzw;                                  // i * cy + j
        int o = off[idx];
        while (m) {
            int bit = __ffs(m) - 1;
            m &= m - 1;
            cells[o++] = (int)(col * cd.cz + (size_t)(32 * w + bit));
        }
    }
}

int classify_cells(mrb_ctx* ctx, const float* F, Dims d, float iso, int** cells_out, int* ncells_out) {
    const CellDims cd{d.nx, d.ny, d.nz, d.nx - 1, d.ny - 1, d.nz - 1};
    const int nzw = (d.nz + 31) / 32;
    const size_t nwords = (size_t)d.nx * d.ny * nzw, ncw = (size_t)cd.cx * cd.cy * nzw;
    uint32_t *sb = nullptr, *aw = nullptr;
    int* cnt = nullptr;
    const bool have_bits = ctx->sign_bits && ctx->sign_F == F && ctx->sign_iso == iso && d.nz % 32 == 0;
    if (have_bits) sb = ctx->sign_bits;         // packed by the kernel that wrote F (mrb_reconstruct)
    else MRB_TRY(arena_alloc(ctx, nwords, &sb));
    ctx->sign_bits = nullptr;
    MRB_TRY(arena_alloc(ctx, ncw, &aw));
    MRB_TRY(arena_alloc(ctx, ncw, &cnt));
    const size_t cap = (size_t)ctx->sm_count * 32;
    size_t want = ((size_t)d.nx * d.ny + 7) / 8;
    if (!have_bits) MRB_LAUNCH(ctx, k_sign_bits, (unsigned)(want < cap ? want : cap), 256, 0, F, d, nzw, iso, sb);
    want = (ncw + 255) / 256;
    MRB_LAUNCH(ctx, k_cell_words, (unsigned)(want < cap ? want : cap), 256, 0, sb, cd, nzw, aw, cnt);
    MRB_TRY(scan_exclusive(ctx, cnt, cnt, ncw, ctx->d_mail));
    int total = 0;
    MRB_TRY(read_int(ctx, ctx->d_mail, &total));
    int* cells = nullptr;
    MRB_TRY(arena_alloc(ctx, (size_t)(total > 0 ? total : 1), &cells));
    if (total > 0) MRB_LAUNCH(ctx, k_cell_list, (unsigned)(want < cap ? want : cap), 256, 0, aw, cnt, cd, nzw, cells);
    *cells_out = cells;
    *ncells_out = total;
    return MRB_OK;
}

int dev_mcubes(mrb_ctx* ctx, const float* F, Dims d, float isolevel, float** verts_dev, int** faces_dev, int* nverts,
               int* nfaces, const McStack* stack, int xorg) {
    *verts_dev = nullptr; *faces_dev = nullptr; *nverts = 0; *nfaces = 0;
    if (stack) {
        for (int o = 0; o <= stack->nobj; o++) { stack->face_off_host[o] = 0; stack->vert_off_host[o] = 0; }
        *stack->seam_err_host = 0;
    }
    if (d.nx < 2 || d.ny < 2 || d.nz < 2) return MRB_OK;
    CellDims cd{d.nx, d.ny, d.nz, d.nx - 1, d.ny - 1, d.nz - 1, 0, 0ull, xorg};
    if (stack) { cd.xpitch = stack->xpitch; cd.kstride = stack->kstride; }
    int* cells = nullptr;
    int ncells = 0;
    MRB_TRY(classify_cells(ctx, F, d, isolevel, &cells, &ncells));
    if (ncells == 0) return MRB_OK;
    int* tri_off = nullptr;
    MRB_TRY(arena_alloc(ctx, (size_t)ncells, &tri_off));
    int* seam = ctx->d_mail + 54;
    MRB_CUDA(cudaMemsetAsync(seam, 0, sizeof(int), ctx->stream));
    MRB_LAUNCH(ctx, k_mc_count, cdiv(ncells, 128), 128, 0, F, cd, cells, ncells, isolevel, tri_off, seam);
    MRB_TRY(scan_exclusive(ctx, tri_off, tri_off, (size_t)ncells, ctx->d_mail));
    MRB_CUDA(cudaMemcpyAsync(ctx->h_mail + 54, seam, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    int ntri = 0;
    MRB_TRY(read_int(ctx, ctx->d_mail, &ntri));
    if (stack && ctx->h_mail[54]) { *stack->seam_err_host = 1; return MRB_OK; }
    if (ntri == 0) return MRB_OK;
    float* soup = nullptr;
    unsigned long long* keys = nullptr;
    int* faces = nullptr;
    MRB_TRY(arena_alloc(ctx, (size_t)ntri * 9, &soup));
    MRB_TRY(arena_alloc(ctx, (size_t)ntri * 3, &keys));
    MRB_TRY(arena_alloc(ctx, (size_t)ntri * 3, &faces));
    MRB_LAUNCH(ctx, k_mc_emit, cdiv(ncells, 128), 128, 0, F, cd, cells, ncells, isolevel, tri_off, soup, keys);
    int* vid = nullptr;
    MRB_TRY(dev_dedup(ctx, soup, keys, ntri * 3, faces, verts_dev, nverts, &vid));
    *faces_dev = faces;
    *nfaces = ntri;
    if (stack) {
        int *foff = nullptr, *voff = nullptr;
        const int n1 = stack->nobj + 1;
        MRB_TRY(arena_alloc(ctx, (size_t)n1, &foff));
        MRB_TRY(arena_alloc(ctx, (size_t)n1, &voff));
        MRB_LAUNCH(ctx, k_stack_face_off, cdiv(ncells + 1, 256), 256, 0, cells, ncells, tri_off, ntri, cd, stack->nobj, foff);
        MRB_LAUNCH(ctx, k_stack_vert_off, cdiv(n1, 256), 256, 0, foff, vid, ntri, *nverts, stack->nobj, voff);
        MRB_LAUNCH(ctx, k_stack_localise, cdiv(ntri, 256), 256, 0, faces, ntri, foff, voff, stack->nobj);
        MRB_CUDA(cudaMemcpyAsync(stack->face_off_host, foff, sizeof(int) * n1, cudaMemcpyDeviceToHost, ctx->stream));
        MRB_CUDA(cudaMemcpyAsync(stack->vert_off_host, voff, sizeof(int) * n1, cudaMemcpyDeviceToHost, ctx->stream));
        MRB_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return MRB_OK;
}

__global__ void k_hash_points(const float* __restrict__ xyz, int n, int ny, int nz, unsigned long long* __restrict__ keys) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    F3 p = {xyz[3 * (size_t)i], xyz[3 * (size_t)i + 1], xyz[3 * (size_t)i + 2]};
    keys[i] = hash_xyz(p, ny, nz);
}

// Copies a device-resident mesh into the caller-visible mrb_mesh (host malloc or device arena pointers).
int export_mesh(mrb_ctx* ctx, float* verts_dev, int* faces_dev, int nverts, int nfaces, int device_out, mrb_mesh* out) {
    out->nverts = nverts;
    out->nfaces = nfaces;
    out->on_device = device_out == 2 ? 2 : (device_out ? 1 : 0);
    if (device_out && device_out != 2) {
        out->verts = verts_dev;
        out->faces = faces_dev;
        return sync(ctx);
    }
    if (device_out == 2) {
        // the context's pinned staging itself: full PCIe rate and no second host copy
        const size_t vb2 = (sizeof(float) * 3 * (size_t)nverts + 255) / 256 * 256, fb2 = sizeof(int) * 3 * (size_t)nfaces;
        char* pin2 = nullptr;
        MRB_TRY(pinned_reserve(ctx, vb2 + fb2 + 256, &pin2));
        if (nverts > 0) MRB_CUDA(cudaMemcpyAsync(pin2, verts_dev, sizeof(float) * 3 * (size_t)nverts, cudaMemcpyDeviceToHost, ctx->stream));
        if (nfaces > 0) MRB_CUDA(cudaMemcpyAsync(pin2 + vb2, faces_dev, fb2, cudaMemcpyDeviceToHost, ctx->stream));
        out->verts = (float*)pin2;
        out->faces = (int*)(pin2 + vb2);
        return sync(ctx);
    }
    out->verts = (float*)malloc(sizeof(float) * 3 * (size_t)(nverts > 0 ? nverts : 1));
    out->faces = (int*)malloc(sizeof(int) * 3 * (size_t)(nfaces > 0 ? nfaces : 1));
    if (!out->verts || !out->faces) { set_error("export_mesh: host malloc failed"); return MRB_ENOMEM; }
    // device -> pinned staging (full PCIe rate) -> the caller's malloc'd arrays
    const size_t vb = sizeof(float) * 3 * (size_t)nverts, fb = sizeof(int) * 3 * (size_t)nfaces;
    char* pin = nullptr;
    MRB_TRY(pinned_reserve(ctx, vb + fb, &pin));
    if (nverts > 0) MRB_CUDA(cudaMemcpyAsync(pin, verts_dev, vb, cudaMemcpyDeviceToHost, ctx->stream));
    if (nfaces > 0) MRB_CUDA(cudaMemcpyAsync(pin + vb, faces_dev, fb, cudaMemcpyDeviceToHost, ctx->stream));
    MRB_TRY(sync(ctx));
    if (nverts > 0) memcpy(out->verts, pin, vb);
    if (nfaces > 0) memcpy(out->faces, pin + vb, fb);
    return MRB_OK;
}

}  // namespace mrb

using namespace mrb;

extern "C" {

int mrb_mcubes(mrb_ctx* ctx, const float* F, const int dims[3], float isolevel, int device_out, mrb_mesh* out) {
    MRB_REQUIRE(ctx && F && dims && out, "NULL argument");
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    memset(out, 0, sizeof(*out));
    Dims d = mkdims(dims);
    const float* dF = nullptr;
    MRB_TRY(stage_in(ctx, F, d.count(), &dF));
    float* verts = nullptr;
    int* faces = nullptr;
    int nv = 0, nf = 0;
    MRB_TRY(dev_mcubes(ctx, dF, d, isolevel, &verts, &faces, &nv, &nf));
    return export_mesh(ctx, verts, faces, nv, nf, device_out, out);
}

int mrb_mcubes_slab(mrb_ctx* ctx, const float* F_slab, const int dims[3], int x_origin, float isolevel, int device_out,
                    mrb_mesh* out) {
    MRB_REQUIRE(ctx && F_slab && dims && out && x_origin >= 0, "bad argument");
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    memset(out, 0, sizeof(*out));
    Dims d = mkdims(dims);
    const float* dF = nullptr;
    MRB_TRY(stage_in(ctx, F_slab, d.count(), &dF));
    float* verts = nullptr;
    int* faces = nullptr;
    int nv = 0, nf = 0;
    MRB_TRY(dev_mcubes(ctx, dF, d, isolevel, &verts, &faces, &nv, &nf, nullptr, x_origin));
    return export_mesh(ctx, verts, faces, nv, nf, device_out, out);
}

void mrb_mesh_free(mrb_mesh* mesh) {
    if (!mesh) return;
    if (mesh->on_device == 0) { free(mesh->verts); free(mesh->faces); }     // 1, 2: owned by the context
    mesh->verts = nullptr; mesh->faces = nullptr; mesh->nverts = 0; mesh->nfaces = 0;
}

int mrb_hash_points(mrb_ctx* ctx, const float* xyz, int n, const int dims[3], uint64_t* keys_out) {
    MRB_REQUIRE(ctx && dims && (n == 0 || (xyz && keys_out)) && n >= 0, "bad argument");
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    if (n == 0) return MRB_OK;
    const float* dx = nullptr;
    MRB_TRY(stage_in(ctx, xyz, (size_t)n * 3, &dx));
    Out<unsigned long long> o;
    MRB_TRY(o.init(ctx, (unsigned long long*)keys_out, (size_t)n));
    MRB_LAUNCH(ctx, k_hash_points, cdiv(n, 256), 256, 0, dx, n, dims[1], dims[2], o.dev);
    MRB_TRY(o.commit(ctx));
    return sync(ctx);
}

}  // extern "C"

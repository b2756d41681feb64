// mrb_compact.cuh -- order-preserving stream compaction (warp ballot + prefix scan).
//
// Pred is a device functor  bool operator()(size_t idx) const.  Indexes with pred true are written
// to pos_out in ascending order.  Tile = 1024 consecutive indexes per 256-thread block; inside a
// tile the order is recovered from warp ballots (popc of the lower lanes) and the per-warp counts,
// across tiles from an exclusive scan of the tile counts -- no atomics, deterministic order.
#pragma once
#include "mrb_internal.cuh"

namespace mrb {

static const int kCompactTile = 1024;

template <typename Pred>
__global__ void __launch_bounds__(256) k_compact_count(Pred pred, size_t n, int* __restrict__ counts) {
    size_t base = (size_t)blockIdx.x * kCompactTile;
    int c = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        size_t idx = base + i * 256 + threadIdx.x;
        bool f = idx < n && pred(idx);
        c += __popc(__ballot_sync(0xffffffffu, f));
    }
    __shared__ int s;
    if (threadIdx.x == 0) s = 0;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) atomicAdd(&s, c);   // integer add: order independent
    __syncthreads();
    if (threadIdx.x == 0) counts[blockIdx.x] = s;
}

template <typename Pred>
__global__ void __launch_bounds__(256) k_compact_write(Pred pred, size_t n, const int* __restrict__ offsets,
                                                       int* __restrict__ pos_out) {
    size_t base = (size_t)blockIdx.x * kCompactTile;
    __shared__ int wcount[32];     // segment s = i*8 + wid
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned bal[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        size_t idx = base + i * 256 + threadIdx.x;
        bool f = idx < n && pred(idx);
        bal[i] = __ballot_sync(0xffffffffu, f);
        if (lane == 0) wcount[i * 8 + wid] = __popc(bal[i]);
    }
    __syncthreads();
    if (wid == 0) {   // exclusive scan of the 32 segment counts
        int v = wcount[lane], incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        wcount[lane] = incl - v;
    }
    __syncthreads();
    const int run = offsets[blockIdx.x];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        if (bal[i] & (1u << lane)) {
            int r = __popc(bal[i] & ((1u << lane) - 1u));
            pos_out[run + wcount[i * 8 + wid] + r] = (int)(base + i * 256 + threadIdx.x);
        }
    }
}

// Compacts [0,n): returns the arena-allocated list and its length (synchronises the stream once).
template <typename Pred>
int compact(mrb_ctx* ctx, Pred pred, size_t n, int** pos_out, int* n_out) {
    unsigned ntiles = cdiv(n, kCompactTile);
    if (ntiles == 0) { *pos_out = nullptr; *n_out = 0; return MRB_OK; }
    int* counts = nullptr;
    MRB_TRY(arena_alloc(ctx, ntiles, &counts));
    MRB_LAUNCH(ctx, (k_compact_count<Pred>), ntiles, 256, 0, pred, n, counts);
    MRB_TRY(scan_exclusive(ctx, counts, counts, ntiles, ctx->d_mail));
    int total = 0;
    MRB_TRY(read_int(ctx, ctx->d_mail, &total));
    int* pos = nullptr;
    MRB_TRY(arena_alloc(ctx, (size_t)(total > 0 ? total : 1), &pos));
    if (total > 0) MRB_LAUNCH(ctx, (k_compact_write<Pred>), ntiles, 256, 0, pred, n, counts, pos);
    *pos_out = pos;
    *n_out = total;
    return MRB_OK;
}

}  // namespace mrb

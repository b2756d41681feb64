// mrb_ctx.cu -- context, device arena, error reporting, device-wide scan.
#include "mrb_internal.cuh"

#include <stdarg.h>
#include <string.h>

namespace mrb {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

bool is_device_ptr(const void* p) {
    cudaPointerAttributes a;
    cudaError_t e = cudaPointerGetAttributes(&a, p);
    if (e != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// ---- arena ----------------------------------------------------------------------------------
static const size_t kAlign = 256;

int Arena::alloc(size_t bytes, void** out) {
    bytes = (bytes + kAlign - 1) / kAlign * kAlign;
    if (bytes == 0) bytes = kAlign;
    // Allocations of at least 1 MiB start on a 2 MiB boundary (the device page size): measured on B200, the sweep kernel
    // is 4 % faster with its iterate buffers placed like that than at the 256-byte granularity (2.79 -> 2.68 ms at 256^3).
    // MRB_ARENA_BIG_ALIGN=<KiB> overrides (0 = off).
    static const size_t big_align = getenv("MRB_ARENA_BIG_ALIGN") ? (size_t)atoi(getenv("MRB_ARENA_BIG_ALIGN")) * 1024 : ((size_t)2 << 20);
    if (!blocks.empty()) {
        Block& b = blocks.back();
        size_t start = b.used;
        if (big_align && bytes >= ((size_t)1 << 20)) start = (start + big_align - 1) / big_align * big_align;
        if (start + bytes <= b.cap) {
            high_water += start - b.used + bytes;
            *out = b.base + start;
            b.used = start + bytes;
            return MRB_OK;
        }
    }
    high_water += bytes + big_align;
    size_t cap = bytes;
    size_t prev = blocks.empty() ? 0 : blocks.back().cap;
    if (cap < prev) cap = prev;                 // grow geometrically-ish
    if (cap < (size_t)(16u << 20)) cap = 16u << 20;
    char* base = nullptr;
    cudaError_t e = cudaMalloc((void**)&base, cap);
    if (e != cudaSuccess) {
        cudaGetLastError();
        set_error("arena: cudaMalloc(%zu) failed: %s", cap, cudaGetErrorString(e));
        return MRB_ENOMEM;
    }
    blocks.push_back(Block{base, cap, bytes});
    *out = base;
    return MRB_OK;
}

int Arena::reset() {
    if (blocks.size() > 1) {
        // Consolidate: the previous call needed several blocks; replace them by one big block so
        // that steady-state calls allocate nothing.
        size_t need = high_water + (high_water >> 3);
        for (auto& b : blocks) cudaFree(b.base);
        blocks.clear();
        char* base = nullptr;
        cudaError_t e = cudaMalloc((void**)&base, need);
        if (e != cudaSuccess) {
            cudaGetLastError();
            set_error("arena: consolidate cudaMalloc(%zu) failed: %s", need, cudaGetErrorString(e));
            return MRB_ENOMEM;
        }
        blocks.push_back(Block{base, need, 0});
    }
    for (auto& b : blocks) b.used = 0;
    high_water = 0;
    return MRB_OK;
}

void Arena::release() {
    for (auto& b : blocks) cudaFree(b.base);
    blocks.clear();
}

int side_reserve(mrb_ctx* ctx, int which, size_t bytes, char** out) {
    if (bytes > ctx->side_cap[which]) {
        if (ctx->side[which]) cudaFree(ctx->side[which]);
        ctx->side[which] = nullptr;
        ctx->side_cap[which] = 0;
        const size_t cap = bytes + bytes / 8;
        cudaError_t e = cudaMalloc((void**)&ctx->side[which], cap);
        if (e != cudaSuccess) { cudaGetLastError(); set_error("side buffer: cudaMalloc(%zu) failed", cap); return MRB_ENOMEM; }
        ctx->side_cap[which] = cap;
    }
    *out = ctx->side[which];
    return MRB_OK;
}

int pinned_reserve(mrb_ctx* ctx, size_t bytes, char** out) {
    if (bytes > ctx->h_pinned_cap) {
        if (ctx->h_pinned) cudaFreeHost(ctx->h_pinned);
        ctx->h_pinned = nullptr;
        size_t cap = bytes + (bytes >> 2) + (1u << 20);
        cudaError_t e = cudaMallocHost((void**)&ctx->h_pinned, cap);
        if (e != cudaSuccess) { cudaGetLastError(); ctx->h_pinned_cap = 0; set_error("pinned staging: cudaMallocHost(%zu) failed", cap); return MRB_ENOMEM; }
        ctx->h_pinned_cap = cap;
    }
    *out = ctx->h_pinned;
    return MRB_OK;
}

int read_int(mrb_ctx* ctx, const int* dev, int* host) {
    MRB_CUDA(cudaMemcpyAsync(ctx->h_mail, dev, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    MRB_CUDA(cudaStreamSynchronize(ctx->stream));
    *host = ctx->h_mail[0];
    return MRB_OK;
}

// ---- exclusive scan ---------------------------------------------------------------------------
// Three small kernels: per-block reduce, single-block scan of the partials, per-block downsweep.
// Block = 256 threads x 8 items = 2048 items.
static const int kScanThreads = 256;
static const int kScanItems = 8;
static const int kScanTile = kScanThreads * kScanItems;

__device__ __forceinline__ int warp_incl_scan(int v) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

// Block-wide exclusive scan of one value per thread (blockDim.x <= 1024); returns the exclusive
// prefix, *total receives the block sum.
__device__ __forceinline__ int block_excl_scan(int v, int* total) {
    __shared__ int wsum[32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int incl = warp_incl_scan(v);
    if (lane == 31) wsum[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int nw = (blockDim.x + 31) >> 5;
        int s = lane < nw ? wsum[lane] : 0;
        int si = warp_incl_scan(s);
        wsum[lane] = si - s;
        if (lane == nw - 1) *total = si;
    }
    __syncthreads();
    int res = incl - v + wsum[wid];
    __syncthreads();
    return res;
}

__global__ void k_scan_reduce(const int* __restrict__ in, int* __restrict__ partial, size_t n) {
    size_t base = (size_t)blockIdx.x * kScanTile;
    int s = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        size_t idx = base + (size_t)i * kScanThreads + threadIdx.x;
        if (idx < n) s += in[idx];
    }
    __shared__ int tot;
    block_excl_scan(s, &tot);
    if (threadIdx.x == 0) partial[blockIdx.x] = tot;
}

__global__ void k_scan_partials(int* __restrict__ partial, int nblocks, int* __restrict__ total_out) {
    __shared__ int tot;
    __shared__ int carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < nblocks; base += blockDim.x) {
        int idx = base + threadIdx.x;
        int v = idx < nblocks ? partial[idx] : 0;
        int ex = block_excl_scan(v, &tot);
        int carry = carry_s;
        if (idx < nblocks) partial[idx] = ex + carry;
        __syncthreads();
        if (threadIdx.x == 0) carry_s = carry + tot;
        __syncthreads();
    }
    if (threadIdx.x == 0 && total_out) *total_out = carry_s;
}

__global__ void k_scan_down(const int* __restrict__ in, int* __restrict__ out, const int* __restrict__ partial,
                            size_t n) {
    // items are laid out blocked per thread so the per-thread running sum is sequential in index
    size_t base = (size_t)blockIdx.x * kScanTile + (size_t)threadIdx.x * kScanItems;
    int v[kScanItems];
    int s = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        size_t idx = base + i;
        v[i] = idx < n ? in[idx] : 0;
        s += v[i];
    }
    __shared__ int tot;
    int ex = block_excl_scan(s, &tot) + partial[blockIdx.x];
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        size_t idx = base + i;
        if (idx < n) out[idx] = ex;
        ex += v[i];
    }
}

// Single-pass scan (decoupled look-back): tiles are handed out by a ticket counter (a tile's predecessors are running or done,
// so waiting for them cannot deadlock); a tile publishes its aggregate, then looks back over its predecessors' words until it
// meets one that already carries an inclusive prefix, and publishes its own.  A word is 64 bits -- status (1 aggregate, 2
// inclusive prefix) above the 32-bit value -- written and read with one relaxed 64-bit access, so value and status always
// arrive together and no fence is needed.  `state` (one word per tile) and `ticket` are zero at launch.
__global__ void __launch_bounds__(kScanThreads) k_scan_lookback(const int* in, int* out, size_t n,
                                                                unsigned long long* state, unsigned* ticket,
                                                                int* __restrict__ total_out, unsigned ntiles) {
    __shared__ unsigned s_tile;
    __shared__ int s_prefix;
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const unsigned tile = s_tile;
    const size_t base = (size_t)tile * kScanTile + (size_t)threadIdx.x * kScanItems;
    int v[kScanItems];
    int s = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        const size_t idx = base + i;
        v[i] = idx < n ? in[idx] : 0;
        s += v[i];
    }
    __shared__ int tot;
    const int ex = block_excl_scan(s, &tot);
    auto put = [&](unsigned t, unsigned status, int value) {
        const unsigned long long w = ((unsigned long long)status << 32) | (unsigned)value;
        asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(state + t), "l"(w) : "memory");
    };
    if (threadIdx.x == 0) {
        if (tile == 0) { put(0, 2u, tot); s_prefix = 0; }
        else put(tile, 1u, tot);
    }
    if (tile > 0 && threadIdx.x < 32) {
        int prefix = 0;
        int j = (int)tile - 1;                      // lane L looks at tile j - L
        for (;;) {
            const int t = j - (int)threadIdx.x;
            unsigned long long w = 0;
            unsigned spins = 0;
            if (t >= 0) {
                do {
                    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w) : "l"(state + t) : "memory");
                    if (++spins == (1u << 26)) __trap();
                } while ((w >> 32) == 0);
            }
            const unsigned has_prefix = __ballot_sync(0xffffffffu, t >= 0 && (w >> 32) == 2u);
            const int first = has_prefix ? __ffs(has_prefix) - 1 : 32;        // nearest predecessor with an inclusive prefix
            int val = (t >= 0 && (int)threadIdx.x <= first) ? (int)(unsigned)w : 0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(0xffffffffu, val, o);
            prefix += val;
            if (has_prefix || j - 32 < 0) break;
            j -= 32;
        }
        if (threadIdx.x == 0) { put(tile, 2u, prefix + tot); s_prefix = prefix; }
    }
    __syncthreads();
    int run = ex + s_prefix;
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        const size_t idx = base + i;
        if (idx < n) out[idx] = run;
        run += v[i];
    }
    if (total_out && tile == ntiles - 1 && threadIdx.x == 0) *total_out = s_prefix + tot;
}

int scan_exclusive(mrb_ctx* ctx, const int* in, int* out, size_t n, int* total_dev) {
    if (n == 0) {
        if (total_dev) MRB_CUDA(cudaMemsetAsync(total_dev, 0, sizeof(int), ctx->stream));
        return MRB_OK;
    }
    unsigned nblocks = cdiv(n, kScanTile);
    if (getenv("MRB_SCAN_THREE_PASS")) {            // the three-kernel scan (reduce, scan of the partials, down-sweep)
        int* partial = nullptr;
        MRB_TRY(arena_alloc(ctx, nblocks, &partial));
        MRB_LAUNCH(ctx, k_scan_reduce, nblocks, kScanThreads, 0, in, partial, n);
        MRB_LAUNCH(ctx, k_scan_partials, 1, 1024, 0, partial, (int)nblocks, total_dev);
        MRB_LAUNCH(ctx, k_scan_down, nblocks, kScanThreads, 0, in, out, partial, n);
        return MRB_OK;
    }
    unsigned long long* state = nullptr;
    MRB_TRY(arena_alloc(ctx, (size_t)nblocks + 1, &state));      // last word: the ticket counter
    MRB_CUDA(cudaMemsetAsync(state, 0, ((size_t)nblocks + 1) * sizeof(unsigned long long), ctx->stream));
    MRB_LAUNCH(ctx, k_scan_lookback, nblocks, kScanThreads, 0, in, out, n, state, (unsigned*)(state + nblocks), total_dev, nblocks);
    return MRB_OK;
}

}  // namespace mrb

// ---- C-ABI: context ---------------------------------------------------------------------------
extern "C" {

const char* mrb_last_error(void) { return mrb::g_err; }
const char* mrb_version(void) { return "mrb 0.1 sm_100a"; }

int mrb_ctx_create(int device, mrb_ctx** out) {
    if (!out) { mrb::set_error("mrb_ctx_create: out is NULL"); return MRB_EINVAL; }
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        mrb::set_error("mrb_ctx_create: no CUDA device available (%s); libmrb has no CPU fallback",
                       e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
        return MRB_ECUDA;
    }
    if (device < 0 || device >= ndev) { mrb::set_error("mrb_ctx_create: bad device %d", device); return MRB_EINVAL; }
    MRB_CUDA(cudaSetDevice(device));
    mrb_ctx* c = new mrb_ctx();
    c->device = device;
    cudaDeviceProp prop;
    MRB_CUDA(cudaGetDeviceProperties(&prop, device));
    c->sm_count = prop.multiProcessorCount;
    MRB_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    for (int i = 0; i < 8; i++) MRB_CUDA(cudaEventCreate(&c->ev[i]));
    MRB_CUDA(cudaMallocHost((void**)&c->h_mail, 64 * sizeof(int)));
    MRB_CUDA(cudaMalloc((void**)&c->d_mail, 64 * sizeof(int)));
    MRB_CUDA(cudaMalloc((void**)&c->d_bar, 64 * sizeof(int)));
    *out = c;
    return MRB_OK;
}

void mrb_ctx_destroy(mrb_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    ctx->arena.release();
    for (auto& a : ctx->plan_arenas) a.release();
    for (int i = 0; i < 8; i++) if (ctx->ev[i]) cudaEventDestroy(ctx->ev[i]);
    if (ctx->h_mail) cudaFreeHost(ctx->h_mail);
    if (ctx->d_mail) cudaFree(ctx->d_mail);
    if (ctx->d_bar) cudaFree(ctx->d_bar);
    if (ctx->h_pinned) cudaFreeHost(ctx->h_pinned);
    for (int w = 0; w < 2; w++) if (ctx->side[w]) cudaFree(ctx->side[w]);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

void* mrb_ctx_stream(mrb_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
uint64_t mrb_ctx_launch_count(mrb_ctx* ctx) { return ctx ? ctx->launches : 0; }

int mrb_copy(mrb_ctx* ctx, void* dst, const void* src, size_t nbytes) {
    if (!ctx || (nbytes && (!dst || !src))) { mrb::set_error("mrb_copy: NULL argument"); return MRB_EINVAL; }
    if (nbytes == 0) return MRB_OK;
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_CUDA(cudaMemcpyAsync(dst, src, nbytes, cudaMemcpyDefault, ctx->stream));
    MRB_CUDA(cudaStreamSynchronize(ctx->stream));
    return MRB_OK;
}

void mrb_options_default(mrb_options* o) {
    if (!o) return;
    o->using_features = 0;
    o->band_size = MRB_BAND_SIZE_DEFAULT;
    o->iters = MRB_QP_ITERS_DEFAULT;
    o->isolevel = 0.0f;
    o->feature_threshold = 0.75f;
    o->corner_threshold = 0.8f;
}

}  // extern "C"

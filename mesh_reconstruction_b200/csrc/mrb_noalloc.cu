// mrb_noalloc.cu -- mrb_find_indexes_noalloc: ordered band compaction that never allocates.
//
// Who this is for.  Some programs that embed this library must not let it touch the CUDA
// allocator at all: a host framework that has already handed the whole card to its own pool
// (PyTorch's caching allocator, a robot's fixed perception arena), where a stray cudaMalloc can fail
// or fragment the pool; a control loop with a hard deadline, where cudaMalloc / cudaFree stall the
// device for an unbounded time; a process running under a fixed memory quota.  Every other entry
// point of libmrb draws scratch from a context arena that grows with cudaMalloc.  This one does
// not: its scratch (one counter per 1024-voxel tile plus a total) lives in a fixed array inside the
// library image, so the call performs no cudaMalloc, no cudaFree and no other allocating routine.
// The caller provides the input grid and the output list, both already on the device.
//
// The price: a hard size limit (MRB_NOALLOC_MAX_VOXELS), 64.25 KiB of device memory held by every
// GPU that loads the library until the process exits, and no concurrency -- every call on a device
// uses the same array (see the notes at the entry point).
#include "mrb_internal.cuh"
#include "mrb_compact.cuh"

namespace mrb {

// One int per tile of kCompactTile (1024) voxels, then one slot for the grand total.
static const int kNoAllocMaxTiles = MRB_NOALLOC_MAX_VOXELS / kCompactTile;     // 16384
static const int kNoAllocScratchInts = kNoAllocMaxTiles + 64;                  // + total, padded

// The scratch itself.  In host code this name is only a handle for the runtime, not an address a
// kernel could use; the device address is looked up once per GPU with cudaGetSymbolAddress below.
static __device__ __align__(256) int g_noalloc_scratch[kNoAllocScratchInts];

static const int kMaxDevices = 64;
static int* g_scratch_addr[kMaxDevices] = {nullptr};   // per-device address of g_noalloc_scratch

// Looks the array's device address up the first time a given GPU is used (that GPU must be current).
// The cache is filled without a lock: two threads making their first call on the same GPU at the same
// time both look the symbol up and both store the same value, so the race is harmless.
static int noalloc_scratch(int device, int** out) {
    if (device < 0 || device >= kMaxDevices) { set_error("mrb_find_indexes_noalloc: device %d not supported", device); return MRB_EINVAL; }
    if (g_scratch_addr[device] == nullptr) {
        int* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
        if (e != cudaSuccess || p == nullptr) {
            cudaGetLastError();
            set_error("mrb_find_indexes_noalloc: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
            return MRB_ECUDA;
        }
        g_scratch_addr[device] = p;
    }
    *out = g_scratch_addr[device];
    return MRB_OK;
}

struct NonZeroPredNA {
    const float* g;
    __device__ bool operator()(size_t idx) const { return g[idx] != 0.0f; }
};

// Exclusive scan of the tile counts in place, by one block; the sum goes to counts[total_slot].
__global__ void __launch_bounds__(1024) k_noalloc_scan(int* __restrict__ counts, int ntiles, int total_slot) {
    __shared__ int wsum[32];
    __shared__ int carry_s;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < ntiles; base += 1024) {
        int idx = base + threadIdx.x;
        int v = idx < ntiles ? counts[idx] : 0;
        int incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) wsum[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            int s = wsum[lane], si = s;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int t = __shfl_up_sync(0xffffffffu, si, o);
                if (lane >= o) si += t;
            }
            wsum[lane] = si - s;
        }
        __syncthreads();
        const int carry = carry_s;
        if (idx < ntiles) counts[idx] = carry + wsum[wid] + incl - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = carry + wsum[31] + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) counts[total_slot] = carry_s;
}

__global__ void k_noalloc_linear(const int* pos, int n, Dims d, int* lin) {   // pos may alias lin
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int p = pos[i];
    int z = p % d.nz;
    int t = p / d.nz;
    int y = t % d.ny;
    int x = t / d.ny;
    lin[i] = z * d.nx * d.ny + y * d.nx + x;     // grid::sub2ind, grid.cpp:172-174
}

}  // namespace mrb

using namespace mrb;

extern "C" {

// See include/mrb.h for the contract.
int mrb_find_indexes_noalloc(const float* grid_dev, const int dims[3], int* idx_out_dev, int capacity, int* n_out,
                             void* stream_v) {
    if (!grid_dev || !dims || !n_out) { set_error("mrb_find_indexes_noalloc: NULL argument"); return MRB_EINVAL; }
    if (dims[0] < 1 || dims[1] < 1 || dims[2] < 1) { set_error("mrb_find_indexes_noalloc: bad dims"); return MRB_EINVAL; }
    const size_t n = (size_t)dims[0] * dims[1] * dims[2];
    if (n > (size_t)MRB_NOALLOC_MAX_VOXELS) {
        set_error("mrb_find_indexes_noalloc: grid has %zu voxels, the built-in scratch covers at most %d "
                  "(MRB_NOALLOC_MAX_VOXELS); use mrb_find_indexes for larger grids", n, MRB_NOALLOC_MAX_VOXELS);
        return MRB_ECAPACITY;
    }
    if (!is_device_ptr(grid_dev) || (idx_out_dev && !is_device_ptr(idx_out_dev))) {
        set_error("mrb_find_indexes_noalloc: grid and idx_out must be device pointers (this call cannot stage host buffers)");
        return MRB_EINVAL;
    }
    cudaStream_t stream = (cudaStream_t)stream_v;
    int device = 0;
    MRB_CUDA(cudaGetDevice(&device));
    // The kernels run on the current GPU with the current GPU's scratch array: buffers that live on another
    // GPU would fault or be reached through peer access with the wrong scratch, so they are refused.
    {
        cudaPointerAttributes pa;
        const void* ptrs[2] = {grid_dev, idx_out_dev};
        for (int i = 0; i < 2; i++) {
            if (!ptrs[i]) continue;
            if (cudaPointerGetAttributes(&pa, ptrs[i]) != cudaSuccess) { cudaGetLastError(); continue; }
            if (pa.type == cudaMemoryTypeDevice && pa.device != device) {
                set_error("mrb_find_indexes_noalloc: %s lives on GPU %d but GPU %d is current",
                          i == 0 ? "grid" : "idx_out", pa.device, device);
                return MRB_EINVAL;
            }
        }
    }
    int* scratch = nullptr;
    MRB_TRY(noalloc_scratch(device, &scratch));     // first call on this GPU resolves the array's address

    Dims d = mkdims(dims);
    const int ntiles = (int)((n + kCompactTile - 1) / kCompactTile);
    const int total_slot = kNoAllocMaxTiles;
    // The array still holds whatever the previous call left; clear the part this call reads back.
    MRB_CUDA(cudaMemsetAsync(scratch, 0, (size_t)ntiles * sizeof(int), stream));
    MRB_CUDA(cudaMemsetAsync(scratch + total_slot, 0, sizeof(int), stream));
    NonZeroPredNA pred{grid_dev};
    k_compact_count<NonZeroPredNA><<<ntiles, 256, 0, stream>>>(pred, n, scratch);
    MRB_CUDA(cudaGetLastError());
    k_noalloc_scan<<<1, 1024, 0, stream>>>(scratch, ntiles, total_slot);
    MRB_CUDA(cudaGetLastError());
    int total = 0;
    MRB_CUDA(cudaMemcpyAsync(&total, scratch + total_slot, sizeof(int), cudaMemcpyDeviceToHost, stream));
    MRB_CUDA(cudaStreamSynchronize(stream));
    *n_out = total;
    if (!idx_out_dev || total == 0) return MRB_OK;
    if (total > capacity) {
        set_error("mrb_find_indexes_noalloc: idx_out holds %d entries, %d needed", capacity, total);
        return MRB_ECAPACITY;
    }
    // dense offsets in reference order, then converted in place to the reference's x-fastest indexes
    k_compact_write<NonZeroPredNA><<<ntiles, 256, 0, stream>>>(pred, n, scratch, idx_out_dev);
    MRB_CUDA(cudaGetLastError());
    k_noalloc_linear<<<(total + 255) / 256, 256, 0, stream>>>(idx_out_dev, total, d, idx_out_dev);
    MRB_CUDA(cudaGetLastError());
    MRB_CUDA(cudaStreamSynchronize(stream));
    return MRB_OK;
}

}  // extern "C"

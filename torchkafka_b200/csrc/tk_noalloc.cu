// tk_noalloc.cu -- filter compaction (K2) for callers that must not allocate.
//
// tk_compact() keeps its inter-kernel scratch (per-tile kept counts) in memory the
// context obtained with cudaMalloc.  Some hosts of this library cannot do that at the
// point where they need the compaction: code running under a memory budget that was
// fixed up front (every byte of HBM already handed to a framework's own pool), plugins
// whose host forbids driver allocations on its worker threads because cudaMalloc
// synchronises the whole device, and callers inside a latency-critical section where an
// allocation's implicit synchronisation is not acceptable.  For them this file offers
// the same computation with the scratch in a fixed array that lives in the library image:
// loading the library reserves it on every GPU that uses the library, and it stays
// reserved until the process exits.
//
// The first call on a device resolves the array's device address (cudaGetSymbolAddress);
// make that first call outside any stream capture.  With CUDA's lazy module loading that
// first call (symbol lookup, first launch) is also when the driver loads the library's
// device code, which may allocate inside the driver: make it during initialisation, not in
// the section that must not allocate.
//
// Limit: TK_NOALLOC_MAX_RECORDS records per call (scratch = one uint32 per tile of 1024
// records, a little over 4 KiB in total).
//
// The array is shared by every call on a device and keeps whatever the last call left in
// it.  Nothing in it has to start at zero -- each call overwrites every word it later
// reads, in stream order -- but two calls whose kernels overlap on the GPU (different
// streams, no dependency) would overwrite each other's tile counts: callers must order
// calls on one device, e.g. by issuing them on one stream.
#include <atomic>
#include "tk_internal.h"

#define NA_THREADS 256
#define NA_ITEMS 4
#define NA_TILE (NA_THREADS * NA_ITEMS)
#define NA_MAX_TILES (TK_NOALLOC_MAX_RECORDS / NA_TILE)

// per-tile kept counts, turned in place into exclusive tile offsets by k_na_scan
static __device__ __align__(256) uint32_t g_na_scratch[NA_MAX_TILES + 64];

// Device address of g_na_scratch per device.  In host code the name g_na_scratch is only a
// handle for the runtime, not an address a kernel can use, so the address is looked up once
// per device with cudaGetSymbolAddress (with that device current) and passed to the kernels.
// Several host threads may make the first call on a device at the same time: the pointer is
// published with release / read with acquire (every thread that loses the race stores the same value).
static std::atomic<uint32_t*> g_na_ptr[64];

static int na_setup(int device, uint32_t** out) {
    TK_REQUIRE(device >= 0 && device < 64, "device index %d not supported", device);
    uint32_t* ptr = g_na_ptr[device].load(std::memory_order_acquire);
    if (!ptr) {
        void* p = nullptr;
        TK_CUDA(cudaGetSymbolAddress(&p, g_na_scratch));
        TK_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null address");
        ptr = static_cast<uint32_t*>(p);
        g_na_ptr[device].store(ptr, std::memory_order_release);
    }
    *out = ptr;
    return 0;
}

__device__ __forceinline__ uint32_t na_warp_scan(uint32_t v) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, v, d);
        if ((threadIdx.x & 31) >= d) v += t;
    }
    return v;
}

// exclusive scan of one value per thread over the block; *total receives the block sum
__device__ __forceinline__ uint32_t na_block_scan(uint32_t v, uint32_t* total) {
    __shared__ uint32_t wsum[32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const uint32_t incl = na_warp_scan(v);
    if (lane == 31) wsum[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        const uint32_t w = lane < (int)((blockDim.x + 31) >> 5) ? wsum[lane] : 0;
        const uint32_t wi = na_warp_scan(w);
        wsum[lane] = wi - w;
        if (lane == 31) *total = wi;
    }
    __syncthreads();
    const uint32_t r = wsum[wid] + incl - v;
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(NA_THREADS) k_na_count(const int32_t* __restrict__ keep, uint32_t n,
                                                          uint32_t* __restrict__ tile_sum) {
    __shared__ uint32_t tot;
    const uint32_t base = blockIdx.x * NA_TILE + threadIdx.x * NA_ITEMS;
    uint32_t c = 0;
#pragma unroll
    for (int k = 0; k < NA_ITEMS; ++k)
        if (base + k < n && keep[base + k] >= 0) ++c;
    na_block_scan(c, &tot);
    if (threadIdx.x == 0) tile_sum[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(1024) k_na_scan(uint32_t* __restrict__ tile_sum, uint32_t ntiles, uint32_t n,
                                                   uint32_t batch_size, uint32_t carry_in,
                                                   tk_result_t* __restrict__ res) {
    __shared__ uint32_t tot;
    uint32_t running = 0;
    for (uint32_t base = 0; base < ntiles; base += blockDim.x) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < ntiles ? tile_sum[i] : 0;
        const uint32_t ex = na_block_scan(v, &tot);
        if (i < ntiles) tile_sum[i] = running + ex;
        running += tot;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const uint32_t kept = running;
        res->n_records = n;
        res->n_kept = kept;
        res->carry_in = carry_in;
        res->carry_out = (carry_in + kept) % batch_size;
        res->n_closed = (carry_in + kept) / batch_size;
        res->batch_size = batch_size;
        res->n_errors = 0;
        res->first_error_rec = 0xFFFFFFFFu;
        res->first_error_status = 0;
    }
}

__global__ void __launch_bounds__(NA_THREADS) k_na_apply(const int32_t* __restrict__ keep, uint32_t n,
                                                          const uint32_t* __restrict__ tile_off,
                                                          int32_t* __restrict__ kidx) {
    __shared__ uint32_t tot;
    const uint32_t base = blockIdx.x * NA_TILE + threadIdx.x * NA_ITEMS;
    int flag[NA_ITEMS];
    uint32_t c = 0;
#pragma unroll
    for (int k = 0; k < NA_ITEMS; ++k) {
        flag[k] = (base + k < n && keep[base + k] >= 0) ? 1 : 0;
        c += flag[k];
    }
    uint32_t ex = na_block_scan(c, &tot) + tile_off[blockIdx.x];
#pragma unroll
    for (int k = 0; k < NA_ITEMS; ++k) {
        const uint32_t i = base + k;
        if (i >= n) break;
        if (flag[k]) kidx[i] = (int32_t)ex++;
        else kidx[i] = -1 - (int32_t)ex;
    }
}

extern "C" int tk_compact_noalloc(const int32_t* d_keep, uint32_t n, uint32_t batch_size, uint32_t carry_in,
                                  int32_t* d_kidx, tk_result_t* d_result, void* stream) {
    TK_REQUIRE(d_keep && d_kidx && d_result, "null device pointer");
    TK_REQUIRE(batch_size > 0, "batch_size must be positive");
    TK_REQUIRE(carry_in < batch_size, "carry_in (%u) must be below batch_size (%u)", carry_in, batch_size);
    TK_REQUIRE(n <= TK_NOALLOC_MAX_RECORDS,
               "tk_compact_noalloc handles at most %u records per call, got %u (use tk_compact or split the run)",
               (unsigned)TK_NOALLOC_MAX_RECORDS, n);
    int device = -1;
    TK_CUDA(cudaGetDevice(&device));
    uint32_t* scratch = nullptr;
    int rc = na_setup(device, &scratch);      // first call on a device resolves the scratch address
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t ntiles = (n + NA_TILE - 1) / NA_TILE;
    if (ntiles) {
        k_na_count<<<ntiles, NA_THREADS, 0, st>>>(d_keep, n, scratch);
        TK_LAUNCH_CHECK();
    }
    k_na_scan<<<1, 1024, 0, st>>>(scratch, ntiles, n, batch_size, carry_in, d_result);
    TK_LAUNCH_CHECK();
    if (ntiles) {
        k_na_apply<<<ntiles, NA_THREADS, 0, st>>>(d_keep, n, scratch, d_kidx);
        TK_LAUNCH_CHECK();
    }
    return 0;
}

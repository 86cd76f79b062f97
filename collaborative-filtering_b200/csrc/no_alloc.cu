// no_alloc.cu — b200als_row_order_static: the row schedule of b200als_row_order (sparse_build.cu) for callers that
// cannot allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
//
// No counterpart in the reference.
//
// The scratch of the schedule (sort keys, sorted keys, row ids, and the temporary storage of the radix sort) lives in
// ONE file-scope `static __device__` array of a fixed, modest size (kScratchBytes = 7 MiB: rows up to
// B200ALS_ROW_ORDER_STATIC_MAX_ROWS = 262 144).  Every GPU that loads the library holds that array until the process
// exits.  The entry point calls no cudaMalloc / cudaFree / allocating library routine: the caller passes inputs and
// outputs, everything in between is this array.
//
// Host code never uses the array's NAME as a pointer — in host code the name is a host-side stand-in, not a device
// address.  The device address is fetched once per device, with that device current, by cudaGetSymbolAddress (below) and
// handed to the kernels / to cub; device code names the array directly.
//
// Every call shares the array and finds the previous call's contents in it.  Nothing in it has to start at zero: the key
// and id segments are written for all n_rows before they are read, and the sort initialises its own temporary storage;
// the one accumulator of the path, *n_nonempty, is the caller's and is cleared here with a memset on the stream.
// Calls that overlap on the GPU (two streams, two host threads) RACE on the array: serialise them.
#include <cub/device/device_radix_sort.cuh>

#include <mutex>

#include "common.cuh"

namespace b200als {

constexpr int64_t kStaticMaxRows = B200ALS_ROW_ORDER_STATIC_MAX_ROWS;
constexpr size_t kSegBytes = (size_t)kStaticMaxRows * 4;       // one uint32 / int32 per row: 1 MiB
constexpr size_t kCubBytes = (size_t)4 << 20;                  // temporary storage of the radix sort (needs ~2 n * 4 B)
constexpr size_t kScratchBytes = 3 * kSegBytes + kCubBytes;    // keys | sorted keys | ids | sort temp = 7 MiB

static __device__ __align__(256) unsigned char g_row_order_scratch[kScratchBytes];

// key = ~count (ascending radix sort = descending count), value = row id; the sort is stable, so ties keep ascending
// row id.  Device code: the array is named directly.
__global__ void __launch_bounds__(256) row_keys_static_kernel(const int64_t* __restrict__ indptr, int64_t n_rows,
                                                              unsigned long long* __restrict__ n_nonempty) {
    uint32_t* keys = reinterpret_cast<uint32_t*>(g_row_order_scratch);
    int32_t* ids = reinterpret_cast<int32_t*>(g_row_order_scratch + 2 * kSegBytes);
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    unsigned local = 0;
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += stride) {
        const int64_t c = indptr[r + 1] - indptr[r];
        const uint32_t c32 = c > 0xffffffffll ? 0xffffffffu : (uint32_t)c;
        keys[r] = ~c32;
        ids[r] = (int32_t)r;
        local += (c > 0);
    }
    local = __reduce_add_sync(0xffffffffu, local);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(n_nonempty, (unsigned long long)local);
}

// Device address of the scratch array on the CURRENT device: cudaGetSymbolAddress once per device, cached.
static int static_scratch_address(unsigned char** out) {
    static std::mutex mu;
    static unsigned char* ptrs[64] = {nullptr};
    int dev = 0;
    ALS_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) {
        set_error("b200als_row_order_static: device ordinal %d outside [0, 64)", dev);
        return B200ALS_E_UNSUPPORTED;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_row_order_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            set_error("b200als_row_order_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
            return B200ALS_E_CUDA;
        }
        ptrs[dev] = static_cast<unsigned char*>(p);
    }
    *out = ptrs[dev];
    return B200ALS_OK;
}

}  // namespace b200als

using namespace b200als;

extern "C" {

int64_t b200als_row_order_static_max_rows(void) { return kStaticMaxRows; }

int b200als_row_order_static(const int64_t* indptr, int64_t n_rows, int32_t* order, int64_t* n_nonempty,
                             b200als_stream_t stream) {
    ALS_CHECK_ARG(indptr && order && n_nonempty && n_rows > 0, "null pointer or empty");
    if (n_rows > kStaticMaxRows) {
        set_error("b200als_row_order_static: n_rows %lld exceeds the static scratch (at most %lld rows); use "
                  "b200als_row_order with a caller-owned workspace",
                  (long long)n_rows, (long long)kStaticMaxRows);
        return B200ALS_E_UNSUPPORTED;
    }
    unsigned char* scratch = nullptr;
    int rc = static_scratch_address(&scratch);
    if (rc) return rc;
    size_t need = 0;                                   // size query: host arithmetic only
    ALS_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, need, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                             (const int32_t*)nullptr, (int32_t*)nullptr, (size_t)n_rows, 0, 32));
    if (need > kCubBytes) {
        set_error("b200als_row_order_static: the sort needs %zu bytes of temporary storage, the static scratch holds %zu",
                  need, kCubBytes);
        return B200ALS_E_UNSUPPORTED;
    }
    cudaStream_t s = as_stream(stream);
    const uint32_t* keys = reinterpret_cast<const uint32_t*>(scratch);
    uint32_t* keys_out = reinterpret_cast<uint32_t*>(scratch + kSegBytes);
    const int32_t* ids = reinterpret_cast<const int32_t*>(scratch + 2 * kSegBytes);
    void* cub_tmp = scratch + 3 * kSegBytes;
    size_t cub_bytes = kCubBytes;
    ALS_CUDA(cudaMemsetAsync(n_nonempty, 0, sizeof(int64_t), s));      // the caller's accumulator starts at zero
    int64_t blocks = ceil_div64(n_rows, 256);
    if (blocks > (int64_t)kNumSMs * 16) blocks = (int64_t)kNumSMs * 16;
    row_keys_static_kernel<<<(unsigned)blocks, 256, 0, s>>>(indptr, n_rows,
                                                            reinterpret_cast<unsigned long long*>(n_nonempty));
    ALS_LAUNCH_CHECK();
    ALS_CUDA(cub::DeviceRadixSort::SortPairs(cub_tmp, cub_bytes, keys, keys_out, ids, order, (size_t)n_rows, 0, 32, s));
    g_launch_count += 4;
    return B200ALS_OK;
}

}  // extern "C"

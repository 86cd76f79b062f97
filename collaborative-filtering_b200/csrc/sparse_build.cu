// sparse_build.cu — A1: CSR / CSC construction, index-bit-exact with
// als.py:335-340 (np.flatnonzero per row / per column, np.where row-major),
// plus the longest-first row schedule used by the half-step kernels.
//
// Two input forms:
//   dense  (m, n) float64 with NaN = missing      — the reference's own input
//   COO    flat_idx = u*n + i (int64) + float32   — create_folds.py convention
// Ordering is produced by construction (ballot compaction in index order) or by
// a stable LSD radix sort (cub::DeviceRadixSort — one-shot setup, library sort),
// never by atomics, so results are deterministic and identical to numpy's.
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace b200als {

// ---- single-CTA in-place inclusive scan with running carry (setup-only) ----
__global__ void __launch_bounds__(1024) scan_inclusive_i64_kernel(int64_t* data, int64_t count) {
    __shared__ int64_t warp_tot[32];
    __shared__ int64_t carry_s;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) carry_s = 0;
    __syncthreads();
    for (int64_t base = 0; base < count; base += 4096) {
        int64_t v[4];
        int64_t i0 = base + (int64_t)tid * 4;
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = (i0 + j < count) ? data[i0 + j] : 0;
        v[1] += v[0]; v[2] += v[1]; v[3] += v[2];
        int64_t incl = v[3];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int64_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) warp_tot[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            int64_t w = warp_tot[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int64_t t = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += t;
            }
            warp_tot[lane] = w;
        }
        __syncthreads();
        int64_t off = carry_s + (wid ? warp_tot[wid - 1] : 0) + (incl - v[3]);
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (i0 + j < count) data[i0 + j] = v[j] + off;
        __syncthreads();
        if (tid == 1023) carry_s = off + v[3];
        __syncthreads();
    }
}

// ---- N2: fold split on the resident observed set (create_folds.py:177-208) ------------------------------
// The observed entries are one ascending flat-index array; a fold's validation indices are looked up in it
// (binary search), flagged in a bitset, and the unflagged entries are compacted in order — the sorted
// `setdiff1d(all_obs, val_idx)` of create_folds.py:200 — with a block count, a single-CTA scan of the block
// counts and an order-preserving scatter.  Pure integer work: bit-exact by construction.
__global__ void __launch_bounds__(256) fold_mark_kernel(const int64_t* __restrict__ flat, const float* __restrict__ vals,
                                                        int64_t nnz, const int64_t* __restrict__ val_idx, int64_t n_val,
                                                        uint32_t* __restrict__ bits, float* __restrict__ val_vals,
                                                        int32_t* __restrict__ status) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n_val; j += stride) {
        const int64_t v = val_idx[j];
        int64_t lo = 0, hi = nnz;                       // lower bound
        while (lo < hi) {
            const int64_t mid = (lo + hi) >> 1;
            if (flat[mid] < v) lo = mid + 1; else hi = mid;
        }
        if (lo < nnz && flat[lo] == v) {
            const uint32_t bit = 1u << (lo & 31);
            const uint32_t old = atomicOr(&bits[lo >> 5], bit);
            if (old & bit) atomicAdd(&status[1], 1);    // the same entry listed twice
            val_vals[j] = vals[lo];
        } else {
            atomicAdd(&status[0], 1);                   // not an observed entry
            val_vals[j] = __int_as_float(0x7fc00000);   // NaN, what the reference would read there
        }
    }
}

constexpr int FOLD_BLOCK_ENTRIES = 8192;                // 8 warps x 32 words x 32 entries

__global__ void __launch_bounds__(256) fold_count_kernel(const uint32_t* __restrict__ bits, int64_t nnz,
                                                         int64_t* __restrict__ block_cnt) {
    __shared__ int wsum[8];
    const int64_t word = (int64_t)blockIdx.x * 256 + threadIdx.x;
    const int64_t first = word * 32;
    int c = 0;
    if (first < nnz) {
        uint32_t keep = ~bits[word];
        const int64_t left = nnz - first;
        if (left < 32) keep &= (1u << left) - 1u;
        c = __popc(keep);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += wsum[w];
        block_cnt[blockIdx.x] = t;
    }
}

// block_off = inclusive scan of block_cnt (scan_inclusive_i64_kernel); warp w of block b compacts its 1024 entries
__global__ void __launch_bounds__(256) fold_scatter_kernel(const int64_t* __restrict__ flat, const float* __restrict__ vals,
                                                           int64_t nnz, const uint32_t* __restrict__ bits,
                                                           const int64_t* __restrict__ block_incl,
                                                           int64_t* __restrict__ out_flat, float* __restrict__ out_vals,
                                                           int64_t* __restrict__ n_train) {
    __shared__ int wsum[8];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int64_t word = (int64_t)blockIdx.x * 256 + threadIdx.x;   // this lane's word of the warp's 32
    const int64_t first = word * 32;
    uint32_t keep = 0;
    if (first < nnz) {
        keep = ~bits[word];
        const int64_t left = nnz - first;
        if (left < 32) keep &= (1u << left) - 1u;
    }
    const int c = __popc(keep);
    int incl = c;                                       // inclusive scan of the word counts inside the warp
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) wsum[wid] = incl;
    __syncthreads();
    int64_t base = blockIdx.x ? block_incl[blockIdx.x - 1] : 0;
    for (int w = 0; w < wid; ++w) base += wsum[w];
    const int excl = incl - c;
    // the warp walks its 32 words together: lane l takes entry l of word w (coalesced reads, in-order writes)
    const int64_t warp_first = ((int64_t)blockIdx.x * 256 + wid * 32) * 32;
    for (int w = 0; w < 32; ++w) {
        const uint32_t kw = __shfl_sync(0xffffffffu, keep, w);
        const int ew = __shfl_sync(0xffffffffu, excl, w);
        if (kw == 0) continue;
        const int64_t e = warp_first + (int64_t)w * 32 + lane;
        if ((kw >> lane) & 1u) {
            const int64_t dst = base + ew + __popc(kw & ((1u << lane) - 1u));
            out_flat[dst] = flat[e];
            out_vals[dst] = vals[e];
        }
    }
    if (n_train && blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) *n_train = block_incl[gridDim.x - 1];
}

// ---- dense input ----------------------------------------------------------
// warp per row: row count → indptr[u+1]; per-column counts via atomics (integer
// adds commute, so the result is exact regardless of order).
__global__ void dense_count_kernel(const double* __restrict__ R, int64_t m, int64_t n,
                                   int64_t* __restrict__ indptr, unsigned long long* __restrict__ ccount) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t u = warp; u < m; u += nwarps) {
        const double* row = R + u * n;
        int cnt = 0;
        for (int64_t i = lane; i < n; i += 32) {
            double v = row[i];
            if (v == v) { ++cnt; atomicAdd(&ccount[i + 1], 1ull); }
        }
        cnt = (int)__reduce_add_sync(0xffffffffu, cnt);
        if (lane == 0) indptr[u + 1] = cnt;
    }
}

__global__ void dense_fill_csr_kernel(const double* __restrict__ R, int64_t m, int64_t n,
                                      const int64_t* __restrict__ indptr, int32_t* __restrict__ indices,
                                      float* __restrict__ vals) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t u = warp; u < m; u += nwarps) {
        const double* row = R + u * n;
        int64_t out = indptr[u];
        for (int64_t i0 = 0; i0 < n; i0 += 32) {
            int64_t i = i0 + lane;
            double v = (i < n) ? row[i] : __longlong_as_double(0x7ff8000000000000ll);
            bool ok = (v == v);
            unsigned bal = __ballot_sync(0xffffffffu, ok);
            if (ok) {
                int64_t p = out + __popc(bal & ((1u << lane) - 1u));
                indices[p] = (int32_t)i;
                vals[p] = (float)v;
            }
            out += __popc(bal);
        }
    }
}

// warp per column; lanes walk rows in ascending order so the compaction keeps
// np.flatnonzero(mask[:, i]) order.  perm by binary search of i in CSR row u.
__global__ void dense_fill_csc_kernel(const double* __restrict__ R, int64_t m, int64_t n,
                                      const int64_t* __restrict__ indptr, const int32_t* __restrict__ indices,
                                      const int64_t* __restrict__ cindptr, int32_t* __restrict__ cindices,
                                      float* __restrict__ cvals, int64_t* __restrict__ perm) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t i = warp; i < n; i += nwarps) {
        int64_t out = cindptr[i];
        for (int64_t u0 = 0; u0 < m; u0 += 32) {
            int64_t u = u0 + lane;
            double v = (u < m) ? R[u * n + i] : __longlong_as_double(0x7ff8000000000000ll);
            bool ok = (v == v);
            unsigned bal = __ballot_sync(0xffffffffu, ok);
            if (ok) {
                int64_t p = out + __popc(bal & ((1u << lane) - 1u));
                cindices[p] = (int32_t)u;
                cvals[p] = (float)v;
                if (perm) {
                    int64_t lo = indptr[u], hi = indptr[u + 1] - 1;
                    while (lo < hi) {
                        int64_t mid = (lo + hi) >> 1;
                        if (indices[mid] < (int32_t)i) lo = mid + 1; else hi = mid;
                    }
                    perm[p] = lo;
                }
            }
            out += __popc(bal);
        }
    }
}

// ---- COO input --------------------------------------------------------------
// sorted flat index → (row32, col32); indptr by boundary detection.
// Input check before anything is built: indices inside [0, m*n), strictly increasing (on the sorted sequence:
// duplicates; on a `sorted` input also disorder), ratings finite.  status[0..2] count the violations.
__global__ void coo_validate_kernel(const int64_t* __restrict__ flat, const float* __restrict__ vals, int64_t nnz,
                                    int64_t total, unsigned long long* __restrict__ status) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    unsigned oob = 0, dis = 0, nonf = 0;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < nnz; e += stride) {
        const int64_t f = flat[e];
        oob += (f < 0 || f >= total);
        if (e > 0) dis += (flat[e - 1] >= f);
        const float v = vals[e];
        nonf += !(fabsf(v) <= 3.402823466e38f);
    }
    oob = __reduce_add_sync(0xffffffffu, oob);
    dis = __reduce_add_sync(0xffffffffu, dis);
    nonf = __reduce_add_sync(0xffffffffu, nonf);
    if ((threadIdx.x & 31) == 0) {
        if (oob) atomicAdd(&status[0], (unsigned long long)oob);
        if (dis) atomicAdd(&status[1], (unsigned long long)dis);
        if (nonf) atomicAdd(&status[2], (unsigned long long)nonf);
    }
}

__device__ __forceinline__ int64_t clamp_flat(int64_t f, int64_t total) { return f < 0 ? 0 : (f >= total ? total - 1 : f); }

__global__ void coo_split_kernel(const int64_t* __restrict__ flat, int64_t nnz, int64_t m, int64_t n,
                                 int32_t* __restrict__ rows32, int32_t* __restrict__ cols32,
                                 int64_t* __restrict__ indptr) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t total = m * n;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < nnz; e += stride) {
        // out-of-range entries are clamped (and reported by coo_validate_kernel): nothing is written outside
        // indptr[0..m] / the nnz-long arrays whatever the input holds
        int64_t f = clamp_flat(flat[e], total);
        int64_t r = f / n;
        int32_t c = (int32_t)(f - r * n);
        rows32[e] = (int32_t)r;
        cols32[e] = c;
        int64_t rprev = (e == 0) ? -1 : clamp_flat(flat[e - 1], total) / n;
        for (int64_t q = rprev + 1; q <= r; ++q) indptr[q] = e;   // rows rprev+1..r start at e
        if (e == nnz - 1)
            for (int64_t q = r + 1; q <= m; ++q) indptr[q] = nnz;
    }
}

__global__ void fill_empty_indptr_kernel(int64_t* indptr, int64_t count) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < count; e += stride) indptr[e] = 0;
}

__global__ void iota_u32_kernel(uint32_t* p, int64_t nnz) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < nnz; e += stride) p[e] = (uint32_t)e;
}

// sorted-by-column (stable) positions → CSC arrays + cindptr by boundary detection
__global__ void csc_gather_kernel(const uint32_t* __restrict__ pos_sorted, const int32_t* __restrict__ col_sorted,
                                  const int32_t* __restrict__ rows32, const float* __restrict__ vals, int64_t nnz,
                                  int64_t n, int32_t* __restrict__ cindices, float* __restrict__ cvals,
                                  int64_t* __restrict__ perm, int64_t* __restrict__ cindptr) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < nnz; e += stride) {
        uint32_t p = pos_sorted[e];
        cindices[e] = rows32[p];
        cvals[e] = vals[p];
        if (perm) perm[e] = (int64_t)p;
        int64_t c = col_sorted[e];
        int64_t cprev = (e == 0) ? -1 : (int64_t)col_sorted[e - 1];
        for (int64_t q = cprev + 1; q <= c; ++q) cindptr[q] = e;
        if (e == nnz - 1)
            for (int64_t q = c + 1; q <= n; ++q) cindptr[q] = nnz;
    }
}

// ---- row schedule -----------------------------------------------------------
// key = ~count (so ascending radix sort = descending count), value = row id;
// the sort is stable, so ties keep ascending row id.
__global__ void row_keys_kernel(const int64_t* __restrict__ indptr, int64_t n_rows, uint32_t* __restrict__ keys,
                                int32_t* __restrict__ ids, unsigned long long* __restrict__ n_nonempty) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    unsigned local = 0;
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += stride) {
        int64_t c = indptr[r + 1] - indptr[r];
        uint32_t c32 = c > 0xffffffffll ? 0xffffffffu : (uint32_t)c;
        keys[r] = ~c32;
        ids[r] = (int32_t)r;
        local += (c > 0);
    }
    local = __reduce_add_sync(0xffffffffu, local);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(n_nonempty, (unsigned long long)local);
}

static int grid_for(int64_t work, int threads) {
    int64_t b = ceil_div64(work, threads);
    int64_t cap = (int64_t)kNumSMs * 16;
    return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

struct CooLayout {
    size_t off_flat_sorted, off_vals_tmp, off_rows32, off_cols32, off_cols_sorted, off_pos, off_pos_sorted, off_cub;
    size_t cub_bytes, total;
};

static int coo_layout(int64_t nnz, CooLayout* L) {
    size_t a = 256, off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, a); return o; };
    size_t nz = (size_t)(nnz > 0 ? nnz : 1);
    L->off_flat_sorted = take(nz * 8);
    L->off_vals_tmp = take(nz * 4);
    L->off_rows32 = take(nz * 4);
    L->off_cols32 = take(nz * 4);
    L->off_cols_sorted = take(nz * 4);
    L->off_pos = take(nz * 4);
    L->off_pos_sorted = take(nz * 4);
    size_t b1 = 0, b2 = 0;
    cudaError_t e1 = cub::DeviceRadixSort::SortPairs(nullptr, b1, (const int64_t*)nullptr, (int64_t*)nullptr,
                                                     (const float*)nullptr, (float*)nullptr, nz, 0, 64);
    cudaError_t e2 = cub::DeviceRadixSort::SortPairs(nullptr, b2, (const int32_t*)nullptr, (int32_t*)nullptr,
                                                     (const uint32_t*)nullptr, (uint32_t*)nullptr, nz, 0, 32);
    if (e1 != cudaSuccess || e2 != cudaSuccess) {
        set_error("cub temp-storage query failed");
        return B200ALS_E_CUDA;
    }
    L->cub_bytes = b1 > b2 ? b1 : b2;
    L->off_cub = take(L->cub_bytes);
    L->total = off;
    return B200ALS_OK;
}

}  // namespace b200als

using namespace b200als;

extern "C" {

int b200als_dense_indptr(const double* R, int64_t m, int64_t n, int64_t* indptr, int64_t* cindptr,
                         b200als_stream_t stream) {
    ALS_CHECK_ARG(m > 0 && n > 0 && R && indptr && cindptr, "null pointer or empty shape");
    ALS_CHECK_ARG(n < (1ll << 31) && m < (1ll << 31), "m and n must fit int32");
    cudaStream_t s = as_stream(stream);
    ALS_CUDA(cudaMemsetAsync(indptr, 0, sizeof(int64_t) * (m + 1), s));
    ALS_CUDA(cudaMemsetAsync(cindptr, 0, sizeof(int64_t) * (n + 1), s));
    dense_count_kernel<<<grid_for(m * 32, 256), 256, 0, s>>>(R, m, n, indptr,
                                                             reinterpret_cast<unsigned long long*>(cindptr));
    ALS_LAUNCH_CHECK();
    scan_inclusive_i64_kernel<<<1, 1024, 0, s>>>(indptr, m + 1);
    ALS_LAUNCH_CHECK();
    scan_inclusive_i64_kernel<<<1, 1024, 0, s>>>(cindptr, n + 1);
    ALS_LAUNCH_CHECK();
    return B200ALS_OK;
}

int b200als_dense_fill(const double* R, int64_t m, int64_t n, const int64_t* indptr, const int64_t* cindptr,
                       int32_t* indices, float* vals, int32_t* cindices, float* cvals, int64_t* perm,
                       b200als_stream_t stream) {
    ALS_CHECK_ARG(m > 0 && n > 0 && R && indptr && cindptr, "null pointer or empty shape");
    ALS_CHECK_ARG(indices && vals && cindices && cvals, "null output");
    cudaStream_t s = as_stream(stream);
    dense_fill_csr_kernel<<<grid_for(m * 32, 256), 256, 0, s>>>(R, m, n, indptr, indices, vals);
    ALS_LAUNCH_CHECK();
    dense_fill_csc_kernel<<<grid_for(n * 32, 256), 256, 0, s>>>(R, m, n, indptr, indices, cindptr, cindices, cvals,
                                                                perm);
    ALS_LAUNCH_CHECK();
    return B200ALS_OK;
}

int b200als_coo_workspace_bytes(int64_t nnz, int64_t m, int64_t n, size_t* bytes_host) {
    ALS_CHECK_ARG(bytes_host && nnz >= 0 && m > 0 && n > 0, "bad shape");
    CooLayout L;
    int rc = coo_layout(nnz, &L);
    if (rc) return rc;
    *bytes_host = L.total;
    return B200ALS_OK;
}

int b200als_coo_to_csr_csc(const int64_t* flat_idx, const float* vals_in, int64_t nnz, int64_t m, int64_t n,
                           int sorted, int64_t* indptr, int32_t* indices, float* vals, int64_t* cindptr,
                           int32_t* cindices, float* cvals, int64_t* perm, void* workspace, size_t workspace_bytes,
                           int64_t* status3, b200als_stream_t stream) {
    ALS_CHECK_ARG(m > 0 && n > 0 && nnz >= 0, "bad shape");
    ALS_CHECK_ARG(n < (1ll << 31) && m < (1ll << 31), "m and n must fit int32");
    ALS_CHECK_ARG(nnz < (1ll << 32), "nnz per call must be < 2^32 (shard larger jobs by rows)");
    ALS_CHECK_ARG(indptr && cindptr, "null indptr");
    cudaStream_t s = as_stream(stream);
    if (status3) ALS_CUDA(cudaMemsetAsync(status3, 0, 3 * sizeof(int64_t), s));
    if (nnz == 0) {
        fill_empty_indptr_kernel<<<grid_for(m + 1, 256), 256, 0, s>>>(indptr, m + 1);
        ALS_LAUNCH_CHECK();
        fill_empty_indptr_kernel<<<grid_for(n + 1, 256), 256, 0, s>>>(cindptr, n + 1);
        ALS_LAUNCH_CHECK();
        return B200ALS_OK;
    }
    ALS_CHECK_ARG(flat_idx && vals_in && indices && vals && cindices && cvals, "null pointer");
    CooLayout L;
    int rc = coo_layout(nnz, &L);
    if (rc) return rc;
    if (!workspace || workspace_bytes < L.total) {
        set_error("b200als_coo_to_csr_csc: workspace %zu < required %zu", workspace_bytes, L.total);
        return B200ALS_E_WORKSPACE;
    }
    char* w = static_cast<char*>(workspace);
    int64_t* flat_sorted = reinterpret_cast<int64_t*>(w + L.off_flat_sorted);
    int32_t* rows32 = reinterpret_cast<int32_t*>(w + L.off_rows32);
    int32_t* cols_sorted = reinterpret_cast<int32_t*>(w + L.off_cols_sorted);
    uint32_t* pos = reinterpret_cast<uint32_t*>(w + L.off_pos);
    uint32_t* pos_sorted = reinterpret_cast<uint32_t*>(w + L.off_pos_sorted);
    void* cub_tmp = w + L.off_cub;
    size_t cub_bytes = L.cub_bytes;

    const int64_t* flat_use = flat_idx;
    if (!sorted) {
        // bits needed for m*n-1
        int end_bit = 1;
        unsigned long long maxkey = (unsigned long long)m * (unsigned long long)n - 1ull;
        while (end_bit < 64 && (maxkey >> end_bit)) ++end_bit;
        ALS_CUDA(cub::DeviceRadixSort::SortPairs(cub_tmp, cub_bytes, flat_idx, flat_sorted, vals_in, vals,
                                                 (size_t)nnz, 0, end_bit, s));
        g_launch_count += 4;
        flat_use = flat_sorted;
    } else {
        ALS_CUDA(cudaMemcpyAsync(vals, vals_in, sizeof(float) * nnz, cudaMemcpyDeviceToDevice, s));
    }
    if (status3) {
        coo_validate_kernel<<<grid_for(nnz, 256), 256, 0, s>>>(flat_use, vals, nnz, m * n,
                                                               reinterpret_cast<unsigned long long*>(status3));
        ALS_LAUNCH_CHECK();
    }
    coo_split_kernel<<<grid_for(nnz, 256), 256, 0, s>>>(flat_use, nnz, m, n, rows32, indices, indptr);
    ALS_LAUNCH_CHECK();
    iota_u32_kernel<<<grid_for(nnz, 256), 256, 0, s>>>(pos, nnz);
    ALS_LAUNCH_CHECK();
    int col_bits = 1;
    while (col_bits < 32 && ((unsigned long long)(n - 1) >> col_bits)) ++col_bits;
    cub_bytes = L.cub_bytes;
    ALS_CUDA(cub::DeviceRadixSort::SortPairs(cub_tmp, cub_bytes, (const int32_t*)indices, cols_sorted,
                                             (const uint32_t*)pos, pos_sorted, (size_t)nnz, 0, col_bits, s));
    g_launch_count += 4;
    csc_gather_kernel<<<grid_for(nnz, 256), 256, 0, s>>>(pos_sorted, cols_sorted, rows32, vals, nnz, n, cindices,
                                                         cvals, perm, cindptr);
    ALS_LAUNCH_CHECK();
    return B200ALS_OK;
}

int b200als_row_order_workspace_bytes(int64_t n_rows, size_t* bytes_host) {
    ALS_CHECK_ARG(bytes_host && n_rows > 0, "bad shape");
    size_t cb = 0;
    cudaError_t e = cub::DeviceRadixSort::SortPairs(nullptr, cb, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                                    (const int32_t*)nullptr, (int32_t*)nullptr, (size_t)n_rows, 0, 32);
    if (e != cudaSuccess) {
        set_error("cub temp-storage query failed");
        return B200ALS_E_CUDA;
    }
    *bytes_host = align_up((size_t)n_rows * 4, 256) * 3 + align_up(cb, 256);
    return B200ALS_OK;
}

int b200als_row_order(const int64_t* indptr, int64_t n_rows, int32_t* order, int64_t* n_nonempty, void* workspace,
                      size_t workspace_bytes, b200als_stream_t stream) {
    ALS_CHECK_ARG(indptr && order && n_nonempty && n_rows > 0, "null pointer or empty");
    ALS_CHECK_ARG(n_rows < (1ll << 31), "n_rows must fit int32");
    size_t need = 0;
    int rc = b200als_row_order_workspace_bytes(n_rows, &need);
    if (rc) return rc;
    if (!workspace || workspace_bytes < need) {
        set_error("b200als_row_order: workspace %zu < required %zu", workspace_bytes, need);
        return B200ALS_E_WORKSPACE;
    }
    cudaStream_t s = as_stream(stream);
    char* w = static_cast<char*>(workspace);
    size_t seg = align_up((size_t)n_rows * 4, 256);
    uint32_t* keys = reinterpret_cast<uint32_t*>(w);
    uint32_t* keys_out = reinterpret_cast<uint32_t*>(w + seg);
    int32_t* ids = reinterpret_cast<int32_t*>(w + 2 * seg);
    void* cub_tmp = w + 3 * seg;
    size_t cub_bytes = need - 3 * seg;
    ALS_CUDA(cudaMemsetAsync(n_nonempty, 0, sizeof(int64_t), s));
    row_keys_kernel<<<grid_for(n_rows, 256), 256, 0, s>>>(indptr, n_rows, keys, ids,
                                                          reinterpret_cast<unsigned long long*>(n_nonempty));
    ALS_LAUNCH_CHECK();
    ALS_CUDA(cub::DeviceRadixSort::SortPairs(cub_tmp, cub_bytes, (const uint32_t*)keys, keys_out, (const int32_t*)ids,
                                             order, (size_t)n_rows, 0, 32, s));
    g_launch_count += 4;
    return B200ALS_OK;
}

int b200als_fold_split_workspace_bytes(int64_t nnz, size_t* bytes_host) {
    ALS_CHECK_ARG(bytes_host && nnz >= 0, "bad args");
    const size_t words = (size_t)ceil_div64(nnz, 32);
    const size_t blocks = (size_t)ceil_div64(nnz, FOLD_BLOCK_ENTRIES);
    *bytes_host = align_up(words * 4, 256) + align_up(blocks * 8, 256) + 256;
    return B200ALS_OK;
}

int b200als_fold_split(const int64_t* flat_sorted, const float* vals, int64_t nnz, const int64_t* val_idx,
                       int64_t n_val, int64_t* train_flat, float* train_vals, float* val_vals, int64_t* n_train,
                       int32_t* status, void* workspace, size_t workspace_bytes, b200als_stream_t stream) {
    ALS_CHECK_ARG(nnz > 0 && n_val >= 0 && flat_sorted && vals && train_flat && train_vals && n_train && status,
                  "null pointer / empty observed set");
    ALS_CHECK_ARG(n_val == 0 || (val_idx && val_vals), "null validation arrays");
    size_t need = 0;
    b200als_fold_split_workspace_bytes(nnz, &need);
    if (!workspace || workspace_bytes < need) {
        set_error("b200als_fold_split: workspace %zu < required %zu", workspace_bytes, need);
        return B200ALS_E_WORKSPACE;
    }
    cudaStream_t s = as_stream(stream);
    const int64_t words = ceil_div64(nnz, 32), blocks = ceil_div64(nnz, FOLD_BLOCK_ENTRIES);
    uint32_t* bits = static_cast<uint32_t*>(workspace);
    int64_t* block_cnt = reinterpret_cast<int64_t*>(static_cast<char*>(workspace) + align_up((size_t)words * 4, 256));
    ALS_CUDA(cudaMemsetAsync(bits, 0, (size_t)words * 4, s));
    ALS_CUDA(cudaMemsetAsync(status, 0, 2 * sizeof(int32_t), s));
    if (n_val > 0) {
        fold_mark_kernel<<<grid_for(n_val, 256), 256, 0, s>>>(flat_sorted, vals, nnz, val_idx, n_val, bits, val_vals,
                                                              status);
        ALS_LAUNCH_CHECK();
    }
    fold_count_kernel<<<(unsigned)blocks, 256, 0, s>>>(bits, nnz, block_cnt);
    ALS_LAUNCH_CHECK();
    scan_inclusive_i64_kernel<<<1, 1024, 0, s>>>(block_cnt, blocks);
    ALS_LAUNCH_CHECK();
    fold_scatter_kernel<<<(unsigned)blocks, 256, 0, s>>>(flat_sorted, vals, nnz, bits, block_cnt, train_flat, train_vals,
                                                         n_train);
    ALS_LAUNCH_CHECK();
    return B200ALS_OK;
}

}  // extern "C"

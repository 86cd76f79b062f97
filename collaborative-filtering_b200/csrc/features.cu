// features.cu — the item-feature passes around the ALS path, on the device.
//
// N4  b200als_normalize_feature: `normalize_feature` of scripts/prepare_features.py:131-191 — column-median
//     imputation of NaN / Inf, then none | row_l1 | row_l2 | col_zscore | col_minmax — in float64 like numpy
//     (the reference computes in the input's dtype and casts at the end, :51).
// N3  b200als_similarity_topk_rows: the cosine similarity + per-row top-k of scripts/als.py:224-236 with a
//     DOCUMENTED, deterministic tie rule (the reference's np.argpartition picks among equal values in an
//     implementation-defined way, SURVEY.md H2, which is why the host build stays the default): row i keeps the
//     topk largest s_ij, j != i; among equal values the SMALLER column index wins.  Opt-in (ALS.graph_build = "device").
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace b200als {

__device__ __forceinline__ double block_sum(double v, double* red) {
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
    return t;
}

// column j of X (n x d, row-major) -> keys[0..n): finite values as they are, everything else +inf; *n_finite
__global__ void __launch_bounds__(256) col_extract_kernel(const double* __restrict__ X, int64_t n, int d, int col,
                                                          double* __restrict__ keys, unsigned long long* n_finite,
                                                          unsigned long long* n_bad_total) {
    unsigned cnt = 0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const double v = X[i * d + col];
        const bool fin = isfinite(v);
        keys[i] = fin ? v : __longlong_as_double(0x7ff0000000000000ll);
        cnt += fin;
    }
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(n_finite, (unsigned long long)cnt);
    (void)n_bad_total;
}

// np.nanmedian of the column from its sorted finite prefix (mean of the two middle values), 0 for an all-NaN column
__global__ void col_median_kernel(const double* __restrict__ sorted, const unsigned long long* n_finite, double* med) {
    const unsigned long long c = *n_finite;
    if (c == 0) { *med = 0.0; return; }
    const double a = sorted[(c - 1) / 2], b = sorted[c / 2];
    *med = (c & 1) ? a : (a + b) * 0.5;
}

__global__ void __launch_bounds__(256) impute_kernel(double* __restrict__ X, int64_t n, int d, const double* __restrict__ med) {
    const int64_t total = n * d, stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride)
        if (!isfinite(X[e])) X[e] = med[e % d];
}

__global__ void __launch_bounds__(256) count_nonfinite_kernel(const double* __restrict__ X, int64_t total,
                                                              unsigned long long* out) {
    unsigned cnt = 0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) cnt += !isfinite(X[e]);
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(out, (unsigned long long)cnt);
}

// method 1 (row_l1) / 2 (row_l2): one warp per row
__global__ void __launch_bounds__(256) row_norm_kernel(double* __restrict__ X, int64_t n, int d, int method, double eps) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t i = warp; i < n; i += nwarps) {
        double s = 0.0;
        for (int c = lane; c < d; c += 32) {
            const double v = X[i * d + c];
            s += method == 1 ? fabs(v) : v * v;
        }
        s = warp_sum(s);
        if (method == 2) s = sqrt(s);
        s = s > eps ? s : eps;                       // np.maximum(s, eps)
        for (int c = lane; c < d; c += 32) X[i * d + c] /= s;
    }
}

// method 3 (col_zscore) / 4 (col_minmax): one block per column
__global__ void __launch_bounds__(256) col_norm_kernel(double* __restrict__ X, int64_t n, int d, int method, double eps) {
    __shared__ double red[8];
    __shared__ double mn_s[8], mx_s[8];
    const int c = blockIdx.x;
    if (method == 3) {
        double s = 0.0;
        for (int64_t i = threadIdx.x; i < n; i += blockDim.x) s += X[i * d + c];
        const double mu = block_sum(s, red) / (double)n;
        double q = 0.0;
        for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
            const double t = X[i * d + c] - mu;
            q += t * t;
        }
        double sd = sqrt(block_sum(q, red) / (double)n);         // np.std: population standard deviation
        if (sd < eps) sd = 1.0;
        for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
            const double v = (X[i * d + c] - mu) / sd;
            X[i * d + c] = isfinite(v) ? v : 0.0;
        }
    } else {
        double mn = __longlong_as_double(0x7ff0000000000000ll), mx = -mn;
        for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
            const double v = X[i * d + c];
            mn = v < mn ? v : mn;
            mx = v > mx ? v : mx;
        }
        for (int o = 16; o > 0; o >>= 1) {
            const double a = __shfl_xor_sync(0xffffffffu, mn, o), b = __shfl_xor_sync(0xffffffffu, mx, o);
            mn = a < mn ? a : mn;
            mx = b > mx ? b : mx;
        }
        if ((threadIdx.x & 31) == 0) { mn_s[threadIdx.x >> 5] = mn; mx_s[threadIdx.x >> 5] = mx; }
        __syncthreads();
        for (int w = 0; w < 8; ++w) {
            mn = mn_s[w] < mn ? mn_s[w] : mn;
            mx = mx_s[w] > mx ? mx_s[w] : mx;
        }
        double rng = mx - mn;
        rng = rng > eps ? rng : eps;
        for (int64_t i = threadIdx.x; i < n; i += blockDim.x) X[i * d + c] = (X[i * d + c] - mn) / rng;
    }
}

// ---- N3 -------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) unit_rows_kernel(const float* __restrict__ X, int64_t n, int d, float eps,
                                                        float* __restrict__ Xn) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t i = warp; i < n; i += nwarps) {
        float s = 0.0f;
        for (int c = lane; c < d; c += 32) s = fmaf(X[i * d + c], X[i * d + c], s);
        s = sqrtf(warp_sum(s)) + eps;                // als.py:224: norms = sqrt(sum(X*X)) + S_eps
        for (int c = lane; c < d; c += 32) Xn[i * d + c] = X[i * d + c] / s;
    }
}

__device__ __forceinline__ uint32_t order_key(float v) {           // larger float -> larger unsigned key
    const uint32_t b = __float_as_uint(v);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

// One CTA per row i (grid-stride): s_j = Xn_i . Xn_j (s_i = 0, als.py:226-227), radix-select the topk-th largest value,
// keep everything above it and, among the entries equal to it, the smallest column indices; the kept (j, s_j) go
// out in ascending j.  scratch: gridDim.x rows of n floats.
__global__ void __launch_bounds__(256) topk_rows_kernel(const float* __restrict__ Xn, int64_t n, int d, int topk,
                                                        float* __restrict__ scratch, int32_t* __restrict__ nbr_idx,
                                                        float* __restrict__ nbr_val) {
    extern __shared__ float xi_s[];                  // [d]
    __shared__ unsigned hist[256];
    __shared__ uint32_t prefix_s, need_s;
    __shared__ unsigned warp_cnt[8][2];
    __shared__ unsigned base_s[2];
    float* srow = scratch + (size_t)blockIdx.x * n;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    for (int64_t i = blockIdx.x; i < n; i += gridDim.x) {
        __syncthreads();
        for (int c = tid; c < d; c += 256) xi_s[c] = Xn[i * d + c];
        __syncthreads();
        for (int64_t j = tid; j < n; j += 256) {
            float s = 0.0f;
            const float* xj = Xn + j * d;
            for (int c = 0; c < d; ++c) s = fmaf(xi_s[c], xj[c], s);
            srow[j] = (j == i) ? 0.0f : s;
        }
        // radix select (MSB first) of the topk-th largest key among j != i ... the diagonal (value 0) takes part, as in
        // the reference, where argpartition runs over the whole row with a zeroed diagonal
        uint32_t prefix = 0, mask = 0;
        uint32_t need = (uint32_t)topk;              // rank (1-based, from the top) still to locate inside the prefix bucket
        for (int shift = 24; shift >= 0; shift -= 8) {
            __syncthreads();
            hist[tid] = 0;
            __syncthreads();
            for (int64_t j = tid; j < n; j += 256) {
                const uint32_t key = order_key(srow[j]);
                if ((key & mask) == prefix) atomicAdd(&hist[(key >> shift) & 255u], 1u);
            }
            __syncthreads();
            if (tid == 0) {
                uint32_t nd = need;
                int b = 255;
                for (; b > 0; --b) {
                    if (hist[b] >= nd) break;
                    nd -= hist[b];
                }
                prefix_s = prefix | ((uint32_t)b << shift);
                need_s = nd;
            }
            __syncthreads();
            prefix = prefix_s;
            need = need_s;
            mask |= 255u << shift;
        }
        // prefix = key of the topk-th largest value; `need` = how many entries EQUAL to it are kept (smallest j first)
        const uint32_t kth = prefix;
        unsigned out_base = 0, tie_base = 0;
        for (int64_t j0 = 0; j0 < n; j0 += 256) {
            const int64_t j = j0 + tid;
            const uint32_t key = j < n ? order_key(srow[j]) : 0u;
            const bool gt = j < n && key > kth, eq = j < n && key == kth;
            const unsigned bg = __ballot_sync(0xffffffffu, gt), be = __ballot_sync(0xffffffffu, eq);
            if (lane == 0) { warp_cnt[wid][0] = __popc(bg); warp_cnt[wid][1] = __popc(be); }
            __syncthreads();
            const unsigned lm = (1u << lane) - 1u;
            unsigned gt_before = __popc(bg & lm), ties_before = __popc(be & lm);      // inside this 256-chunk
            for (int w = 0; w < wid; ++w) { gt_before += warp_cnt[w][0]; ties_before += warp_cnt[w][1]; }
            const unsigned room = need > tie_base ? need - tie_base : 0u;             // ties this chunk may still keep
            const bool keep_eq = eq && ties_before < room;
            const unsigned pos = out_base + gt_before + (ties_before < room ? ties_before : room);
            if ((gt || keep_eq) && pos < (unsigned)topk) {
                nbr_idx[i * topk + pos] = (int32_t)j;
                nbr_val[i * topk + pos] = srow[j];
            }
            __syncthreads();
            if (tid == 0) {
                unsigned g = 0, e = 0;
                for (int w = 0; w < 8; ++w) { g += warp_cnt[w][0]; e += warp_cnt[w][1]; }
                const unsigned room = need > tie_base ? need - tie_base : 0u;
                base_s[0] = out_base + g + (e < room ? e : room);
                base_s[1] = tie_base + e;
            }
            __syncthreads();
            out_base = base_s[0];
            tie_base = base_s[1];
        }
    }
}

}  // namespace b200als

using namespace b200als;

extern "C" {

int b200als_normalize_workspace_bytes(int64_t n, size_t* bytes_host) {
    ALS_CHECK_ARG(bytes_host && n > 0, "bad shape");
    size_t cb = 0;
    if (cub::DeviceRadixSort::SortKeys(nullptr, cb, (const double*)nullptr, (double*)nullptr, (size_t)n) != cudaSuccess) {
        set_error("cub temp-storage query failed");
        return B200ALS_E_CUDA;
    }
    *bytes_host = align_up((size_t)n * 8, 256) * 2 + align_up(cb, 256) + 256;
    return B200ALS_OK;
}

int b200als_normalize_feature(double* X, int64_t n, int32_t d, int32_t method, int32_t impute, double eps,
                              int64_t* n_nonfinite, double* med_work, void* workspace, size_t workspace_bytes,
                              b200als_stream_t stream) {
    ALS_CHECK_ARG(X && n > 0 && d >= 1 && n_nonfinite, "bad arguments");
    ALS_CHECK_ARG(method >= 0 && method <= 4 && (impute == 0 || impute == 1), "unknown method / impute");
    cudaStream_t s = as_stream(stream);
    auto* bad = reinterpret_cast<unsigned long long*>(n_nonfinite);
    ALS_CUDA(cudaMemsetAsync(bad, 0, sizeof(unsigned long long), s));
    const int grid = (int)(ceil_div64(n * d, 256) < (int64_t)kNumSMs * 8 ? ceil_div64(n * d, 256) : (int64_t)kNumSMs * 8);
    count_nonfinite_kernel<<<grid, 256, 0, s>>>(X, n * d, bad);
    ALS_LAUNCH_CHECK();
    if (impute == 1) {
        size_t need = 0;
        int rc = b200als_normalize_workspace_bytes(n, &need);
        if (rc) return rc;
        ALS_CHECK_ARG(med_work, "imputation needs med_work (d + 1 doubles)");
        if (!workspace || workspace_bytes < need) {
            set_error("b200als_normalize_feature: workspace %zu < required %zu", workspace_bytes, need);
            return B200ALS_E_WORKSPACE;
        }
        char* w = static_cast<char*>(workspace);
        double* keys = reinterpret_cast<double*>(w);
        double* sorted = reinterpret_cast<double*>(w + align_up((size_t)n * 8, 256));
        auto* cnt = reinterpret_cast<unsigned long long*>(w + 2 * align_up((size_t)n * 8, 256));
        void* cub_tmp = w + 2 * align_up((size_t)n * 8, 256) + 256;
        size_t cb = workspace_bytes - (2 * align_up((size_t)n * 8, 256) + 256);
        const int g2 = (int)(ceil_div64(n, 256) < (int64_t)kNumSMs * 8 ? ceil_div64(n, 256) : (int64_t)kNumSMs * 8);
        for (int c = 0; c < d; ++c) {
            ALS_CUDA(cudaMemsetAsync(cnt, 0, sizeof(unsigned long long), s));
            col_extract_kernel<<<g2, 256, 0, s>>>(X, n, d, c, keys, cnt, nullptr);
            ALS_LAUNCH_CHECK();
            ALS_CUDA(cub::DeviceRadixSort::SortKeys(cub_tmp, cb, (const double*)keys, sorted, (size_t)n, 0, 64, s));
            col_median_kernel<<<1, 1, 0, s>>>(sorted, cnt, med_work + c);
            ALS_LAUNCH_CHECK();
        }
        impute_kernel<<<grid, 256, 0, s>>>(X, n, d, med_work);
        ALS_LAUNCH_CHECK();
    }
    if (method == 1 || method == 2) {
        const int g3 = (int)(ceil_div64(n * 32, 256) < (int64_t)kNumSMs * 8 ? ceil_div64(n * 32, 256) : (int64_t)kNumSMs * 8);
        row_norm_kernel<<<g3, 256, 0, s>>>(X, n, d, method, eps);
        ALS_LAUNCH_CHECK();
    } else if (method == 3 || method == 4) {
        col_norm_kernel<<<d, 256, 0, s>>>(X, n, d, method, eps);
        ALS_LAUNCH_CHECK();
    }
    return B200ALS_OK;
}

int b200als_similarity_topk_rows(const float* X, int64_t n, int32_t d, int32_t topk, float eps, float* Xn_work,
                                 float* scratch, int64_t scratch_rows, int32_t* nbr_idx, float* nbr_val,
                                 b200als_stream_t stream) {
    ALS_CHECK_ARG(X && Xn_work && scratch && nbr_idx && nbr_val, "null pointer");
    ALS_CHECK_ARG(n > 1 && d >= 1 && d <= 4096 && topk >= 1 && topk < n && scratch_rows >= 1, "bad shape (needs 1 <= topk < n)");
    cudaStream_t s = as_stream(stream);
    const int g1 = (int)(ceil_div64(n * 32, 256) < (int64_t)kNumSMs * 8 ? ceil_div64(n * 32, 256) : (int64_t)kNumSMs * 8);
    unit_rows_kernel<<<g1, 256, 0, s>>>(X, n, d, eps, Xn_work);
    ALS_LAUNCH_CHECK();
    int64_t grid = n < scratch_rows ? n : scratch_rows;
    topk_rows_kernel<<<(unsigned)grid, 256, sizeof(float) * d, s>>>(Xn_work, n, d, topk, scratch, nbr_idx, nbr_val);
    ALS_LAUNCH_CHECK();
    return B200ALS_OK;
}

}  // extern "C"

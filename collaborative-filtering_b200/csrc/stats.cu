// stats.cu — A5 compose Z, A10 residual / mu / RMSE / norm reductions,
// A13 predictions.  All streaming, HBM-bound kernels: coalesced 16-byte loads,
// float64 accumulators for every reduction (early stopping compares RMSE
// differences to 1e-4, als.py:283-297).
#include "common.cuh"

namespace b200als {

struct ComposeParams {
    int nf;
    int skip;
    const float* X[B200ALS_MAX_FEATURES];
    const float* W[B200ALS_MAX_FEATURES];
    int d[B200ALS_MAX_FEATURES];
};

// Z[i, c] = V[i, c] + sum_f sum_a X_f[i, a] * W_f[a, c]; one thread per (i, c4) float4.
__global__ void compose_z_kernel(const float* __restrict__ V, int64_t ldv, int64_t n, int k, ComposeParams cp,
                                 int64_t ldw, float* __restrict__ Z, int64_t ldz) {
    const int kq = (k + 3) >> 2;
    const int64_t total = n * kq;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
        const int64_t i = e / kq;
        const int c = (int)(e - i * kq) * 4;
        float4 z = *reinterpret_cast<const float4*>(V + i * ldv + c);
        for (int f = 0; f < cp.nf; ++f) {
            if (f == cp.skip) continue;
            const float* x = cp.X[f] + i * cp.d[f];
            const float* w = cp.W[f] + c;
            for (int a = 0; a < cp.d[f]; ++a) {
                const float xa = x[a];
                if (xa != 0.0f) {
                    const float4 w4 = *reinterpret_cast<const float4*>(w + (size_t)a * ldw);
                    z.x = fmaf(xa, w4.x, z.x); z.y = fmaf(xa, w4.y, z.y);
                    z.z = fmaf(xa, w4.z, z.z); z.w = fmaf(xa, w4.w, z.w);
                }
            }
        }
        *reinterpret_cast<float4*>(Z + i * ldz + c) = z;
    }
}

__device__ __forceinline__ void block_atomic_add2(double a, double b, double* out0, double* out1) {
    __shared__ double sh[2][32];
    a = warp_sum(a);
    b = warp_sum(b);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) { sh[0][w] = a; sh[1][w] = b; }
    __syncthreads();
    if (w == 0) {
        const int nw = (blockDim.x + 31) >> 5;
        a = lane < nw ? sh[0][lane] : 0.0;
        b = lane < nw ? sh[1][lane] : 0.0;
        a = warp_sum(a);
        b = warp_sum(b);
        if (lane == 0) {
            atomicAdd(out0, a);
            if (out1) atomicAdd(out1, b);
        }
    }
}

// warp per user row; 4 groups of 8 lanes, each group one rating at a time; lane g8
// holds float4 slices g8, g8+8, ... of U[u] in registers.
template <int KQ_MAX>   // max float4 slices per lane (k <= 32*KQ_MAX)
__global__ void __launch_bounds__(256) residual_stats_kernel(int64_t m, const int64_t* __restrict__ indptr,
                                                             const int32_t* __restrict__ indices,
                                                             const float* __restrict__ vals, int k,
                                                             const float* __restrict__ U, int64_t ldu,
                                                             const float* __restrict__ Z, int64_t ldz,
                                                             const float* __restrict__ b_u,
                                                             const float* __restrict__ b_i, double* acc) {
    const int lane = threadIdx.x & 31, g8 = lane & 7, grp = lane >> 3;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int kq = (k + 3) >> 2;
    double sd = 0.0, sd2 = 0.0;
    for (int64_t u = warp; u < m; u += nwarps) {
        const int64_t beg = indptr[u], end = indptr[u + 1];
        if (end == beg) continue;
        float4 uu[KQ_MAX];
#pragma unroll
        for (int s = 0; s < KQ_MAX; ++s) {
            const int q = g8 + s * 8;
            uu[s] = (q < kq) ? *reinterpret_cast<const float4*>(U + u * ldu + q * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        const float bu = b_u[u];
        for (int64_t e0 = beg; e0 < end; e0 += 4) {   // warp-uniform trip count (shuffles below)
            const int64_t e = e0 + grp;
            const bool live = e < end;
            const int32_t i = live ? ld_stream_s32(indices + e) : 0;
            const float r = live ? ld_stream_f32(vals + e) : 0.0f;
            float dot = 0.0f;
#pragma unroll
            for (int s = 0; s < KQ_MAX; ++s) {
                const int q = g8 + s * 8;
                if (q < kq) {
                    const float4 z = *reinterpret_cast<const float4*>(Z + (size_t)i * ldz + q * 4);
                    dot = fmaf(uu[s].x, z.x, dot); dot = fmaf(uu[s].y, z.y, dot);
                    dot = fmaf(uu[s].z, z.z, dot); dot = fmaf(uu[s].w, z.w, dot);
                }
            }
            dot += __shfl_xor_sync(0xffffffffu, dot, 4);
            dot += __shfl_xor_sync(0xffffffffu, dot, 2);
            dot += __shfl_xor_sync(0xffffffffu, dot, 1);
            if (live && g8 == 0) {
                const float d = r - (dot + bu + b_i[i]);
                sd += (double)d;
                sd2 += (double)d * (double)d;
            }
        }
    }
    block_atomic_add2(sd, sd2, acc, acc + 1);
}

__global__ void __launch_bounds__(256) sumsq_kernel(const float* __restrict__ x, int64_t rows, int64_t cols, int64_t ld,
                                                    double* acc) {
    const int64_t total = rows * cols;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    double s = 0.0;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
        const int64_t r = e / cols;
        const float v = x[r * ld + (e - r * cols)];
        s += (double)v * (double)v;
    }
    block_atomic_add2(s, 0.0, acc, nullptr);
}

__global__ void finalize_iter_kernel(const double* __restrict__ acc, int64_t nnz, float* mu_f32, double* hist) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        const double N = (double)nnz;
        const double mu = acc[0] / N;                       // als.py:507
        double sse = acc[1] - acc[0] * acc[0] / N;          // sum (d - mu)^2
        if (sse < 0.0) sse = 0.0;
        hist[0] = sqrt(sse / N);                            // als.py:512
        hist[1] = sqrt(acc[2]);
        hist[2] = sqrt(acc[3]);
        hist[3] = sqrt(acc[4]);
        hist[4] = sqrt(acc[5]);
        hist[5] = mu;
        *mu_f32 = (float)mu;
    }
}

// acc[0..1] are sums of d' = d - mu_old (the item step ran with mu_old, als.py:447): shift back
__global__ void finalize_iter_fused_kernel(double* __restrict__ acc, const double* __restrict__ reset, int64_t nnz,
                                           float* mu_f32, double* hist) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        const double N = (double)nnz;
        const double delta = acc[0] / N;
        const double mu = (double)*mu_f32 + delta;          // als.py:507
        double sse = acc[1] - acc[0] * delta;               // sum (d' - delta)^2
        if (sse < 0.0) sse = 0.0;
        hist[0] = sqrt(sse / N);                            // als.py:512
        hist[1] = sqrt(acc[2]);
        hist[2] = sqrt(acc[3]);
        hist[3] = sqrt(acc[4]);
        hist[4] = sqrt(acc[5]);
        hist[5] = mu;
        *mu_f32 = (float)mu;
        if (reset)                                          // start values of the next iteration's accumulators
            for (int i = 0; i < 6; ++i) acc[i] = reset[i];
    }
}

__global__ void f64_to_f32_kernel(const double* __restrict__ src, float* __restrict__ dst, int64_t count) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < count; e += stride) dst[e] = (float)src[e];
}

// out (m, n) float64 = U Z^T + mu + b_u + b_i.  64x64 tile / CTA, 4x4 / thread.
__global__ void __launch_bounds__(256) predict_dense_kernel(const float* __restrict__ U, int64_t ldu,
                                                            const float* __restrict__ Z, int64_t ldz,
                                                            const float* __restrict__ b_u,
                                                            const float* __restrict__ b_i,
                                                            const float* __restrict__ mu_p, int64_t m, int64_t n, int k,
                                                            double* __restrict__ out) {
    __shared__ float As[16][64 + 4];
    __shared__ float Bs[16][64 + 4];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int64_t u0 = (int64_t)blockIdx.y * 64, i0 = (int64_t)blockIdx.x * 64;
    float acc[4][4] = {};
    for (int kk = 0; kk < k; kk += 16) {
        // 64 rows x 16 cols each: 1024 elements / 256 threads = 4
        for (int e = threadIdx.x; e < 64 * 16; e += 256) {
            const int r = e >> 4, c = e & 15;
            const int64_t uu = u0 + r, ii = i0 + r;
            As[c][r] = (uu < m && kk + c < k) ? U[uu * ldu + kk + c] : 0.0f;
            Bs[c][r] = (ii < n && kk + c < k) ? Z[ii * ldz + kk + c] : 0.0f;
        }
        __syncthreads();
#pragma unroll
        for (int c = 0; c < 16; ++c) {
            float a[4], b[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) { a[j] = As[c][ty * 4 + j]; b[j] = Bs[c][tx * 4 + j]; }
#pragma unroll
            for (int x = 0; x < 4; ++x)
#pragma unroll
                for (int y = 0; y < 4; ++y) acc[x][y] = fmaf(a[x], b[y], acc[x][y]);
        }
        __syncthreads();
    }
    const float mu = *mu_p;
#pragma unroll
    for (int x = 0; x < 4; ++x) {
        const int64_t uu = u0 + ty * 4 + x;
        if (uu >= m) continue;
        const float bu = b_u[uu];
#pragma unroll
        for (int y = 0; y < 4; ++y) {
            const int64_t ii = i0 + tx * 4 + y;
            if (ii < n) out[uu * n + ii] = (double)(acc[x][y] + mu + bu + b_i[ii]);
        }
    }
}

// 8 lanes per pair
__global__ void __launch_bounds__(256) predict_pairs_kernel(const float* __restrict__ U, int64_t ldu,
                                                            const float* __restrict__ Z, int64_t ldz,
                                                            const float* __restrict__ b_u,
                                                            const float* __restrict__ b_i,
                                                            const float* __restrict__ mu_p,
                                                            const int32_t* __restrict__ rows,
                                                            const int32_t* __restrict__ cols, int64_t count, int k,
                                                            float* __restrict__ out) {
    const int g8 = threadIdx.x & 7;
    const int64_t grp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    const int64_t ngrp = ((int64_t)gridDim.x * blockDim.x) >> 3;
    const int kq = (k + 3) >> 2;
    const float mu = *mu_p;
    // every lane of a warp runs the same trip count (tail groups compute a dummy pair)
    const int64_t iters = (count + ngrp - 1) / ngrp;
    for (int64_t it = 0; it < iters; ++it) {
        const int64_t e = grp + it * ngrp;
        const bool live = e < count;
        const int32_t u = live ? rows[e] : 0, i = live ? cols[e] : 0;
        float dot = 0.0f;
        for (int q = g8; q < kq; q += 8) {
            const float4 a = *reinterpret_cast<const float4*>(U + (size_t)u * ldu + q * 4);
            const float4 z = *reinterpret_cast<const float4*>(Z + (size_t)i * ldz + q * 4);
            dot = fmaf(a.x, z.x, dot); dot = fmaf(a.y, z.y, dot);
            dot = fmaf(a.z, z.z, dot); dot = fmaf(a.w, z.w, dot);
        }
        dot += __shfl_xor_sync(0xffffffffu, dot, 4);
        dot += __shfl_xor_sync(0xffffffffu, dot, 2);
        dot += __shfl_xor_sync(0xffffffffu, dot, 1);
        if (live && g8 == 0) out[e] = dot + mu + b_u[u] + b_i[i];
    }
}

// N1 — squared error over flat indices u*n+i without the dense prediction (tune_params.py:147-167), split by an
// item bin when given (evaluate_models.py:173-191, 263-274).  8 lanes per entry (16-byte factor loads), float64
// per-bin sums per warp in shared memory, one atomic per bin and block.
constexpr int SSE_MAX_BINS = 32;
__global__ void __launch_bounds__(256) sse_on_flat_kernel(const float* __restrict__ U, int64_t ldu,
                                                          const float* __restrict__ Z, int64_t ldz,
                                                          const float* __restrict__ b_u, const float* __restrict__ b_i,
                                                          const float* __restrict__ mu_p,
                                                          const int64_t* __restrict__ flat, const float* __restrict__ y,
                                                          int64_t count, int64_t n, int k,
                                                          const int32_t* __restrict__ item_bin, int n_bins,
                                                          double* __restrict__ sse, long long* __restrict__ cnt) {
    __shared__ double sh_sse[SSE_MAX_BINS];
    __shared__ unsigned long long sh_cnt[SSE_MAX_BINS];
    if (threadIdx.x < SSE_MAX_BINS) { sh_sse[threadIdx.x] = 0.0; sh_cnt[threadIdx.x] = 0ull; }
    __syncthreads();
    const int g8 = threadIdx.x & 7;
    const int64_t grp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    const int64_t ngrp = ((int64_t)gridDim.x * blockDim.x) >> 3;
    const int kq = (k + 3) >> 2;
    const float mu = *mu_p;
    const int64_t iters = (count + ngrp - 1) / ngrp;   // same trip count for every lane of a warp
    for (int64_t it = 0; it < iters; ++it) {
        const int64_t e = grp + it * ngrp;
        const bool live = e < count;
        const int64_t f = live ? flat[e] : 0;
        const int64_t u = f / n, i = f - u * n;        // als.py's row-major flat convention (create_folds.py:73-84)
        float dot = 0.0f;
        for (int q = g8; q < kq; q += 8) {
            const float4 a = *reinterpret_cast<const float4*>(U + (size_t)u * ldu + q * 4);
            const float4 z = *reinterpret_cast<const float4*>(Z + (size_t)i * ldz + q * 4);
            dot = fmaf(a.x, z.x, dot); dot = fmaf(a.y, z.y, dot);
            dot = fmaf(a.z, z.z, dot); dot = fmaf(a.w, z.w, dot);
        }
        dot += __shfl_xor_sync(0xffffffffu, dot, 4);
        dot += __shfl_xor_sync(0xffffffffu, dot, 2);
        dot += __shfl_xor_sync(0xffffffffu, dot, 1);
        if (live && g8 == 0) {
            const double d = (double)y[e] - (double)(dot + mu + b_u[u] + b_i[i]);
            const int b = item_bin ? item_bin[i] : 0;
            if (b >= 0 && b < n_bins) {
                atomicAdd(&sh_sse[b], d * d);
                atomicAdd(&sh_cnt[b], 1ull);
            }
        }
    }
    __syncthreads();
    if (threadIdx.x < n_bins && sh_cnt[threadIdx.x]) {
        atomicAdd(&sse[threadIdx.x], sh_sse[threadIdx.x]);
        atomicAdd(reinterpret_cast<unsigned long long*>(cnt) + threadIdx.x, sh_cnt[threadIdx.x]);
    }
}

static int grid_for(int64_t work, int threads, int per_sm = 8) {
    int64_t b = ceil_div64(work, threads);
    int64_t cap = (int64_t)kNumSMs * per_sm;
    return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace b200als

using namespace b200als;

extern "C" {

int b200als_compose_z(const float* V, int64_t ldv, int64_t n, int32_t k, int32_t n_features,
                      const float* const* X, const int32_t* d, const float* const* W, int64_t ldw, int32_t skip,
                      float* Z, int64_t ldz, b200als_stream_t stream) {
    ALS_CHECK_ARG(V && Z && n > 0 && k >= 1 && k <= B200ALS_MAX_K, "bad V/Z/shape");
    ALS_CHECK_ARG(n_features >= 0 && n_features <= B200ALS_MAX_FEATURES, "too many features (max 8)");
    ALS_CHECK_ARG((ldv & 3) == 0 && (ldz & 3) == 0 && ldv >= ((k + 3) & ~3) && ldz >= ((k + 3) & ~3), "ld not 16B aligned");
    ALS_CHECK_ARG(n_features == 0 || ((ldw & 3) == 0 && ldw >= ((k + 3) & ~3)), "ldw not 16B aligned");
    ComposeParams cp;
    memset(&cp, 0, sizeof(cp));
    cp.nf = n_features;
    cp.skip = skip;
    for (int f = 0; f < n_features; ++f) {
        ALS_CHECK_ARG(X && W && d && X[f] && W[f] && d[f] > 0, "null feature pointer");
        cp.X[f] = X[f]; cp.W[f] = W[f]; cp.d[f] = d[f];
    }
    cudaStream_t s = as_stream(stream);
    compose_z_kernel<<<grid_for(n * ((k + 3) >> 2), 256), 256, 0, s>>>(V, ldv, n, k, cp, ldw, Z, ldz);
    ALS_LAUNCH_CHECK();
    return B200ALS_OK;
}

int b200als_residual_stats(int64_t m, const int64_t* indptr, const int32_t* indices, const float* vals, int32_t k,
                           const float* U, int64_t ldu, const float* Z, int64_t ldz, const float* b_u,
                           const float* b_i, double* acc, b200als_stream_t stream) {
    ALS_CHECK_ARG(m > 0 && indptr && indices && vals && U && Z && b_u && b_i && acc, "null pointer");
    ALS_CHECK_ARG(k >= 1 && k <= B200ALS_MAX_K, "k out of range");
    ALS_CHECK_ARG((ldu & 3) == 0 && (ldz & 3) == 0, "ld not 16B aligned");
    cudaStream_t s = as_stream(stream);
    const int grid = grid_for(m * 32, 256, 8);
    if (k <= 32)
        residual_stats_kernel<1><<<grid, 256, 0, s>>>(m, indptr, indices, vals, k, U, ldu, Z, ldz, b_u, b_i, acc);
    else if (k <= 64)
        residual_stats_kernel<2><<<grid, 256, 0, s>>>(m, indptr, indices, vals, k, U, ldu, Z, ldz, b_u, b_i, acc);
    else if (k <= 128)
        residual_stats_kernel<4><<<grid, 256, 0, s>>>(m, indptr, indices, vals, k, U, ldu, Z, ldz, b_u, b_i, acc);
    else
        residual_stats_kernel<5><<<grid, 256, 0, s>>>(m, indptr, indices, vals, k, U, ldu, Z, ldz, b_u, b_i, acc);
    ALS_LAUNCH_CHECK();
    return B200ALS_OK;
}

int b200als_sumsq(const float* x, int64_t rows, int64_t cols, int64_t ld, double* acc, b200als_stream_t stream) {
    ALS_CHECK_ARG(x && acc && rows >= 0 && cols > 0 && ld >= cols, "bad args");
    if (rows == 0) return B200ALS_OK;
    sumsq_kernel<<<grid_for(rows * cols, 256, 4), 256, 0, as_stream(stream)>>>(x, rows, cols, ld, acc);
    ALS_LAUNCH_CHECK();
    return B200ALS_OK;
}

int b200als_finalize_iter(const double* acc6, int64_t nnz_total, float* mu_f32, double* hist6,
                          b200als_stream_t stream) {
    ALS_CHECK_ARG(acc6 && mu_f32 && hist6 && nnz_total > 0, "bad args");
    finalize_iter_kernel<<<1, 32, 0, as_stream(stream)>>>(acc6, nnz_total, mu_f32, hist6);
    ALS_LAUNCH_CHECK();
    return B200ALS_OK;
}

int b200als_finalize_iter_fused(double* acc6, const double* reset6, int64_t nnz_total, float* mu_f32, double* hist6,
                                b200als_stream_t stream) {
    ALS_CHECK_ARG(acc6 && mu_f32 && hist6 && nnz_total > 0, "bad args");
    finalize_iter_fused_kernel<<<1, 32, 0, as_stream(stream)>>>(acc6, reset6, nnz_total, mu_f32, hist6);
    ALS_LAUNCH_CHECK();
    return B200ALS_OK;
}

int b200als_f64_to_f32(const double* src, float* dst, int64_t count, b200als_stream_t stream) {
    ALS_CHECK_ARG(src && dst && count >= 0, "bad args");
    if (count == 0) return B200ALS_OK;
    f64_to_f32_kernel<<<grid_for(count, 256, 4), 256, 0, as_stream(stream)>>>(src, dst, count);
    ALS_LAUNCH_CHECK();
    return B200ALS_OK;
}

int b200als_predict_dense(const float* U, int64_t ldu, const float* Z, int64_t ldz, const float* b_u,
                          const float* b_i, const float* mu, int64_t m, int64_t n, int32_t k, double* out,
                          b200als_stream_t stream) {
    ALS_CHECK_ARG(U && Z && b_u && b_i && mu && out && m > 0 && n > 0, "null pointer");
    ALS_CHECK_ARG(k >= 1 && k <= B200ALS_MAX_K, "k out of range");
    dim3 grid((unsigned)ceil_div64(n, 64), (unsigned)ceil_div64(m, 64));
    ALS_CHECK_ARG(grid.y <= 65535, "m too large for the dense predictor (use b200als_predict_pairs)");
    predict_dense_kernel<<<grid, 256, 0, as_stream(stream)>>>(U, ldu, Z, ldz, b_u, b_i, mu, m, n, k, out);
    ALS_LAUNCH_CHECK();
    return B200ALS_OK;
}

int b200als_predict_pairs(const float* U, int64_t ldu, const float* Z, int64_t ldz, const float* b_u,
                          const float* b_i, const float* mu, const int32_t* rows, const int32_t* cols,
                          int64_t count, int32_t k, float* out, b200als_stream_t stream) {
    ALS_CHECK_ARG(U && Z && b_u && b_i && mu && rows && cols && out, "null pointer");
    ALS_CHECK_ARG(k >= 1 && k <= B200ALS_MAX_K && (ldu & 3) == 0 && (ldz & 3) == 0, "bad k / ld");
    if (count <= 0) return B200ALS_OK;
    predict_pairs_kernel<<<grid_for(count * 8, 256, 8), 256, 0, as_stream(stream)>>>(U, ldu, Z, ldz, b_u, b_i, mu,
                                                                                    rows, cols, count, k, out);
    ALS_LAUNCH_CHECK();
    return B200ALS_OK;
}

int b200als_sse_on_flat(const float* U, int64_t ldu, const float* Z, int64_t ldz, const float* b_u,
                        const float* b_i, const float* mu, const int64_t* flat_idx, const float* y_true,
                        int64_t count, int64_t n, int32_t k, const int32_t* item_bin, int32_t n_bins,
                        double* sse, int64_t* cnt, b200als_stream_t stream) {
    ALS_CHECK_ARG(n > 0 && n_bins >= 1 && n_bins <= SSE_MAX_BINS, "n_bins outside [1, 32]");
    if (count <= 0) return B200ALS_OK;                  // an empty fold: nothing to add (the RMSE is NaN, tune_params.py:162)
    ALS_CHECK_ARG(U && Z && b_u && b_i && mu && flat_idx && y_true && sse && cnt, "null pointer");
    ALS_CHECK_ARG(k >= 1 && k <= B200ALS_MAX_K && (ldu & 3) == 0 && (ldz & 3) == 0, "bad k / ld");
    sse_on_flat_kernel<<<grid_for(count * 8, 256, 8), 256, 0, as_stream(stream)>>>(
        U, ldu, Z, ldz, b_u, b_i, mu, flat_idx, y_true, count, n, k, item_bin, n_bins, sse,
        reinterpret_cast<long long*>(cnt));
    ALS_LAUNCH_CHECK();
    return B200ALS_OK;
}

}  // extern "C"

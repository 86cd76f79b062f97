// wstep.cu — A9: feature-projection update W_f (als.py:468-501) without the
// (N_obs, d*k) design matrix.  With X_design[e, a*k+c] = X[i_e, a] * U[u_e, c]
// (als.py:495) the normal equations factor per item:
//   A[(a,c),(a',c')] = sum_i X[i,a] X[i,a'] G_i[c,c'] + (lam + EPS) delta
//   b[(a,c)]         = sum_i X[i,a] h_i[c]
//   G_i = sum_{u in Omega_i} U_u U_u^T,   h_i = sum_u y_ui U_u = c_i - G_i zmf_i
// where y = residual_plus_self (als.py:486-489) = R - mu - b_u - b_i - <U_u, zmf_i>
// and zmf_i = V_i + sum_{g != f} x_i^g W_g (all W old: the reference never
// refreshes `residual` inside the feature loop, als.py:474-489).
// Everything is float64: lambda_w may be 0 (evaluate_models.py "no_features"),
// leaving the (d k)^2 system regularised by EPS alone (als.py:497-498).
#include <cooperative_groups.h>

#include "common.cuh"

namespace b200als {

constexpr int GR_ACC = 8;      // accumulators per thread per pass
constexpr int GR_TR = 32;      // user rows staged per tile

// CTA per item; pass p covers entries [p*2048, (p+1)*2048) of the k x k Gram.
__global__ void __launch_bounds__(256) item_gram_f64_kernel(int64_t n, const int64_t* __restrict__ cindptr,
                                                            const int32_t* __restrict__ cindices,
                                                            const float* __restrict__ cvals, int k,
                                                            const float* __restrict__ U, int64_t ldu,
                                                            const float* __restrict__ b_u,
                                                            const float* __restrict__ b_i,
                                                            const float* __restrict__ mu_p, double* __restrict__ G,
                                                            double* __restrict__ cvec) {
    extern __shared__ float sm[];
    float* rows_s = sm;                      // [GR_TR][k]
    float* q_s = sm + GR_TR * k;             // [GR_TR]
    const int tid = threadIdx.x;
    const float mu = *mu_p;
    const int kk2 = k * k;
    const int npass = (kk2 + 256 * GR_ACC - 1) / (256 * GR_ACC);
    for (int64_t i = blockIdx.x; i < n; i += gridDim.x) {
        const int64_t beg = cindptr[i], end = cindptr[i + 1];
        const float bi = b_i[i];
        double* Gi = G + (size_t)i * kk2;
        for (int p = 0; p <= npass; ++p) {   // pass npass computes c_i
            double acc[GR_ACC];
#pragma unroll
            for (int j = 0; j < GR_ACC; ++j) acc[j] = 0.0;
            for (int64_t t0 = beg; t0 < end; t0 += GR_TR) {
                const int nr = (int)((end - t0) < GR_TR ? (end - t0) : GR_TR);
                __syncthreads();
                for (int e = tid; e < nr * k; e += 256) {
                    const int rr = e / k, c = e - rr * k;
                    rows_s[rr * k + c] = U[(size_t)cindices[t0 + rr] * ldu + c];
                }
                if (tid < nr) {
                    const int32_t u = cindices[t0 + tid];
                    q_s[tid] = cvals[t0 + tid] - (mu + b_u[u] + bi);
                }
                __syncthreads();
                if (p < npass) {
#pragma unroll
                    for (int j = 0; j < GR_ACC; ++j) {
                        const int e = p * 256 * GR_ACC + j * 256 + tid;
                        if (e < kk2) {
                            const int c = e / k, c2 = e - c * k;
                            double a = acc[j];
                            for (int rr = 0; rr < nr; ++rr)
                                a = fma((double)rows_s[rr * k + c], (double)rows_s[rr * k + c2], a);
                            acc[j] = a;
                        }
                    }
                } else if (tid < k) {
                    double a = acc[0];
                    for (int rr = 0; rr < nr; ++rr) a = fma((double)rows_s[rr * k + tid], (double)q_s[rr], a);
                    acc[0] = a;
                }
            }
            if (p < npass) {
#pragma unroll
                for (int j = 0; j < GR_ACC; ++j) {
                    const int e = p * 256 * GR_ACC + j * 256 + tid;
                    if (e < kk2) Gi[e] = acc[j];
                }
            } else if (tid < k) {
                cvec[(size_t)i * k + tid] = acc[0];
            }
        }
    }
}

// h[i, c] = cvec[i, c] - sum_c' G_i[c, c'] * zmf[i, c']      (warp per (i, c))
__global__ void __launch_bounds__(256) wstep_h_kernel(int64_t n, int k, const double* __restrict__ G,
                                                      const double* __restrict__ cvec,
                                                      const float* __restrict__ Zmf, int64_t ldz,
                                                      double* __restrict__ h) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t e = warp; e < n * k; e += nwarps) {
        const int64_t i = e / k;
        const int c = (int)(e - i * k);
        const double* g = G + ((size_t)i * k + c) * k;
        double a = 0.0;
        for (int c2 = lane; c2 < k; c2 += 32) a = fma(g[c2], (double)Zmf[(size_t)i * ldz + c2], a);
        a = warp_sum(a);
        if (lane == 0) h[e] = cvec[e] - a;
    }
}

// b[a*k + c] = sum_i X[i, a] h[i, c]; block per a: lane = c (32 columns per pass), the 8 warps stride the items and
// their partial sums are combined in shared memory (a thread-per-c loop over all n items was 1.5 ms per call at the
// bundled shape: 4980 dependent loads per thread on 32 threads per block)
__global__ void __launch_bounds__(256) wstep_b_kernel(int64_t n, int k, int d, const float* __restrict__ X,
                                                      const double* __restrict__ h, double* __restrict__ b) {
    __shared__ double part[8][32];
    const int a = blockIdx.x, lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int c0 = 0; c0 < k; c0 += 32) {
        const int c = c0 + lane;
        double s0 = 0.0, s1 = 0.0;
        int64_t i = w;
        for (; i + 8 < n; i += 16) {                 // two independent chains per thread
            const float x0 = X[i * d + a], x1 = X[(i + 8) * d + a];
            if (c < k) {
                if (x0 != 0.0f) s0 = fma((double)x0, h[i * k + c], s0);
                if (x1 != 0.0f) s1 = fma((double)x1, h[(i + 8) * k + c], s1);
            }
        }
        for (; i < n; i += 8) {
            const float x0 = X[i * d + a];
            if (c < k && x0 != 0.0f) s0 = fma((double)x0, h[i * k + c], s0);
        }
        __syncthreads();
        part[w][lane] = s0 + s1;
        __syncthreads();
        if (w == 0 && c < k) {
            double t = 0.0;
            for (int ww = 0; ww < 8; ++ww) t += part[ww][lane];
            b[(size_t)a * k + c] = t;
        }
    }
}

// A[(a,c),(a',c')] += sum_{i in split} X[i,a] X[i,a'] G_i[c,c']
// grid = (ceil(k*k/256), d, nsplit); thread = (c,c'); DCH accumulators over a'.
template <int DCH>
__global__ void __launch_bounds__(256) wstep_A_kernel(int64_t n, int k, int d, const float* __restrict__ X,
                                                      const double* __restrict__ G, double* __restrict__ A) {
    const int e = blockIdx.x * 256 + threadIdx.x;
    const int kk2 = k * k;
    const int a = blockIdx.y;
    const int64_t per = (n + gridDim.z - 1) / gridDim.z;
    const int64_t i0 = (int64_t)blockIdx.z * per, i1 = (i0 + per < n) ? i0 + per : n;
    const int c = e / k, c2 = e - c * k;
    const int N = d * k;
    for (int a0 = 0; a0 <= a; a0 += DCH) {       // lower block triangle: a' <= a
        double acc[DCH];
#pragma unroll
        for (int j = 0; j < DCH; ++j) acc[j] = 0.0;
        for (int64_t i = i0; i < i1; ++i) {
            const float xa = X[i * d + a];       // block-uniform
            if (xa == 0.0f) continue;
            if (e < kk2) {
                const double g = (double)xa * G[(size_t)i * kk2 + e];
#pragma unroll
                for (int j = 0; j < DCH; ++j) {
                    const int a2 = a0 + j;
                    if (a2 <= a) acc[j] = fma(g, (double)X[i * d + a2], acc[j]);
                }
            }
        }
        if (e < kk2) {
#pragma unroll
            for (int j = 0; j < DCH; ++j) {
                const int a2 = a0 + j;
                if (a2 <= a && acc[j] != 0.0) atomicAdd(&A[(size_t)(a * k + c) * N + a2 * k + c2], acc[j]);
            }
        }
    }
}

// mirror the lower block triangle into the upper one and add lam to the diagonal
__global__ void wstep_finish_A_kernel(int k, int d, double lam, double* __restrict__ A) {
    const int N = d * k;
    const int64_t total = (int64_t)N * N;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
        const int r = (int)(e / N), cc = (int)(e - (int64_t)r * N);
        const int a = r / k, a2 = cc / k;
        if (a2 > a) {
            // A[(a,c),(a2,c2)] = A[(a2,c2),(a,c)]
            A[e] = A[(size_t)cc * N + r];
        } else if (r == cc) {
            A[e] += lam;
        }
    }
}

// Single-CTA right-looking Cholesky + solve, float64, matrix in global (L2-resident).
// b rides along as an extra row (forward substitution for free).  Uses the
// lower triangle only.  x overwrites b.
__global__ void __launch_bounds__(1024) cholesky_solve_f64_kernel(double* __restrict__ A, double* __restrict__ b,
                                                                  int N, int32_t* status) {
    extern __shared__ double col[];     // [N+1]
    __shared__ int bad;
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int lane = tid & 31, wid = tid >> 5, nw = nthr >> 5;
    if (tid == 0) bad = 0;
    __syncthreads();
    for (int j = 0; j < N; ++j) {
        const double dj = A[(size_t)j * N + j];
        if (!(dj > 0.0)) {
            if (tid == 0) bad = 1;
            break;
        }
        const double sd = sqrt(dj), inv = 1.0 / sd;
        for (int i = j + 1 + tid; i <= N; i += nthr) {
            const double v = ((i == N) ? b[j] : A[(size_t)i * N + j]) * inv;
            col[i] = v;
            if (i == N) b[j] = v; else A[(size_t)i * N + j] = v;
        }
        __syncthreads();                 // everyone has read dj; column j is in col[]
        if (tid == 0) A[(size_t)j * N + j] = sd;
        for (int i = j + 1 + wid; i <= N; i += nw) {
            const double li = col[i];
            if (i == N) {
                for (int c = j + 1 + lane; c < N; c += 32) b[c] -= li * col[c];
            } else {
                double* row = A + (size_t)i * N;
                for (int c = j + 1 + lane; c <= i; c += 32) row[c] -= li * col[c];
            }
        }
        __syncthreads();
    }
    __syncthreads();
    if (bad) {
        if (tid == 0 && status) {
            if (atomicAdd(&status[0], 1) == 0) { status[1] = -1; status[2] = 9; }
        }
        return;
    }
    for (int j = N - 1; j >= 0; --j) {
        const double xj = b[j] / A[(size_t)j * N + j];
        __syncthreads();
        const double* row = A + (size_t)j * N;
        for (int i = tid; i < j; i += nthr) b[i] -= row[i] * xj;
        if (tid == 0) b[j] = xj;
        __syncthreads();
    }
}

// Blocked right-looking Cholesky + solve across the whole GPU, float64, one cooperative launch.
// The single-CTA kernel above walks N columns with the trailing matrix going through L2 once per column
// (measured: ~20 ms at N = 608, the genres system of the bundled shape — a fifth of a C2 fit).  Here the matrix is
// cut into 32 x 32 tiles; per block column: (1) CTA 0 factors the diagonal tile in shared memory, (2) every row
// below — and b, carried as row N — is forward-substituted against it, one thread per row, (3) the trailing tiles
// are updated C -= P_I P_J^T, one tile per CTA at a time, operands staged in shared memory; grid barriers in between.
// The back substitution runs block-wise on CTA 0.  Lower triangle only; x overwrites b.
constexpr int CB = 32;

__global__ void __launch_bounds__(256) cholesky_solve_f64_coop_kernel(double* __restrict__ A, double* __restrict__ b,
                                                                      int N, int32_t* status, int* bad_flag) {
    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();
    __shared__ double T[CB][CB + 1];
    __shared__ double Pj[CB][CB + 1];
    __shared__ double xs[CB];
    __shared__ int bad_s;
    const int tid = threadIdx.x;
    const int nblk = (N + CB - 1) / CB;
    auto row_ptr = [&](int i) -> double* { return i == N ? b : A + (size_t)i * N; };   // b rides along as row N

    for (int jb = 0; jb < nblk; ++jb) {
        const int j0 = jb * CB, w = (N - j0) < CB ? (N - j0) : CB;
        // ---- (1) diagonal tile ----
        if (blockIdx.x == 0) {
            if (tid == 0) bad_s = 0;
            for (int e = tid; e < CB * CB; e += 256) {
                const int r = e / CB, c = e - r * CB;
                T[r][c] = (r < w && c <= r) ? A[(size_t)(j0 + r) * N + j0 + c] : 0.0;
            }
            __syncthreads();
            for (int j = 0; j < w; ++j) {
                const double d = T[j][j];
                if (!(d > 0.0)) {
                    if (tid == 0) bad_s = 1;
                    break;
                }
                const double sd = sqrt(d), inv = 1.0 / sd;
                __syncthreads();
                if (tid > j && tid < w) T[tid][j] *= inv;
                if (tid == j) T[j][j] = sd;
                __syncthreads();
                for (int e = tid; e < CB * CB; e += 256) {
                    const int r = e / CB, c = e - r * CB;
                    if (c > j && c <= r && r < w) T[r][c] = fma(-T[r][j], T[c][j], T[r][c]);
                }
                __syncthreads();
            }
            __syncthreads();
            if (bad_s) {
                if (tid == 0) *bad_flag = 1;
            } else {
                for (int e = tid; e < CB * CB; e += 256) {
                    const int r = e / CB, c = e - r * CB;
                    if (r < w && c <= r) A[(size_t)(j0 + r) * N + j0 + c] = T[r][c];
                }
            }
            __threadfence();
        }
        grid.sync();
        if (*reinterpret_cast<volatile int*>(bad_flag)) break;
        // ---- (2) panel: rows below (and b) against the factored tile ----
        const int first = j0 + w;                            // first row below the tile; row N is b
        for (int e = tid; e < CB * CB; e += 256) {
            const int r = e / CB, c = e - r * CB;
            T[r][c] = (r < w && c <= r) ? A[(size_t)(j0 + r) * N + j0 + c] : 0.0;
        }
        __syncthreads();
        for (int64_t i = first + (int64_t)blockIdx.x * 256 + tid; i <= N; i += (int64_t)gridDim.x * 256) {
            double* rp = row_ptr((int)i) + j0;
            double a[CB];
#pragma unroll
            for (int c = 0; c < CB; ++c) a[c] = c < w ? rp[c] : 0.0;
#pragma unroll
            for (int c = 0; c < CB; ++c) {
                if (c < w) {
                    double v = a[c];
#pragma unroll
                    for (int t = 0; t < CB; ++t)
                        if (t < c) v = fma(-a[t], T[c][t], v);
                    a[c] = v / T[c][c];
                }
            }
#pragma unroll
            for (int c = 0; c < CB; ++c)
                if (c < w) rp[c] = a[c];
        }
        __threadfence();
        grid.sync();
        // ---- (3) trailing update: tile (I, J), jb < J <= I, I == nblk is the single row b ----
        const int nI = nblk - jb - 1;                        // block rows below
        if (nI > 0 || first <= N) {
            const int64_t tiles_sq = (int64_t)nI * (nI + 1) / 2;
            const int64_t tiles = tiles_sq + nI;             // + one row-N strip per block column
            for (int64_t tl = blockIdx.x; tl < tiles; tl += gridDim.x) {
                int I, J;
                if (tl < tiles_sq) {
                    I = (int)((sqrt(8.0 * (double)tl + 1.0) - 1.0) * 0.5);
                    while ((int64_t)(I + 1) * (I + 2) / 2 <= tl) ++I;
                    while ((int64_t)I * (I + 1) / 2 > tl) --I;
                    J = (int)(tl - (int64_t)I * (I + 1) / 2);
                } else {
                    I = nI;                                   // the b strip
                    J = (int)(tl - tiles_sq);
                }
                const int ri0 = (I == nI) ? N : first + I * CB, rj0 = first + J * CB;
                const int hi = (I == nI) ? 1 : ((N - ri0) < CB ? (N - ri0) : CB);
                const int hj = (N - rj0) < CB ? (N - rj0) : CB;
                __syncthreads();
                for (int e = tid; e < CB * CB; e += 256) {
                    const int r = e / CB, c = e - r * CB;
                    T[r][c] = (r < hi && c < w) ? row_ptr(ri0 + r)[j0 + c] : 0.0;
                    Pj[r][c] = (r < hj && c < w) ? A[(size_t)(rj0 + r) * N + j0 + c] : 0.0;
                }
                __syncthreads();
                for (int e = tid; e < CB * CB; e += 256) {
                    const int r = e / CB, c = e - r * CB;    // output (ri0 + r, rj0 + c)
                    if (r < hi && c < hj && (I == nI || rj0 + c <= ri0 + r)) {
                        double acc = 0.0;
#pragma unroll
                        for (int t = 0; t < CB; ++t) acc = fma(T[r][t], Pj[c][t], acc);
                        row_ptr(ri0 + r)[rj0 + c] -= acc;
                    }
                }
            }
        }
        __threadfence();
        grid.sync();
    }
    if (*reinterpret_cast<volatile int*>(bad_flag)) {
        if (blockIdx.x == 0 && tid == 0 && status) {
            if (atomicAdd(&status[0], 1) == 0) { status[1] = -1; status[2] = 9; }
        }
        return;
    }
    // ---- back substitution L^T x = y on CTA 0, block by block from the end ----
    if (blockIdx.x != 0) return;
    for (int jb = nblk - 1; jb >= 0; --jb) {
        const int j0 = jb * CB, w = (N - j0) < CB ? (N - j0) : CB;
        __syncthreads();
        for (int e = tid; e < CB * CB; e += 256) {
            const int r = e / CB, c = e - r * CB;
            T[r][c] = (r < w && c <= r) ? A[(size_t)(j0 + r) * N + j0 + c] : 0.0;
        }
        if (tid < CB) xs[tid] = tid < w ? b[j0 + tid] : 0.0;
        __syncthreads();
        if (tid < 32) {                                      // one warp: w dependent steps
            double y = xs[tid];
            for (int j = w - 1; j >= 0; --j) {
                const double xj = __shfl_sync(0xffffffffu, y, j) / T[j][j];
                if (tid == j) y = xj;
                else if (tid < j) y = fma(-T[j][tid], xj, y);
            }
            xs[tid] = y;
            if (tid < w) b[j0 + tid] = y;
        }
        __syncthreads();
        for (int i = tid; i < j0; i += 256) {
            double acc = b[i];
            for (int r = 0; r < w; ++r) acc = fma(-A[(size_t)(j0 + r) * N + i], xs[r], acc);
            b[i] = acc;
        }
    }
}

}  // namespace b200als

using namespace b200als;

extern "C" {

int b200als_item_gram_f64(int64_t n, const int64_t* cindptr, const int32_t* cindices, const float* cvals, int32_t k,
                          const float* U, int64_t ldu, const float* b_u, const float* b_i, const float* mu,
                          double* G, double* c, b200als_stream_t stream) {
    ALS_CHECK_ARG(n > 0 && cindptr && cindices && cvals && U && b_u && b_i && mu && G && c, "null pointer");
    ALS_CHECK_ARG(k >= 1 && k <= B200ALS_MAX_K, "k out of range");
    const size_t smem = sizeof(float) * ((size_t)GR_TR * k + GR_TR);
    int grid = (int)(n < (int64_t)kNumSMs * 8 ? n : (int64_t)kNumSMs * 8);
    item_gram_f64_kernel<<<grid, 256, smem, as_stream(stream)>>>(n, cindptr, cindices, cvals, k, U, ldu, b_u, b_i, mu,
                                                                 G, c);
    ALS_LAUNCH_CHECK();
    return B200ALS_OK;
}

int b200als_wstep_system(int64_t n, int32_t k, int32_t d, const float* X, const double* G, const double* c,
                         const float* Zmf, int64_t ldz, double lam, double* A, double* b, b200als_stream_t stream) {
    ALS_CHECK_ARG(n > 0 && k >= 1 && k <= B200ALS_MAX_K && d >= 1, "bad shape");
    ALS_CHECK_ARG(X && G && c && Zmf && A && b, "null pointer");
    cudaStream_t s = as_stream(stream);
    const int64_t N = (int64_t)d * k;
    ALS_CUDA(cudaMemsetAsync(A, 0, sizeof(double) * N * N, s));
    // h lives in b's tail?  no: use a stream-ordered scratch (setup-sized, freed on the stream)
    double* h = nullptr;
    ALS_CUDA(cudaMallocAsync(&h, sizeof(double) * n * k, s));
    {
        int64_t warps = n * k;
        int64_t blocks = ceil_div64(warps * 32, 256);
        if (blocks > (int64_t)kNumSMs * 16) blocks = (int64_t)kNumSMs * 16;
        wstep_h_kernel<<<(unsigned)blocks, 256, 0, s>>>(n, k, G, c, Zmf, ldz, h);
        ALS_LAUNCH_CHECK();
    }
    wstep_b_kernel<<<d, 256, 0, s>>>(n, k, d, X, h, b);
    ALS_LAUNCH_CHECK();
    ALS_CUDA(cudaFreeAsync(h, s));
    {
        const int bx = (k * k + 255) / 256;
        int nsplit = (int)((int64_t)kNumSMs * 4 / ((int64_t)bx * d));
        if (nsplit < 1) nsplit = 1;
        if (nsplit > 64) nsplit = 64;
        if (nsplit > n) nsplit = (int)n;
        dim3 grid(bx, d, nsplit);
        wstep_A_kernel<8><<<grid, 256, 0, s>>>(n, k, d, X, G, A);
        ALS_LAUNCH_CHECK();
    }
    {
        int64_t blocks = ceil_div64(N * N, 256);
        if (blocks > (int64_t)kNumSMs * 8) blocks = (int64_t)kNumSMs * 8;
        wstep_finish_A_kernel<<<(unsigned)blocks, 256, 0, s>>>(k, d, lam, A);
        ALS_LAUNCH_CHECK();
    }
    return B200ALS_OK;
}

int b200als_cholesky_solve_f64(double* A, double* b, int32_t N, int32_t* status, b200als_stream_t stream) {
    ALS_CHECK_ARG(A && b && N >= 1, "bad args");
    if (N > 64) {
        // blocked, all SMs, one cooperative launch; its 'not positive definite' flag lives in the caller's status words
        cudaStream_t s = as_stream(stream);
        int* flag = nullptr;
        ALS_CUDA(cudaMallocAsync(&flag, sizeof(int), s));
        ALS_CUDA(cudaMemsetAsync(flag, 0, sizeof(int), s));
        int per_sm = 0;
        ALS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, cholesky_solve_f64_coop_kernel, 256, 0));
        const int nblk = (N + CB - 1) / CB;
        int64_t want = (int64_t)nblk * (nblk + 1) / 2;
        int64_t cap = (int64_t)kNumSMs * (per_sm < 1 ? 1 : (per_sm > 2 ? 2 : per_sm));
        int grid = (int)(want < cap ? want : cap);
        if (grid < 1) grid = 1;
        void* args[] = {&A, &b, &N, &status, &flag};
        ALS_CUDA(cudaLaunchCooperativeKernel((const void*)cholesky_solve_f64_coop_kernel, dim3(grid), dim3(256), args, 0, s));
        g_launch_count++;
        ALS_CUDA(cudaFreeAsync(flag, s));
        return B200ALS_OK;
    }
    const size_t smem = sizeof(double) * ((size_t)N + 1);
    ALS_CHECK_ARG(smem <= 200 * 1024, "system too large for the single-CTA float64 solver (d*k <= 25599)");
    if (smem > 48 * 1024)
        ALS_CUDA(cudaFuncSetAttribute(cholesky_solve_f64_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)smem));
    cholesky_solve_f64_kernel<<<1, 1024, smem, as_stream(stream)>>>(A, b, N, status);
    ALS_LAUNCH_CHECK();
    return B200ALS_OK;
}

}  // extern "C"

// row_step_f64.cuh — the float64 half-step (B200ALS_PATH_F64): the same row update as the FFMA / tcgen05 kernels
// (als.py:414-433 / 436-466, helpers.py:19-20) with the normal matrix accumulated, factored and solved in DOUBLE.
//
// Why it exists.  The reference is float64 end to end and its tuner draws lambda_u, lambda_v log-uniformly from
// [1e-4, 1e4] (tune_params.py:102-103).  A row with fewer ratings than factors has a rank-deficient Gram that only
// the ridge keeps positive definite; the solution then amplifies any perturbation of the Gram by sigma_max / lambda.
// Measured on B200 against the oracle (rows of ~20 ratings): float32 holds 1e-3 on the factors down to lambda ~ 0.03
// on the FFMA path and ~ 0.1 on the tcgen05 path (bf16 hi/lo operands), fails it at lambda = 1e-2 and loses positive
// definiteness at 1e-4 — a quarter of the tuner's range.  Those rows (small ridge AND few ratings — the host decides,
// als.py) take this kernel; everything else stays on the fast paths.  It is also the whole fit when the caller asks for
// `precision = "f64"`.
//
// One CTA per row, rows from the shared atomic cursor.  Shared memory: packed lower triangle of G (k (k + 1) / 2
// doubles: 103 KB at k = 160), c = F^T q, s = F^T 1, a 16-rating float tile of gathered rows, and a table of the
// (i, j) of every packed entry (built once per CTA) so that the threads stride the triangle without index arithmetic.
// Cholesky is right-looking on the packed triangle; the solves are column sweeps.  Throughput is not the point here
// (DFMA stream from shared memory, a few barriers per column); exactness against the float64 reference is.
#pragma once
#include "row_step_ffma.cuh"

namespace b200als {

constexpr int F64_THREADS = 256;
constexpr int F64_TR = 16;        // ratings per staged tile

__host__ __device__ constexpr int f64_tri(int k) { return k * (k + 1) / 2; }
__host__ __device__ constexpr int f64_tri_pad(int k) { return (f64_tri(k) + 1) & ~1; }   // keeps the float4 tile 16-byte aligned
static inline size_t f64_smem_bytes(int k) {
    const int k4 = (k + 3) & ~3;
    return sizeof(double) * ((size_t)f64_tri_pad(k) + 4 * (size_t)k4 + 8 + F64_TR)   // G | c | s | b / y / x | - | scalars | q
           + sizeof(float) * ((size_t)F64_TR * k4)                       // tile
           + sizeof(int32_t) * (F64_TR + 8)                              // idx | flags
           + sizeof(uint16_t) * 2 * (size_t)f64_tri(k);                  // (i, j) table
}

__global__ void __launch_bounds__(F64_THREADS) row_step_f64_kernel(const RowStepParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int k = p.k, k4 = (k + 3) & ~3, tri = f64_tri(k);
    double* G = reinterpret_cast<double*>(smem_raw);
    double* c_s = G + f64_tri_pad(k);
    double* s_s = c_s + k4;
    double* b_s = s_s + k4;          // right-hand side, then y, then x
    double* x_s = b_s + k4;
    double* sc = x_s + k4;           // [0] t, [1] qq, [2] pivot-ok flag
    double* q_t = sc + 8;            // q = rating - mu - bias_other of the staged ratings, in double
    float* tile = reinterpret_cast<float*>(q_t + F64_TR);
    int32_t* idx_t = reinterpret_cast<int32_t*>(tile + F64_TR * k4);
    int32_t* misc = idx_t + F64_TR;
    uint16_t* tab = reinterpret_cast<uint16_t*>(misc + 8);
    const int tid = threadIdx.x;
    const double mu = (double)*p.mu;

    // (i, j) of every packed entry, built once per CTA
    for (int i = 0; i < k; ++i) {
        const int base = i * (i + 1) / 2;
        for (int j = tid; j <= i; j += F64_THREADS) {
            tab[2 * (base + j)] = (uint16_t)i;
            tab[2 * (base + j) + 1] = (uint16_t)j;
        }
    }
    double st_d = 0.0, st_d2 = 0.0, st_xx = 0.0, st_bb = 0.0;     // thread 0: fused statistics / norms

    for (;;) {
        __syncthreads();
        if (tid == 0) {
            const unsigned long long claimed = atomicAdd(p.cursor, 1ull);
            misc[0] = claimed < (unsigned long long)p.n_list ? (int32_t)claimed : -1;
        }
        __syncthreads();
        const int64_t slot = misc[0];
        if (slot < 0) break;
        const int64_t r = p.rows ? (int64_t)p.rows[slot] : slot;
        const int64_t gr = p.row_ids ? (int64_t)p.row_ids[r] : r;
        const int64_t beg = p.indptr[r], end = p.indptr[r + 1];
        const int64_t cnt = end - beg;
        if (cnt <= 0) continue;
        for (int e = tid; e < tri; e += F64_THREADS) G[e] = 0.0;
        for (int i = tid; i < 4 * k4 + 8; i += F64_THREADS) c_s[i] = 0.0;    // c, s, b, x, scalars are contiguous

        for (int64_t t0 = beg; t0 < end; t0 += F64_TR) {
            const int live = (int)((end - t0) < F64_TR ? (end - t0) : F64_TR);
            __syncthreads();                                   // previous tile consumed (and the zeroing done)
            if (tid < F64_TR) {
                int32_t ix = 0;
                double q = 0.0;
                if (tid < live) {
                    ix = p.indices[t0 + tid];
                    q = (double)p.vals[t0 + tid] - (mu + (double)p.bias_other[ix]);
                }
                idx_t[tid] = ix;
                q_t[tid] = q;
            }
            __syncthreads();
            for (int e = tid; e < F64_TR * (k4 >> 2); e += F64_THREADS) {
                const int rr = e / (k4 >> 2), cc = (e - rr * (k4 >> 2)) * 4;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (rr < live) v = *reinterpret_cast<const float4*>(p.F + (size_t)idx_t[rr] * p.ldf + cc);
                *reinterpret_cast<float4*>(tile + rr * k4 + cc) = v;
            }
            __syncthreads();
            for (int e = tid; e < tri; e += F64_THREADS) {
                const int i = tab[2 * e], j = tab[2 * e + 1];
                double acc = G[e];
#pragma unroll
                for (int rr = 0; rr < F64_TR; ++rr) acc = fma((double)tile[rr * k4 + i], (double)tile[rr * k4 + j], acc);
                G[e] = acc;
            }
            if (tid < k) {
                double ca = c_s[tid], sa = s_s[tid];
#pragma unroll
                for (int rr = 0; rr < F64_TR; ++rr) {
                    const double f = (double)tile[rr * k4 + tid];
                    ca = fma(f, q_t[rr], ca);
                    sa += f;
                }
                c_s[tid] = ca;
                s_s[tid] = sa;
            }
            if (tid == F64_THREADS - 1) {
                double ta = sc[0], qa = sc[1];
                for (int rr = 0; rr < live; ++rr) {
                    const double q = q_t[rr];
                    ta += q;
                    qa = fma(q, q, qa);
                }
                sc[0] = ta;
                sc[1] = qa;
            }
        }
        __syncthreads();

        // ---- A = G + ridge I, b = c - bias_old s ----
        const double ridge = (double)(p.reg_row ? p.reg_row[gr] : p.reg);
        const double lam_bias = (double)(p.lam_bias_row ? p.lam_bias_row[gr] : p.lam_bias);
        const double bias_old = (double)p.bias_own[gr];
        for (int i = tid; i < k; i += F64_THREADS) {
            G[i * (i + 1) / 2 + i] += ridge;
            b_s[i] = c_s[i] - bias_old * s_s[i];
        }
        if (tid == 0) sc[2] = 1.0;
        __syncthreads();
        // ---- Cholesky, right-looking on the packed lower triangle (helpers.py:19: cho_factor) ----
        for (int j = 0; j < k; ++j) {
            const double d = G[j * (j + 1) / 2 + j];
            if (!(d > 0.0)) {                                  // uniform: every thread reads the same pivot
                if (tid == 0) sc[2] = 0.0;
                break;
            }
            const double ljj = sqrt(d), inv = 1.0 / ljj;
            __syncthreads();                                   // everyone has read the pivot
            for (int i = j + tid; i < k; i += F64_THREADS) {
                const int e = i * (i + 1) / 2 + j;
                G[e] = (i == j) ? ljj : G[e] * inv;
            }
            __syncthreads();
            const int first = (j + 1) * (j + 2) / 2;           // entries of rows > j
            for (int e = first + tid; e < tri; e += F64_THREADS) {
                const int i = tab[2 * e], l = tab[2 * e + 1];
                if (l > j) G[e] = fma(-G[i * (i + 1) / 2 + j], G[l * (l + 1) / 2 + j], G[e]);
            }
            __syncthreads();
        }
        __syncthreads();
        if (sc[2] == 0.0) {
            if (tid == 0 && p.status) {
                if (atomicAdd(&p.status[0], 1) == 0) p.status[1] = (int32_t)gr;
            }
            continue;
        }
        if (p.mode == B200ALS_STEP_FACTOR) {
            float* Lo = p.chol_out + (size_t)gr * k * k;
            for (int e = tid; e < k * k; e += F64_THREADS) {
                const int i = e / k, j = e - i * k;
                Lo[e] = (j <= i) ? (float)G[i * (i + 1) / 2 + j] : 0.0f;
            }
            float* ao = p.aux_out + (size_t)gr * (2 * k + 2);
            for (int i = tid; i < k; i += F64_THREADS) {
                ao[i] = (float)c_s[i];
                ao[k + i] = (float)s_s[i];
            }
            if (tid == 0) {
                ao[2 * k] = (float)sc[0];
                ao[2 * k + 1] = (float)cnt;
            }
            continue;
        }
        // ---- L y = b, L^T x = y (cho_solve) ----
        for (int j = 0; j < k; ++j) {
            const double yj = b_s[j] / G[j * (j + 1) / 2 + j];
            __syncthreads();
            if (tid == 0) b_s[j] = yj;
            for (int i = j + 1 + tid; i < k; i += F64_THREADS) b_s[i] = fma(-G[i * (i + 1) / 2 + j], yj, b_s[i]);
            __syncthreads();
        }
        for (int j = k - 1; j >= 0; --j) {
            const double xj = b_s[j] / G[j * (j + 1) / 2 + j];
            __syncthreads();
            if (tid == 0) b_s[j] = xj;
            for (int i = tid; i < j; i += F64_THREADS) b_s[i] = fma(-G[j * (j + 1) / 2 + i], xj, b_s[i]);
            __syncthreads();
        }
        // ---- bias, statistics, stores ----
        if (tid < 32) {
            double sx = 0.0, cx = 0.0, xx = 0.0;
            for (int i = tid; i < k; i += 32) {
                const double x = b_s[i];
                sx = fma(s_s[i], x, sx);
                cx = fma(c_s[i], x, cx);
                xx = fma(x, x, xx);
            }
            sx = warp_sum(sx);
            cx = warp_sum(cx);
            xx = warp_sum(xx);
            if (tid == 0) {
                const double t = sc[0], qq = sc[1], cd = (double)cnt;
                const double beta = (t - sx) / (cd + lam_bias);                    // als.py:431-433 / 464-466
                store_bias_all(p, gr, (float)beta);
                st_xx += xx;
                st_bb += beta * beta;
                st_d += t - cd * beta - sx;
                st_d2 += qq - 2.0 * beta * t + cd * beta * beta - cx + (2.0 * beta - bias_old) * sx - ridge * xx;
            }
        }
        for (int i = tid; i < k; i += F64_THREADS) store_x_all(p, (size_t)gr * p.ldx + i, (float)b_s[i]);
    }
    if (tid == 0) {
        if (p.stats) {
            atomicAdd(&p.stats[0], st_d);
            atomicAdd(&p.stats[1], st_d2);
        }
        if (p.norms) {
            atomicAdd(&p.norms[0], st_xx);
            atomicAdd(&p.norms[2], st_bb);
        }
    }
}

static int launch_f64(const RowStepParams& p, cudaStream_t s) {
    const size_t smem = f64_smem_bytes(p.k);
    if (smem > 227 * 1024) {
        set_error("row_step (float64): k = %d needs %zu bytes of shared memory", p.k, smem);
        return B200ALS_E_UNSUPPORTED;
    }
    ALS_CUDA(cudaFuncSetAttribute(row_step_f64_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = (int)((227 * 1024) / (smem + 1024));
    if (per_sm < 1) per_sm = 1;
    if (per_sm > 4) per_sm = 4;
    int64_t grid = (int64_t)kNumSMs * per_sm;
    if (grid > p.n_list) grid = p.n_list;
    row_step_f64_kernel<<<(unsigned)grid, F64_THREADS, smem, s>>>(p);
    ALS_LAUNCH_CHECK();
    return B200ALS_OK;
}

}  // namespace b200als

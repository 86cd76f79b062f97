// row_solve.cuh — the per-row epilogue shared by every Gram path (A6/A7/A11):
// in-shared-memory Cholesky of the k x k normal matrix, forward substitution
// folded in as an augmented row, back substitution, closed-form bias.
//
// Layout: G is (KP+1) rows x LD floats, LD = KP+1 (odd → column walks are
// bank-conflict free).  Rows 0..KP-1 hold the lower triangle of A (diagonal
// already regularised; padded columns k..KP-1 carry a unit diagonal so they
// factor to the identity).  Row KP holds the right-hand side b, which the
// right-looking factorisation turns into y = L^-1 b for free.
#pragma once
#include "common.cuh"

namespace b200als {

// Right-looking Cholesky on the lower triangle, all `nthr` threads of the CTA.
// Returns false (uniformly) on a non-positive / NaN pivot.
template <int KP>
__device__ __forceinline__ bool chol_factor_smem(float* __restrict__ G, int kk, int tid, int nthr,
                                                 int* __restrict__ flag_s) {
    constexpr int LD = KP + 1;
    const int lane = tid & 31, wid = tid >> 5, nw = nthr >> 5;
    if (tid == 0) *flag_s = 0;
    for (int j = 0; j < kk; ++j) {
        __syncthreads();
        const float d = G[j * LD + j];
        if (!(d > 0.0f)) {           // uniform: every thread reads the same word
            if (tid == 0) *flag_s = 1;
            break;
        }
        const float sd = sqrtf(d);
        const float inv = 1.0f / sd;
        // scale column j (rows j+1..kk-1 and the rhs row KP)
        for (int i = j + 1 + tid; i <= kk; i += nthr) {
            const int row = (i == kk) ? KP : i;
            G[row * LD + j] *= inv;
        }
        __syncthreads();
        if (tid == 0) G[j * LD + j] = sd;
        // trailing update: rows i in (j, kk) plus rhs row; cols c in (j, min(i, kk-1)]
        for (int i = j + 1 + wid; i <= kk; i += nw) {
            const int row = (i == kk) ? KP : i;
            const float lij = G[row * LD + j];
            const int cmax = (i == kk) ? kk - 1 : i;
            for (int c = j + 1 + lane; c <= cmax; c += 32) G[row * LD + c] -= lij * G[c * LD + j];
        }
    }
    __syncthreads();
    return *flag_s == 0;
}

// Back substitution L^T x = y with y in row KP; x overwrites row KP.
template <int KP>
__device__ __forceinline__ void chol_backsolve_smem(float* __restrict__ G, int kk, int tid, int nthr) {
    constexpr int LD = KP + 1;
    float* y = G + KP * LD;
    for (int j = kk - 1; j >= 0; --j) {
        __syncthreads();
        const float xj = y[j] / G[j * LD + j];
        // y[i] -= L[j][i] * x[j] for i < j   (row j of L is contiguous)
        for (int i = tid; i < j; i += nthr) y[i] -= G[j * LD + i] * xj;
        __syncthreads();
        if (tid == 0) y[j] = xj;
    }
    __syncthreads();
}

struct RowEpilogueArgs {
    int k;
    int mode;
    float reg, lam_bias;
    float* X; int64_t ldx;
    float* bias_own;
    float* chol_out; float* aux_out;
    int32_t* status;
};

// G lower triangle + c (F^T q), s (F^T 1) in smem, t = sum q, cnt.
// Applies regularisation, builds b = c - bias_old * s, solves, writes outputs.
template <int KP>
__device__ __forceinline__ void row_epilogue(float* __restrict__ G, const float* __restrict__ c_s,
                                             const float* __restrict__ s_s, float t_sum, int64_t cnt, int64_t r,
                                             const RowEpilogueArgs& a, int tid, int nthr, int* flag_s,
                                             float* red_s) {
    constexpr int LD = KP + 1;
    const int k = a.k;
    const float bias_old = a.bias_own[r];
    __syncthreads();
    for (int i = tid; i < KP; i += nthr) {
        G[i * LD + i] = (i < k) ? G[i * LD + i] + a.reg : 1.0f;
        G[KP * LD + i] = (i < k) ? c_s[i] - bias_old * s_s[i] : 0.0f;
    }
    // (chol_factor_smem starts with a barrier)
    const bool ok = chol_factor_smem<KP>(G, k, tid, nthr, flag_s);
    if (!ok) {
        if (tid == 0 && a.status) {
            if (atomicAdd(&a.status[0], 1) == 0) a.status[1] = (int32_t)r;
        }
        return;
    }
    if (a.mode == B200ALS_STEP_FACTOR) {
        float* Lo = a.chol_out + (size_t)r * k * k;
        for (int e = tid; e < k * k; e += nthr) {
            int i = e / k, j = e - i * k;
            Lo[e] = (j <= i) ? G[i * LD + j] : 0.0f;
        }
        float* ao = a.aux_out + (size_t)r * (2 * k + 2);
        for (int i = tid; i < k; i += nthr) {
            ao[i] = c_s[i];
            ao[k + i] = s_s[i];
        }
        if (tid == 0) {
            ao[2 * k] = t_sum;
            ao[2 * k + 1] = (float)cnt;
        }
        return;
    }
    chol_backsolve_smem<KP>(G, k, tid, nthr);
    const float* x = G + KP * LD;
    // bias_own = (t - s·x) / (cnt + lam_bias)        (als.py:431-433, 464-466)
    float part = 0.0f;
    for (int i = tid; i < k; i += nthr) part += s_s[i] * x[i];
    part = warp_sum(part);
    if ((tid & 31) == 0) red_s[tid >> 5] = part;
    __syncthreads();
    if (tid == 0) {
        float dot = 0.0f;
        for (int w = 0; w < (nthr >> 5); ++w) dot += red_s[w];
        a.bias_own[r] = (t_sum - dot) / ((float)cnt + a.lam_bias);
    }
    float* xo = a.X + (size_t)r * a.ldx;
    for (int i = tid; i < k; i += nthr) xo[i] = x[i];
}

}  // namespace b200als

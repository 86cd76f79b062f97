// row_ldl.cuh — register-resident, thread-per-row LDL^T solve fed straight from TMEM
// (A6/A7/A11: the per-row normal equations, als.py:426-428 / 455-461, helpers.py:19-20).
//
// The tcgen05 Gram kernel leaves row i of the k x k normal matrix in TMEM lane i, so a
// tcgen05.ld hands every solver thread one full matrix row in registers.  The
// factorisation never copies the matrix to shared memory:
//
//   for column j (fully unrolled, register indices static):
//     every thread r >= j publishes its entry a_r[j] (still unscaled: l_rj * d_j) into
//     the packed column store; thread j also publishes z_j (forward-substituted rhs);
//     ONE named barrier;
//     t = a_r[j] / d_j;  rhs_r -= t * z_j;  a_r[c] -= t * col_j[c]  for c in (j, warp max row]
//     (col_j read as broadcast LDS.128: one shared-memory wavefront per four FFMAs).
//
// The column store doubles as L^T for the back substitution (thread j owns column j:
// x_j = (z_j - sum_{i>j} col_j[i] x_i) / d_j), done block by block: the owning warp
// resolves its RPW rows with shuffles, the warps below fold the block in with per-thread
// LDS.128.  Shared memory per matrix: k^2/2 + 2k floats of columns + z + x.
//
// A solver group is 4 warps = the 4 TMEM lane quadrants.  k = 128: one matrix per group,
// RPW = 32 rows per warp.  k = 64: the Gram MMAs run with M = 64, whose accumulator uses
// 16 lanes of every quadrant, and two rows' matrices are interleaved (lanes 0-15 / 16-31
// of each quadrant), so RPW = 16 and a group factors two matrices in lock step.
#pragma once
#include "common.cuh"

namespace b200als {

// element (i, j), i >= j, of the packed column store sits at ldl_col_base(j) + i
// (column j is allocated from row 4*(j/4), so every column keeps i % 4 alignment for LDS.128)
template <int K>
__host__ __device__ constexpr int ldl_col_base(int j) {
    const int g = j >> 2, jj = j & 3;
    return 4 * g * K - 8 * g * (g - 1) + jj * (K - 4 * g) - 4 * g;
}
template <int K>
__host__ __device__ constexpr int ldl_cols_floats() { return K * K / 2 + 2 * K; }
// per-matrix buffer: columns | z[K] | x[K] | pad — size = 16 (mod 32) floats so that the two
// interleaved matrices of a k=64 group sit 16 banks apart
template <int K>
__host__ __device__ constexpr int ldl_mat_floats() {
    const int raw = ldl_cols_floats<K>() + 2 * K;
    return ((raw + 31) / 32) * 32 + 16;
}

__device__ __forceinline__ float fast_rcp(float d) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d));
    return r * fmaf(-d, r, 2.0f);             // one Newton step: < 1 ulp
}

__device__ __forceinline__ void named_bar_sync(int id, int nthr) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthr) : "memory");
}

template <int K, int RPW>
struct Ldl {
    static constexpr int NW = K / RPW;       // warps per group (4)
    static_assert(NW == 4, "a solver group is the four TMEM lane quadrants");

    // One column step (J is a template parameter so that every register index is static even
    // where nvcc would not fully unroll a 128-trip loop).
    template <int J>
    static __device__ __forceinline__ void col_step(float (&a)[K], float& rhs, bool& ok, const float ridge,
                                                    const int r, const int q, float* __restrict__ cb,
                                                    const int bar_id) {
        float* __restrict__ zb = cb + ldl_cols_floats<K>();
        float* __restrict__ colj = cb + ldl_col_base<K>(J);
        const bool warp_active = q >= (J / RPW);              // warp still owns a row >= J
        if (warp_active) {
            if (r == J) {
                const float d = a[J] + ridge;
                ok = d > 0.0f;
                colj[J] = d;
                zb[J] = rhs;
            } else if (r > J) {
                colj[r] = a[J];
            }
        }
        named_bar_sync(bar_id, NW * 32);
        if (warp_active) {
            const float inv = fast_rcp(colj[J]);
            const float t = a[J] * inv;
            rhs = fmaf(-t, zb[J], rhs);
            // columns J+1 .. : scalar head up to the next multiple of 4, then LDS.128 groups;
            // whole RPW-blocks above this warp's rows are skipped (warp-uniform)
            constexpr int HEAD_END = ((J + 1 + 3) & ~3) < K ? ((J + 1 + 3) & ~3) : K;
#pragma unroll
            for (int c = J + 1; c < HEAD_END; ++c) a[c] = fmaf(-t, colj[c], a[c]);
#pragma unroll
            for (int cb0 = (HEAD_END / RPW) * RPW; cb0 < K; cb0 += RPW) {
                if (q >= cb0 / RPW) {
#pragma unroll
                    for (int c4 = (cb0 > HEAD_END ? cb0 : HEAD_END); c4 < cb0 + RPW; c4 += 4) {
                        const float4 v = *reinterpret_cast<const float4*>(colj + c4);
                        a[c4 + 0] = fmaf(-t, v.x, a[c4 + 0]);
                        a[c4 + 1] = fmaf(-t, v.y, a[c4 + 1]);
                        a[c4 + 2] = fmaf(-t, v.z, a[c4 + 2]);
                        a[c4 + 3] = fmaf(-t, v.w, a[c4 + 3]);
                    }
                }
            }
        }
    }
    template <int J0, int J1>
    static __device__ __forceinline__ void col_range(float (&a)[K], float& rhs, bool& ok, const float ridge,
                                                     const int r, const int q, float* __restrict__ cb,
                                                     const int bar_id) {
        if constexpr (J1 - J0 == 1) {
            col_step<J0>(a, rhs, ok, ridge, r, q, cb, bar_id);
        } else {
            constexpr int JM = (J0 + J1) / 2;
            col_range<J0, JM>(a, rhs, ok, ridge, r, q, cb, bar_id);
            col_range<JM, J1>(a, rhs, ok, ridge, r, q, cb, bar_id);
        }
    }

    // Factor + forward substitution.  a[] = this thread's matrix row (row position r in
    // warp quadrant q), rhs = its right-hand-side entry, ridge is added on the pivot.
    // cb = this thread's matrix buffer (columns, then z at cb + ldl_cols_floats).
    // Returns true if thread r's pivot was positive.
    static __device__ __forceinline__ bool factor(float (&a)[K], float rhs, const float ridge, const int r,
                                                  const int q, float* __restrict__ cb, const int bar_id) {
        bool ok = true;
        col_range<0, K>(a, rhs, ok, ridge, r, q, cb, bar_id);
        return ok;
    }

    // Back substitution.  Needs a group barrier (phase_bar) between blocks; ends with x_r
    // returned to thread r and stored in xb[r].  `seg` = lane % RPW, `full` = shuffle mask.
    static __device__ __forceinline__ float backsolve(const int r, const int q, const int seg,
                                                      const float* __restrict__ cb, float* __restrict__ xb,
                                                      const int phase_bar) {
        const float* __restrict__ zb = cb + ldl_cols_floats<K>();
        const int g = r >> 2, jj = r & 3;
        const float* __restrict__ mycol = cb + (4 * g * K - 8 * g * (g - 1) + jj * (K - 4 * g) - 4 * g);
        const float dinv = fast_rcp(mycol[r]);
        float acc = zb[r];
        float x = 0.0f;
#pragma unroll
        for (int blk = NW - 1; blk >= 0; --blk) {
            if (q == blk) {
                float lt[RPW];
#pragma unroll
                for (int i4 = 0; i4 < RPW; i4 += 4) {
                    const float4 v = *reinterpret_cast<const float4*>(mycol + blk * RPW + i4);
                    lt[i4] = v.x; lt[i4 + 1] = v.y; lt[i4 + 2] = v.z; lt[i4 + 3] = v.w;
                }
#pragma unroll
                for (int ii = RPW - 1; ii >= 0; --ii) {
                    const float mine = acc * dinv;
                    const float xi = __shfl_sync(0xffffffffu, mine, ii, RPW);
                    if (seg < ii) acc = fmaf(-lt[ii], xi, acc);
                    if (seg == ii) x = mine;
                }
                xb[r] = x;
            }
            named_bar_sync(phase_bar, NW * 32);
            if (q < blk) {
#pragma unroll
                for (int i4 = 0; i4 < RPW; i4 += 4) {
                    const float4 c4 = *reinterpret_cast<const float4*>(mycol + blk * RPW + i4);
                    const float4 x4 = *reinterpret_cast<const float4*>(xb + blk * RPW + i4);
                    acc = fmaf(-c4.x, x4.x, acc);
                    acc = fmaf(-c4.y, x4.y, acc);
                    acc = fmaf(-c4.z, x4.z, acc);
                    acc = fmaf(-c4.w, x4.w, acc);
                }
            }
        }
        return x;
    }
};

}  // namespace b200als

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
//     ONE named barrier among the warps that still own a row >= j;
//     t = a_r[j] / d_j;  rhs_r -= t * z_j;  a_r[c] -= t * col_j[c]  for c in (j, warp max row]
//     (col_j read as broadcast LDS.128: one shared-memory wavefront per four FFMAs).
//
// The column loop is ROLLED: four columns per iteration, the fourth writing its update four
// registers down so that the window slides and every register index stays static.  Warp q
// owns rows [RPW q, RPW (q+1)): it runs phases p = 0..q of RPW columns on a window of
// RPW (q+1-p) registers, then leaves the column barrier to the warps above it.
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
//
// All shared-memory traffic uses explicit .shared addresses (32-bit): generic LD/ST on
// the carved dynamic buffer would cost the long-scoreboard path.
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
__device__ __forceinline__ void named_bar_arrive(int id, int nthr) {
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(nthr) : "memory");
}
__device__ __forceinline__ float lds_f32(uint32_t a) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ float4 lds_f32x4(uint32_t a) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts_f32(uint32_t a, float v) {
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory");
}

template <int K, int RPW>
struct Ldl {
    static constexpr int NW = K / RPW;       // warps per group (4)
    static_assert(NW == 4, "a solver group is the four TMEM lane quadrants");

    // Four column steps J = j0 .. j0+3 (j0 % 4 == 0) on a row window of W registers:
    // a[c] holds the row's entry for column j0 + c.  The last step writes its update four
    // registers down, so after the call a[c] holds column j0 + 4 + c: the code of a step
    // never depends on J, every register index is static, and the loop stays rolled — the
    // whole factorisation is a few KB of instructions that live in the instruction cache
    // (a fully unrolled column loop is >100 KB of straight-line code and ran
    // instruction-fetch bound).  nthr = threads of the warps that still own a row >= j0.
    // `last` (warp-uniform): j0 + 3 is this warp's last row — publish, arrive, no update.
    template <int W>
    static __device__ __forceinline__ void body4(float (&a)[K], float& rhs, bool& ok, const float ridge,
                                                 const int r, const uint32_t cbs, const int j0, const int bar_id,
                                                 const int nthr, const bool last) {
        const int g = j0 >> 2;
        const uint32_t zb = cbs + 4u * ldl_cols_floats<K>();
        const uint32_t S4 = 4u * (uint32_t)(K - 4 * g);                                  // column stride (bytes)
        const uint32_t col0 = cbs + 4u * (uint32_t)(4 * g * K - 8 * g * (g - 1) - 4 * g);   // element (i, j0) at col0 + 4 i
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            const int J = j0 + s;
            const uint32_t cj = col0 + (uint32_t)s * S4;
            {
                const float d = a[s] + ridge;
                if (r >= J) sts_f32(cj + 4u * r, r == J ? d : a[s]);
                if (r == J) {
                    sts_f32(zb + 4u * J, rhs);
                    ok = d > 0.0f;
                }
            }
            if (s == 3 && last) {
                if (nthr > 32) named_bar_arrive(bar_id, nthr);
                return;
            }
            named_bar_sync(bar_id, nthr);
            const float inv = fast_rcp(lds_f32(cj + 4u * J));
            const float t = a[s] * inv;
            rhs = fmaf(-t, lds_f32(zb + 4u * J), rhs);
            const uint32_t cw = cj + 4u * j0;                  // window start: element (j0 + c, J) at cw + 4 c
            if (s < 3) {
#pragma unroll
                for (int c = s + 1; c < 4; ++c) a[c] = fmaf(-t, lds_f32(cw + 4u * c), a[c]);
#pragma unroll
                for (int c4 = 4; c4 < W; c4 += 4) {
                    const float4 v = lds_f32x4(cw + 4u * c4);
                    a[c4 + 0] = fmaf(-t, v.x, a[c4 + 0]);
                    a[c4 + 1] = fmaf(-t, v.y, a[c4 + 1]);
                    a[c4 + 2] = fmaf(-t, v.z, a[c4 + 2]);
                    a[c4 + 3] = fmaf(-t, v.w, a[c4 + 3]);
                }
            } else {
#pragma unroll
                for (int c4 = 4; c4 < W; c4 += 4) {
                    const float4 v = lds_f32x4(cw + 4u * c4);
                    a[c4 - 4] = fmaf(-t, v.x, a[c4 + 0]);
                    a[c4 - 3] = fmaf(-t, v.y, a[c4 + 1]);
                    a[c4 - 2] = fmaf(-t, v.z, a[c4 + 2]);
                    a[c4 - 1] = fmaf(-t, v.w, a[c4 + 3]);
                }
            }
        }
    }

    // phase p = columns [RPW p, RPW (p+1)) on a window of W = RPW (q + 1 - p) registers
    template <int W>
    static __device__ __forceinline__ void phase(float (&a)[K], float& rhs, bool& ok, const float ridge, const int r,
                                                 const uint32_t cbs, const int p, const int bar_id) {
        const int nthr = 32 * (NW - p);
#pragma unroll 1
        for (int it = 0; it < RPW / 4; ++it)
            body4<W>(a, rhs, ok, ridge, r, cbs, p * RPW + 4 * it, bar_id, nthr, W == RPW && it == RPW / 4 - 1);
    }

    // Factor + forward substitution.  a[] = this thread's matrix row (row position r in
    // warp quadrant q), rhs = its right-hand-side entry, ridge is added on the pivot.
    // Returns true if thread r's pivot was positive.
    static __device__ __forceinline__ bool factor(float (&a)[K], float rhs, const float ridge, const int r,
                                                  const int q, const uint32_t cbs, const int bar_id) {
        bool ok = true;
#pragma unroll 1
        for (int p = 0; p <= q; ++p) {
            const int w = q + 1 - p;
            if (w == 4) phase<4 * RPW>(a, rhs, ok, ridge, r, cbs, p, bar_id);
            else if (w == 3) phase<3 * RPW>(a, rhs, ok, ridge, r, cbs, p, bar_id);
            else if (w == 2) phase<2 * RPW>(a, rhs, ok, ridge, r, cbs, p, bar_id);
            else phase<RPW>(a, rhs, ok, ridge, r, cbs, p, bar_id);
        }
        return ok;
    }

    // Back substitution.  Needs a group barrier (phase_bar) between blocks; ends with x_r
    // returned to thread r and stored in xb[r].  `seg` = lane % RPW.
    static __device__ __forceinline__ float backsolve(const int r, const int q, const int seg, const uint32_t cbs,
                                                      const int phase_bar) {
        const uint32_t zb = cbs + 4u * ldl_cols_floats<K>();
        const uint32_t xb = zb + 4u * K;
        const int g = r >> 2, jj = r & 3;
        const uint32_t mycol = cbs + 4u * (uint32_t)(4 * g * K - 8 * g * (g - 1) + jj * (K - 4 * g) - 4 * g);
        const float dinv = fast_rcp(lds_f32(mycol + 4u * r));
        float acc = lds_f32(zb + 4u * r);
        float x = 0.0f;
#pragma unroll
        for (int blk = NW - 1; blk >= 0; --blk) {
            if (q == blk) {
                float lt[RPW];
#pragma unroll
                for (int i4 = 0; i4 < RPW; i4 += 4) {
                    const float4 v = lds_f32x4(mycol + 4u * (blk * RPW + i4));
                    lt[i4] = v.x; lt[i4 + 1] = v.y; lt[i4 + 2] = v.z; lt[i4 + 3] = v.w;
                }
#pragma unroll
                for (int ii = RPW - 1; ii >= 0; --ii) {
                    const float mine = acc * dinv;
                    const float xi = __shfl_sync(0xffffffffu, mine, ii, RPW);
                    if (seg < ii) acc = fmaf(-lt[ii], xi, acc);
                    if (seg == ii) x = mine;
                }
                sts_f32(xb + 4u * r, x);
            }
            if (blk > 0) {
                named_bar_sync(phase_bar, NW * 32);
                if (q < blk) {
#pragma unroll
                    for (int i4 = 0; i4 < RPW; i4 += 4) {
                        const float4 c4 = lds_f32x4(mycol + 4u * (blk * RPW + i4));
                        const float4 x4 = lds_f32x4(xb + 4u * (blk * RPW + i4));
                        acc = fmaf(-c4.x, x4.x, acc);
                        acc = fmaf(-c4.y, x4.y, acc);
                        acc = fmaf(-c4.z, x4.z, acc);
                        acc = fmaf(-c4.w, x4.w, acc);
                    }
                }
            }
        }
        return x;
    }
};

}  // namespace b200als

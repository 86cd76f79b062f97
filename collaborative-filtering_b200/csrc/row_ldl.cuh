// row_ldl.cuh — thread-per-row LDL^T solve fed straight from TMEM
// (A6/A7/A11: the per-row normal equations, als.py:426-428 / 455-461, helpers.py:19-20).
//
// The tcgen05 Gram kernel leaves row i of the k x k normal matrix in TMEM lane i.  The matrix
// STAYS in TMEM while it is factored: warp q owns rows [RPW q, RPW (q+1)) and runs phases
// p = 0..q; in phase p a thread pulls only the RPW columns [RPW p, RPW (p+1)) of its row out of
// TMEM (tcgen05.ld), so the register footprint is one RPW-wide window, not a k-wide row.
//
//   phase p, per thread (row r >= RPW p):
//     left-looking: window -= sum over finished columns j < RPW p of l_rj d_j l_cj, four
//       columns per step: own entries u_rj (LDS.32), 1/d_j (broadcast LDS.128), column
//       entries u_cj for the window rows c (broadcast LDS.128) — no barrier, pure FFMA stream;
//     in-block: RPW/4 "block pivot" steps of four columns (body4): publish the raw entries,
//       ONE named barrier among the warps that still own a row >= j, redundant 4 x 4 LDL^T,
//       rank-4 update of the rest of the window written four registers down (the window slides,
//       every register index is static, the loop stays rolled and instruction-cache resident).
//
// The packed column store (unscaled u_rj = l_rj d_j) doubles as L^T for the back substitution
// (thread j owns column j: x_j = (z_j - sum_{i>j} col_j[i] x_i) / d_j), done block by block: the
// owning warp resolves its RPW rows with shuffles, the warps below fold the block in with
// per-thread LDS.128.  Shared memory per matrix: k^2/2 + 2k floats of columns + z + x + 1/d.
//
// A solver group is 4 warps = the 4 TMEM lane quadrants.  k = 128: one matrix per group,
// RPW = 32 rows per warp.  k = 64: the Gram MMAs run with M = 64, whose accumulator uses
// 16 lanes of every quadrant, and two rows' matrices are interleaved (lanes 0-15 / 16-31
// of each quadrant), so RPW = 16 and a group factors two matrices in lock step.
//
// All shared-memory traffic uses explicit .shared addresses (32-bit): generic LD/ST on
// the carved dynamic buffer would cost the long-scoreboard path.
#pragma once
#include <cuda_bf16.h>

#include "common.cuh"
#include "umma_ptx.cuh"

namespace b200als {

#ifdef B200ALS_TIMING
__device__ float g_dbg[64];
__device__ unsigned long long g_timing[16];
#define LDL_TICK(i) do { if (blockIdx.x == 0 && threadIdx.x == 96) { const long long t__ = clock64(); g_timing[i] += (unsigned long long)(t__ - tl__); tl__ = t__; } } while (0)
#else
#define LDL_TICK(i) do { } while (0)
#endif

// element (i, j), i >= j, of the packed column store sits at ldl_col_base(j) + i
// (column j is allocated from row 4*(j/4), so every column keeps i % 4 alignment for LDS.128)
template <int K>
__host__ __device__ constexpr int ldl_col_base(int j) {
    const int g = j >> 2, jj = j & 3;
    return 4 * g * K - 8 * g * (g - 1) + jj * (K - 4 * g) - 4 * g;
}
template <int K>
__host__ __device__ constexpr int ldl_cols_floats() { return K * K / 2 + 2 * K; }
// per-matrix buffer: columns | z[K] | x[K] | 1/d[K] | raw scratch | pad — size = 16 (mod 32) floats
// so that the two interleaved matrices of a k=64 group sit 16 banks apart
template <int K>
__host__ __device__ constexpr int ldl_mat_floats() {
    const int raw = ldl_cols_floats<K>() + 3 * K + 2 * (4 * K + 8);   // columns | z | x | 1/d | raw-row scratch (2 buffers)
    return ((raw + 31) / 32) * 32 + 16;
}

__device__ __forceinline__ float fast_rcp(float d) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d));
    return r * fmaf(-d, r, 2.0f);             // one Newton step: < 1 ulp
}

__device__ __forceinline__ float rcp_approx(float d) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d));
    return r;
}

__device__ __forceinline__ float rsqrt_approx(float d) {
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d));
    return r;
}

__device__ __forceinline__ void named_bar_sync(int id, int nthr) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthr) : "memory");
}
__device__ __forceinline__ void named_bar_arrive(int id, int nthr) {
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(nthr) : "memory");
}
// Shared-memory accessors on 32-bit .shared addresses.  The pointer is rebuilt with
// cvta.shared->generic, which NVVM's address-space inference folds back into LDS/STS, so
// the compiler can batch the loads ahead of the FFMAs (volatile asm loads would pin each
// LDS right in front of its consumers and expose its latency sixteen times per column).
__device__ __forceinline__ const float* smem_ptr(uint32_t a) {
    return reinterpret_cast<const float*>(__cvta_shared_to_generic((size_t)a));
}
__device__ __forceinline__ float lds_f32(uint32_t a) { return *smem_ptr(a); }
__device__ __forceinline__ float4 lds_f32x4(uint32_t a) { return *reinterpret_cast<const float4*>(smem_ptr(a)); }
__device__ __forceinline__ void sts_f32(uint32_t a, float v) { *const_cast<float*>(smem_ptr(a)) = v; }

// predicated store: one STS under a predicate, never a branch (lanes of a warp differ in `on`)
__device__ __forceinline__ void sts_f32_if(uint32_t a, float v, bool on) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %2, 0;\n"
        "@p st.shared.f32 [%0], %1;\n"
        "}\n" ::"r"(a), "f"(v), "r"((int)on)
        : "memory");
}

__device__ __forceinline__ void sts_b64_if(uint32_t a, uint32_t x, uint32_t y, bool on) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %3, 0;\n"
        "@p st.shared.v2.b32 [%0], {%1, %2};\n"
        "}\n" ::"r"(a), "r"(x), "r"(y), "r"((int)on)
        : "memory");
}
// Stores WITHOUT a memory clobber: the compiler may batch later (non-aliasing) loads ahead of them; they stay
// ordered before every later volatile asm (barriers, fences).  Only for addresses the thread does not read back.
__device__ __forceinline__ void sts_f32_nc(uint32_t a, float v) {
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v));
}
__device__ __forceinline__ void sts_b64_nc(uint32_t a, uint32_t x, uint32_t y) {
    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(a), "r"(x), "r"(y));
}
__device__ __forceinline__ uint32_t bf162_bits(__nv_bfloat162 v) { return *reinterpret_cast<uint32_t*>(&v); }

// (o0..o3) = (i0..i3) - t0 v0 - t1 v1 - t2 v2 - t3 v3, entry by entry in that order (t0 first)
__device__ __forceinline__ void rank4_sub(float& o0, float& o1, float& o2, float& o3, const float i0, const float i1,
                                          const float i2, const float i3, const float t0, const float t1, const float t2,
                                          const float t3, const float4 v0, const float4 v1, const float4 v2,
                                          const float4 v3) {
#if B200ALS_FFMA2
    const uint64_t T0 = pack_f32x2(-t0, -t0), T1 = pack_f32x2(-t1, -t1), T2 = pack_f32x2(-t2, -t2), T3 = pack_f32x2(-t3, -t3);
    uint64_t x = pack_f32x2(i0, i1), y = pack_f32x2(i2, i3);
    x = fma_f32x2(T0, pack_f32x2(v0.x, v0.y), x); y = fma_f32x2(T0, pack_f32x2(v0.z, v0.w), y);
    x = fma_f32x2(T1, pack_f32x2(v1.x, v1.y), x); y = fma_f32x2(T1, pack_f32x2(v1.z, v1.w), y);
    x = fma_f32x2(T2, pack_f32x2(v2.x, v2.y), x); y = fma_f32x2(T2, pack_f32x2(v2.z, v2.w), y);
    x = fma_f32x2(T3, pack_f32x2(v3.x, v3.y), x); y = fma_f32x2(T3, pack_f32x2(v3.z, v3.w), y);
    unpack_f32x2(x, o0, o1);
    unpack_f32x2(y, o2, o3);
#else
    o0 = fmaf(-t3, v3.x, fmaf(-t2, v2.x, fmaf(-t1, v1.x, fmaf(-t0, v0.x, i0))));
    o1 = fmaf(-t3, v3.y, fmaf(-t2, v2.y, fmaf(-t1, v1.y, fmaf(-t0, v0.y, i1))));
    o2 = fmaf(-t3, v3.z, fmaf(-t2, v2.z, fmaf(-t1, v1.z, fmaf(-t0, v0.z, i2))));
    o3 = fmaf(-t3, v3.w, fmaf(-t2, v2.w, fmaf(-t1, v1.w, fmaf(-t0, v0.w, i3))));
#endif
}

// 0 (default): diag4; 1: diag4p, the software-pipelined variant — correct (parity suite green) but measured NOT faster
// on B200: C4 user half-step 35.3 ms (unchanged), k = 64 shapes ~5 % slower (its 96-register solver starts to spill)
#ifndef B200ALS_DIAG_PIPELINE
#define B200ALS_DIAG_PIPELINE 0
#endif
// 1: the second half of a diagonal block's steps runs a copy of diag4 that updates only the window entries still inside
// the block (3 groups of four instead of 7 at k = 128, 1 instead of 3 at k = 64)
// 1: the forward substitution of the rows below a diagonal block runs in right-looking order (subst_tri_rl)
#ifndef B200ALS_SUBST_RL
#define B200ALS_SUBST_RL 1
#endif
// 1: at k = 128 the right-looking substitution drops the dead writes of the multipliers into the window registers and
// takes the panel flag as a compile-time constant (C4 user half-step -2.3 %; at k = 64 the same change costs 0.3 %
// — its 80-register solver starts to spill — so it is not applied there)
#ifndef B200ALS_SUBST_LEAN
#define B200ALS_SUBST_LEAN 1
#endif
#ifndef B200ALS_DIAG_TAPER
#define B200ALS_DIAG_TAPER 1
#endif

// 1: the panel entries of the tensor-core update are scaled and split into bf16 hi + lo with packed FMUL2 / FADD2 and
// two-instruction bf16 -> fp32 widenings (same IEEE operations per entry: identical bits)
#ifndef B200ALS_PANEL2
#define B200ALS_PANEL2 1
#endif
__device__ __forceinline__ uint64_t mul_f32x2(uint64_t a, uint64_t b) {
    uint64_t d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ uint64_t sub_f32x2(uint64_t a, uint64_t b) {
    uint64_t d;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
// w_k = u_k rs_k (k = 0..3) as bf16 hi + lo: four panel columns = 8 bytes of the row's core-matrix line at `ta` (hi)
// and `ta + tile_lo` (lo)
__device__ __forceinline__ void panel_store4(const float u0, const float u1, const float u2, const float u3,
                                             const float4 rs, const uint32_t ta, const uint32_t tile_lo) {
#ifdef B200ALS_MMA_NULL
    const __nv_bfloat162 z = __floats2bfloat162_rn(0.0f, 0.0f);
    sts_b64_nc(ta, bf162_bits(z), bf162_bits(z));
    sts_b64_nc(ta + tile_lo, bf162_bits(z), bf162_bits(z));
#elif B200ALS_PANEL2
    const uint64_t w01 = mul_f32x2(pack_f32x2(u0, u1), pack_f32x2(rs.x, rs.y));
    const uint64_t w23 = mul_f32x2(pack_f32x2(u2, u3), pack_f32x2(rs.z, rs.w));
    float w0, w1, w2, w3;
    unpack_f32x2(w01, w0, w1);
    unpack_f32x2(w23, w2, w3);
    const uint32_t h01 = bf162_bits(__floats2bfloat162_rn(w0, w1)), h23 = bf162_bits(__floats2bfloat162_rn(w2, w3));
    // bf16 -> fp32 is the same bits in the upper half of the word
    const uint64_t f01 = pack_f32x2(__uint_as_float(h01 << 16), __uint_as_float(h01 & 0xffff0000u));
    const uint64_t f23 = pack_f32x2(__uint_as_float(h23 << 16), __uint_as_float(h23 & 0xffff0000u));
    float l0, l1, l2, l3;
    unpack_f32x2(sub_f32x2(w01, f01), l0, l1);
    unpack_f32x2(sub_f32x2(w23, f23), l2, l3);
    sts_b64_nc(ta, h01, h23);
    sts_b64_nc(ta + tile_lo, bf162_bits(__floats2bfloat162_rn(l0, l1)), bf162_bits(__floats2bfloat162_rn(l2, l3)));
#else
    const float w0 = u0 * rs.x, w1 = u1 * rs.y, w2 = u2 * rs.z, w3 = u3 * rs.w;
    const __nv_bfloat162 h01 = __floats2bfloat162_rn(w0, w1), h23 = __floats2bfloat162_rn(w2, w3);
    const float2 f01 = __bfloat1622float2(h01), f23 = __bfloat1622float2(h23);
    const __nv_bfloat162 l01 = __floats2bfloat162_rn(w0 - f01.x, w1 - f01.y);
    const __nv_bfloat162 l23 = __floats2bfloat162_rn(w2 - f23.x, w3 - f23.y);
    sts_b64_nc(ta, bf162_bits(h01), bf162_bits(h23));
    sts_b64_nc(ta + tile_lo, bf162_bits(l01), bf162_bits(l23));
#endif
}

template <int K, int RPW>
struct Ldl {
    static constexpr int NW = K / RPW;       // warps per group (4)
    static constexpr bool SUBST_LEAN = B200ALS_SUBST_LEAN && K >= 128;
    static constexpr int RAWSZ = 4 * K + 8;  // floats of one raw-row scratch buffer: rho[4][K] | rhs[4] | 4 dummy words (stores of lanes that have nothing to store)
    static_assert(NW == 4, "a solver group is the four TMEM lane quadrants");

    // byte offsets inside a matrix buffer
    static constexpr uint32_t Z_OFF = 4u * ldl_cols_floats<K>();
    static constexpr uint32_t X_OFF = Z_OFF + 4u * K;
    static constexpr uint32_t INV_OFF = X_OFF + 4u * K;
    static constexpr uint32_t RAW_OFF = INV_OFF + 4u * K;

    // ---- trailing update on the tensor core ----------------------------------------------------
    // After phase p (p < MMA_PHASES) the rank-RPW update of everything below/right of the panel,
    //   A[m][n] -= sum_{j in phase p} w_mj w_nj,   w_mj = u_mj / sqrt(d_j),
    // is ONE tcgen05.mma group accumulating (negated) into the matrix's own TMEM accumulator.  The panel
    // W is written by the row threads as bf16 hi + lo (three products, like the Gram) in the K-major
    // no-swizzle canonical layout: 8-row x 16-byte core matrices, 128 B apart along the panel columns,
    // TILE_SBO apart along the rows.  It lives in the part of the packed column store that only later
    // phases fill (columns >= RPW (p+1)), which is free until this MMA has completed.  The last phase's
    // panel would not fit there, so phase NW-1 folds phase NW-2 in with FFMA (left4).
#ifndef B200ALS_MMA_PHASES
#define B200ALS_MMA_PHASES (NW - 1)
#endif
    static constexpr int MMA_PHASES = B200ALS_MMA_PHASES;
    static constexpr uint32_t TILE_SBO = (RPW / 8) * 128u;
    static __host__ __device__ constexpr uint32_t tile_off(int p) {          // bytes from the matrix buffer
        // the panel of phase NW-2 (one warp's rows) lives in the raw-row scratch, idle between the diagonal blocks
        if (p == NW - 2) return RAW_OFF;
        // first stored element of column J = RPW (p+1) is (J, J) at ldl_col_base(J) + J
        return ((4u * (uint32_t)(ldl_col_base<K>(RPW * (p + 1)) + RPW * (p + 1))) + 127u) & ~127u;
    }
    static __host__ __device__ constexpr uint32_t tile_half(int p) {         // bytes of the hi (or lo) panel
        return (uint32_t)((K - RPW * (p + 1)) / 8) * TILE_SBO;
    }
    static_assert(tile_off(0) + 2 * tile_half(0) <= Z_OFF, "phase-0 panel must fit behind the phase-0 columns");
    static_assert(tile_off(1) + 2 * tile_half(1) <= Z_OFF, "phase-1 panel must fit behind the phase-1 columns");
    static_assert(2 * tile_half(NW - 2) <= 2u * 4u * RAWSZ && RAW_OFF % 16 == 0, "phase-(NW-2) panel must fit in the raw-row scratch");

    // element (i, 4 g) of the column store at col_block(g) + 4 i; the block's columns are S4(g) bytes apart
    static __device__ __forceinline__ uint32_t col_block(const uint32_t cbs, const int g) {
        return cbs + 4u * (uint32_t)(4 * g * K - 8 * g * (g - 1) - 4 * g);
    }

    // Left-looking step: fold the finished columns 4 jb .. 4 jb + 3 into the window
    // a[c] = A[r][c0 + c], c0 % 4 == 0, c0 > 4 jb + 3:  a[c] -= sum_k (u_rk / d_k) u_(c0+c)k.
    static __device__ __forceinline__ void left4(float (&a)[RPW], const int r, const uint32_t cbs, const int jb,
                                                 const int c0) {
        const uint32_t S4 = 4u * (uint32_t)(K - 4 * jb);
        const uint32_t col0 = col_block(cbs, jb);
        const float4 inv = lds_f32x4(cbs + INV_OFF + 16u * (uint32_t)jb);
        const float t0 = lds_f32(col0 + 4u * r) * inv.x;
        const float t1 = lds_f32(col0 + S4 + 4u * r) * inv.y;
        const float t2 = lds_f32(col0 + 2u * S4 + 4u * r) * inv.z;
        const float t3 = lds_f32(col0 + 3u * S4 + 4u * r) * inv.w;
        const uint32_t cw = col0 + 4u * (uint32_t)c0;
#pragma unroll
        for (int c4 = 0; c4 < RPW; c4 += 4) {
            const float4 v0 = lds_f32x4(cw + 4u * c4);
            const float4 v1 = lds_f32x4(cw + S4 + 4u * c4);
            const float4 v2 = lds_f32x4(cw + 2u * S4 + 4u * c4);
            const float4 v3 = lds_f32x4(cw + 3u * S4 + 4u * c4);
            rank4_sub(a[c4 + 0], a[c4 + 1], a[c4 + 2], a[c4 + 3], a[c4 + 0], a[c4 + 1], a[c4 + 2], a[c4 + 3],
                      t0, t1, t2, t3, v0, v1, v2, v3);
        }
    }

    // Phase p has two parts.  (1) The warp that owns the diagonal block (rows RPW p .. RPW (p+1) - 1) factors it
    // ALONE, four columns per step (diag4), synchronising with __syncwarp only.  (2) One named barrier later
    // every row below forward-substitutes its RPW window entries against the finished block (subst): no
    // barrier, no redundant pivot work, a pure LDS.128 + FFMA stream.  (The first version ran all rows through
    // the four-column steps in lock step: RPW/4 block-wide barriers per phase and a ~1.4 k-cycle chain each.)
    //
    // diag4: four columns J = j0 .. j0+3 (j0 % 4 == 0) on the row window, a[c] = entry for column j0 + c.
    //   every row r >= j0 of the warp publishes its four raw entries rho_r = a[0..3] (the four pivot rows also
    //   their rhs); __syncwarp; every thread reads the 4 x 4 pivot block P (+ its rhs), factors it redundantly
    //   in registers (P = L4 D4 L4^T), forward-substitutes its own four entries (u_k = l_rk d_k go to the
    //   column store, t_k = l_rk) and applies the rank-4 update in terms of the RAW published rows:
    //   a[c] -= (t L4^-1) . rho_c.  The update writes four registers down, so the window slides and every
    //   register index is static while the loop over the steps stays rolled.
    // tile_on (warp-uniform): the phase feeds a tensor-core update, so 1/sqrt(d) of the pivots is kept (in the
    // x array, unused until the back substitution) for the rows below.  final_step: the block's last four
    // columns — nothing of this warp is left to update.
    // NG: groups of four window columns the rank-4 update writes.  After the step at position `it` of the block only
    // RPW - 4 (it + 1) window entries are still inside the block (the window slides), so NG >= RPW/4 - 1 - it is enough;
    // the columns behind them are never read again by this warp.
    template <int NG = RPW / 4 - 1>
    static __device__ __forceinline__ void diag4(float (&a)[RPW], float& rhs, bool& ok, const float ridge,
                                                 const int r, const uint32_t cbs, const int j0, const bool tile_on,
                                                 const bool final_step) {
        const int g = j0 >> 2;
        const uint32_t zb = cbs + Z_OFF;
        const uint32_t raw = cbs + RAW_OFF + (uint32_t)(g & 1) * (4u * RAWSZ);           // raw[parity][4][K] | zraw[4]
        const uint32_t S4 = 4u * (uint32_t)(K - 4 * g);                                  // column stride (bytes)
        const uint32_t col0 = col_block(cbs, g);                                         // element (i, j0) at col0 + 4 i
        const int dr = r - j0;                                                           // row position relative to the block
        // The ridge is added where the pivot block is read: its diagonal is the only place a pivot row's own
        // diagonal entry is used (the column store's (j, j) slot is never read: 1/d lives in the inv array).
        // every store below is unconditional: a lane with nothing to store writes a dummy word instead (a predicated
        // store compiles to a divergent branch region here, and this loop is one long dependent chain).  Finished
        // lanes (dr < 0) may publish their stale window: nobody reads raw rows above the block.
        const uint32_t dmy = raw + 4u * (uint32_t)(4 * K + 4);
#pragma unroll
        for (int s = 0; s < 4; ++s) sts_f32(raw + 4u * (uint32_t)(s * K + r), a[s]);
        sts_f32(raw + 4u * (uint32_t)(4 * K) + 4u * min((uint32_t)dr, 4u), rhs);
        __syncwarp();
        const float4 p0 = lds_f32x4(raw + 4u * (uint32_t)(0 * K + j0));     // P[0..3][0]
        const float4 p1 = lds_f32x4(raw + 4u * (uint32_t)(1 * K + j0));     // P[0..3][1]
        const float4 p2 = lds_f32x4(raw + 4u * (uint32_t)(2 * K + j0));
        const float4 p3 = lds_f32x4(raw + 4u * (uint32_t)(3 * K + j0));
        const float4 zr = lds_f32x4(raw + 4u * (uint32_t)(4 * K));
        // 4 x 4 LDL^T of the pivot block (lower part of P; MUFU.RCP is good to 1 ulp)
        // Two levels of two reciprocals each instead of four in a row: for a 2 x 2 block [a b; b c],
        // d0 = a, d1 = det / a with det = a c - b^2, so 1/d0 = rcp(a) and 1/d1 = a rcp(det) are independent MUFUs.
        const float d0 = p0.x + ridge, c11 = p1.y + ridge;
        const float det1 = fmaf(d0, c11, -p0.y * p0.y);
        const float i0 = rcp_approx(d0), i1 = d0 * rcp_approx(det1);
        const float d1 = det1 * i0;
        const float l10 = p0.y * i0, l20 = p0.z * i0, l30 = p0.w * i0;
        const float u21 = fmaf(-l20, p0.y, p1.z), u31 = fmaf(-l30, p0.y, p1.w);
        const float l21 = u21 * i1, l31 = u31 * i1;
        const float d2 = fmaf(-l21, u21, fmaf(-l20, p0.z, p2.z + ridge));
        const float u32 = fmaf(-l31, u21, fmaf(-l30, p0.z, p2.w));
        const float s33 = fmaf(-l31, u31, fmaf(-l30, p0.w, p3.w + ridge));
        const float det2 = fmaf(d2, s33, -u32 * u32);
        const float i2 = rcp_approx(d2), i3 = d2 * rcp_approx(det2);
        const float d3 = det2 * i2;
        const float l32 = u32 * i2;
        // own row: unscaled entries u_k = l_rk d_k and multipliers t_k = l_rk
        const float u0 = a[0], t0 = u0 * i0;
        const float u1 = fmaf(-t0, p0.y, a[1]), t1 = u1 * i1;
        const float u2 = fmaf(-t1, u21, fmaf(-t0, p0.z, a[2])), t2 = u2 * i2;
        const float u3 = fmaf(-t2, u32, fmaf(-t1, u31, fmaf(-t0, p0.w, a[3]))), t3 = u3 * i3;
        sts_f32(dr >= 0 ? col0 + 4u * r : dmy, u0);
        sts_f32(dr >= 1 ? col0 + S4 + 4u * r : dmy, u1);
        sts_f32(dr >= 2 ? col0 + 2u * S4 + 4u * r : dmy, u2);
        sts_f32(dr >= 3 ? col0 + 3u * S4 + 4u * r : dmy, u3);
        // rhs of the block rows, forward-substituted; the pivot rows keep theirs in the z array
        const float z0 = zr.x;
        const float z1 = fmaf(-l10, z0, zr.y);
        const float z2 = fmaf(-l21, z1, fmaf(-l20, z0, zr.z));
        const float z3 = fmaf(-l32, z2, fmaf(-l31, z1, fmaf(-l30, z0, zr.w)));
        {
            const bool piv = (unsigned)dr < 4u;
            // flat select chains (FSEL): a nested ?: became a divergent branch region
            const bool e1 = dr == 1, e2 = dr == 2, e3 = dr == 3;
            float zsel = e1 ? z1 : z0, dsel = e1 ? d1 : d0, isel = e1 ? i1 : i0;
            zsel = e2 ? z2 : zsel; dsel = e2 ? d2 : dsel; isel = e2 ? i2 : isel;
            zsel = e3 ? z3 : zsel; dsel = e3 ? d3 : dsel; isel = e3 ? i3 : isel;
            sts_f32(piv ? zb + 4u * r : dmy, zsel);
            sts_f32(piv ? cbs + INV_OFF + 4u * r : dmy, isel);
            if (tile_on) sts_f32(piv ? cbs + X_OFF + 4u * r : dmy, rsqrt_approx(dsel));   // 1/sqrt(d) for the panel of the rows below
            ok = piv ? (dsel > 0.0f) : ok;
        }
        if (final_step) return;
        rhs = fmaf(-t3, z3, fmaf(-t2, z2, fmaf(-t1, z1, fmaf(-t0, z0, rhs))));
        // tau = t L4^-1, so that l_r D4 l_c^T = tau . rho_c
        const float w3 = t3;
        const float w2 = fmaf(-w3, l32, t2);
        const float w1 = fmaf(-w3, l31, fmaf(-w2, l21, t1));
        const float w0 = fmaf(-w3, l30, fmaf(-w2, l20, fmaf(-w1, l10, t0)));
        const uint32_t rw = raw + 4u * (uint32_t)j0;           // rho_{j0+c}[s] at rw + 4 (s K + c)
        static_assert(NG >= 1 && NG <= RPW / 4 - 1, "update groups");
#pragma unroll
        for (int c4 = 4; c4 <= 4 * NG; c4 += 4) {
            const float4 v0 = lds_f32x4(rw + 4u * (uint32_t)(0 * K + c4));
            const float4 v1 = lds_f32x4(rw + 4u * (uint32_t)(1 * K + c4));
            const float4 v2 = lds_f32x4(rw + 4u * (uint32_t)(2 * K + c4));
            const float4 v3 = lds_f32x4(rw + 4u * (uint32_t)(3 * K + c4));
            rank4_sub(a[c4 - 4], a[c4 - 3], a[c4 - 2], a[c4 - 1], a[c4 + 0], a[c4 + 1], a[c4 + 2], a[c4 + 3],
                      w0, w1, w2, w3, v0, v1, v2, v3);
        }
    }

    // B200ALS_DIAG_TAPER == 2: the block's steps two at a time, each pair with the update width its first step needs
    // (7, 5, 3, 1 groups at k = 128): RPW/8 copies of the step
    template <int IT0>
    static __device__ __forceinline__ void diag_pairs(float (&a)[RPW], float& rhs, bool& ok, const float ridge,
                                                      const int r, const uint32_t cbs, const int jb, const bool tile_on) {
        constexpr int STEPS = RPW / 4;
        if constexpr (IT0 < STEPS) {
            constexpr int NGP = (STEPS - 1 - IT0) > 0 ? (STEPS - 1 - IT0) : 1;
#pragma unroll 1
            for (int it = IT0; it < IT0 + 2; ++it)
                diag4<NGP>(a, rhs, ok, ridge, r, cbs, jb + 4 * it, tile_on, it == STEPS - 1);
            diag_pairs<IT0 + 2>(a, rhs, ok, ridge, r, cbs, jb, tile_on);
        }
    }

    // diag4 software-pipelined (B200ALS_DIAG_PIPELINE=1; kept as a measured negative, off by default).  A step is a long dependent pivot chain followed by
    // the rank-4 update of the window's remaining columns; the NEXT step's chain needs only the first four of them.  So a
    // step applies only the first update group (columns 4..7 -> the next pivots) and hands its multipliers w and its
    // raw-row buffer to the next step, which applies the other groups (DEFER) in the same basic block as its own pivot
    // chain: independent LDS.128 + FFMA work that the scheduler interleaves with the chain's dependent instructions.
    // The raw-row scratch is double buffered by step parity, so the previous step's rows are intact while this step
    // publishes its own.  Same operations per entry, in the same order, as diag4.
    template <bool DEFER>
    static __device__ __forceinline__ void diag4p(float (&a)[RPW], float& rhs, bool& ok, float (&wp)[4], uint32_t& rwp,
                                                  const float ridge, const int r, const uint32_t cbs, const int j0,
                                                  const bool final_step) {
        const int g = j0 >> 2;
        const uint32_t zb = cbs + Z_OFF;
        const uint32_t raw = cbs + RAW_OFF + (uint32_t)(g & 1) * (4u * RAWSZ);
        const uint32_t S4 = 4u * (uint32_t)(K - 4 * g);
        const uint32_t col0 = col_block(cbs, g);
        const int dr = r - j0;
        const uint32_t dmy = raw + 4u * (uint32_t)(4 * K + 4);
#pragma unroll
        for (int s = 0; s < 4; ++s) sts_f32(raw + 4u * (uint32_t)(s * K + r), a[s]);
        sts_f32(raw + 4u * (uint32_t)(4 * K) + 4u * min((uint32_t)dr, 4u), rhs);
        __syncwarp();
        const float4 p0 = lds_f32x4(raw + 4u * (uint32_t)(0 * K + j0));
        const float4 p1 = lds_f32x4(raw + 4u * (uint32_t)(1 * K + j0));
        const float4 p2 = lds_f32x4(raw + 4u * (uint32_t)(2 * K + j0));
        const float4 p3 = lds_f32x4(raw + 4u * (uint32_t)(3 * K + j0));
        const float4 zr = lds_f32x4(raw + 4u * (uint32_t)(4 * K));
        if (DEFER) {
            // the previous step's update of its window columns 8 .. RPW-1 (that step only slid its first group)
            const float q0 = wp[0], q1 = wp[1], q2 = wp[2], q3 = wp[3];
#pragma unroll
            for (int c4 = 8; c4 < RPW; c4 += 4) {
                const float4 v0 = lds_f32x4(rwp + 4u * (uint32_t)(0 * K + c4));
                const float4 v1 = lds_f32x4(rwp + 4u * (uint32_t)(1 * K + c4));
                const float4 v2 = lds_f32x4(rwp + 4u * (uint32_t)(2 * K + c4));
                const float4 v3 = lds_f32x4(rwp + 4u * (uint32_t)(3 * K + c4));
                a[c4 - 4] = fmaf(-q3, v3.x, fmaf(-q2, v2.x, fmaf(-q1, v1.x, fmaf(-q0, v0.x, a[c4 + 0]))));
                a[c4 - 3] = fmaf(-q3, v3.y, fmaf(-q2, v2.y, fmaf(-q1, v1.y, fmaf(-q0, v0.y, a[c4 + 1]))));
                a[c4 - 2] = fmaf(-q3, v3.z, fmaf(-q2, v2.z, fmaf(-q1, v1.z, fmaf(-q0, v0.z, a[c4 + 2]))));
                a[c4 - 1] = fmaf(-q3, v3.w, fmaf(-q2, v2.w, fmaf(-q1, v1.w, fmaf(-q0, v0.w, a[c4 + 3]))));
            }
        }
        const float d0 = p0.x + ridge, c11 = p1.y + ridge;
        const float det1 = fmaf(d0, c11, -p0.y * p0.y);
        const float i0 = rcp_approx(d0), i1 = d0 * rcp_approx(det1);
        const float d1 = det1 * i0;
        const float l10 = p0.y * i0, l20 = p0.z * i0, l30 = p0.w * i0;
        const float u21 = fmaf(-l20, p0.y, p1.z), u31 = fmaf(-l30, p0.y, p1.w);
        const float l21 = u21 * i1, l31 = u31 * i1;
        const float d2 = fmaf(-l21, u21, fmaf(-l20, p0.z, p2.z + ridge));
        const float u32 = fmaf(-l31, u21, fmaf(-l30, p0.z, p2.w));
        const float s33 = fmaf(-l31, u31, fmaf(-l30, p0.w, p3.w + ridge));
        const float det2 = fmaf(d2, s33, -u32 * u32);
        const float i2 = rcp_approx(d2), i3 = d2 * rcp_approx(det2);
        const float d3 = det2 * i2;
        const float l32 = u32 * i2;
        const float u0 = a[0], t0 = u0 * i0;
        const float u1 = fmaf(-t0, p0.y, a[1]), t1 = u1 * i1;
        const float u2 = fmaf(-t1, u21, fmaf(-t0, p0.z, a[2])), t2 = u2 * i2;
        const float u3 = fmaf(-t2, u32, fmaf(-t1, u31, fmaf(-t0, p0.w, a[3]))), t3 = u3 * i3;
        sts_f32(dr >= 0 ? col0 + 4u * r : dmy, u0);
        sts_f32(dr >= 1 ? col0 + S4 + 4u * r : dmy, u1);
        sts_f32(dr >= 2 ? col0 + 2u * S4 + 4u * r : dmy, u2);
        sts_f32(dr >= 3 ? col0 + 3u * S4 + 4u * r : dmy, u3);
        const float z0 = zr.x;
        const float z1 = fmaf(-l10, z0, zr.y);
        const float z2 = fmaf(-l21, z1, fmaf(-l20, z0, zr.z));
        const float z3 = fmaf(-l32, z2, fmaf(-l31, z1, fmaf(-l30, z0, zr.w)));
        {
            const bool piv = (unsigned)dr < 4u;
            const bool e1 = dr == 1, e2 = dr == 2, e3 = dr == 3;
            float zsel = e1 ? z1 : z0, dsel = e1 ? d1 : d0, isel = e1 ? i1 : i0;
            zsel = e2 ? z2 : zsel; dsel = e2 ? d2 : dsel; isel = e2 ? i2 : isel;
            zsel = e3 ? z3 : zsel; dsel = e3 ? d3 : dsel; isel = e3 ? i3 : isel;
            sts_f32(piv ? zb + 4u * r : dmy, zsel);
            sts_f32(piv ? cbs + INV_OFF + 4u * r : dmy, isel);
            // 1/sqrt(d) for the tensor-core panel of the rows below: stored whether or not this phase feeds an update
            // (the x array is free until the back substitution) — no branch inside the step
            sts_f32(piv ? cbs + X_OFF + 4u * r : dmy, rsqrt_approx(dsel));
            ok = piv ? (dsel > 0.0f) : ok;
        }
        if (final_step) return;
        rhs = fmaf(-t3, z3, fmaf(-t2, z2, fmaf(-t1, z1, fmaf(-t0, z0, rhs))));
        const float w3 = t3;
        const float w2 = fmaf(-w3, l32, t2);
        const float w1 = fmaf(-w3, l31, fmaf(-w2, l21, t1));
        const float w0 = fmaf(-w3, l30, fmaf(-w2, l20, fmaf(-w1, l10, t0)));
        const uint32_t rw = raw + 4u * (uint32_t)j0;
        {
            // first update group only: the next step's pivots
            const float4 v0 = lds_f32x4(rw + 4u * (uint32_t)(0 * K + 4));
            const float4 v1 = lds_f32x4(rw + 4u * (uint32_t)(1 * K + 4));
            const float4 v2 = lds_f32x4(rw + 4u * (uint32_t)(2 * K + 4));
            const float4 v3 = lds_f32x4(rw + 4u * (uint32_t)(3 * K + 4));
            a[0] = fmaf(-w3, v3.x, fmaf(-w2, v2.x, fmaf(-w1, v1.x, fmaf(-w0, v0.x, a[4]))));
            a[1] = fmaf(-w3, v3.y, fmaf(-w2, v2.y, fmaf(-w1, v1.y, fmaf(-w0, v0.y, a[5]))));
            a[2] = fmaf(-w3, v3.z, fmaf(-w2, v2.z, fmaf(-w1, v1.z, fmaf(-w0, v0.z, a[6]))));
            a[3] = fmaf(-w3, v3.w, fmaf(-w2, v2.w, fmaf(-w1, v1.w, fmaf(-w0, v0.w, a[7]))));
        }
        wp[0] = w0; wp[1] = w1; wp[2] = w2; wp[3] = w3;
        rwp = rw;
    }

    // Rows below the diagonal block J0 .. J0 + RPW - 1 (finished, in the column store): forward substitution of the
    // row's window a[c] = A[r][J0 + c] (already updated for every earlier phase),
    //   u_j = a_j - sum_{i < j} t_i U[J0+j][J0+i],   t_j = u_j / d_j,   rhs -= sum_j t_j z_j,
    // four columns at a time: the earlier column groups fold in with four broadcast LDS.128 + 16 FFMA each, the
    // group's own 4 x 4 triangle sits in three more.  Every index is static (fully unrolled: RPW/4 (RPW/4 - 1)/2
    // group pairs), the multipliers t_j replace a[j] in place.  u_j go to the column store; tile_on: w_j = u_j / sqrt(d_j)
    // as bf16 hi + lo into this row's line of the tensor-core panel (tile_row: .shared address of the hi-panel row,
    // tile_lo: bytes from there to the lo panel).
    template <int NC>
    static __device__ __forceinline__ void subst_tri(float (&a)[RPW], float& rhs, const int r, const uint32_t cbs,
                                                     const int J0, const uint32_t tile_row, const uint32_t tile_lo,
                                                     const bool tile_on) {
        const int g0 = J0 >> 2;
#pragma unroll
        for (int gb = 0; gb < NC / 4; ++gb) {
            const int g = g0 + gb;
            const uint32_t S4 = 4u * (uint32_t)(K - 4 * g);
            const uint32_t colb = col_block(cbs, g);                 // element (i, 4 g + s) at colb + s S4 + 4 i
            const uint32_t rowoff = 16u * (uint32_t)g;               // rows 4 g .. 4 g + 3
#pragma unroll
            for (int ib = 0; ib < gb; ++ib) {
                const int gi = g0 + ib;
                const uint32_t S4i = 4u * (uint32_t)(K - 4 * gi);
                const uint32_t ci = col_block(cbs, gi) + rowoff;     // U[4g .. 4g+3][4 gi + s] at ci + s S4i
                const float4 v0 = lds_f32x4(ci);
                const float4 v1 = lds_f32x4(ci + S4i);
                const float4 v2 = lds_f32x4(ci + 2u * S4i);
                const float4 v3 = lds_f32x4(ci + 3u * S4i);
                const float t0 = a[4 * ib], t1 = a[4 * ib + 1], t2 = a[4 * ib + 2], t3 = a[4 * ib + 3];
                rank4_sub(a[4 * gb + 0], a[4 * gb + 1], a[4 * gb + 2], a[4 * gb + 3], a[4 * gb + 0], a[4 * gb + 1],
                          a[4 * gb + 2], a[4 * gb + 3], t0, t1, t2, t3, v0, v1, v2, v3);
            }
            const float4 c0 = lds_f32x4(colb + rowoff);              // d, U10, U20, U30 of the group's triangle
            const float4 c1 = lds_f32x4(colb + S4 + rowoff);         // -, -,   U21, U31
            const float4 c2 = lds_f32x4(colb + 2u * S4 + rowoff);    // -, -,   -,   U32
            const float4 inv = lds_f32x4(cbs + INV_OFF + rowoff);
            const float4 zz = lds_f32x4(cbs + Z_OFF + rowoff);
            const float u0 = a[4 * gb], t0 = u0 * inv.x;
            const float u1 = fmaf(-t0, c0.y, a[4 * gb + 1]), t1 = u1 * inv.y;
            const float u2 = fmaf(-t1, c1.z, fmaf(-t0, c0.z, a[4 * gb + 2])), t2 = u2 * inv.z;
            const float u3 = fmaf(-t2, c2.w, fmaf(-t1, c1.w, fmaf(-t0, c0.w, a[4 * gb + 3]))), t3 = u3 * inv.w;
            rhs = fmaf(-t3, zz.w, fmaf(-t2, zz.z, fmaf(-t1, zz.y, fmaf(-t0, zz.x, rhs))));
            sts_f32_nc(colb + 4u * r, u0);
            sts_f32_nc(colb + S4 + 4u * r, u1);
            sts_f32_nc(colb + 2u * S4 + 4u * r, u2);
            sts_f32_nc(colb + 3u * S4 + 4u * r, u3);
            a[4 * gb] = t0; a[4 * gb + 1] = t1; a[4 * gb + 2] = t2; a[4 * gb + 3] = t3;
            if (tile_on) {                                   // warp-uniform: this phase feeds the tensor-core update
                // w_k = u_k / sqrt(d_k) as bf16 hi + lo, four panel columns = 8 bytes of the row's core-matrix line
                const float4 rs = lds_f32x4(cbs + X_OFF + rowoff);
                const int jj = 4 * gb;                       // column inside the panel
                panel_store4(u0, u1, u2, u3, rs, tile_row + (uint32_t)((jj >> 3) * 128 + (jj & 7) * 2), tile_lo);
            }
        }
    }

    // The same substitution in RIGHT-LOOKING order (B200ALS_SUBST_RL): as soon as the triangle of column group gb is
    // solved, its four multipliers are folded into every later group of the window — the rest of this half and, when
    // `ext` is set (warp-uniform: the first half), the NC entries behind it.  Every entry still receives its column
    // groups in ascending order, one fused multiply-add per column, so the results are bit-identical to subst_tri +
    // the rectangle of subst; what changes is the address arithmetic: all loads of column group gb sit at
    // (column s of the group) + 16 (g0 + gc) with gc a compile-time constant, i.e. FOUR base registers per group and
    // immediates, where the left-looking order computes four addresses per (gb, ib) pair.
    template <int NC>
    static __device__ __forceinline__ void subst_tri_rl(float (&a)[RPW], float& rhs, const int r, const uint32_t cbs,
                                                        const int J0, const uint32_t tile_row, const uint32_t tile_lo,
                                                        const bool tile_on, const bool ext) {
        static_assert(2 * NC <= RPW, "the extension is the second half of the window");
        const int g0 = J0 >> 2;
        const uint32_t row16 = 16u * (uint32_t)g0;
#pragma unroll
        for (int gb = 0; gb < NC / 4; ++gb) {
            const int g = g0 + gb;
            const uint32_t S4 = 4u * (uint32_t)(K - 4 * g);
            const uint32_t colb = col_block(cbs, g);                 // element (i, 4 g + s) at colb + s S4 + 4 i
            const uint32_t b0 = colb + row16, b1 = b0 + S4, b2 = b1 + S4, b3 = b2 + S4;   // rows 4 g0 .. of the group's columns
            const float4 c0 = lds_f32x4(b0 + 16u * gb);              // d, U10, U20, U30 of the group's triangle
            const float4 c1 = lds_f32x4(b1 + 16u * gb);              // -, -,   U21, U31
            const float4 c2 = lds_f32x4(b2 + 16u * gb);              // -, -,   -,   U32
            const float4 inv = lds_f32x4(cbs + INV_OFF + row16 + 16u * gb);
            const float4 zz = lds_f32x4(cbs + Z_OFF + row16 + 16u * gb);
            const float u0 = a[4 * gb], t0 = u0 * inv.x;
            const float u1 = fmaf(-t0, c0.y, a[4 * gb + 1]), t1 = u1 * inv.y;
            const float u2 = fmaf(-t1, c1.z, fmaf(-t0, c0.z, a[4 * gb + 2])), t2 = u2 * inv.z;
            const float u3 = fmaf(-t2, c2.w, fmaf(-t1, c1.w, fmaf(-t0, c0.w, a[4 * gb + 3]))), t3 = u3 * inv.w;
            rhs = fmaf(-t3, zz.w, fmaf(-t2, zz.z, fmaf(-t1, zz.y, fmaf(-t0, zz.x, rhs))));
            const uint32_t own = 4u * (uint32_t)r - row16;           // this row inside the group's columns
            sts_f32_nc(b0 + own, u0);
            sts_f32_nc(b1 + own, u1);
            sts_f32_nc(b2 + own, u2);
            sts_f32_nc(b3 + own, u3);
            // the multipliers are consumed right here: nothing reads a[4 gb ..] again (the caller reloads the window),
            // and the lean form drops these register writes
            if constexpr (!SUBST_LEAN) { a[4 * gb] = t0; a[4 * gb + 1] = t1; a[4 * gb + 2] = t2; a[4 * gb + 3] = t3; }
            // fold the group into the later groups of this half ...
#pragma unroll
            for (int gc = gb + 1; gc < NC / 4; ++gc) {
                const float4 v0 = lds_f32x4(b0 + 16u * gc), v1 = lds_f32x4(b1 + 16u * gc);
                const float4 v2 = lds_f32x4(b2 + 16u * gc), v3 = lds_f32x4(b3 + 16u * gc);
                rank4_sub(a[4 * gc + 0], a[4 * gc + 1], a[4 * gc + 2], a[4 * gc + 3], a[4 * gc + 0], a[4 * gc + 1],
                          a[4 * gc + 2], a[4 * gc + 3], t0, t1, t2, t3, v0, v1, v2, v3);
            }
            // ... and, in the first half, into the half behind it
            if (ext) {
#pragma unroll
                for (int gc = NC / 4; gc < 2 * (NC / 4); ++gc) {
                    const float4 v0 = lds_f32x4(b0 + 16u * gc), v1 = lds_f32x4(b1 + 16u * gc);
                    const float4 v2 = lds_f32x4(b2 + 16u * gc), v3 = lds_f32x4(b3 + 16u * gc);
                    rank4_sub(a[4 * gc + 0], a[4 * gc + 1], a[4 * gc + 2], a[4 * gc + 3], a[4 * gc + 0], a[4 * gc + 1],
                              a[4 * gc + 2], a[4 * gc + 3], t0, t1, t2, t3, v0, v1, v2, v3);
                }
            }
            if (tile_on) {                                   // warp-uniform: this phase feeds the tensor-core update
                const float4 rs = lds_f32x4(cbs + X_OFF + row16 + 16u * gb);
                const int jj = 4 * gb;                       // column inside the panel
                panel_store4(u0, u1, u2, u3, rs, tile_row + (uint32_t)((jj >> 3) * 128 + (jj & 7) * 2), tile_lo);
            }
        }
    }

    // The window goes through ONE copy of the half-width code as two halves (a flat 32-column substitution is 1400
    // instructions, and instruction fetch was a quarter of its stall samples): half 0, then the rectangle
    // a[H..2H) -= t[0..H) . U, then the second half moves down into a[0..H) and the loop runs the same code on it.
    // Same operations as the flat form, in the same order per entry.
    static __device__ __forceinline__ void subst(float (&a)[RPW], float& rhs, const int r, const uint32_t cbs,
                                                 const int J0, const uint32_t tile_row, const uint32_t tile_lo,
                                                 const bool tile_on) {
        constexpr int H = RPW / 2;
        static_assert(H % 8 == 0, "a half must be whole 8-column core matrices of the panel");
#if B200ALS_SUBST_RL
#pragma unroll 1
        for (int hf = 0; hf < 2; ++hf) {
            subst_tri_rl<H>(a, rhs, r, cbs, J0 + H * hf, tile_row + (uint32_t)(H / 8) * 128u * (uint32_t)hf, tile_lo, tile_on,
                            hf == 0);
            if (hf == 0) {
#pragma unroll
                for (int i = 0; i < H; ++i) a[i] = a[H + i];
            }
        }
#else
#pragma unroll 1
        for (int hf = 0; hf < 2; ++hf) {
            const int Jh = J0 + H * hf;
            subst_tri<H>(a, rhs, r, cbs, Jh, tile_row + (uint32_t)(H / 8) * 128u * (uint32_t)hf, tile_lo, tile_on);
            if (hf == 0) {
                const int g0 = Jh >> 2;
#pragma unroll
                for (int jb = 0; jb < H / 4; ++jb) {
                    const uint32_t rowoff = 16u * (uint32_t)(g0 + H / 4 + jb);       // rows Jh + H + 4 jb .. + 3
#pragma unroll
                    for (int ib = 0; ib < H / 4; ++ib) {
                        const int gi = g0 + ib;
                        const uint32_t S4i = 4u * (uint32_t)(K - 4 * gi);
                        const uint32_t ci = col_block(cbs, gi) + rowoff;
                        const float4 v0 = lds_f32x4(ci);
                        const float4 v1 = lds_f32x4(ci + S4i);
                        const float4 v2 = lds_f32x4(ci + 2u * S4i);
                        const float4 v3 = lds_f32x4(ci + 3u * S4i);
                        const float t0 = a[4 * ib], t1 = a[4 * ib + 1], t2 = a[4 * ib + 2], t3 = a[4 * ib + 3];
                        const int c = H + 4 * jb;
                        rank4_sub(a[c + 0], a[c + 1], a[c + 2], a[c + 3], a[c + 0], a[c + 1], a[c + 2], a[c + 3],
                                  t0, t1, t2, t3, v0, v1, v2, v3);
                    }
                }
#pragma unroll
                for (int i = 0; i < H; ++i) a[i] = a[H + i];
            }
        }
#endif
    }

    // Factor + forward substitution.  load_window(p, a, is_last) fills a[c] = A[r][RPW p + c] from
    // TMEM (and releases the accumulator after this thread's last load); rhs = this row's
    // right-hand-side entry, ridge is added on the pivot.  r / q = row and warp in ELIMINATION order,
    // r_phys = the row's TMEM lane row (K-1-r for the reversed groups), d_tmem = the matrix's accumulator,
    // upd_bar = an mbarrier (count 1) of this matrix for the tensor-core updates, upd_base = updates it has completed so far
    // (MMA_PHASES per factorisation; the wait parity follows the running count).
    // Returns true if thread r's pivot was positive.
    template <class LoadWindow>
    static __device__ __forceinline__ bool factor(float rhs, const float ridge, const int r, const int q,
                                                  const int r_phys, const uint32_t cbs, const int bar_id,
                                                  const int start_bar, const uint32_t d_tmem, uint64_t* upd_bar, const uint32_t upd_base,
                                                  LoadWindow&& load_window) {
        bool ok = true;
        float a[RPW];
        const bool rev = r_phys != r;
#ifdef B200ALS_TIMING
        long long tl__ = clock64();
#endif
#pragma unroll 1
        for (int p = 0; p <= q; ++p) {
            const int nthr = 32 * (NW - p);
#ifdef B200ALS_TIMING
            float before[RPW];
#endif
            if (p > 0) {
                ptx::fence_proxy_async();                      // this thread's panel rows -> async proxy
                ptx::tc_fence_before();                        // its window loads precede the update MMA
                named_bar_sync(start_bar, nthr + 32);          // phase p-1 fully in the column store / panel
                LDL_TICK(9);
#ifdef B200ALS_TIMING
                if (p == 1) load_window(1, before, false);
#endif
                if (p - 1 < MMA_PHASES) {
                    if (r == RPW * p) {                        // one thread per matrix issues the update of phase p-1
                        ptx::tc_fence_after();
                        const int R0 = RPW * p;
                        // physical rows needed: [R0, K) forward, [0, K-R0) reversed; the descriptor base is
                        // moved so that row 0 of the MMA lands R0 rows before the panel (garbage rows only
                        // touch rows / columns of the accumulator that are finished)
                        const uint32_t base = cbs + tile_off(p - 1) - (rev ? 0u : (uint32_t)(R0 / 8) * TILE_SBO);
                        const uint32_t lo = tile_half(p - 1);
#ifdef B200ALS_MMA_NONEG
                        constexpr uint32_t IDESC = ptx::idesc_bf16_kk_nega(K, K) & ~(1u << 13);
#else
                        constexpr uint32_t IDESC = ptx::idesc_bf16_kk_nega(K, K);
#endif
#pragma unroll
                        for (int ks = 0; ks < RPW / 16; ++ks) {
#ifdef B200ALS_MMA_ZDESC
                            // debug: every row / chunk reads the first (zeroed) 16 bytes of the written panel
                            const uint32_t zb = cbs + tile_off(p - 1) + (rev ? 0u : 0u);
                            const uint64_t hi_d = ptx::smem_desc(zb, 0u, 0u, ptx::SWZ_NONE);
                            const uint64_t lo_d = hi_d;
                            (void)base; (void)lo;
#else
                            const uint64_t hi_d = ptx::smem_desc(base + ks * 256u, 128u, TILE_SBO, ptx::SWZ_NONE);
                            const uint64_t lo_d = ptx::smem_desc(base + lo + ks * 256u, 128u, TILE_SBO, ptx::SWZ_NONE);
#endif
#ifndef B200ALS_MMA_SKIP
                            ptx::umma_bf16(d_tmem, hi_d, hi_d, IDESC, 1u);
                            ptx::umma_bf16(d_tmem, hi_d, lo_d, IDESC, 1u);
                            ptx::umma_bf16(d_tmem, lo_d, hi_d, IDESC, 1u);
#endif
                        }
                        ptx::umma_commit(upd_bar);
                    }
                    ptx::mbar_wait(upd_bar, (upd_base + (uint32_t)(p - 1)) & 1u);
                    ptx::tc_fence_after();
                }
            }
            load_window(p, a, p == q);
#ifdef B200ALS_TIMING
            if (p == 1 && blockIdx.x == 0 && start_bar == 3 && r == RPW && g_dbg[63] == 0.0f) {
                for (int i = 0; i < 16; ++i) { g_dbg[i] = before[i]; g_dbg[16 + i] = a[i]; }
                g_dbg[63] = 1.0f;
            }
#endif
            LDL_TICK(8);
#ifdef B200ALS_MMA_NULL
            if (p > 0) {
#else
            if (p > MMA_PHASES) {
#endif
                // columns of the phases the tensor core did not fold in
#pragma unroll 1
#ifdef B200ALS_MMA_NULL
                for (int jb = 0; jb < p * (RPW / 4); ++jb) left4(a, r, cbs, jb, p * RPW);
#else
                for (int jb = MMA_PHASES * (RPW / 4); jb < p * (RPW / 4); ++jb) left4(a, r, cbs, jb, p * RPW);
#endif
                LDL_TICK(10);
            }
            // this thread's panel row for the update after this phase
            const bool tile_on = p < MMA_PHASES;
            if (p == q) {
                // the diagonal block: this warp alone, four columns per step
#if B200ALS_DIAG_PIPELINE
                {
                    float wp[4] = {0.0f, 0.0f, 0.0f, 0.0f};
                    uint32_t rwp = 0;
                    diag4p<false>(a, rhs, ok, wp, rwp, ridge, r, cbs, p * RPW, false);
#pragma unroll 1
                    for (int it = 1; it < RPW / 4; ++it)
                        diag4p<true>(a, rhs, ok, wp, rwp, ridge, r, cbs, p * RPW + 4 * it, it == RPW / 4 - 1);
                }
#elif B200ALS_DIAG_TAPER == 2
                diag_pairs<0>(a, rhs, ok, ridge, r, cbs, p * RPW, tile_on);
#elif B200ALS_DIAG_TAPER
                // the first half of the block's steps updates the whole window, the second half only the
                // RPW/2 - 4 entries that are still inside the block: two copies of the step instead of one
                {
                    constexpr int STEPS = RPW / 4, HALF = STEPS / 2;
#pragma unroll 1
                    for (int it = 0; it < HALF; ++it)
                        diag4<STEPS - 1>(a, rhs, ok, ridge, r, cbs, p * RPW + 4 * it, tile_on, false);
#pragma unroll 1
                    for (int it = HALF; it < STEPS; ++it)
                        diag4<HALF - 1>(a, rhs, ok, ridge, r, cbs, p * RPW + 4 * it, tile_on, it == STEPS - 1);
                }
#else
#pragma unroll 1
                for (int it = 0; it < RPW / 4; ++it)
                    diag4(a, rhs, ok, ridge, r, cbs, p * RPW + 4 * it, tile_on, it == RPW / 4 - 1);
#endif
                if (nthr > 32) {
                    named_bar_arrive(bar_id, nthr);        // block, z, 1/d, 1/sqrt(d) published: the rows below substitute
                    named_bar_arrive(start_bar, nthr);     // and this warp signs off on the phase
                }
                LDL_TICK(11);
            } else {
                const int trow = rev ? r_phys : r - RPW * (p + 1);
                const uint32_t tile_row = cbs + tile_off(tile_on ? p : 0) + (uint32_t)((trow >> 3) * (int)TILE_SBO + (trow & 7) * 16);
                const uint32_t tile_lo = tile_half(tile_on ? p : 0);
                named_bar_sync(bar_id, nthr);
                LDL_TICK(12);
                // rows below exist only in phases p <= NW - 2: when every such phase feeds a tensor-core update the flag
                // is a compile-time constant and the four branches per pass around the panel stores go away
                if constexpr (SUBST_LEAN && MMA_PHASES >= NW - 1) subst(a, rhs, r, cbs, p * RPW, tile_row, tile_lo, true);
                else subst(a, rhs, r, cbs, p * RPW, tile_row, tile_lo, tile_on);
                LDL_TICK(13);
            }
        }
        return ok;
    }

    // Back substitution.  Needs a group barrier (phase_bar) between blocks; ends with x_r
    // returned to thread r and stored in xb[r].  `seg` = lane % RPW.
    // lane_xor: 0, or RPW-1 when the lanes of a segment hold their rows in reverse order.
    static __device__ __forceinline__ float backsolve(const int r, const int q, const int seg, const int lane_xor,
                                                      const uint32_t cbs, const int phase_bar) {
        const uint32_t zb = cbs + Z_OFF;
        const uint32_t xb = cbs + X_OFF;
        const int g = r >> 2, jj = r & 3;
        const uint32_t mycol = cbs + 4u * (uint32_t)(4 * g * K - 8 * g * (g - 1) + jj * (K - 4 * g) - 4 * g);
        const float dinv = lds_f32(cbs + INV_OFF + 4u * r);       // 1/d_r as the factorisation used it
        float acc = lds_f32(zb + 4u * r);
        float x = 0.0f;
        // the block loop stays rolled (one copy of the RPW-step in-warp solve in the instruction cache: the
        // solver warps of a CTA sit at different places of a ~35 KB loop body and instruction fetch was 20 % of
        // their issue stalls)
#pragma unroll 1
        for (int blk = NW - 1; blk >= 0; --blk) {
            if (q == blk) {
                float lt[RPW];
#pragma unroll
                for (int i4 = 0; i4 < RPW; i4 += 4) {
                    const float4 v = lds_f32x4(mycol + 4u * (uint32_t)(blk * RPW + i4));
                    lt[i4] = v.x; lt[i4 + 1] = v.y; lt[i4 + 2] = v.z; lt[i4 + 3] = v.w;
                }
#pragma unroll
                for (int ii = RPW - 1; ii >= 0; --ii) {
                    const float mine = acc * dinv;
                    const float xi = __shfl_sync(0xffffffffu, mine, ii ^ lane_xor, RPW);
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
                        const float4 c4 = lds_f32x4(mycol + 4u * (uint32_t)(blk * RPW + i4));
                        const float4 x4 = lds_f32x4(xb + 4u * (uint32_t)(blk * RPW + i4));
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

// row_step_umma.cuh — tcgen05 / TMEM Gram path for k in {64, 128}.
//
// One persistent CTA per SM, warp-specialised, rows dealt round-robin from the
// longest-first schedule.  Gathered operand: the factor matrix pre-split into
// bf16 (hi | lo) records, H[row] = [hi(k) | lo(k)], 4k bytes — the same bytes per
// rating as the float32 row.  hi + lo carries ~17 mantissa bits, and
//     F^T F  ~=  hi^T hi + hi^T lo + (hi^T lo)^T            (lo^T lo ~ 2^-18 dropped)
// so two bf16 MMAs give a float32-quality Gram at TF32 cost:
//   k = 64 : one MMA  D[128 x 64]  = [hi;lo]^T hi      (rows 0-63 = P, rows 64-127 = Q^T)
//   k = 128: one MMA  D[128 x 256] = hi^T [hi | lo]    (cols 0-127 = P, cols 128-255 = Q)
// plus a 16-column MMA against [q_hi, q_lo, 1] per rating for c = F^T q and s = F^T 1.
// Ratings are the MMA K dimension, so both operands are MN-major: a gathered record
// is one 128-byte line per 64 factors, written straight into the 128B-swizzled
// canonical layout by 16-byte cp.async (no register pass, no transposition).
//
// Warp roles:  [solver warps: 4 per group] [meta] [gather x4] [mma]
//   meta   : streams (index, rating) of each 32-rating stage into smem (cp.async → mbarrier)
//   gather : per stage, 128 threads cp.async the 32 records (and the 32 opposite biases)
//   mma    : builds the [q_hi,q_lo,1] operand, issues tcgen05.mma, commits stage / accumulator
//   solver : drains the accumulator from TMEM into 4x4 lower-triangle tiles, then runs
//            the blocked Cholesky epilogue (row_solve.cuh) while the next rows accumulate.
//            k=64: a group drains 4 rows together, then each warp solves one of them;
//            k=128: the 4 warps of a group solve one row together.
#pragma once
#include "row_step_ffma.cuh"
#include "umma_ptx.cuh"

namespace b200als {

constexpr int UM_R = 32;          // ratings per stage
constexpr int UM_RD = 8;          // row-descriptor ring (t_sum hand-off)
constexpr int UM_GATHER_WARPS = 4;

template <int K>
struct UmmaCfg {
    static_assert(K == 64 || K == 128, "tcgen05 path covers k = 64 and k = 128");
    static constexpr int REC_BYTES = 4 * K;                 // [hi | lo] bf16
    static constexpr int N_MB = REC_BYTES / 128;            // 128-byte atoms per record
    static constexpr int STAGE_DATA = UM_R * REC_BYTES;
    static constexpr int STAGE_STRIDE = STAGE_DATA;         // data ring (1024-byte aligned atoms)
    static constexpr int B2_BYTES = 1024;                   // per stage: [q_hi,q_lo,1] operand, 32 ratings x 32 B
    static constexpr int META = 32;                         // (index, rating) ring: runs far ahead of the data ring
    static constexpr int META_BYTES = 256;                  // idx[32] int32 + val[32] float
    static constexpr int NG = 3;                            // solver groups (4 warps each)
    static constexpr int STAGES = 5;
    static constexpr int MATS = (K == 64) ? NG * 4 : NG;    // matrices being solved at once
    static constexpr int ACC_BUFS = (K == 64) ? 4 : 1;
    static constexpr int ACC_STRIDE = (K == 64) ? 128 : 512;  // TMEM columns per accumulator
    static constexpr int D2_COL = (K == 64) ? 64 : 256;
    static constexpr int MAIN_N = (K == 64) ? 64 : 256;
    static constexpr int SOLVER_WARPS = NG * 4;
    static constexpr int WARP_META = SOLVER_WARPS;
    static constexpr int WARP_GATHER0 = SOLVER_WARPS + 1;
    static constexpr int WARP_MMA = SOLVER_WARPS + 1 + UM_GATHER_WARPS;
    static constexpr int THREADS = (WARP_MMA + 1) * 32;
    static constexpr size_t MAT_FLOATS = solve_smem_floats(K) + 2 * K;
    static constexpr int NBAR = 2 * STAGES + 2 * META + MATS + ACC_BUFS + 2 * UM_RD;
    static constexpr size_t smem_bytes = 1024 /*align slack*/ + (size_t)STAGES * (STAGE_STRIDE + B2_BYTES + 128) +
                                         (size_t)META * META_BYTES +
                                         (size_t)MATS * MAT_FLOATS * 4 + (size_t)NBAR * 8 + UM_RD * 4 + 128;
    static_assert(smem_bytes <= 232448, "tcgen05 Gram kernel exceeds 227 KB of shared memory");
};

// Warp-level walker over this CTA's rows: row(n) = rows[b + n*G]; bounds prefetched 32 rows at a time.
struct RowWalker {
    const int64_t* indptr;
    const int32_t* rows;
    int64_t N, n, base;
    int G, b;
    int64_t beg_l, end_l;
    int32_t row_l;
    __device__ __forceinline__ void init(const RowStepParams& p, int lane) {
        indptr = p.indptr; rows = p.rows;
        G = gridDim.x; b = blockIdx.x;
        N = (p.n_list > b) ? (p.n_list - b + G - 1) / G : 0;
        n = 0; base = -64;
        beg_l = end_l = 0; row_l = 0;
        (void)lane;
    }
    __device__ __forceinline__ bool next(int lane, int64_t& beg, int64_t& end, int32_t& row) {
        if (n >= N) return false;
        if (n - base >= 32) {
            base = n;
            const int64_t seq = n + lane;
            if (seq < N) {
                const int64_t slot = (int64_t)b + seq * G;
                row_l = rows ? rows[slot] : (int32_t)slot;
                beg_l = indptr[row_l];
                end_l = indptr[row_l + 1];
            }
        }
        const int src = (int)(n - base);
        beg = __shfl_sync(0xffffffffu, beg_l, src);
        end = __shfl_sync(0xffffffffu, end_l, src);
        row = __shfl_sync(0xffffffffu, row_l, src);
        ++n;
        return true;
    }
};

// float32 → [hi | lo] bf16 records
__global__ void __launch_bounds__(256) split_bf16_kernel(const float* __restrict__ F, int64_t ldf, int64_t n_rows, int k,
                                                         __nv_bfloat16* __restrict__ H) {
    const int kq = k >> 2;
    const int64_t total = n_rows * kq;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
        const int64_t r = e / kq;
        const int c = (int)(e - r * kq) * 4;
        const float4 f = *reinterpret_cast<const float4*>(F + r * ldf + c);
        const __nv_bfloat16 h0 = __float2bfloat16_rn(f.x), h1 = __float2bfloat16_rn(f.y),
                            h2 = __float2bfloat16_rn(f.z), h3 = __float2bfloat16_rn(f.w);
        const __nv_bfloat16 l0 = __float2bfloat16_rn(f.x - __bfloat162float(h0)),
                            l1 = __float2bfloat16_rn(f.y - __bfloat162float(h1)),
                            l2 = __float2bfloat16_rn(f.z - __bfloat162float(h2)),
                            l3 = __float2bfloat16_rn(f.w - __bfloat162float(h3));
        __nv_bfloat16* hp = H + r * 2 * k + c;
        __nv_bfloat162* hv = reinterpret_cast<__nv_bfloat162*>(hp);
        hv[0] = __halves2bfloat162(h0, h1);
        hv[1] = __halves2bfloat162(h2, h3);
        __nv_bfloat162* lv = reinterpret_cast<__nv_bfloat162*>(hp + k);
        lv[0] = __halves2bfloat162(l0, l1);
        lv[1] = __halves2bfloat162(l2, l3);
    }
}

struct UmmaSmem {
    unsigned char* stages;   // [STAGES][STAGE_DATA]
    unsigned char* b2;       // [STAGES][1024]
    unsigned char* bias;     // [STAGES][128]
    unsigned char* meta;     // [META][256]
    float* mats;
    uint64_t *meta_full, *meta_empty, *full, *empty, *acc_full, *acc_empty, *rd_full, *rd_empty;
    float* tsum;
    uint32_t* tmem_base;
};

template <int K>
__device__ __forceinline__ void umma_carve(unsigned char* raw, UmmaSmem& s) {
    using C = UmmaCfg<K>;
    unsigned char* base = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
    s.stages = base;
    s.b2 = s.stages + (size_t)C::STAGES * C::STAGE_STRIDE;
    s.bias = s.b2 + (size_t)C::STAGES * C::B2_BYTES;
    s.meta = s.bias + (size_t)C::STAGES * 128;
    s.mats = reinterpret_cast<float*>(s.meta + (size_t)C::META * C::META_BYTES);
    uint64_t* bars = reinterpret_cast<uint64_t*>(s.mats + (size_t)C::MATS * C::MAT_FLOATS);
    s.meta_full = bars;
    s.meta_empty = s.meta_full + C::META;
    s.full = s.meta_empty + C::META;
    s.empty = s.full + C::STAGES;
    s.acc_full = s.empty + C::STAGES;
    s.acc_empty = s.acc_full + C::MATS;
    s.rd_full = s.acc_empty + C::ACC_BUFS;
    s.rd_empty = s.rd_full + UM_RD;
    s.tsum = reinterpret_cast<float*>(s.rd_empty + UM_RD);
    s.tmem_base = reinterpret_cast<uint32_t*>(s.tsum + UM_RD);
}

// acc_full is indexed by CONSUMER (solver group, matrix slot), not by TMEM buffer: every barrier then
// has exactly one waiting group that sees its phases strictly in order, so the parity wait is exact
// (indexing by buffer lets a group that is two uses behind mistake an old phase for the new one).
template <int K>
__device__ __forceinline__ void acc_full_slot(uint32_t me, uint32_t& bar_idx, uint32_t& parity) {
    using C = UmmaCfg<K>;
    if (K == 128) {
        bar_idx = me % C::NG;
        parity = (me / C::NG) & 1;
    } else {
        bar_idx = me % (4 * C::NG);           // group (me%12)/4, matrix (me%12)%4
        parity = (me / (4 * C::NG)) & 1;
    }
}

// ---- meta warp: (index, rating) streams of every stage → smem, fully asynchronous ----------
template <int K>
__device__ __forceinline__ void umma_role_meta(const RowStepParams& p, const UmmaSmem& s, int lane) {
    using C = UmmaCfg<K>;
    RowWalker w;
    w.init(p, lane);
    int64_t beg, end;
    int32_t row;
    uint32_t sg = 0;                                   // global stage counter of this CTA
    while (w.next(lane, beg, end, row)) {
        if (beg >= end) continue;
        for (int64_t t0 = beg; t0 < end; t0 += UM_R, ++sg) {
            const uint32_t ms = sg % C::META, mph = (sg / C::META) & 1;
            ptx::mbar_wait(&s.meta_empty[ms], mph ^ 1, 1, (int)sg);
            unsigned char* mt = s.meta + (size_t)ms * C::META_BYTES;
            const int64_t e = t0 + lane;
            const bool live = e < end;
            const int64_t es = live ? e : beg;         // any valid address for the zero-fill form
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(ptx::smem_u32(mt + lane * 4)),
                         "l"(p.indices + es), "r"(live ? 4u : 0u)
                         : "memory");
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(ptx::smem_u32(mt + 128 + lane * 4)),
                         "l"(p.vals + es), "r"(live ? 4u : 0u)
                         : "memory");
            ptx::cp_async_mbar_arrive_noinc(&s.meta_full[ms]);
        }
    }
    cp_async_wait<0>();
}

// ---- gather warps: 32 records (+ opposite biases) per stage into the swizzled MN-major tile ----
template <int K>
__device__ __forceinline__ void umma_role_gather(const RowStepParams& p, const UmmaSmem& s, const unsigned char* H,
                                                 int gtid, int lane) {
    using C = UmmaCfg<K>;
    RowWalker w;
    w.init(p, lane);
    int64_t beg, end;
    int32_t row;
    uint32_t sg = 0;
    const int rr = (gtid >> 3) & 7;        // rating inside its group of 8 == swizzle row
    const int ch = gtid & 7;               // 16-byte chunk inside the 128-byte line
    const int half = gtid >> 6;            // 0 / 1
    while (w.next(lane, beg, end, row)) {
        if (beg >= end) continue;
        for (int64_t t0 = beg; t0 < end; t0 += UM_R, ++sg) {
            const uint32_t slot = sg % C::STAGES, ph = (sg / C::STAGES) & 1;
            const uint32_t ms = sg % C::META, mph = (sg / C::META) & 1;
            unsigned char* st = s.stages + (size_t)slot * C::STAGE_STRIDE;
            ptx::mbar_wait(&s.meta_full[ms], mph, 2, (int)sg);
            ptx::mbar_wait(&s.empty[slot], ph ^ 1, 11, (int)sg);
            const int32_t* idx_s = reinterpret_cast<const int32_t*>(s.meta + (size_t)ms * C::META_BYTES);
            const int live_n = (int)((end - t0) < UM_R ? (end - t0) : UM_R);
            const uint32_t st_addr = ptx::smem_u32(st);
#pragma unroll
            for (int j = 0; j < C::N_MB * 2; ++j) {
                // chunk id c = gtid + 128 j  →  (kb, mb) atom, line rr, chunk ch
                const int atom = half + 2 * j;             // = c / 64
                const int kb = atom / C::N_MB, mb = atom % C::N_MB;
                const int rating = kb * 8 + rr;
                const bool live = rating < live_n;
                const int32_t ix = idx_s[rating];
                const unsigned char* src = H + (size_t)ix * C::REC_BYTES + mb * 128 + ch * 16;
                const uint32_t dst = st_addr + (uint32_t)((kb * C::N_MB + mb) * 1024 + rr * 128 + ((ch ^ rr) << 4));
                ptx::cp_async16_zfill(dst, src, live ? 16u : 0u);
            }
            if (gtid < UM_R) {
                const bool live = gtid < live_n;
                const int32_t ix = idx_s[gtid];
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(ptx::smem_u32(s.bias + slot * 128 + gtid * 4)),
                             "l"(p.bias_other + ix), "r"(live ? 4u : 0u)
                             : "memory");
            }
            ptx::cp_async_mbar_arrive_noinc(&s.full[slot]);
        }
    }
    cp_async_wait<0>();
}

// ---- mma warp ------------------------------------------------------------------------------
template <int K>
__device__ __forceinline__ void umma_role_mma(const RowStepParams& p, const UmmaSmem& s, int lane) {
    using C = UmmaCfg<K>;
    RowWalker w;
    w.init(p, lane);
    int64_t beg, end;
    int32_t row;
    uint32_t sg = 0;
    uint32_t nrow = 0;
    const float mu = *p.mu;
    const uint32_t tmem = *s.tmem_base;
    constexpr uint32_t IDESC_MAIN = ptx::idesc_bf16_mn(128, C::MAIN_N);
    constexpr uint32_t IDESC_D2 = ptx::idesc_bf16_mn(128, 16);
    constexpr uint32_t SBO = C::N_MB * 1024;           // K-atom (8 ratings) stride
    constexpr uint32_t LBO = 1024;                     // MN-atom (64 factors) stride
    while (w.next(lane, beg, end, row)) {
        if (beg >= end) continue;
        const uint32_t buf = nrow % C::ACC_BUFS, aph = (nrow / C::ACC_BUFS) & 1;
        ptx::mbar_wait(&s.acc_empty[buf], aph ^ 1, 3, (int)nrow);     // accumulator drained by its solver group
        ptx::tc_fence_after();
        const uint32_t d_main = tmem + buf * C::ACC_STRIDE;
        const uint32_t d_two = d_main + C::D2_COL;
        float tsum = 0.0f;
        bool first = true;
        for (int64_t t0 = beg; t0 < end; t0 += UM_R, ++sg) {
            const uint32_t slot = sg % C::STAGES, ph = (sg / C::STAGES) & 1;
            unsigned char* st = s.stages + (size_t)slot * C::STAGE_STRIDE;
            const uint32_t ms = sg % C::META, mph = (sg / C::META) & 1;
            ptx::mbar_wait(&s.meta_full[ms], mph, 4, (int)sg);
            ptx::mbar_wait(&s.full[slot], ph, 5, (int)sg);
            // the [q_hi, q_lo, 1] operand of this stage (no-swizzle MN-major: 8 ratings x 16 B core matrices)
            const bool live = (t0 + lane) < end;
            const float v = reinterpret_cast<const float*>(s.meta + (size_t)ms * C::META_BYTES + 128)[lane];
            const float bo = reinterpret_cast<const float*>(s.bias + slot * 128)[lane];
            const float q = live ? v - (mu + bo) : 0.0f;
            tsum += q;
            const __nv_bfloat16 qh = __float2bfloat16_rn(q);
            const __nv_bfloat16 ql = __float2bfloat16_rn(q - __bfloat162float(qh));
            const __nv_bfloat162 w0 = __halves2bfloat162(qh, ql);
            const __nv_bfloat162 w1 = __halves2bfloat162(__float2bfloat16_rn(live ? 1.0f : 0.0f), __float2bfloat16_rn(0.0f));
            uint4 c0;
            c0.x = *reinterpret_cast<const uint32_t*>(&w0);
            c0.y = *reinterpret_cast<const uint32_t*>(&w1);
            c0.z = 0u; c0.w = 0u;
            unsigned char* b2 = s.b2 + (size_t)slot * C::B2_BYTES + (lane >> 3) * 256 + (lane & 7) * 16;
            *reinterpret_cast<uint4*>(b2) = c0;                          // columns 0..7
            *reinterpret_cast<uint4*>(b2 + 128) = make_uint4(0u, 0u, 0u, 0u);   // columns 8..15
            ptx::fence_proxy_async();
            __syncwarp();
            ptx::tc_fence_after();
            if (lane == 0) {
                ptx::mbar_arrive(&s.meta_empty[ms]);                     // idx (gather) and val (this warp) consumed
                const uint32_t sa = ptx::smem_u32(st);
                const uint32_t sb2 = ptx::smem_u32(s.b2 + (size_t)slot * C::B2_BYTES);
#pragma unroll
                for (int ks = 0; ks < UM_R / 16; ++ks) {
                    const uint32_t a_addr = sa + ks * 2 * SBO;
                    const uint64_t a_desc = ptx::smem_desc(a_addr, LBO, SBO, ptx::SWZ_128B);
                    const uint64_t b2_desc = ptx::smem_desc(sb2 + ks * 512, 256, 128, ptx::SWZ_NONE);
                    const uint32_t acc = (first && ks == 0) ? 0u : 1u;
                    // main Gram MMA: B starts at atom 0 of the same tile (k=64: hi only; k=128: hi|lo)
                    ptx::umma_bf16(d_main, a_desc, a_desc, IDESC_MAIN, acc);
                    ptx::umma_bf16(d_two, a_desc, b2_desc, IDESC_D2, acc);
                    if (K == 128) {
                        const uint64_t alo_desc = ptx::smem_desc(a_addr + 2 * 1024, LBO, SBO, ptx::SWZ_128B);
                        ptx::umma_bf16(d_two, alo_desc, b2_desc, IDESC_D2, 1u);
                    }
                }
                ptx::umma_commit(&s.empty[slot]);                        // stage reusable once these MMAs finish
                if (t0 + UM_R >= end) {
                    uint32_t fb, fp;
                    acc_full_slot<K>(nrow, fb, fp);
                    ptx::umma_commit(&s.acc_full[fb]);
                }
            }
            __syncwarp();
            first = false;
        }
        // t_sum = sum q of the row → solver (explicit release/acquire pair, not the async commit)
        tsum = warp_sum(tsum);
        const uint32_t rd = nrow % UM_RD, rph = (nrow / UM_RD) & 1;
        ptx::mbar_wait(&s.rd_empty[rd], rph ^ 1, 6, (int)nrow);
        if (lane == 0) {
            s.tsum[rd] = tsum;
            ptx::mbar_arrive(&s.rd_full[rd]);
        }
        ++nrow;
    }
}

// ---- solver groups ----------------------------------------------------------------------------

// k = 64: accumulator rows 0..63 = P = hi^T hi, rows 64..127 = Q^T = lo^T hi.  A = P + Q^T + Q.
// Called by the 4 warps of a group (wq = warp in group = TMEM lane quadrant).
__device__ __forceinline__ void umma_drain64(uint32_t taddr_buf, float* Gt, float* c_s, float* s_s, int wq, int lane,
                                             int bar) {
    const uint32_t trow = taddr_buf + ((uint32_t)(32 * wq) << 16);
    const int i = (32 * wq + lane) & 63;
    const bool isQ = wq >= 2;
    const int imax = ((32 * wq) & 63) + 31;            // warp-uniform bounds for the collective tcgen05.ld
    const int imin = (32 * wq) & 63;
    float d2[4];
    ptx::tmem_ld4(trow + 64, d2);
    const float cpart = d2[0] + d2[1], spart = d2[2];
    float v[16];
    if (!isQ) {
        c_s[i] = cpart; s_s[i] = spart;
        for (int jc = 0; jc * 16 <= imax; ++jc) {
            ptx::tmem_ld16(trow + jc * 16, v);
#pragma unroll
            for (int jj = 0; jj < 16; ++jj) {
                const int j = jc * 16 + jj;
                if (j <= i) *tile_elem(Gt, i, j) = v[jj];
            }
        }
    }
    bar_sync(bar, 128);
    if (isQ) {
        c_s[i] += cpart; s_s[i] += spart;
        for (int jc = 0; jc * 16 <= imax; ++jc) {      // row part: A[i][j] += Q^T[i][j], j <= i
            ptx::tmem_ld16(trow + jc * 16, v);
#pragma unroll
            for (int jj = 0; jj < 16; ++jj) {
                const int j = jc * 16 + jj;
                if (j <= i) *tile_elem(Gt, i, j) += v[jj];
            }
        }
    }
    bar_sync(bar, 128);
    if (isQ) {
        for (int jc = imin / 16; jc < 4; ++jc) {       // column part: A[j][i] += Q^T[i][j], j >= i
            ptx::tmem_ld16(trow + jc * 16, v);
#pragma unroll
            for (int jj = 0; jj < 16; ++jj) {
                const int j = jc * 16 + jj;
                if (j >= i) *tile_elem(Gt, j, i) += v[jj];
            }
        }
    }
}

// k = 128: accumulator row i = [P[i][0..127] | Q[i][0..127]].  A = P + Q + Q^T.
__device__ __forceinline__ void umma_drain128(uint32_t taddr_buf, float* Gt, float* c_s, float* s_s, int wq, int lane,
                                              int bar) {
    const uint32_t trow = taddr_buf + ((uint32_t)(32 * wq) << 16);
    const int i = 32 * wq + lane;
    const int imax = 32 * wq + 31, imin = 32 * wq;
    float d2[4];
    ptx::tmem_ld4(trow + 256, d2);
    c_s[i] = d2[0] + d2[1];
    s_s[i] = d2[2];
    float v[16], u[16];
    for (int jc = 0; jc * 16 <= imax; ++jc) {
        ptx::tmem_ld16(trow + jc * 16, v);
        ptx::tmem_ld16(trow + 128 + jc * 16, u);
#pragma unroll
        for (int jj = 0; jj < 16; ++jj) {
            const int j = jc * 16 + jj;
            if (j < i) *tile_elem(Gt, i, j) = v[jj] + u[jj];
            else if (j == i) *tile_elem(Gt, i, j) = v[jj] + 2.0f * u[jj];
        }
    }
    bar_sync(bar, 128);
    for (int jc = imin / 16; jc < 8; ++jc) {
        ptx::tmem_ld16(trow + 128 + jc * 16, u);
#pragma unroll
        for (int jj = 0; jj < 16; ++jj) {
            const int j = jc * 16 + jj;
            if (j > i) *tile_elem(Gt, j, i) += u[jj];
        }
    }
}

template <int K>
__device__ __forceinline__ void umma_role_solver(const RowStepParams& p, const UmmaSmem& s, int warp, int lane) {
    using C = UmmaCfg<K>;
    const int g = warp >> 2, wq = warp & 3;            // group, warp in group (== TMEM lane quadrant)
    const int gtid = wq * 32 + lane;
    const int gbar = 1 + g;                            // named barrier of the group
    const uint32_t tmem = *s.tmem_base;
    RowWalker w;
    w.init(p, lane);
    int64_t beg, end;
    int32_t row;
    uint32_t nrow = 0;
    RowEpilogueArgs ea;
    ea.k = K; ea.mode = p.mode;
    ea.X = p.X; ea.ldx = p.ldx; ea.bias_own = p.bias_own;
    ea.chol_out = p.chol_out; ea.aux_out = p.aux_out; ea.status = p.status;

    if (K == 128) {
        // rows nrow with nrow % NG == g; the 4 warps drain and solve together
        float* mat = s.mats + (size_t)g * C::MAT_FLOATS;
        SolveSmem S;
        S.carve<K>(mat);
        float* c_s = mat + solve_smem_floats(K);
        float* s_s = c_s + K;
        while (w.next(lane, beg, end, row)) {
            if (beg >= end) continue;
            const uint32_t me = nrow++;
            if ((int)(me % C::NG) != g) continue;
            const uint32_t buf = me % C::ACC_BUFS;
            {
                uint32_t fb, fp;
                acc_full_slot<K>(me, fb, fp);
                ptx::mbar_wait(&s.acc_full[fb], fp, 7, (int)me);
            }
            ptx::tc_fence_after();
            umma_drain128(tmem + buf * C::ACC_STRIDE, S.Gt, c_s, s_s, wq, lane, gbar);
            ptx::tc_fence_before();
            ptx::mbar_arrive(&s.acc_empty[buf]);
            const uint32_t rd = me % UM_RD, rph = (me / UM_RD) & 1;
            ptx::mbar_wait(&s.rd_full[rd], rph, 8, (int)me);
            const float t_sum = s.tsum[rd];
            bar_sync(gbar, 128);
            if (gtid == 0) ptx::mbar_arrive(&s.rd_empty[rd]);
            ea.reg = p.reg_row ? p.reg_row[row] : p.reg;
            ea.lam_bias = p.lam_bias_row ? p.lam_bias_row[row] : p.lam_bias;
            row_epilogue_tiles(S, c_s, s_s, t_sum, end - beg, row, ea, gtid, 128, gbar);
            bar_sync(gbar, 128);
        }
    } else {
        // batches of 4 consecutive rows per group: drained together, then one warp solves each
        float* mat = s.mats + (size_t)(g * 4 + wq) * C::MAT_FLOATS;     // this warp's own matrix
        SolveSmem S;
        S.carve<K>(mat);
        float* c_s = mat + solve_smem_floats(K);
        float* s_s = c_s + K;
        bool more = true;
        while (more) {
            int64_t my_beg = 0, my_end = 0;
            int32_t my_row = -1;
            uint32_t my_seq = 0;
            for (int slot4 = 0; slot4 < 4 * C::NG; ++slot4) {
                // every group walks all rows; a batch is 4*NG consecutive rows, group g owns rows [4g, 4g+4)
                do { more = w.next(lane, beg, end, row); } while (more && beg >= end);
                if (!more) break;
                const uint32_t me = nrow++;
                if (slot4 / 4 != g) continue;
                const int j = slot4 & 3;                                 // matrix of warp j of this group
                float* matj = s.mats + (size_t)(g * 4 + j) * C::MAT_FLOATS;
                const uint32_t buf = me % C::ACC_BUFS;
                {
                    uint32_t fb, fp;
                    acc_full_slot<K>(me, fb, fp);
                    ptx::mbar_wait(&s.acc_full[fb], fp, 9, (int)me);
                }
                ptx::tc_fence_after();
                umma_drain64(tmem + buf * C::ACC_STRIDE, matj, matj + solve_smem_floats(K),
                             matj + solve_smem_floats(K) + K, wq, lane, gbar);
                ptx::tc_fence_before();
                ptx::mbar_arrive(&s.acc_empty[buf]);
                if (j == wq) { my_beg = beg; my_end = end; my_row = row; my_seq = me; }
            }
            bar_sync(gbar, 128);                                         // all drains of the batch visible
            if (my_row >= 0) {
                const uint32_t rd = my_seq % UM_RD, rph = (my_seq / UM_RD) & 1;
                ptx::mbar_wait(&s.rd_full[rd], rph, 10, (int)my_seq);
                const float t_sum = s.tsum[rd];
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&s.rd_empty[rd]);
                ea.reg = p.reg_row ? p.reg_row[my_row] : p.reg;
                ea.lam_bias = p.lam_bias_row ? p.lam_bias_row[my_row] : p.lam_bias;
                row_epilogue_tiles(S, c_s, s_s, t_sum, my_end - my_beg, my_row, ea, lane, 32, 0);
            }
            bar_sync(gbar, 128);                                         // matrices free for the next batch
        }
    }
}

template <int K>
__global__ void __launch_bounds__(UmmaCfg<K>::THREADS, 1) row_step_umma_kernel(const RowStepParams p,
                                                                               const unsigned char* __restrict__ H) {
    using C = UmmaCfg<K>;
    extern __shared__ unsigned char smem_raw[];
    UmmaSmem s;
    umma_carve<K>(smem_raw, s);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int i = 0; i < C::META; ++i) {
            ptx::mbar_init(&s.meta_full[i], 32);
            ptx::mbar_init(&s.meta_empty[i], 1);
        }
        for (int i = 0; i < C::STAGES; ++i) {
            ptx::mbar_init(&s.full[i], UM_GATHER_WARPS * 32);
            ptx::mbar_init(&s.empty[i], 1);
        }
        for (int i = 0; i < C::MATS; ++i) ptx::mbar_init(&s.acc_full[i], 1);
        for (int i = 0; i < C::ACC_BUFS; ++i) ptx::mbar_init(&s.acc_empty[i], 128);
        for (int i = 0; i < UM_RD; ++i) {
            ptx::mbar_init(&s.rd_full[i], 1);
            ptx::mbar_init(&s.rd_empty[i], 1);
        }
        ptx::fence_barrier_init();
    }
    if (warp == C::WARP_MMA) ptx::tmem_alloc(s.tmem_base, 512);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();

    if (warp < C::SOLVER_WARPS) {
        umma_role_solver<K>(p, s, warp, lane);
    } else if (warp == C::WARP_META) {
        umma_role_meta<K>(p, s, lane);
    } else if (warp < C::WARP_MMA) {
        umma_role_gather<K>(p, s, H, (warp - C::WARP_GATHER0) * 32 + lane, lane);
    } else {
        umma_role_mma<K>(p, s, lane);
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == C::WARP_MMA) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(*s.tmem_base, 512);
    }
}

static inline bool umma_supported(int k, int64_t ldf) { return (k == 64 || k == 128) && ldf == k; }

static inline size_t umma_workspace_bytes(int k, int64_t n_f_rows) {
    if (k != 64 && k != 128) return 256;
    return 256 + align_up((size_t)(n_f_rows > 0 ? n_f_rows : 0) * 4 * k, 256);
}

template <int K>
static int launch_umma_k(const RowStepParams& p, int64_t n_f_rows, void* ws, size_t ws_bytes, cudaStream_t st) {
    using C = UmmaCfg<K>;
    if (ws_bytes < (size_t)n_f_rows * C::REC_BYTES) {
        set_error("row_step (tcgen05): workspace too small for the bf16 split of %lld rows", (long long)n_f_rows);
        return B200ALS_E_WORKSPACE;
    }
    __nv_bfloat16* H = static_cast<__nv_bfloat16*>(ws);
    {
        int64_t work = n_f_rows * (K / 4);
        int64_t blocks = ceil_div64(work, 256);
        if (blocks > (int64_t)kNumSMs * 16) blocks = (int64_t)kNumSMs * 16;
        split_bf16_kernel<<<(unsigned)blocks, 256, 0, st>>>(p.F, p.ldf, n_f_rows, K, H);
        ALS_LAUNCH_CHECK();
    }
    auto kern = row_step_umma_kernel<K>;
    static thread_local bool configured = false;
    if (!configured) {
        ALS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::smem_bytes));
        configured = true;
    }
    int64_t grid = p.n_list < kNumSMs ? p.n_list : kNumSMs;
    kern<<<(unsigned)grid, C::THREADS, C::smem_bytes, st>>>(p, reinterpret_cast<const unsigned char*>(H));
    ALS_LAUNCH_CHECK();
    return B200ALS_OK;
}

static inline int launch_umma(const RowStepParams& p, int64_t n_f_rows, void* ws, size_t ws_bytes, cudaStream_t s) {
    if (p.k == 64) return launch_umma_k<64>(p, n_f_rows, ws, ws_bytes, s);
    return launch_umma_k<128>(p, n_f_rows, ws, ws_bytes, s);
}

}  // namespace b200als

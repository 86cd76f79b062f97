// row_step_umma.cuh — tcgen05 / TMEM Gram path for k in {64, 128} (A6/A7, als.py:414-466).
//
// One persistent CTA per SM, warp-specialised; rows are claimed longest-first from a global
// atomic cursor (guided chunks), so the power-law tail cannot unbalance the SMs.
//
// Gathered operand: the factor matrix pre-split into bf16 records H[row] = [hi(k) | lo(k)],
// 4k bytes — the same bytes per rating as the float32 row.  hi + lo carries ~17 mantissa bits:
//     F^T F ~= hi^T hi + hi^T lo + lo^T hi                       (lo^T lo ~ 2^-18 dropped)
// All three products are accumulated by the tensor core INTO THE SAME k x k accumulator
// (three tcgen05.mma per 16 ratings, M = N = k), so TMEM lane i ends up holding the whole
// symmetric row i of the normal matrix and the solver threads read it straight into
// registers (row_ldl.cuh) — the Gram never touches shared memory.  A 16-column MMA against
// [q_hi, q_lo, 1] per rating gives c = F^T q and s = F^T 1 next to it.
// Ratings are the MMA K dimension, so both operands are MN-major: a gathered record is one
// 128-byte line per 64 factors, written straight into the 128B-swizzled canonical layout by
// 16-byte cp.async (no register pass, no transposition).
//
// k = 64 uses M = 64 MMAs: the accumulator occupies 16 lanes of each TMEM lane quadrant, and
// consecutive rows alternate between lanes 0-15 and 16-31 ("interleaved" accumulators), so a
// solver group (4 warps = 4 quadrants) factors two rows' matrices in lock step.
//
// Warp roles:  [solver groups: 4 warps each] [meta] [gather x4] [prep] [mma]
//   meta   : claims rows, streams (index, rating) + a stage header of each 32-rating stage
//   gather : per stage, 128 threads cp.async the 32 records (and the 32 opposite biases)
//   prep   : q = rating - mu - bias, writes the [q_hi,q_lo,1] atom of the stage, hands (row, cnt,
//            sum q, sum q^2) to the solver through a descriptor ring
//   mma    : issues tcgen05.mma (3 per 16 ratings), commits stage / accumulator
//   solver : tcgen05.ld the accumulator rows, LDL^T factor + solve in registers, write x, bias,
//            and the fused residual statistics (A10) of the row.
#pragma once
#include "row_ldl.cuh"
#include "row_step_ffma.cuh"
#include "split_plan.cuh"
#include "umma_ptx.cuh"

namespace b200als {

constexpr int UM_RD = 16;         // row-descriptor ring

constexpr int HDR_FIRST = 1, HDR_LAST = 2, HDR_END = 4;

struct StageHdr {                 // 16 bytes at the end of a meta slot
    int32_t live_n;
    int32_t row;
    int32_t flags;
    int32_t cnt;
};
struct RowDesc {                  // 32 bytes: one 16-byte half per prep warp
    int32_t row;                  // -1: no row (odd tail / end of work), <= -2: piece of a split row
    int32_t cnt;
    float tsum;                   // sum q   over the stages of prep warp 0
    float qq;                     // sum q^2
    float tsum1, qq1;             // the same over the stages of prep warp 1
    int32_t pad[2];
};

template <int K>
struct UmmaCfg {
    static_assert(K == 64 || K == 128, "tcgen05 path covers k = 64 and k = 128");
    static constexpr int REC_BYTES = 4 * K;                 // [hi | lo] bf16
    static constexpr int N_MB = REC_BYTES / 128;            // 128-byte atoms per record
    static constexpr int HI_ATOMS = N_MB / 2;
    static constexpr int NA = N_MB + 1;                     // atoms per 8-rating K-atom: hi.. | q | lo..
#ifndef B200ALS_R64
#define B200ALS_R64 64
#endif
    // ratings per stage: every producer warp pays a fixed latency chain per stage (barrier waits, header,
    // commit), so k = 64 — whose stages are half the bytes — takes 64 ratings per stage
    static constexpr int R = (K == 64) ? B200ALS_R64 : 32;
    static constexpr int RL = R / 32;                       // ratings per lane of a single-warp role
    static constexpr int STAGE_DATA = (R / 8) * NA * 1024;
    static constexpr int STAGE_STRIDE = STAGE_DATA;         // data ring (1024-byte aligned atoms)
#ifndef B200ALS_NG128
#define B200ALS_NG128 3
#endif
    static constexpr int META = (K == 64) ? 32 : (B200ALS_NG128 > 3 ? 16 : 32);   // (index, rating) ring: runs far ahead of the data ring
    static constexpr int META_BYTES = 8 * R + 32;           // idx[R] int32 | val[R] float | StageHdr | pad
    static constexpr int BIAS_BYTES = 4 * R;
    static constexpr int STAGES = (K == 64) ? (R == 64 ? 4 : 6) : (B200ALS_NG128 > 3 ? 3 : 4);
#ifndef B200ALS_NG64
#define B200ALS_NG64 4
#endif
    static constexpr int NG = (K == 64) ? B200ALS_NG64 : B200ALS_NG128; // solver groups (4 warps each); 3 named barriers per group
    static constexpr int RPB = (K == 64) ? 2 : 1;           // rows per TMEM accumulator buffer
    static constexpr int RPW = (K == 64) ? 16 : 32;         // matrix rows per solver warp
    static constexpr int NBUF = (K == 64) ? 6 : 3;          // TMEM accumulator buffers
    static constexpr int ACC_STRIDE = K + 16;               // TMEM columns per buffer: main | [c_hi,c_lo,s,..]
    static constexpr int D2_COL = K;
    // 'accumulator ready' barriers, indexed pair % ACCF.  Per buffer when buffers outnumber groups (the MMA warp
    // may finish two pairs of one group before it looks), per group when groups outnumber buffers (a group
    // may wait for a buffer's next use before the previous one completed) — either way a waiter never
    // sees the phase parity alias.
    static constexpr int ACCF = NG > NBUF ? NG : NBUF;
    static constexpr int SOLVER_WARPS = NG * 4;
    static constexpr int WARP_META = SOLVER_WARPS;
    static constexpr int WARP_GATHER0 = SOLVER_WARPS + 1;
    // gather warps stride the stage sequence; more warps than ring slots would let a warp wait on a slot two
    // uses ahead, where the mbarrier phase parity aliases
    static constexpr int GATHER_WARPS = STAGES < 4 ? STAGES : 4;
    static constexpr int WARP_PREP = SOLVER_WARPS + 1 + GATHER_WARPS;
    static constexpr int PREP_WARPS = 2;                     // stage sg is prepared by warp sg % 2
    static_assert(STAGES % PREP_WARPS == 0, "a ring slot must always be prepared by the same warp (barrier phase parity)");
    static constexpr int WARP_MMA = WARP_PREP + PREP_WARPS;
    static constexpr int THREADS = (SOLVER_WARPS + 8) * 32;  // whole warpgroups: solvers | 7 producer warps + 1 idle
    // setmaxnreg split of the CTA's register allocation (launch: 65536 / THREADS rounded down to 8)
    static constexpr int REG_PRODUCER = 48;
    static constexpr int REG_LAUNCH = (65536 / THREADS) & ~7;
    static constexpr int REG_SOLVER_RAW = ((REG_LAUNCH * THREADS - 256 * REG_PRODUCER) / (SOLVER_WARPS * 32)) & ~7;
    static constexpr int REG_SOLVER = REG_SOLVER_RAW > 232 ? 232 : REG_SOLVER_RAW;
    static_assert(3 * NG + 1 <= 16, "named barriers");
    static constexpr int MATS = NG * RPB;
    static constexpr int MAT_FLOATS = ldl_mat_floats<K>();
    static constexpr int RED_FLOATS = NG * 4 * RPB * 4;     // per group: [warp][matrix][4]
    static constexpr int NBAR = (2 * META + 3 * STAGES + NBUF + ACCF + 2 * UM_RD + MATS + 1) & ~1;   // even: RowDesc ring stays 16-byte aligned
    static constexpr size_t smem_bytes = 1024 /*align slack*/ + (size_t)STAGES * (STAGE_STRIDE + BIAS_BYTES + 16) +
                                         (size_t)META * META_BYTES + (size_t)MATS * MAT_FLOATS * 4 +
                                         (size_t)RED_FLOATS * 4 + (size_t)NBAR * 8 + (size_t)UM_RD * 32 + 128 + 128;
    static_assert(smem_bytes <= 232448, "tcgen05 Gram kernel exceeds 227 KB of shared memory");
    static_assert(NBUF * ACC_STRIDE <= 512, "TMEM columns");
};

// float32 -> [hi | lo] bf16 records
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
    unsigned char* bias;     // [STAGES][128]
    int32_t* sflags;         // [STAGES] stage flags, prep -> mma
    unsigned char* meta;     // [META][META_BYTES]
    float* mats;             // [MATS][MAT_FLOATS]
    float* red;              // [NG][4][RPB][4]
    uint64_t *meta_full, *meta_empty, *full, *prep_full, *empty, *acc_full, *acc_empty, *rd_full, *rd_empty, *upd;
    RowDesc* rd;             // [UM_RD]
    int32_t* bad;            // [16] non-SPD flags, one per matrix slot | [16] split-row 'last piece' flags
    uint32_t* tmem_base;
};

template <int K>
__device__ __forceinline__ void umma_carve(unsigned char* raw, UmmaSmem& s) {
    using C = UmmaCfg<K>;
    unsigned char* base = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
    s.stages = base;
    s.bias = s.stages + (size_t)C::STAGES * C::STAGE_STRIDE;
    s.sflags = reinterpret_cast<int32_t*>(s.bias + (size_t)C::STAGES * C::BIAS_BYTES);
    s.meta = s.bias + (size_t)C::STAGES * (C::BIAS_BYTES + 16);
    s.mats = reinterpret_cast<float*>(s.meta + (size_t)C::META * C::META_BYTES);
    s.red = s.mats + (size_t)C::MATS * C::MAT_FLOATS;
    uint64_t* bars = reinterpret_cast<uint64_t*>(s.red + C::RED_FLOATS);
    s.meta_full = bars;
    s.meta_empty = s.meta_full + C::META;
    s.full = s.meta_empty + C::META;
    s.prep_full = s.full + C::STAGES;
    s.empty = s.prep_full + C::STAGES;
    s.acc_full = s.empty + C::STAGES;
    s.acc_empty = s.acc_full + C::ACCF;
    s.rd_full = s.acc_empty + C::NBUF;
    s.rd_empty = s.rd_full + UM_RD;
    s.upd = s.rd_empty + UM_RD;                        // [MATS] tensor-core trailing updates of the solver
    s.rd = reinterpret_cast<RowDesc*>(bars + C::NBAR);
    s.bad = reinterpret_cast<int32_t*>(s.rd + UM_RD);
    s.tmem_base = reinterpret_cast<uint32_t*>(s.bad + 32);
}

// shared-space loads (explicit .shared addressing: generic LD on the carved buffer takes the long-scoreboard path)
__device__ __forceinline__ int4 lds_s32x4(uint32_t a) {
    int4 v;
    asm volatile("ld.shared.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ int32_t lds_s32(uint32_t a) {
    int32_t v;
    asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ float lds_f32v(uint32_t a) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts_s32(uint32_t a, int32_t v) {
    asm volatile("st.shared.s32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}
__device__ __forceinline__ void sts_s32x4(uint32_t a, int4 v) {
    asm volatile("st.shared.v4.s32 [%0], {%1,%2,%3,%4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
template <int K>
__device__ __forceinline__ StageHdr stage_hdr(const UmmaSmem& s, uint32_t ms) {
    const int4 v = lds_s32x4(ptx::smem_u32(s.meta + (size_t)ms * UmmaCfg<K>::META_BYTES + 8 * UmmaCfg<K>::R));
    StageHdr h;
    h.live_n = v.x; h.row = v.y; h.flags = v.z; h.cnt = v.w;
    return h;
}

// ---- meta warp: claims rows, streams (index, rating) + header of every stage --------------
template <int K>
__device__ __forceinline__ void umma_role_meta(const RowStepParams& p, const UmmaSmem& s, int lane) {
    using C = UmmaCfg<K>;
    uint32_t sg = 0;                                   // global stage counter of this CTA
    // work list: the items of the split plan (rows cut into pieces) or the plain row list
    const long long n_list = p.items ? (long long)p.split_hdr[0] : (long long)p.n_list;
    const long long G = gridDim.x;
    long long seen = 0;

    // one batch of claimed rows lives in registers: lane l holds entry l
    // (row_l: the row, or -2 - item for a piece of a split row — the solver looks the item up)
    int nb = 0;
    int32_t row_l = 0;
    int64_t beg_l = 0, end_l = 0;
    auto claim = [&](int& nb_o, int32_t& row_o, int64_t& beg_o, int64_t& end_o) {
        // guided chunks: single rows while the list is in its long-row head, up to 32 later
        long long want = (seen < 16 * G) ? 1 : (n_list - seen) / (G * 8);
        want = want < 1 ? 1 : (want > 32 ? 32 : want);
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(p.cursor, (unsigned long long)want);
        base = __shfl_sync(0xffffffffu, base, 0);
        seen = (long long)base + want;
        long long avail = n_list - (long long)base;
        nb_o = (int)(avail < 0 ? 0 : (avail < want ? avail : want));
        row_o = 0; beg_o = 0; end_o = 0;
        if (lane < nb_o) {
            const long long slot = (long long)base + lane;
            if (p.items) {
                const int4 it = p.items[slot];
                beg_o = p.indptr[it.x] + it.y;
                end_o = beg_o + it.z;
                row_o = it.w < 0 ? it.x : (int32_t)(-2 - slot);
            } else {
                row_o = p.rows ? p.rows[slot] : (int32_t)slot;
                beg_o = p.indptr[row_o];
                end_o = p.indptr[row_o + 1];
            }
        }
    };
    claim(nb, row_l, beg_l, end_l);
    while (nb > 0) {
        // prefetch the next batch while this one streams
        int nb_n; int32_t row_n; int64_t beg_n, end_n;
        claim(nb_n, row_n, beg_n, end_n);
        for (int e = 0; e < nb; ++e) {
            const int64_t beg = __shfl_sync(0xffffffffu, beg_l, e);
            const int64_t end = __shfl_sync(0xffffffffu, end_l, e);
            const int32_t row = __shfl_sync(0xffffffffu, row_l, e);
            for (int64_t t0 = beg; t0 < end; t0 += C::R, ++sg) {
                const uint32_t ms = sg % C::META, mph = (sg / C::META) & 1;
                ptx::mbar_wait(&s.meta_empty[ms], mph ^ 1, 1, (int)sg);
                unsigned char* mt = s.meta + (size_t)ms * C::META_BYTES;
#pragma unroll
                for (int j = 0; j < C::RL; ++j) {
                    const int64_t ee = t0 + lane + 32 * j;
                    const bool live = ee < end;
                    const int64_t es = live ? ee : beg;    // any valid address for the zero-fill form
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(ptx::smem_u32(mt + (lane + 32 * j) * 4)),
                                 "l"(p.indices + es), "r"(live ? 4u : 0u)
                                 : "memory");
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(ptx::smem_u32(mt + 4 * C::R + (lane + 32 * j) * 4)),
                                 "l"(p.vals + es), "r"(live ? 4u : 0u)
                                 : "memory");
                }
                ptx::cp_async_mbar_arrive_noinc(&s.meta_full[ms]);
                if (lane == 0) {
                    StageHdr h;
                    h.live_n = (int32_t)((end - t0) < C::R ? (end - t0) : C::R);
                    h.row = row;
                    h.flags = (t0 == beg ? HDR_FIRST : 0) | (t0 + C::R >= end ? HDR_LAST : 0);
                    h.cnt = (int32_t)(end - beg);
                    sts_s32x4(ptx::smem_u32(mt + 8 * C::R), make_int4(h.live_n, h.row, h.flags, h.cnt));
                    ptx::mbar_arrive(&s.meta_full[ms]);    // release: header visible to the waiters
                }
            }
        }
        nb = nb_n; row_l = row_n; beg_l = beg_n; end_l = end_n;
    }
    // end-of-work stages: one per gather warp (each walks its own stride of the stage sequence);
    // the prep and mma warps stop at the first
    for (int e = 0; e < C::GATHER_WARPS; ++e, ++sg) {
        const uint32_t ms = sg % C::META, mph = (sg / C::META) & 1;
        ptx::mbar_wait(&s.meta_empty[ms], mph ^ 1, 1, (int)sg);
        unsigned char* mt = s.meta + (size_t)ms * C::META_BYTES;
        ptx::cp_async_mbar_arrive_noinc(&s.meta_full[ms]);
        if (lane == 0) {
            sts_s32x4(ptx::smem_u32(mt + 8 * C::R), make_int4(0, -1, HDR_END, 0));
            ptx::mbar_arrive(&s.meta_full[ms]);
        }
    }
    cp_async_wait<0>();
}

// ---- gather warps: 32 records (+ opposite biases) per stage into the swizzled MN-major tile ----
// Warp w takes stages w, w + 4, ... whole: the per-stage fixed cost (two barrier waits, header) is a
// latency chain, and four warps walking it in lock step on the same stage would bound the pipeline.
template <int K>
__device__ __forceinline__ void umma_role_gather(const RowStepParams& p, const UmmaSmem& s, const unsigned char* H,
                                                 int w, int lane) {
    using C = UmmaCfg<K>;
    const int ch = lane & 7;               // 16-byte chunk inside the 128-byte line
    const int r4 = lane >> 3;              // rating inside its group of 4
    for (uint32_t sg = (uint32_t)w;; sg += C::GATHER_WARPS) {
        const uint32_t slot = sg % C::STAGES, ph = (sg / C::STAGES) & 1;
        const uint32_t ms = sg % C::META, mph = (sg / C::META) & 1;
        ptx::mbar_wait(&s.meta_full[ms], mph, 2, (int)sg);
        const StageHdr hd = stage_hdr<K>(s, ms);
        if (hd.flags & HDR_END) break;
        const int live_n = hd.live_n;
        ptx::mbar_wait(&s.empty[slot], ph ^ 1, 11, (int)sg);
        const uint32_t idx_a = ptx::smem_u32(s.meta + (size_t)ms * C::META_BYTES);
        const uint32_t st_addr = ptx::smem_u32(s.stages + (size_t)slot * C::STAGE_STRIDE);
        int32_t ixb[C::RL];
#pragma unroll
        for (int j = 0; j < C::RL; ++j) ixb[j] = lds_s32(idx_a + 4u * (uint32_t)(lane + 32 * j));
#pragma unroll
        for (int t = 0; t < C::R / 4; ++t) {
            if ((t & ~3) * 4 >= live_n) break;         // 16-rating MMA steps past the ragged end are not issued
            const int rating = 4 * t + r4;             // = 8 kb + rr
            const int kb = t >> 1, rr = (t & 1) * 4 + r4;
            const bool live = rating < live_n;
            const int32_t ix = __shfl_sync(0xffffffffu, ixb[t / 8], rating & 31);
            const unsigned char* src = H + (size_t)ix * C::REC_BYTES + ch * 16;
            const uint32_t dst = st_addr + (uint32_t)(kb * C::NA * 1024 + rr * 128 + ((ch ^ rr) << 4));
#pragma unroll
            for (int mb = 0; mb < C::N_MB; ++mb) {
                const int at = mb < C::HI_ATOMS ? mb : mb + 1;     // the q atom sits between the hi and lo atoms
                ptx::cp_async16_zfill(dst + at * 1024, src + mb * 128, live ? 16u : 0u);
            }
        }
#pragma unroll
        for (int j = 0; j < C::RL; ++j) {
            const bool live = lane + 32 * j < live_n;
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(ptx::smem_u32(s.bias + slot * C::BIAS_BYTES + (lane + 32 * j) * 4)),
                         "l"(p.bias_other + ixb[j]), "r"(live ? 4u : 0u)
                         : "memory");
        }
        ptx::cp_async_mbar_arrive_noinc(&s.full[slot]);
    }
    cp_async_wait<0>();
}

// ---- prep warp: q = rating - mu - bias, the [q_hi, q_lo, 1] atom of the stage, row descriptors ----
template <int K>
__device__ __forceinline__ void umma_publish_desc(const UmmaSmem& s, uint32_t n, int lane, int pw, bool owner,
                                                  int32_t row, int32_t cnt, float tsum, float qq) {
    // both prep warps write their half of descriptor n (the owner of the row's last stage also the row and count)
    const uint32_t rd = n % UM_RD, rph = (n / UM_RD) & 1;
    ptx::mbar_wait(&s.rd_empty[rd], rph ^ 1, 6, (int)n);
    if (lane == 0) {
        const uint32_t a = ptx::smem_u32(&s.rd[rd]);
        if (pw == 0) {
            if (owner) {
                sts_s32x4(a, make_int4(row, cnt, __float_as_int(tsum), __float_as_int(qq)));
            } else {
                sts_s32(a + 8, __float_as_int(tsum));
                sts_s32(a + 12, __float_as_int(qq));
            }
        } else {
            if (owner) {
                sts_s32(a, row);
                sts_s32(a + 4, cnt);
            }
            sts_s32(a + 16, __float_as_int(tsum));
            sts_s32(a + 20, __float_as_int(qq));
        }
        ptx::mbar_arrive(&s.rd_full[rd]);                  // release (count 2: one per prep warp)
    }
    __syncwarp();
}

// Two prep warps; warp pw prepares the stages sg % 2 == pw and only reads the headers of the others (row
// boundaries).  Each keeps the row's sum q / sum q^2 over ITS stages and publishes its half of the row descriptor.
template <int K>
__device__ __forceinline__ void umma_role_prep(const RowStepParams& p, const UmmaSmem& s, int pw, int lane) {
    using C = UmmaCfg<K>;
    const float mu = *p.mu;
    uint32_t nrow = 0;                                 // rows finished so far
    float tsum = 0.0f, qq = 0.0f;
    // this lane's rating line inside the stage's q atoms (128B-swizzled MN-major, like the records)
    const uint32_t kb = (uint32_t)lane >> 3, rr = (uint32_t)lane & 7;
    const uint32_t q_off = (kb * C::NA + C::HI_ATOMS) * 1024 + rr * 128;
    for (uint32_t sg = 0;; ++sg) {
        const uint32_t slot = sg % C::STAGES, ph = (sg / C::STAGES) & 1;
        const uint32_t ms = sg % C::META, mph = (sg / C::META) & 1;
        const bool mine = (int)(sg & 1) == pw;
        ptx::mbar_wait(&s.meta_full[ms], mph, 4, (int)sg);
        const StageHdr hd = stage_hdr<K>(s, ms);
        const int flags = hd.flags, live_n = hd.live_n, hrow = hd.row, hcnt = hd.cnt;
        if (flags & HDR_END) {
            if (mine) {
                ptx::mbar_wait(&s.empty[slot], ph ^ 1, 12, (int)sg);  // the slot's previous MMAs (and flag read) are done
                if (lane == 0) {
                    sts_s32(ptx::smem_u32(s.sflags + slot), HDR_END);
                    ptx::mbar_arrive(&s.prep_full[slot]);
                }
            }
            break;
        }
        if (flags & HDR_FIRST) { tsum = 0.0f; qq = 0.0f; }
        if (mine) {
            ptx::mbar_wait(&s.full[slot], ph, 5, (int)sg);
#pragma unroll
            for (int j = 0; j < C::RL; ++j) {
                const uint32_t rating = (uint32_t)lane + 32u * j;
                const bool live = (int)rating < live_n;
                const float v = lds_f32v(ptx::smem_u32(s.meta + (size_t)ms * C::META_BYTES + 4 * C::R) + 4u * rating);
                const float bo = lds_f32v(ptx::smem_u32(s.bias + slot * C::BIAS_BYTES) + 4u * rating);
                const float q = live ? v - (mu + bo) : 0.0f;
                tsum += q;
                qq = fmaf(q, q, qq);
                const __nv_bfloat16 qh = __float2bfloat16_rn(q);
                const __nv_bfloat16 ql = __float2bfloat16_rn(q - __bfloat162float(qh));
                const __nv_bfloat162 w0 = __halves2bfloat162(qh, ql);
                const __nv_bfloat162 w1 = __halves2bfloat162(__float2bfloat16_rn(live ? 1.0f : 0.0f), __float2bfloat16_rn(0.0f));
                const uint32_t qa = ptx::smem_u32(s.stages + (size_t)slot * C::STAGE_STRIDE + q_off) + 4u * j * C::NA * 1024u;
                sts_s32x4(qa + (rr << 4), make_int4(*reinterpret_cast<const int*>(&w0), *reinterpret_cast<const int*>(&w1), 0, 0));   // columns 0..7  (chunk 0 ^ rr)
                sts_s32x4(qa + ((rr ^ 1) << 4), make_int4(0, 0, 0, 0));                                                              // columns 8..15 (chunk 1 ^ rr)
            }
            ptx::fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
                sts_s32(ptx::smem_u32(s.sflags + slot), flags | (live_n << 8));
                ptx::mbar_arrive(&s.prep_full[slot]);                // release: atom + flags visible to the mma warp
                ptx::mbar_arrive(&s.meta_empty[ms]);                 // header and ratings consumed
            }
        } else {
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&s.meta_empty[ms]);      // header consumed
        }
        if (flags & HDR_LAST) {
            // (row, cnt, sum q, sum q^2) -> solver
            const float ts = warp_sum(tsum), qs = warp_sum(qq);
            umma_publish_desc<K>(s, nrow, lane, pw, mine, hrow, hcnt, ts, qs);
            ++nrow;
        }
    }
    // odd tail: the second half of the last accumulator pair holds no row
    if (nrow % C::RPB != 0) {
        umma_publish_desc<K>(s, nrow, lane, pw, pw == 0, -1, 0, 0.0f, 0.0f);
        ++nrow;
    }
    // end-of-work descriptors: one empty pair per solver group
    for (int i = 0; i < C::NG * C::RPB; ++i, ++nrow) umma_publish_desc<K>(s, nrow, lane, pw, pw == 0, -1, 0, 0.0f, 0.0f);
}

// ---- mma warp: nothing but tcgen05 issue ------------------------------------------------------
__device__ __forceinline__ uint64_t desc_from_words(uint32_t lo, uint32_t hi) {
    uint64_t d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(hi));
    return d;
}

template <int K>
__device__ __forceinline__ void umma_role_mma(const RowStepParams& p, const UmmaSmem& s, int lane) {
    using C = UmmaCfg<K>;
    (void)p;
    const uint32_t tmem = *s.tmem_base;
    constexpr uint32_t IDESC_HQ = ptx::idesc_bf16_mn(K, K + 16);   // B = [hi | q atom's 16 columns]
    constexpr uint32_t IDESC_LO = ptx::idesc_bf16_mn(K, K);        // B = lo
    constexpr uint32_t SBO = C::NA * 1024;             // K-atom (8 ratings) stride
    constexpr uint32_t LBO = 1024;                     // MN-atom (64 factors) stride
    constexpr uint32_t LO_OFF = (C::HI_ATOMS + 1) * 1024;
    // descriptor template: everything but the 14-bit start address
    const uint64_t tmpl = ptx::smem_desc(0u, LBO, SBO, ptx::SWZ_128B);
    const uint32_t stage0 = ptx::smem_u32(s.stages) >> 4;
    uint32_t nrow = 0;                                 // rows started so far
    uint32_t buf = 0, grp = 0, d_main = 0;             // accumulator buffer / solver group of the current pair
    bool first = true;
    for (uint32_t sg = 0;; ++sg) {
        const uint32_t slot = sg % C::STAGES, ph = (sg / C::STAGES) & 1;
        ptx::mbar_wait(&s.prep_full[slot], ph, 13, (int)sg);
        const int sflag = lds_s32(ptx::smem_u32(s.sflags + slot));
        const int flags = sflag & 0xff, live_n = sflag >> 8;
        if (flags & HDR_END) break;
        if (flags & HDR_FIRST) {
            const uint32_t pr = nrow / C::RPB, h = nrow % C::RPB;
            buf = pr % C::NBUF;
            grp = pr % C::ACCF;
            if (h == 0) ptx::mbar_wait(&s.acc_empty[buf], ((pr / C::NBUF) & 1) ^ 1, 3, (int)nrow);   // drained by its solver group
            d_main = tmem + buf * C::ACC_STRIDE + ((h * 16u) << 16);
            first = true;
        }
        ptx::tc_fence_after();
        if (ptx::elect_one()) {
            const uint64_t d0 = tmpl + (uint64_t)(stage0 + slot * (uint32_t)(C::STAGE_STRIDE >> 4));
#pragma unroll
            for (int ks = 0; ks < C::R / 16; ++ks) {
                if (ks * 16 >= live_n) break;                       // 16-rating steps past the ragged end hold nothing
                const uint64_t hi_desc = d0 + (uint64_t)(ks * (2 * SBO >> 4));
                const uint64_t lo_desc = hi_desc + (uint64_t)(LO_OFF >> 4);
                // [A | c s] += hi^T [hi | q] + lo^T [hi | q] + hi^T lo : TMEM lane i = row i of A, then (F^T q, F^T 1)
                ptx::umma_bf16(d_main, hi_desc, hi_desc, IDESC_HQ, (first && ks == 0) ? 0u : 1u);
                ptx::umma_bf16(d_main, lo_desc, hi_desc, IDESC_HQ, 1u);
                ptx::umma_bf16(d_main, hi_desc, lo_desc, IDESC_LO, 1u);
            }
            ptx::umma_commit(&s.empty[slot]);                        // stage reusable once these MMAs finish
            if (flags & HDR_LAST) ptx::umma_commit(&s.acc_full[grp]);
        }
        __syncwarp();
        first = false;
        if (flags & HDR_LAST) ++nrow;
    }
    // odd tail: the second half of the last accumulator pair holds no row
    if (nrow % C::RPB != 0) {
        if (lane == 0) ptx::mbar_arrive(&s.acc_full[grp]);
    }
}

// ---- solver groups ----------------------------------------------------------------------------
#ifdef B200ALS_TIMING
#define TICK(i) do { if (blockIdx.x == 0 && threadIdx.x == 96) { const long long t__ = clock64(); tacc[i] += t__ - tlast; tlast = t__; } } while (0)
#else
#define TICK(i) do { } while (0)
#endif
template <int K>
__device__ __forceinline__ void umma_role_solver(const RowStepParams& p, const UmmaSmem& s, int warp, int lane) {
    using C = UmmaCfg<K>;
    using L = Ldl<K, C::RPW>;
    const int g = warp >> 2, q = warp & 3;             // group, warp in group (== TMEM lane quadrant)
    const int h = lane / C::RPW;                       // matrix of the accumulator pair (k=64), else 0
    const int seg = lane % C::RPW;
    const int r = C::RPW * q + seg;                    // matrix row owned by this thread
    // Odd groups eliminate in reverse order (the permuted system P A P^T, P = index reversal): the
    // thread-per-row factorisation loads the LAST rows most, so alternating the direction spreads
    // the heavy warps over quadrants 0 and 3 instead of piling them on one SM sub-partition.
    // k = 128: the last group stays forward whatever its parity — a reversed group's tensor-core update of phase NW-2
    // reads (K - RPW)/8 row groups of garbage PAST its panel in the raw-row scratch at the end of the matrix buffer:
    // 1.5 KB at k = 64 (stays inside the scratch words behind the matrices), 6 KB at k = 128, which leaves the CTA's
    // shared memory when the LAST matrix buffer is a reversed group's (measured: illegal access with NG = 4)
    const bool rev = (g & 1) != 0 && (K == 64 || g != C::NG - 1);
    const int qe = rev ? 3 - q : q;                    // elimination-order block / row / lane segment
    const int re = rev ? K - 1 - r : r;
    const int sege = rev ? C::RPW - 1 - seg : seg;
    const int col_bar = 1 + 3 * g, phase_bar = 2 + 3 * g, start_bar = 3 + 3 * g;
    const uint32_t tmem = *s.tmem_base;
    const uint32_t cbs = ptx::smem_u32(s.mats + (size_t)(g * C::RPB + h) * C::MAT_FLOATS);
    // explicit .shared addresses for the group's scratch words (a generic LD/ST takes the long-scoreboard path)
    const uint32_t red_a = ptx::smem_u32(s.red + (size_t)g * 4 * C::RPB * 4);
    const uint32_t bad_a = ptx::smem_u32(s.bad + g * C::RPB + h);
    double acc_d = 0.0, acc_d2 = 0.0;                  // fused residual statistics (thread r == 0 of each matrix)
    double acc_xx = 0.0, acc_bb = 0.0;                 // fused norms: sum |x|^2, sum bias^2 over the rows solved
#ifdef B200ALS_TIMING
    long long tacc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    long long tlast = clock64();
    if (blockIdx.x == 0 && threadIdx.x == 96)
        for (int i = 8; i < 16; ++i) g_timing[i] = 0;
#endif

    uint32_t nfac = 0;                                 // factorisations so far (parity of the tensor-core update barrier)
    const uint32_t lastf_a = ptx::smem_u32(s.bad + 16 + g * C::RPB + h);   // split rows: "this group added the last piece"
    constexpr int SLOT_LD = K + 4;                     // slot row: k matrix entries | c, s, -, -

    for (uint32_t pr = g;; pr += C::NG) {
        // descriptors of the rows in this accumulator buffer
        const uint32_t n0 = pr * C::RPB;
#pragma unroll
        for (int hh = 0; hh < C::RPB; ++hh) {
            const uint32_t n = n0 + hh;
            ptx::mbar_wait(&s.rd_full[n % UM_RD], (n / UM_RD) & 1, 8, (int)n);
        }
        TICK(0);
        const int4 d0v = lds_s32x4(ptx::smem_u32(&s.rd[n0 % UM_RD]));
        const int4 dmv = lds_s32x4(ptx::smem_u32(&s.rd[(n0 + h) % UM_RD]));
        const int4 dm1 = lds_s32x4(ptx::smem_u32(&s.rd[(n0 + h) % UM_RD]) + 16u);
        const int drow = dmv.x;                        // row | -1: none | <= -2: piece (-2 - item) of a split row
        int cnt = dmv.y;
        float tsum = __int_as_float(dmv.z) + __int_as_float(dm1.x), qq = __int_as_float(dmv.w) + __int_as_float(dm1.y);
        named_bar_sync(phase_bar, 128);                // everyone holds its descriptor
        if (q == 0 && seg == 0) ptx::mbar_arrive(&s.rd_empty[(n0 + h) % UM_RD]);
        if (d0v.x == -1) break;                        // end of work (rows fill buffers in order)
        const bool part = drow <= -2;
        int64_t row = drow >= 0 ? drow : d0v.x;        // absent half: shadow a valid row, write nothing
        int slot = 0, npieces = 0;
        if (part) {
            const int4 it = p.items[-2 - drow];
            row = it.x;
            slot = it.w;
            npieces = p.slot_np[slot];
        } else if (row < 0) {                          // first half is a piece, this half is absent
            row = p.items[-2 - d0v.x].x;
        }
        const bool any_part = __any_sync(0xffffffffu, part);   // both halves sit in every warp: group-uniform
        const int64_t grow = p.row_ids ? (int64_t)p.row_ids[row] : row;   // the row's id in X / bias_own / reg_row
        const float ridge = p.reg_row ? p.reg_row[grow] : p.reg;
        const float lam_bias = p.lam_bias_row ? p.lam_bias_row[grow] : p.lam_bias;
        const float bias_old = p.bias_own[grow];
        if (q == 0 && seg == 0) { sts_s32(bad_a, 0); sts_s32(lastf_a, 0); }

        const uint32_t buf = pr % C::NBUF;
        ptx::mbar_wait(&s.acc_full[pr % C::ACCF], (pr / C::ACCF) & 1, 7, (int)pr);
        ptx::tc_fence_after();
        TICK(1);
        const uint32_t trow = tmem + buf * C::ACC_STRIDE + ((uint32_t)(32 * q) << 16);
        float d2[4];
        ptx::tmem_ld4(trow + C::D2_COL, d2);
        float c_r = d2[0] + d2[1], s_r = d2[2];
        TICK(2);

        // ---- pieces of split rows: add the partial matrix to the row's slot; the last piece solves ----
        bool from_slot = false;
        float* gs = nullptr;
        if (any_part) {
            // The slot is accumulated TRANSPOSED: thread r adds its entry (r, c) at [c][r], so the lanes of a warp
            // (consecutive rows r) hit consecutive addresses and each warp-level red.add is one 128-byte (k = 64:
            // two 64-byte) L2 transaction — adding at [r][c] was 32 scattered 4-byte atomics per instruction and
            // made a piece's flush cost ~57 k cycles.  The normal matrix is symmetric, so reading the slot back
            // row-wise ([r][c..c+3], float4) returns the sums of the entries (c, r) = (r, c).  c and s are rows K
            // and K + 1 of the slot, (sum q, sum q^2) follow.
            float* sb = p.slot_acc + (size_t)slot * split_slot_floats(K);
            if (part) gs = sb + (size_t)r * SLOT_LD;
#pragma unroll 1
            for (int c0 = 0; c0 < K; c0 += 16) {
                float v[16];
                ptx::tmem_ld16(trow + c0, v);
                if (part) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) atomicAdd(sb + (size_t)(c0 + i) * SLOT_LD + r, v[i]);
                }
            }
            if (part) {
                atomicAdd(sb + (size_t)K * SLOT_LD + r, c_r);
                atomicAdd(sb + (size_t)(K + 1) * SLOT_LD + r, s_r);
                if (q == 0 && seg == 0) {
                    float* tq = sb + (size_t)(K + 2) * SLOT_LD;
                    atomicAdd(tq, tsum);
                    atomicAdd(tq + 1, qq);
                }
                __threadfence();
            }
            named_bar_sync(phase_bar, 128);            // every thread's adds are fenced
            if (part && q == 0 && seg == 0) {
                const int old = atomicAdd(p.slot_ctr + slot, 1);
                if (old == npieces - 1) sts_s32(lastf_a, 1);
            }
            named_bar_sync(phase_bar, 128);
            from_slot = part && (lds_s32(lastf_a) != 0);
            if (from_slot) {
                __threadfence();
                c_r = __ldcg(sb + (size_t)K * SLOT_LD + r);
                s_r = __ldcg(sb + (size_t)(K + 1) * SLOT_LD + r);
                const float* tq = sb + (size_t)(K + 2) * SLOT_LD;
                tsum = __ldcg(tq);
                qq = __ldcg(tq + 1);
                cnt = (int)(p.indptr[row + 1] - p.indptr[row]);
            }
            // the last piece solves from the slot: bring the summed matrix back into the accumulator, where the
            // factorisation (window loads and tensor-core trailing updates) works on it
            if (__any_sync(0xffffffffu, from_slot)) {
#pragma unroll 1
                for (int c0 = 0; c0 < K; c0 += 16) {
                    float v[16];
                    ptx::tmem_ld16(trow + c0, v);
                    if (from_slot) {
#pragma unroll
                        for (int i4 = 0; i4 < 16; i4 += 4) {
                            const float4 gv = __ldcg(reinterpret_cast<const float4*>(gs + c0 + i4));
                            v[i4] = gv.x; v[i4 + 1] = gv.y; v[i4 + 2] = gv.z; v[i4 + 3] = gv.w;
                        }
                    }
                    ptx::tmem_st16(trow + c0, v);
                }
                ptx::tc_fence_before();
                named_bar_sync(phase_bar, 128);
                ptx::tc_fence_after();
            }
            // nothing to solve in this buffer: hand the accumulator back and move on
            const bool solve_any = __any_sync(0xffffffffu, from_slot || drow >= 0);
            if (!solve_any) {
                ptx::tc_fence_before();
                ptx::mbar_arrive(&s.acc_empty[buf]);
                named_bar_sync(phase_bar, 128);        // lastf / bad free for the next pair
                continue;
            }
        }
        const bool valid = drow >= 0 || from_slot;

        // the matrix stays in TMEM; every phase pulls one RPW-column window of this thread's row
        auto load_window = [&](const int ph, float (&w)[C::RPW], const bool is_last) {
            const uint32_t c0 = (uint32_t)(rev ? K - C::RPW * (ph + 1) : C::RPW * ph);
#pragma unroll
            for (int cc = 0; cc < C::RPW; cc += 16) {
                float v[16];
                ptx::tmem_ld16(trow + c0 + cc, v);
                if (!rev) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) w[cc + i] = v[i];
                } else {
#pragma unroll
                    for (int i = 0; i < 16; ++i) w[C::RPW - 1 - cc - i] = v[i];
                }
            }
            if (is_last) {
                ptx::tc_fence_before();
                ptx::mbar_arrive(&s.acc_empty[buf]);   // this thread is done with the accumulator
            }
        };
        const bool ok = L::factor(c_r - bias_old * s_r, ridge, re, qe, r, cbs, col_bar, start_bar,
                                  tmem + buf * C::ACC_STRIDE + ((uint32_t)(16 * h) << 16), &s.upd[g * C::RPB + h],
                                  (uint32_t)L::MMA_PHASES * nfac, load_window);
        ++nfac;
        if (!ok && valid) sts_s32(bad_a, 1);
        TICK(3);
        const float x = L::backsolve(re, qe, sege, rev ? C::RPW - 1 : 0, cbs, phase_bar);   // barriers inside publish *bad too

        TICK(4);
        // s.x, c.x, x.x over the matrix: segment shuffle, then the 4 warps through shared memory
        float v0 = s_r * x, v1 = c_r * x, v2 = x * x;
#pragma unroll
        for (int o = C::RPW / 2; o > 0; o >>= 1) {
            v0 += __shfl_xor_sync(0xffffffffu, v0, o);
            v1 += __shfl_xor_sync(0xffffffffu, v1, o);
            v2 += __shfl_xor_sync(0xffffffffu, v2, o);
        }
        if (seg == 0)
            sts_s32x4(red_a + 16u * (uint32_t)(q * C::RPB + h),
                      make_int4(__float_as_int(v0), __float_as_int(v1), __float_as_int(v2), 0));
        named_bar_sync(phase_bar, 128);
        const bool good = valid && (lds_s32(bad_a) == 0);
        if (good) store_x_all(p, (size_t)grow * p.ldx + r, x);
        if (q == 0 && seg == 0) {
            if (good) {
                float sx = 0.0f, cx = 0.0f, xx = 0.0f;
#pragma unroll
                for (int w = 0; w < 4; ++w) {
                    const int4 rr = lds_s32x4(red_a + 16u * (uint32_t)(w * C::RPB + h));
                    sx += __int_as_float(rr.x); cx += __int_as_float(rr.y); xx += __int_as_float(rr.z);
                }
                const float cntf = (float)cnt;
                const float beta = (tsum - sx) / (cntf + lam_bias);            // als.py:431-433 / 464-466
                store_bias_all(p, grow, beta);
                // residuals of the row with the fresh x, beta:  d_e = q_e - beta - f_e.x
                //   sum d   = t - cnt beta - s.x
                //   sum d^2 = qq - 2 beta t + cnt beta^2 - 2 (c.x - beta s.x) + x^T G x,
                //   x^T G x = x.(c - beta_old s) - ridge x.x      (x solves (G + ridge I) x = c - beta_old s)
                // (terms of O(cnt) cancelling to ~0.4 cnt: combined in float64, once per row on one thread)
                const double bd = (double)beta, td = (double)tsum, cd = (double)cnt, sxd = (double)sx;
                acc_d += td - cd * bd - sxd;
                acc_d2 += (double)qq - 2.0 * bd * td + cd * bd * bd - (double)cx +
                          (2.0 * bd - (double)bias_old) * sxd - (double)ridge * (double)xx;
                acc_xx += (double)xx;
                acc_bb += (double)beta * (double)beta;
            } else if (valid && p.status) {
                if (atomicAdd(&p.status[0], 1) == 0) p.status[1] = (int32_t)grow;
            }
        }
        // no barrier here: the next pair's descriptor barrier stands between these reads (red, bad, x, columns:
        // everything after the barrier above) and the next writes to them
        TICK(5);
#ifdef B200ALS_TIMING
        if (blockIdx.x == 0 && threadIdx.x == 96) tacc[6] += 1;
#endif
    }
#ifdef B200ALS_TIMING
    if (blockIdx.x == 0 && threadIdx.x == 96)
        for (int i = 0; i < 8; ++i) g_timing[i] = (unsigned long long)tacc[i];
    // g_timing[8..11] (window load / start barrier / left-looking / in-block) accumulate inside Ldl::factor
#endif
    if (p.stats && q == 0 && seg == 0) {
        atomicAdd(&p.stats[0], acc_d);
        atomicAdd(&p.stats[1], acc_d2);
    }
    if (p.norms && q == 0 && seg == 0) {
        atomicAdd(&p.norms[0], acc_xx);
        atomicAdd(&p.norms[2], acc_bb);
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
            ptx::mbar_init(&s.meta_full[i], 33);
            ptx::mbar_init(&s.meta_empty[i], C::PREP_WARPS);
        }
        for (int i = 0; i < C::STAGES; ++i) {
            ptx::mbar_init(&s.full[i], 32);
            ptx::mbar_init(&s.prep_full[i], 1);
            ptx::mbar_init(&s.empty[i], 1);
        }
        for (int i = 0; i < C::ACCF; ++i) ptx::mbar_init(&s.acc_full[i], C::RPB);
        for (int i = 0; i < C::NBUF; ++i) ptx::mbar_init(&s.acc_empty[i], 128);
        for (int i = 0; i < C::MATS; ++i) ptx::mbar_init(&s.upd[i], 1);
        for (int i = 0; i < UM_RD; ++i) {
            ptx::mbar_init(&s.rd_full[i], C::PREP_WARPS);
            ptx::mbar_init(&s.rd_empty[i], 1);
        }
        ptx::fence_barrier_init();
    }
    if (warp == C::WARP_MMA) ptx::tmem_alloc(s.tmem_base, 512);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();

    if (warp < C::SOLVER_WARPS) {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(C::REG_SOLVER));
        umma_role_solver<K>(p, s, warp, lane);
    } else {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(C::REG_PRODUCER));
        if (warp == C::WARP_META) {
            umma_role_meta<K>(p, s, lane);
        } else if (warp < C::WARP_PREP) {
            umma_role_gather<K>(p, s, H, warp - C::WARP_GATHER0, lane);
        } else if (warp < C::WARP_MMA) {
            umma_role_prep<K>(p, s, warp - C::WARP_PREP, lane);
        } else if (warp == C::WARP_MMA) {
            umma_role_mma<K>(p, s, lane);
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == C::WARP_MMA) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(*s.tmem_base, 512);
    }
}

// The kernel is instantiated for K = 64 and K = 128.  Any k <= K runs on it through zero padding: the caller
// keeps the factor rows with a row stride of exactly K floats and zeros in columns [k, K) (b200als.h), so the
// padded block of the normal matrix is ridge * I with a zero right-hand side, decoupled from the rest — the
// k leading entries of the solution are unchanged and the padded ones come out as exact zeros.
static inline int umma_padded_k(int k) { return k <= 64 ? 64 : (k <= 128 ? 128 : 0); }
static inline bool umma_supported(int k, int64_t ldf) { return (ldf == 64 || ldf == 128) && k >= 1 && k <= ldf; }
// AUTO takes the tensor-core path when more than half of the padded width is real work
static inline bool umma_preferred(int k, int64_t ldf) { return umma_supported(k, ldf) && 2 * k > ldf; }

// workspace: [256 B cursor] [bf16 copy of F]   (kp: padded width, 64 or 128)
static inline size_t umma_workspace_bytes(int kp, int64_t n_f_rows) {
    if (kp != 64 && kp != 128) return 256;
    return 256 + align_up((size_t)(n_f_rows > 0 ? n_f_rows : 0) * 4 * kp, 256);
}

template <int K>
static int launch_umma_k(RowStepParams& p, int64_t n_f_rows, const b200als_split_plan* plan, char* ws, size_t ws_bytes,
                         cudaStream_t st) {
    using C = UmmaCfg<K>;
    if (ws_bytes < umma_workspace_bytes(K, n_f_rows)) {
        set_error("row_step (tcgen05): workspace too small for the bf16 split of %lld rows", (long long)n_f_rows);
        return B200ALS_E_WORKSPACE;
    }
    __nv_bfloat16* H = reinterpret_cast<__nv_bfloat16*>(ws + 256);
    {
        int64_t work = n_f_rows * (K / 4);
        int64_t blocks = ceil_div64(work, 256);
        if (blocks > (int64_t)kNumSMs * 16) blocks = (int64_t)kNumSMs * 16;
        split_bf16_kernel<<<(unsigned)blocks, 256, 0, st>>>(p.F, p.ldf, n_f_rows, K, H);
        ALS_LAUNCH_CHECK();
    }
    p.items = nullptr; p.split_hdr = nullptr; p.slot_np = nullptr; p.slot_ctr = nullptr; p.slot_acc = nullptr;
    const bool split = plan && plan->items;
    if (split) {
        if (plan->n_list != p.n_list || plan->slot_floats != split_slot_floats(K)) {
            set_error("row_step (tcgen05): split plan was built for another list / k (n_list %lld vs %lld)",
                      (long long)plan->n_list, (long long)p.n_list);
            return B200ALS_E_INVALID;
        }
        split_zero_kernel<<<kNumSMs * 2, 256, 0, st>>>(plan->hdr, plan->slot_floats, plan->slot_acc, plan->slot_ctr);
        ALS_LAUNCH_CHECK();
        p.items = static_cast<const int4*>(plan->items);
        p.split_hdr = plan->hdr;
        p.slot_np = plan->slot_np;
        p.slot_ctr = plan->slot_ctr;
        p.slot_acc = plan->slot_acc;
    }
    auto kern = row_step_umma_kernel<K>;
    // function attributes are per device: set on every launch (cheap), not cached per thread
    ALS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::smem_bytes));
    int64_t grid = p.n_list < kNumSMs ? p.n_list : kNumSMs;
    if (split) grid = kNumSMs;                         // pieces: the item count is only known on the device
    kern<<<(unsigned)grid, C::THREADS, C::smem_bytes, st>>>(p, reinterpret_cast<const unsigned char*>(H));
    ALS_LAUNCH_CHECK();
    return B200ALS_OK;
}

static inline int launch_umma(RowStepParams& p, int64_t n_f_rows, const b200als_split_plan* plan, char* ws,
                              size_t ws_bytes, cudaStream_t s) {
    if (p.ldf == 64) return launch_umma_k<64>(p, n_f_rows, plan, ws, ws_bytes, s);
    return launch_umma_k<128>(p, n_f_rows, plan, ws, ws_bytes, s);
}

}  // namespace b200als

// row_step_ffma.cuh — generic Gram + RHS accumulation on FFMA (any k <= 160).
//
// One CTA per row, persistent, rows claimed longest-first from an atomic
// cursor.  The row's ratings are walked in chunks of FF_CHUNK: the chunk's
// (index, q = rating - mu - bias_other[index]) pairs are staged in shared
// memory, then the gathered factor rows stream through a 3-stage cp.async
// ring of FF_TR-row tiles (16-byte vectorised, coalesced per factor row).
// Thread (ti, tj), ti >= tj, owns a T x T register block of the lower
// triangle of F^T F; threads 0..KP-1 also own one column of c = F^T q and
// s = F^T 1.  The epilogue (row_solve.cuh) factors and solves in place.
#pragma once
#include "row_solve.cuh"

// the Gram register tile on packed FFMA2 (common.cuh); own switch so that it can be measured apart from the LDL^T's
#ifndef B200ALS_FFMA2_GRAM
#define B200ALS_FFMA2_GRAM B200ALS_FFMA2
#endif

namespace b200als {

constexpr int FF_TR = 16;       // ratings per pipeline tile
constexpr int FF_STAGES = 3;
constexpr int FF_CHUNK = 256;   // ratings whose (idx, q) are staged at once

struct RowStepParams {
    int k;
    int mode;
    int64_t n_list;
    const int64_t* indptr;
    const int32_t* indices;
    const float* vals;
    const int32_t* rows;
    const float* F; int64_t ldf;
    const float* bias_other;
    const float* mu;
    float reg; const float* reg_row;
    float lam_bias; const float* lam_bias_row;
    float* X; int64_t ldx;
    float* bias_own;
    float* chol_out; float* aux_out;
    int32_t* status;
    double* stats;                // optional: [0] += sum d, [1] += sum d^2 over the rows solved (A10 fused)
    double* norms;                // optional: [0] += sum |x|^2, [2] += sum bias^2 over the rows solved (als.py:513-516 fused)
    unsigned long long* cursor;   // zeroed before launch
    // tcgen05 path only: work items of the schedule with long rows cut into pieces (split_plan.cuh)
    const int4* items;            // {row, first rating (offset in the row), ratings, slot or -1}; null: rows[] as is
    const int32_t* split_hdr;     // [0] = number of items
    const int32_t* slot_np;       // pieces of split row `slot`
    int32_t* slot_ctr;            // pieces of the slot accumulated so far (zeroed)
    float* slot_acc;              // per slot: k rows x (k + 4) [normal matrix | c, s, -, -] | t, qq, -, -  (zeroed)
    // slab storage: indptr / rows[] are local to this rank's slab, row_ids[r] is the row's id in X / bias_own / reg_row
    const int32_t* row_ids;
    // fused exchange: the solved row and its bias are also stored into the replicas of the other ranks (peer-mapped
    // device memory, NVLink) — the half-step's all-gather without a separate collective
    int n_peers;
    float* peer_X[B200ALS_MAX_PEERS];
    float* peer_bias[B200ALS_MAX_PEERS];
};

// x and bias of a solved row go to every replica (thread-level: the caller's lanes hold consecutive entries, so the
// peer stores are 128-byte coalesced NVLink writes)
__device__ __forceinline__ void store_x_all(const RowStepParams& p, size_t off, float x) {
    p.X[off] = x;
    for (int j = 0; j < p.n_peers; ++j) p.peer_X[j][off] = x;
}
__device__ __forceinline__ void store_bias_all(const RowStepParams& p, int64_t row, float b) {
    p.bias_own[row] = b;
    for (int j = 0; j < p.n_peers; ++j) p.peer_bias[j][row] = b;
}

template <int KP, int T>
struct FfmaCfg {
    static constexpr int NT = KP / T;
    static constexpr int NLOW = NT * (NT + 1) / 2;
    static constexpr int THREADS = (round_up(NLOW, 32) > round_up(KP, 32)) ? round_up(NLOW, 32) : round_up(KP, 32);
    static constexpr int LDT = KP + 4;   // tile row stride (floats): 16B aligned, staggers banks
    static constexpr size_t smem_bytes =
        sizeof(float) * ((size_t)FF_STAGES * FF_TR * LDT     // gather ring
                         + solve_smem_floats(KP)             // tile-format Gram + solver scratch
                         + 2 * KP                            // c, s
                         + 2 * FF_CHUNK                      // q ring (2 chunks)
                         + 32)                               // reductions
        + sizeof(int32_t) * (2 * FF_CHUNK + 8);              // idx ring + flags
};

template <int KP, int T>
__global__ void __launch_bounds__(FfmaCfg<KP, T>::THREADS) row_step_ffma_kernel(const RowStepParams p) {
    using C = FfmaCfg<KP, T>;
    constexpr int NTHR = C::THREADS;
    constexpr int LDT = C::LDT;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* tiles = reinterpret_cast<float*>(smem_raw);
    SolveSmem S;
    S.carve<KP>(tiles + FF_STAGES * FF_TR * LDT);
    float* c_s = tiles + FF_STAGES * FF_TR * LDT + solve_smem_floats(KP);
    float* s_s = c_s + KP;
    float* q_s = s_s + KP;                 // [2][FF_CHUNK]
    float* red_s = q_s + 2 * FF_CHUNK;     // [32]
    int32_t* idx_s = reinterpret_cast<int32_t*>(red_s + 32);   // [2][FF_CHUNK]
    int32_t* misc_s = idx_s + 2 * FF_CHUNK;                    // [0]=flag, [2]=claimed row slot

    const int tid = threadIdx.x;
    const int k = p.k;
    const int k4 = (k + 3) & ~3;           // floats copied per gathered row (F is zero padded to ld)
    const int cpr = k4 >> 2;               // 16B chunks per row
    const float mu = *p.mu;

    // lower-triangle tile owned by this thread
    int ti = 0, tj = 0;
    const bool has_tile = tid < C::NLOW;
    if (has_tile) {
        ti = (int)((sqrtf(8.0f * tid + 1.0f) - 1.0f) * 0.5f);
        while ((ti + 1) * (ti + 2) / 2 <= tid) ++ti;
        while (ti * (ti + 1) / 2 > tid) --ti;
        tj = tid - ti * (ti + 1) / 2;
    }
    double stat_d = 0.0, stat_d2 = 0.0;    // fused residual statistics (thread 0)
    double stat_xx = 0.0, stat_bb = 0.0;   // fused norms (thread 0)
    // zero the tile ring once: columns [k4, KP) are never written by cp.async
    for (int e = tid; e < FF_STAGES * FF_TR * LDT; e += NTHR) tiles[e] = 0.0f;

    for (;;) {
        __syncthreads();
        if (tid == 0) {
            unsigned long long claimed = atomicAdd(p.cursor, 1ull);
            misc_s[2] = claimed < (unsigned long long)p.n_list ? (int32_t)claimed : -1;   // n_list < 2^31
        }
        __syncthreads();
        const int64_t slot = misc_s[2];
        if (slot < 0) break;
        const int64_t r = p.rows ? (int64_t)p.rows[slot] : slot;
        const int64_t gr = p.row_ids ? (int64_t)p.row_ids[r] : r;        // the row's id in X / bias_own / reg_row
        const int64_t beg = p.indptr[r], end = p.indptr[r + 1];
        const int64_t cnt = end - beg;
        if (cnt <= 0) continue;

        float acc[T][T];
#pragma unroll
        for (int a = 0; a < T; ++a)
#pragma unroll
            for (int b = 0; b < T; ++b) acc[a][b] = 0.0f;
        float c_acc = 0.0f, s_acc = 0.0f, t_acc = 0.0f, qq_acc = 0.0f;

        const int64_t nchunks = (cnt + FF_CHUNK - 1) / FF_CHUNK;
        // stage chunk 0 meta
        for (int e = tid; e < FF_CHUNK; e += NTHR) {
            int32_t ix = 0; float q = 0.0f;
            if (e < cnt) {
                ix = ld_stream_s32(p.indices + beg + e);
                q = ld_stream_f32(p.vals + beg + e) - (mu + p.bias_other[ix]);
            }
            idx_s[e] = ix; q_s[e] = q;
        }
        for (int64_t ch = 0; ch < nchunks; ++ch) {
            const int cur = (int)(ch & 1);
            const int64_t cbeg = ch * FF_CHUNK;
            const int clen = (int)((cnt - cbeg) < FF_CHUNK ? (cnt - cbeg) : FF_CHUNK);
            const int ntiles = (clen + FF_TR - 1) / FF_TR;
            const int32_t* idx_c = idx_s + cur * FF_CHUNK;
            const float* q_c = q_s + cur * FF_CHUNK;
            __syncthreads();   // meta of this chunk visible; previous chunk's tiles fully consumed

            // issue the next chunk's meta loads now; they land in registers while we compute
            constexpr int MPT = (FF_CHUNK + NTHR - 1) / NTHR;
            int32_t nix[MPT]; float nq[MPT];
            const bool have_next = (ch + 1 < nchunks);
            if (have_next) {
                const int64_t nb = beg + cbeg + FF_CHUNK;
#pragma unroll
                for (int j = 0; j < MPT; ++j) {
                    int e = tid + j * NTHR;
                    nix[j] = 0; nq[j] = 0.0f;
                    if (e < FF_CHUNK && nb + e < end) {
                        nix[j] = ld_stream_s32(p.indices + nb + e);
                        nq[j] = ld_stream_f32(p.vals + nb + e);
                    }
                }
            }

            auto issue_tile = [&](int t) {
                float* dst = tiles + (t % FF_STAGES) * FF_TR * LDT;
                const int rows_here = (clen - t * FF_TR) < FF_TR ? (clen - t * FF_TR) : FF_TR;
                for (int cch = tid; cch < FF_TR * cpr; cch += NTHR) {
                    const int rr = cch / cpr, cc = cch - rr * cpr;
                    if (rr < rows_here) {
                        const float* src = p.F + (size_t)idx_c[t * FF_TR + rr] * p.ldf + cc * 4;
                        cp_async16(dst + rr * LDT + cc * 4, src);
                    } else {
                        *reinterpret_cast<float4*>(dst + rr * LDT + cc * 4) = make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                }
            };
#pragma unroll
            for (int s = 0; s < FF_STAGES - 1; ++s) {
                if (s < ntiles) issue_tile(s);
                cp_async_commit();
            }
            for (int t = 0; t < ntiles; ++t) {
                cp_async_wait<FF_STAGES - 2>();
                __syncthreads();
                // prefetch tile t + STAGES-1 into the slot consumed at iteration t-1
                if (t + FF_STAGES - 1 < ntiles) issue_tile(t + FF_STAGES - 1);
                cp_async_commit();
                const float* tl = tiles + (t % FF_STAGES) * FF_TR * LDT;
                const float* qt = q_c + t * FF_TR;
                if (has_tile) {
#pragma unroll 4
                    for (int rr = 0; rr < FF_TR; ++rr) {
                        float av[T], bv[T];
#pragma unroll
                        for (int v = 0; v < T; v += 4) {
                            float4 a4 = *reinterpret_cast<const float4*>(tl + rr * LDT + ti * T + v);
                            float4 b4 = *reinterpret_cast<const float4*>(tl + rr * LDT + tj * T + v);
                            av[v] = a4.x; av[v + 1] = a4.y; av[v + 2] = a4.z; av[v + 3] = a4.w;
                            bv[v] = b4.x; bv[v + 1] = b4.y; bv[v + 2] = b4.z; bv[v + 3] = b4.w;
                        }
#if B200ALS_FFMA2_GRAM
#pragma unroll
                        for (int a = 0; a < T; ++a) {
                            const uint64_t aa = pack_f32x2(av[a], av[a]);
#pragma unroll
                            for (int b = 0; b < T; b += 2) {
                                const uint64_t x = fma_f32x2(aa, pack_f32x2(bv[b], bv[b + 1]), pack_f32x2(acc[a][b], acc[a][b + 1]));
                                unpack_f32x2(x, acc[a][b], acc[a][b + 1]);
                            }
                        }
#else
#pragma unroll
                        for (int a = 0; a < T; ++a)
#pragma unroll
                            for (int b = 0; b < T; ++b) acc[a][b] = fmaf(av[a], bv[b], acc[a][b]);
#endif
                    }
                }
                if (tid < KP) {
#pragma unroll 4
                    for (int rr = 0; rr < FF_TR; ++rr) {
                        const float f = tl[rr * LDT + tid];
                        const float q = qt[rr];   // zero beyond the chunk (tile rows are zero there too)
                        c_acc = fmaf(f, q, c_acc);
                        s_acc += f;
                        t_acc += q;
                        qq_acc = fmaf(q, q, qq_acc);
                    }
                }
            }
            cp_async_wait<0>();
            // finish next chunk's meta: dependent bias gather, then publish to the other ring slot
            if (have_next) {
                int32_t* idx_n = idx_s + (cur ^ 1) * FF_CHUNK;
                float* q_n = q_s + (cur ^ 1) * FF_CHUNK;
                const int64_t nb = beg + cbeg + FF_CHUNK;
#pragma unroll
                for (int j = 0; j < MPT; ++j) {
                    int e = tid + j * NTHR;
                    if (e < FF_CHUNK) {
                        float q = 0.0f;
                        if (nb + e < end) q = nq[j] - (mu + p.bias_other[nix[j]]);
                        idx_n[e] = nix[j];
                        q_n[e] = q;
                    }
                }
            }
        }
        // note: q ring entries past cnt in the last chunk are zero only for chunk 0's
        // staging and the publish above (both write all FF_CHUNK entries).

        // ---- epilogue: lower triangle → 4x4 tiles in smem, solve ---------------
        __syncthreads();
        const int kt = (k + 3) >> 2;
        if (has_tile) {
#pragma unroll
            for (int a = 0; a < T / 4; ++a)
#pragma unroll
                for (int b = 0; b < T / 4; ++b) {
                    const int sti = ti * (T / 4) + a, stj = tj * (T / 4) + b;
                    if (stj <= sti && sti < kt) {
#pragma unroll
                        for (int rr = 0; rr < 4; ++rr)
                            *tile_row(S.Gt, sti, stj, rr) = make_float4(acc[a * 4 + rr][b * 4 + 0], acc[a * 4 + rr][b * 4 + 1],
                                                  acc[a * 4 + rr][b * 4 + 2], acc[a * 4 + rr][b * 4 + 3]);
                    }
                }
        }
        if (tid < KP) { c_s[tid] = c_acc; s_s[tid] = s_acc; }
        if (tid == 0) { red_s[31] = t_acc; red_s[30] = qq_acc; }
        __syncthreads();
        RowEpilogueArgs ea;
        ea.k = k; ea.mode = p.mode;
        ea.reg = p.reg_row ? p.reg_row[gr] : p.reg;
        ea.lam_bias = p.lam_bias_row ? p.lam_bias_row[gr] : p.lam_bias;
        ea.X = p.X; ea.ldx = p.ldx; ea.bias_own = p.bias_own;
        ea.n_peers = p.n_peers;
        for (int j = 0; j < B200ALS_MAX_PEERS; ++j) { ea.peer_X[j] = p.peer_X[j]; ea.peer_bias[j] = p.peer_bias[j]; }
        ea.chol_out = p.chol_out; ea.aux_out = p.aux_out; ea.status = p.status;
        ea.stats_d = &stat_d; ea.stats_d2 = &stat_d2;
        ea.stats_xx = &stat_xx; ea.stats_bb = &stat_bb;
        const float t_sum = red_s[31];
        ea.qq = red_s[30];
        row_epilogue_tiles(S, c_s, s_s, t_sum, cnt, gr, ea, tid, NTHR, 1);
    }
    if (p.stats && tid == 0) {
        atomicAdd(&p.stats[0], stat_d);
        atomicAdd(&p.stats[1], stat_d2);
    }
    if (p.norms && tid == 0) {
        atomicAdd(&p.norms[0], stat_xx);
        atomicAdd(&p.norms[2], stat_bb);
    }
}

}  // namespace b200als

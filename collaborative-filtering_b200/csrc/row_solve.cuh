// row_solve.cuh — the per-row epilogue shared by every Gram path (A6/A7/A11):
// blocked in-shared-memory Cholesky of the k x k normal matrix, forward
// substitution folded in as an augmented row, blocked back substitution,
// closed-form bias.
//
// Storage: 4x4 tiles of the LOWER triangle only.  Tile (ti, tj), ti >= tj,
// lives at Gt + (ti(ti+1)/2 + tj) * 16, row-major inside the tile, so a tile
// is four 16-byte vectors and consecutive tiles are consecutive in memory
// (conflict-free LDS.128 / STS.128).  kt = ceil(k/4) tile rows hold A; tile
// row kt holds the right-hand side in its first row (other rows zero), so the
// right-looking update turns it into y = L^-1 b.  k=128 needs 36 KB (vs 66 KB
// dense), which is what lets 3 CTAs share an SM and hide each other's
// pivot chains.
//
// Algorithm: panels of 8 columns (two tile columns).  Per panel:
//   (a) one thread factors the 8x8 diagonal block in registers (rsqrt+Newton);
//   (b) every remaining row solves its 8 entries against it (thread per row);
//   (c) all threads apply the rank-8 update to the trailing tiles (thread per
//       4x4 tile, 128 FMA per 12 vector loads; rows of the panel are shared
//       by neighbouring threads, so the loads broadcast).
// Three barriers per panel instead of two per column.
#pragma once
#include "common.cuh"

namespace b200als {

constexpr int SOLVE_NB = 8;

__host__ __device__ constexpr int tri_tiles(int kt) { return (kt + 1) * (kt + 2) / 2; }   // incl. rhs tile row
// tiles are stored in groups of 8: [group][row 0..3][tile 0..7][4 floats], so that row r of 8
// consecutive tiles is one 128-byte line (conflict-free LDS.128 for thread-per-tile loops)
__host__ __device__ constexpr int tile_store_floats(int kt) { return ((tri_tiles(kt) + 7) / 8) * 128; }
__host__ __device__ constexpr size_t solve_smem_floats(int KP) {
    // tiles (+rhs tile row) + panel buffer Lp[8][KP/4+1][4] + Ld[64] + invd[KP+4] + x[KP+4] + 8 misc
    return (size_t)tile_store_floats(KP / 4) + (size_t)(KP + 4) * SOLVE_NB + 64 + 2 * (KP + 4) + 8;
}

__device__ __forceinline__ int tile_base(int ti, int tj) {
    const int T = (ti * (ti + 1)) / 2 + tj;
    return ((T >> 3) << 7) + ((T & 7) << 2);
}
__device__ __forceinline__ float4* tile_row(float* Gt, int ti, int tj, int r) {
    return reinterpret_cast<float4*>(Gt + tile_base(ti, tj) + (r << 5));
}
__device__ __forceinline__ float* tile_elem(float* Gt, int i, int j) {     // i >= j
    return Gt + tile_base(i >> 2, j >> 2) + ((i & 3) << 5) + (j & 3);
}
// panel buffer: Lp[h][r][t] (float4) = entries 4h..4h+3 of the panel row 4t+r
__device__ __forceinline__ float4* lp_vec(float* Lp, int NTP, int t, int r, int h) {
    return reinterpret_cast<float4*>(Lp) + ((h * 4 + r) * NTP + t);
}

// barrier among the `nthr` threads that solve one matrix: a named barrier, or just the warp
__device__ __forceinline__ void bar_sync(int id, int nthr) {
    if (nthr == 32) __syncwarp();
    else asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthr) : "memory");
}

// decode a linear index over the lower triangle: q -> (a, b), a >= b
__device__ __forceinline__ void tri_decode(int q, int& a, int& b) {
    a = (int)((sqrtf(8.0f * (float)q + 1.0f) - 1.0f) * 0.5f);
    while ((a + 1) * (a + 2) / 2 <= q) ++a;
    while (a * (a + 1) / 2 > q) --a;
    b = q - a * (a + 1) / 2;
}

struct SolveSmem {
    float* Gt;     // tiles
    float* Lp;     // current panel, [half][row in tile][tile row] float4 (see lp_vec)
    int NTP;       // tile rows in Lp (= KP/4 + 1)
    float* Ld;     // [8][8] diagonal block of the current panel
    float* invd;   // [KP+4] 1 / L[j][j]
    float* x;      // [KP+4] solution / scratch
    int* flag;     // [0] = bad pivot
    template <int KP>
    __device__ __forceinline__ void carve(float* base) {
        Gt = base;
        NTP = KP / 4 + 1;
        Lp = Gt + tile_store_floats(KP / 4);
        Ld = Lp + (KP + 4) * SOLVE_NB;
        invd = Ld + 64;
        x = invd + (KP + 4);
        flag = reinterpret_cast<int*>(x + (KP + 4));
    }
};

// Factor + forward substitution.  kt = ceil(k/4).  All `nthr` threads of barrier `bar`.
// Returns false (uniformly) on a non-positive / NaN pivot.
__device__ __forceinline__ bool chol_factor_tiles(const SolveSmem& S, int kt, int tid, int nthr, int bar) {
    float* __restrict__ Gt = S.Gt;
    float* __restrict__ Lp = S.Lp;
    float* __restrict__ Ld = S.Ld;
    if (tid == 0) *S.flag = 0;
    for (int tc0 = 0; tc0 < kt; tc0 += 2) {
        const int ntc = (kt - tc0) >= 2 ? 2 : 1;
        const int nb = ntc * 4;
        bar_sync(bar, nthr);
        // ---- (a) diagonal block, one thread, registers ----------------------------------
        if (tid == 0) {
            float a[8][8];
#pragma unroll
            for (int r = 0; r < 8; ++r)
#pragma unroll
                for (int c = 0; c < 8; ++c) a[r][c] = (r == c) ? 1.0f : 0.0f;
            {
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    float4 v = *tile_row(Gt, tc0, tc0, r);
                    a[r][0] = v.x; a[r][1] = v.y; a[r][2] = v.z; a[r][3] = v.w;
                }
                if (ntc == 2) {
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        float4 v = *tile_row(Gt, tc0 + 1, tc0, r), w = *tile_row(Gt, tc0 + 1, tc0 + 1, r);
                        a[4 + r][0] = v.x; a[4 + r][1] = v.y; a[4 + r][2] = v.z; a[4 + r][3] = v.w;
                        a[4 + r][4] = w.x; a[4 + r][5] = w.y; a[4 + r][6] = w.z; a[4 + r][7] = w.w;
                    }
                }
            }
            bool ok = true;
            float inv[8];
#pragma unroll
            for (int p = 0; p < 8; ++p) {
                const float d = a[p][p];
                ok = ok && (d > 0.0f);
                float iv = rsqrtf(d);
                iv = iv * (1.5f - 0.5f * d * iv * iv);      // one Newton step: ~1 ulp
                inv[p] = iv;
                a[p][p] = d * iv;
#pragma unroll
                for (int r = p + 1; r < 8; ++r) a[r][p] *= iv;
#pragma unroll
                for (int r = p + 1; r < 8; ++r)
#pragma unroll
                    for (int c = p + 1; c <= r; ++c) a[r][c] = fmaf(-a[r][p], a[c][p], a[r][c]);
            }
            if (!ok) *S.flag = 1;
            // write back: tiles, Ld (row-major 8x8, strictly-lower + diag), invd
#pragma unroll
            for (int r = 0; r < 4; ++r) *tile_row(Gt, tc0, tc0, r) = make_float4(a[r][0], a[r][1], a[r][2], a[r][3]);
            if (ntc == 2) {
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    *tile_row(Gt, tc0 + 1, tc0, r) = make_float4(a[4 + r][0], a[4 + r][1], a[4 + r][2], a[4 + r][3]);
                    *tile_row(Gt, tc0 + 1, tc0 + 1, r) = make_float4(a[4 + r][4], a[4 + r][5], a[4 + r][6], a[4 + r][7]);
                }
            }
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                reinterpret_cast<float4*>(Ld + r * 8)[0] = make_float4(a[r][0], a[r][1], a[r][2], a[r][3]);
                reinterpret_cast<float4*>(Ld + r * 8)[1] = make_float4(a[r][4], a[r][5], a[r][6], a[r][7]);
                S.invd[tc0 * 4 + r] = inv[r];
            }
        }
        bar_sync(bar, nthr);
        if (*S.flag) break;
        // ---- (b) panel rows below the block (+ rhs row): a <- a * Ld^-T ------------------
        const int tr0 = tc0 + ntc;                        // first tile row below the block
        const int ntr = kt - tr0 + 1;                      // tile rows incl. the rhs tile row (only its row 0 is real)
        for (int t = tid; t < ntr * 4; t += nthr) {
            // lanes walk tile rows first (same row-in-tile), so panel-buffer stores are contiguous
            const int ti = tr0 + (t % ntr), li = t / ntr;
            if (ti == kt && li != 0) continue;
            float4* p0 = tile_row(Gt, ti, tc0, li);
            float4* p1 = tile_row(Gt, ti, tc0 + 1, li);
            float a[8];
            {
                float4 v = *p0;
                a[0] = v.x; a[1] = v.y; a[2] = v.z; a[3] = v.w;
                if (ntc == 2) { float4 w = *p1; a[4] = w.x; a[5] = w.y; a[6] = w.z; a[7] = w.w; }
                else { a[4] = a[5] = a[6] = a[7] = 0.0f; }
            }
#pragma unroll
            for (int p = 0; p < 8; ++p) {
                float s = a[p];
#pragma unroll
                for (int q = 0; q < p; ++q) s = fmaf(-a[q], Ld[p * 8 + q], s);
                a[p] = s * S.invd[tc0 * 4 + p];
            }
            *p0 = make_float4(a[0], a[1], a[2], a[3]);
            if (ntc == 2) *p1 = make_float4(a[4], a[5], a[6], a[7]);
            *lp_vec(Lp, S.NTP, ti, li, 0) = make_float4(a[0], a[1], a[2], a[3]);
            *lp_vec(Lp, S.NTP, ti, li, 1) = make_float4(a[4], a[5], a[6], a[7]);
        }
        bar_sync(bar, nthr);
        // ---- (c) trailing update: tile(ti,tj) -= Lp[ti rows] * Lp[tj rows]^T ---------------
        const int ta0 = tc0 + ntc;
        const int na = kt - ta0;
        const int ntri = na * (na + 1) / 2;
        const int ntiles = ntri + na;                      // + rhs tile row
        for (int q = tid; q < ntiles; q += nthr) {
            int ti, tj;
            if (q < ntri) {
                int aa, bb;
                tri_decode(q, aa, bb);
                ti = ta0 + aa; tj = ta0 + bb;
            } else {
                ti = kt; tj = ta0 + (q - ntri);
            }
            float* tb = Gt + tile_base(ti, tj);
            float4 acc[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) acc[r] = *reinterpret_cast<float4*>(tb + (r << 5));
#pragma unroll
            for (int h = 0; h < 2; ++h) {                 // two halves of the 8-wide panel
                float4 rr[4], cc[4];
#pragma unroll
                for (int r = 0; r < 4; ++r) { rr[r] = *lp_vec(Lp, S.NTP, ti, r, h); cc[r] = *lp_vec(Lp, S.NTP, tj, r, h); }
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const float4 a4 = rr[r];
                    acc[r].x = fmaf(-a4.x, cc[0].x, fmaf(-a4.y, cc[0].y, fmaf(-a4.z, cc[0].z, fmaf(-a4.w, cc[0].w, acc[r].x))));
                    acc[r].y = fmaf(-a4.x, cc[1].x, fmaf(-a4.y, cc[1].y, fmaf(-a4.z, cc[1].z, fmaf(-a4.w, cc[1].w, acc[r].y))));
                    acc[r].z = fmaf(-a4.x, cc[2].x, fmaf(-a4.y, cc[2].y, fmaf(-a4.z, cc[2].z, fmaf(-a4.w, cc[2].w, acc[r].z))));
                    acc[r].w = fmaf(-a4.x, cc[3].x, fmaf(-a4.y, cc[3].y, fmaf(-a4.z, cc[3].z, fmaf(-a4.w, cc[3].w, acc[r].w))));
                }
            }
#pragma unroll
            for (int r = 0; r < 4; ++r) *reinterpret_cast<float4*>(tb + (r << 5)) = acc[r];
        }
    }
    bar_sync(bar, nthr);
    return *S.flag == 0;
}

// Back substitution L^T x = y.  y sits in the rhs tile row; x ends in S.x[0..kt*4).
__device__ __forceinline__ void chol_backsolve_tiles(const SolveSmem& S, int kt, int tid, int nthr, int bar) {
    float* __restrict__ Gt = S.Gt;
    float* __restrict__ x = S.x;
    for (int j = tid; j < kt * 4; j += nthr) x[j] = Gt[tile_base(kt, j >> 2) + (j & 3)];
    const int last_tc0 = ((kt - 1) >> 1) << 1;
    for (int tc0 = last_tc0; tc0 >= 0; tc0 -= 2) {
        const int ntc = (kt - tc0) >= 2 ? 2 : 1;
        const int nb = ntc * 4;
        bar_sync(bar, nthr);
        if (tid == 0) {
            // 8x8 (or 4x4) lower block, transposed solve from the bottom row up
            float l[8][8];
#pragma unroll
            for (int r = 0; r < 8; ++r)
#pragma unroll
                for (int c = 0; c < 8; ++c) l[r][c] = 0.0f;
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                float4 v = *tile_row(Gt, tc0, tc0, r);
                l[r][0] = v.x; l[r][1] = v.y; l[r][2] = v.z; l[r][3] = v.w;
            }
            float y[8];
#pragma unroll
            for (int r = 0; r < 8; ++r) y[r] = 0.0f;
#pragma unroll
            for (int r = 0; r < 4; ++r) y[r] = x[tc0 * 4 + r];
            if (ntc == 2) {
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    float4 v = *tile_row(Gt, tc0 + 1, tc0, r), w = *tile_row(Gt, tc0 + 1, tc0 + 1, r);
                    l[4 + r][0] = v.x; l[4 + r][1] = v.y; l[4 + r][2] = v.z; l[4 + r][3] = v.w;
                    l[4 + r][4] = w.x; l[4 + r][5] = w.y; l[4 + r][6] = w.z; l[4 + r][7] = w.w;
                    y[4 + r] = x[tc0 * 4 + 4 + r];
                }
            }
#pragma unroll
            for (int j = 7; j >= 0; --j) {
                if (j < nb) {
                    const float xj = y[j] * S.invd[tc0 * 4 + j];
                    y[j] = xj;
#pragma unroll
                    for (int i = 0; i < j; ++i) y[i] = fmaf(-l[j][i], xj, y[i]);
                }
            }
#pragma unroll
            for (int r = 0; r < 8; ++r)
                if (r < nb) x[tc0 * 4 + r] = y[r];
        }
        bar_sync(bar, nthr);
        // y_i -= sum_{c in panel} L[c][i] * x_c  for i < tc0*4  (row c of L is contiguous in tiles)
        for (int i = tid; i < tc0 * 4; i += nthr) {
            const int tj = i >> 2, lj = i & 3;
            float acc = x[i];
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                if (c < nb) {
                    const int row = tc0 * 4 + c;
                    acc = fmaf(-Gt[tile_base(row >> 2, tj) + ((row & 3) << 5) + lj], x[row], acc);
                }
            }
            x[i] = acc;
        }
    }
    bar_sync(bar, nthr);
}

struct RowEpilogueArgs {
    int k;
    int mode;
    float reg, lam_bias;
    float* X; int64_t ldx;
    float* bias_own;
    float* chol_out; float* aux_out;
    int32_t* status;
    float qq;                     // sum q^2 of the row (fused residual statistics)
    double* stats_d; double* stats_d2;   // thread-0 accumulators, may be null
    double* stats_xx; double* stats_bb;  // thread-0 accumulators of |x|^2 and bias^2
    int n_peers;                         // fused exchange: replicas of the other ranks (row_step_ffma.cuh)
    float* peer_X[B200ALS_MAX_PEERS];
    float* peer_bias[B200ALS_MAX_PEERS];
};

// Preconditions: the lower-triangle tiles of F^T F are in S.Gt (tile rows 0..kt-1), c_s = F^T q
// and s_s = F^T 1 hold k entries, all visible to the `nthr` threads (barrier done by caller).
// Adds the ridge, builds b = c - bias_old * s, factors, solves, writes X[r], bias_own[r]
// (or the factor + aux in FACTOR mode).
__device__ __forceinline__ void row_epilogue_tiles(const SolveSmem& S, const float* __restrict__ c_s,
                                                   const float* __restrict__ s_s, float t_sum, int64_t cnt,
                                                   int64_t r, const RowEpilogueArgs& a, int tid, int nthr, int bar) {
    const int k = a.k;
    const int kt = (k + 3) >> 2;
    float* __restrict__ Gt = S.Gt;
    const float bias_old = a.bias_own[r];
    // ridge on the diagonal; identity on padded rows; rhs tile row
    for (int i = tid; i < kt * 4; i += nthr) {
        float* d = tile_elem(Gt, i, i);
        *d = (i < k) ? *d + a.reg : 1.0f;
        float* rt = Gt + tile_base(kt, i >> 2) + (i & 3);
        rt[0] = (i < k) ? c_s[i] - bias_old * s_s[i] : 0.0f;
        rt[32] = 0.0f; rt[64] = 0.0f; rt[96] = 0.0f;
    }
    if (tid < 6) *lp_vec(S.Lp, S.NTP, kt, 1 + (tid >> 1), tid & 1) = make_float4(0.f, 0.f, 0.f, 0.f);   // rows 1..3 of the rhs tile row stay zero
    const bool ok = chol_factor_tiles(S, kt, tid, nthr, bar);   // starts with a barrier
    if (!ok) {
        if (tid == 0 && a.status) {
            if (atomicAdd(&a.status[0], 1) == 0) a.status[1] = (int32_t)r;
        }
        return;
    }
    if (a.mode == B200ALS_STEP_FACTOR) {
        float* Lo = a.chol_out + (size_t)r * k * k;
        for (int e = tid; e < k * k; e += nthr) {
            const int i = e / k, j = e - i * k;
            Lo[e] = (j <= i) ? *tile_elem(Gt, i, j) : 0.0f;
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
    chol_backsolve_tiles(S, kt, tid, nthr, bar);
    const float* x = S.x;
    // bias_own = (t - s·x) / (cnt + lam_bias)        (als.py:431-433, 464-466); first warp reduces
    if (tid < 32) {
        float part = 0.0f, cx = 0.0f, xx = 0.0f;
        for (int i = tid; i < k; i += 32) {
            part = fmaf(s_s[i], x[i], part);
            cx = fmaf(c_s[i], x[i], cx);
            xx = fmaf(x[i], x[i], xx);
        }
        part = warp_sum(part);
        cx = warp_sum(cx);
        xx = warp_sum(xx);
        if (tid == 0) {
            const float fc = (float)cnt;
            const float beta = (t_sum - part) / (fc + a.lam_bias);
            a.bias_own[r] = beta;
            for (int j = 0; j < a.n_peers; ++j) a.peer_bias[j][r] = beta;
            *a.stats_xx += (double)xx;
            *a.stats_bb += (double)beta * (double)beta;
            if (a.stats_d) {
                // d_e = q_e - beta - f_e.x over the row, from the quantities already at hand:
                //   sum d^2 = qq - 2 beta t + cnt beta^2 - 2 (c.x - beta s.x) + x^T G x,
                //   x^T G x = x.(c - beta_old s) - ridge x.x
                // The terms are O(cnt) and cancel to ~0.4 cnt: combined in float64 (once per row, one thread).
                const double bd = (double)beta, td = (double)t_sum, cd = (double)cnt, pd = (double)part;
                *a.stats_d += td - cd * bd - pd;
                *a.stats_d2 += (double)a.qq - 2.0 * bd * td + cd * bd * bd - (double)cx +
                               (2.0 * bd - (double)bias_old) * pd - (double)a.reg * (double)xx;
            }
        }
    }
    float* xo = a.X + (size_t)r * a.ldx;
    for (int i = tid; i < k; i += nthr) xo[i] = x[i];
    for (int j = 0; j < a.n_peers; ++j) {
        float* xp = a.peer_X[j] + (size_t)r * a.ldx;
        for (int i = tid; i < k; i += nthr) xp[i] = x[i];
    }
}

}  // namespace b200als

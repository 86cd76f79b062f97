// row_step.cu — b200als_row_step: dispatch of one ALS half-step (A6/A7) to the
// Gram path that fits k: tcgen05 (row_step_umma.cuh) for k in {64, 128},
// register-tiled FFMA (row_step_ffma.cuh) otherwise; and the Gauss–Seidel
// Laplacian level solve (A8).
#include <cooperative_groups.h>

#include "row_step_f64.cuh"
#include "row_step_ffma.cuh"
#include "row_step_umma.cuh"

namespace b200als {

static thread_local int g_last_path = 0;   // Gram path the last b200als_row_step on this thread took

template <int KP, int T>
static int launch_ffma(const RowStepParams& p, cudaStream_t s) {
    using C = FfmaCfg<KP, T>;
    auto kern = row_step_ffma_kernel<KP, T>;
    // per (KP,T) instantiation and per device (function attributes belong to the device's context)
    static thread_local int blocks_per_dev[64] = {0};
    int dev = 0;
    ALS_CUDA(cudaGetDevice(&dev));
    int& blocks_per_sm = blocks_per_dev[dev & 63];
    if (blocks_per_sm == 0) {
        ALS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::smem_bytes));
        int b = 0;
        ALS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, kern, C::THREADS, C::smem_bytes));
        if (b < 1) {
            set_error("row_step_ffma<%d,%d>: kernel does not fit on an SM (smem %zu)", KP, T, C::smem_bytes);
            return B200ALS_E_UNSUPPORTED;
        }
        blocks_per_sm = b;
    }
    int64_t grid = (int64_t)kNumSMs * blocks_per_sm;
    if (grid > p.n_list) grid = p.n_list;
    kern<<<(unsigned)grid, C::THREADS, C::smem_bytes, s>>>(p);
    ALS_LAUNCH_CHECK();
    return B200ALS_OK;
}

static int dispatch_ffma(const RowStepParams& p, cudaStream_t s) {
    const int k = p.k;
    if (k <= 8) return launch_ffma<8, 4>(p, s);
    if (k <= 16) return launch_ffma<16, 4>(p, s);
    if (k <= 32) return launch_ffma<32, 4>(p, s);
    if (k <= 64) return launch_ffma<64, 4>(p, s);
    if (k <= 96) return launch_ffma<96, 8>(p, s);
    if (k <= 128) return launch_ffma<128, 8>(p, s);
    return launch_ffma<160, 8>(p, s);
}

// ---- A8: one Gauss–Seidel level ------------------------------------------------
// One warp per item of the level.  b = c - b_i*s + alpha * S[i,:] V (CSR row, live V),
// forward/back substitution against the stored lower factor, closed-form bias.
template <int KMAX>
__global__ void __launch_bounds__(128) graph_level_kernel(int k, const int32_t* __restrict__ level_rows,
                                                          int64_t n_level, const float* __restrict__ chol,
                                                          const float* __restrict__ aux,
                                                          const int64_t* __restrict__ sptr,
                                                          const int32_t* __restrict__ sidx,
                                                          const float* __restrict__ sval, float alpha, float lam_bias,
                                                          const float* __restrict__ lam_bias_row, float* V,
                                                          int64_t ldv, float* b_i) {
    __shared__ float ysm[4][KMAX];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t item_slot = (int64_t)blockIdx.x * 4 + w;
    if (item_slot >= n_level) return;
    const int64_t i = level_rows[item_slot];
    float* y = ysm[w];
    const float* L = chol + (size_t)i * k * k;
    const float* ax = aux + (size_t)i * (2 * k + 2);
    const float bold = b_i[i];
    // rhs
    for (int c = lane; c < k; c += 32) y[c] = ax[c] - bold * ax[k + c];
    __syncwarp();
    for (int64_t e = sptr[i]; e < sptr[i + 1]; ++e) {
        const float sv = alpha * sval[e];
        const float* vj = V + (size_t)sidx[e] * ldv;
        for (int c = lane; c < k; c += 32) y[c] = fmaf(sv, vj[c], y[c]);
    }
    __syncwarp();
    // forward: L y' = y
    for (int j = 0; j < k; ++j) {
        const float yj = y[j] / L[j * k + j];
        __syncwarp();
        if (lane == 0) y[j] = yj;
        for (int r = j + 1 + lane; r < k; r += 32) y[r] -= L[r * k + j] * yj;
        __syncwarp();
    }
    // backward: L^T x = y'
    for (int j = k - 1; j >= 0; --j) {
        const float xj = y[j] / L[j * k + j];
        __syncwarp();
        if (lane == 0) y[j] = xj;
        for (int r = lane; r < j; r += 32) y[r] -= L[j * k + r] * xj;
        __syncwarp();
    }
    float dot = 0.0f;
    for (int c = lane; c < k; c += 32) {
        const float x = y[c];
        V[(size_t)i * ldv + c] = x;
        dot = fmaf(ax[k + c], x, dot);
    }
    dot = warp_sum(dot);
    if (lane == 0) {
        const float lb = lam_bias_row ? lam_bias_row[i] : lam_bias;
        b_i[i] = (ax[2 * k] - dot) / (ax[2 * k + 1] + lb);
    }
}

// ---- A8 at speed: the whole Gauss–Seidel sweep in ONE cooperative launch ------------------------
// All levels are walked inside the kernel with a grid barrier between them (the per-level launch above costs a
// host call + a launch per level: 140 of them per iteration on the bundled genres graph).  One warp per item of a
// level: the item's Cholesky factor (k x k float32, stored by the FACTOR step) is staged into the warp's shared
// memory with coalesced 16-byte loads, the right-hand side b = c - b_i s + alpha S[i,:] V is built with the live V,
// the forward substitution runs row-wise (one coalesced shared-memory row dot + a warp reduction per unknown, no
// column-strided global reads), the backward substitution column-wise as rank-1 updates with a shuffle broadcast.
// y stays in registers (lane l holds entries l, l + 32, ...).  Same arithmetic order across levels as als.py:453-461:
// S is symmetric, so an item's neighbours are always in other levels.
template <int NCH>
__global__ void __launch_bounds__(256) graph_sweep_kernel(int k, const int64_t* __restrict__ level_ptr, int n_levels,
                                                          const int32_t* __restrict__ level_rows,
                                                          const float* __restrict__ chol, const float* __restrict__ aux,
                                                          const int64_t* __restrict__ sptr,
                                                          const int32_t* __restrict__ sidx,
                                                          const float* __restrict__ sval, float alpha, float lam_bias,
                                                          const float* __restrict__ lam_bias_row, float* V,
                                                          int64_t ldv, float* b_i, int warps_per_cta) {
    extern __shared__ __align__(16) float sweep_sm[];
    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const bool active = w < warps_per_cta;
    // the factor is staged with an ODD row stride: rows (backward substitution) and columns (forward substitution) of L
    // are then both conflict-free for the 32 lanes
    const int ks = k | 1;
    float* Ls = sweep_sm + (size_t)w * k * ks;
    const int64_t gw = (int64_t)blockIdx.x * warps_per_cta + w, nw = (int64_t)gridDim.x * warps_per_cta;
    const int kk = k * k;
    for (int lv = 0; lv < n_levels; ++lv) {
        const int64_t lo = level_ptr[lv], hi = level_ptr[lv + 1];
        if (active) {
            for (int64_t slot = lo + gw; slot < hi; slot += nw) {
                const int64_t i = level_rows[slot];
                const float* L = chol + (size_t)i * kk;
                const float* ax = aux + (size_t)i * (2 * k + 2);
                // stage the factor: coalesced 16-byte global loads, scalar shared stores (the odd stride breaks the
                // 16-byte alignment of the rows)
                const int kk4 = kk & ~3;
                if ((((uintptr_t)L) & 15) == 0) {
                    for (int e = lane * 4; e < kk4; e += 128) {
                        const float4 v = *reinterpret_cast<const float4*>(L + e);
                        const float vv4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const int ee = e + u, rr = ee / k;
                            Ls[rr * ks + (ee - rr * k)] = vv4[u];
                        }
                    }
                    for (int e = kk4 + lane; e < kk; e += 32) { const int rr = e / k; Ls[rr * ks + (e - rr * k)] = L[e]; }
                } else {
                    for (int e = lane; e < kk; e += 32) { const int rr = e / k; Ls[rr * ks + (e - rr * k)] = L[e]; }
                }
                const float bold = b_i[i];
                float y[NCH];
#pragma unroll
                for (int t = 0; t < NCH; ++t) {
                    const int c = lane + 32 * t;
                    y[t] = c < k ? ax[c] - bold * ax[k + c] : 0.0f;
                }
                // b += alpha S[i,:] V: the lanes fetch 32 (column, weight) pairs at once, then the 32 rows of V are read
                // with independent loads (eight in flight) — a row of S holds up to ~500 neighbours on the bundled
                // genres graph, and one dependent L2 round trip per neighbour made the widest row the whole level's time.
                // Same summation order as the edge-by-edge loop.
                const int64_t e1 = sptr[i + 1];
                for (int64_t e0 = sptr[i]; e0 < e1; e0 += 32) {
                    const int64_t eme = e0 + lane;
                    const int32_t jl = eme < e1 ? sidx[eme] : 0;
                    const float sl = eme < e1 ? alpha * sval[eme] : 0.0f;
                    const int cnt_e = (int)((e1 - e0) < 32 ? (e1 - e0) : 32);
                    // eight rows of V in flight: all loads of a group are issued before the first is consumed (lanes
                    // past the end of the row carry weight 0 and column 0: no branch inside the group)
#pragma unroll 1
                    for (int u0 = 0; u0 < 32; u0 += 8) {
                        if (u0 >= cnt_e) break;
                        float vv[8][NCH];
#pragma unroll
                        for (int q = 0; q < 8; ++q) {
                            const int32_t j = __shfl_sync(0xffffffffu, jl, u0 + q);
                            const float* vj = V + (size_t)j * ldv;
#pragma unroll
                            for (int t = 0; t < NCH; ++t) {
                                const int c = lane + 32 * t;
                                vv[q][t] = c < k ? __ldcg(vj + c) : 0.0f;
                            }
                        }
#pragma unroll
                        for (int q = 0; q < 8; ++q) {
                            const float sv = __shfl_sync(0xffffffffu, sl, u0 + q);
#pragma unroll
                            for (int t = 0; t < NCH; ++t) y[t] = fmaf(sv, vv[q][t], y[t]);
                        }
                    }
                }
                __syncwarp();
                // reciprocals of the diagonal, once (the divisions leave the substitution chains)
                float dinv[NCH];
#pragma unroll
                for (int t = 0; t < NCH; ++t) {
                    const int c = lane + 32 * t;
                    dinv[t] = c < k ? 1.0f / Ls[c * ks + c] : 0.0f;
                }
                // forward, column-oriented: y_r = y_r / L_rr is final once the columns before r are folded in; its
                // owner broadcasts it and every lane c > r folds column r in (L[c][r]: stride ks, conflict-free).
                // Chain per unknown: multiply, shuffle, multiply-add (the row-oriented form cost a five-shuffle
                // reduction and a division per unknown).
                for (int r = 0; r < k; ++r) {
                    float mine = 0.0f;
#pragma unroll
                    for (int t = 0; t < NCH; ++t)
                        if (t == (r >> 5)) mine = y[t] * dinv[t];
                    const float yr = __shfl_sync(0xffffffffu, mine, r & 31);
#pragma unroll
                    for (int t = 0; t < NCH; ++t) {
                        const int c = lane + 32 * t;
                        if (c == r) y[t] = yr;
                        else if (c > r && c < k) y[t] = fmaf(-Ls[c * ks + r], yr, y[t]);
                    }
                }
                // backward: L^T x = y', column-oriented as well (rank-1 updates with row r of L)
                for (int r = k - 1; r >= 0; --r) {
                    const float* Lr = Ls + r * ks;
                    float mine = 0.0f;
#pragma unroll
                    for (int t = 0; t < NCH; ++t)
                        if (t == (r >> 5)) mine = y[t] * dinv[t];
                    const float xr = __shfl_sync(0xffffffffu, mine, r & 31);
#pragma unroll
                    for (int t = 0; t < NCH; ++t) {
                        const int c = lane + 32 * t;
                        if (c == r) y[t] = xr;
                        else if (c < r) y[t] = fmaf(-Lr[c], xr, y[t]);
                    }
                }
                float dot = 0.0f;
#pragma unroll
                for (int t = 0; t < NCH; ++t) {
                    const int c = lane + 32 * t;
                    if (c < k) {
                        V[(size_t)i * ldv + c] = y[t];
                        dot = fmaf(ax[k + c], y[t], dot);
                    }
                }
                dot = warp_sum(dot);
                if (lane == 0) {
                    const float lb = lam_bias_row ? lam_bias_row[i] : lam_bias;
                    b_i[i] = (ax[2 * k] - dot) / (ax[2 * k + 1] + lb);
                }
                __syncwarp();                       // Ls is reused by this warp's next item
            }
        }
        __threadfence();
        grid.sync();                                // level lv's rows of V are visible to level lv + 1
    }
}

template <int NCH>
static int launch_graph_sweep(int k, const int64_t* level_ptr, int n_levels, int64_t max_level, const int32_t* level_rows,
                              const float* chol, const float* aux, const int64_t* sptr, const int32_t* sidx,
                              const float* sval, float alpha, float lam_bias, const float* lam_bias_row, float* V,
                              int64_t ldv, float* b_i, cudaStream_t s) {
    auto kern = graph_sweep_kernel<NCH>;
    const size_t per_warp = sizeof(float) * (size_t)k * (size_t)(k | 1);
    int wpc = (int)((200 * 1024) / per_warp);
    if (wpc > 8) wpc = 8;
    if (wpc < 1) {
        set_error("graph_sweep: k = %d does not fit a factor in shared memory", k);
        return B200ALS_E_UNSUPPORTED;
    }
    const size_t smem = per_warp * wpc;
    ALS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    ALS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 256, smem));
    if (per_sm < 1) {
        set_error("graph_sweep: kernel does not fit on an SM");
        return B200ALS_E_UNSUPPORTED;
    }
    int64_t grid = ceil_div64(max_level, wpc);            // no more CTAs than the widest level can use
    if (grid > (int64_t)kNumSMs * per_sm) grid = (int64_t)kNumSMs * per_sm;
    if (grid < 1) grid = 1;
    void* args[] = {&k, &level_ptr, &n_levels, &level_rows, &chol, &aux, &sptr, &sidx, &sval, &alpha, &lam_bias,
                    &lam_bias_row, &V, &ldv, &b_i, &wpc};
    ALS_CUDA(cudaLaunchCooperativeKernel((const void*)kern, dim3((unsigned)grid), dim3(256), args, smem, s));
    g_launch_count++;
    return B200ALS_OK;
}

}  // namespace b200als

using namespace b200als;

extern "C" {

int b200als_step_workspace_bytes(int32_t k, int32_t path, int64_t n_f_rows, size_t* bytes_host) {
    ALS_CHECK_ARG(bytes_host, "null");
    ALS_CHECK_ARG(k >= 1 && k <= B200ALS_MAX_K, "k out of range");
    ALS_CHECK_ARG(n_f_rows >= 0, "negative row count");
    *bytes_host = (path == B200ALS_PATH_FFMA) ? 256 : umma_workspace_bytes(umma_padded_k(k), n_f_rows);
    return B200ALS_OK;
}

int b200als_last_path(void) { return g_last_path; }

int b200als_split_default(int32_t which) { return which == 0 ? B200ALS_SPLIT_ROWS_DEFAULT : B200ALS_SPLIT_PIECE_DEFAULT; }

int b200als_split_plan_bytes(int32_t k, int64_t n_list, int64_t nnz, int32_t split_rows, int32_t split_piece,
                             int64_t n_long, size_t* bytes_host) {
    ALS_CHECK_ARG(bytes_host, "null");
    ALS_CHECK_ARG(k >= 1 && k <= 128, "k out of range (the tcgen05 path covers k <= 128)");
    ALS_CHECK_ARG(n_list >= 0 && n_list < (1ll << 31) && nnz >= 0, "count out of range");
    k = umma_padded_k(k);
    SplitLayout L;
    split_layout(k, n_list, nnz, split_rows, split_piece, n_long, 256, L);
    *bytes_host = L.total;
    return B200ALS_OK;
}

int b200als_split_plan_build(const int64_t* indptr, const int32_t* rows, int64_t n_list, int64_t nnz, int32_t k,
                             int32_t split_rows, int32_t split_piece, int64_t n_long, void* plan_mem, size_t plan_bytes,
                             b200als_split_plan* plan_host, b200als_stream_t stream) {
    ALS_CHECK_ARG(indptr && plan_host, "null");
    ALS_CHECK_ARG(k >= 1 && k <= 128, "k out of range (the tcgen05 path covers k <= 128)");
    ALS_CHECK_ARG(n_list >= 0 && n_list < (1ll << 31) && nnz >= 0, "count out of range");
    k = umma_padded_k(k);
    memset(plan_host, 0, sizeof(*plan_host));
    plan_host->n_list = n_list;
    SplitLayout L;
    if (!split_layout(k, n_list, nnz, split_rows, split_piece, n_long, 256, L)) return B200ALS_OK;   // nothing can be split
    if (!plan_mem || plan_bytes < L.total) {
        set_error("b200als_split_plan_build: plan memory %zu < required %zu", plan_bytes, L.total);
        return B200ALS_E_WORKSPACE;
    }
    return split_build(indptr, rows, n_list, L, static_cast<char*>(plan_mem), plan_host, as_stream(stream));
}

int b200als_row_step(const b200als_step_args* a, b200als_stream_t stream) {
    g_launch_count = 0;
    ALS_CHECK_ARG(a && a->struct_bytes == sizeof(b200als_step_args), "args struct size mismatch (ABI)");
    ALS_CHECK_ARG(a->k >= 1 && a->k <= B200ALS_MAX_K, "k out of range [1,160]");
    ALS_CHECK_ARG(a->indptr && a->indices && a->vals && a->F && a->bias_other && a->mu, "null input");
    ALS_CHECK_ARG(a->bias_own, "null bias_own");
    ALS_CHECK_ARG((a->ldf & 3) == 0 && a->ldf >= ((a->k + 3) & ~3), "ldf must be a multiple of 4 and >= round4(k)");
    ALS_CHECK_ARG(a->n_list >= 0 && a->n_list < (1ll << 31), "n_list out of range");
    if (a->mode == B200ALS_STEP_SOLVE) {
        ALS_CHECK_ARG(a->X && a->ldx >= a->k, "null X / ldx < k");
    } else if (a->mode == B200ALS_STEP_FACTOR) {
        ALS_CHECK_ARG(a->chol_out && a->aux_out, "FACTOR mode needs chol_out and aux_out");
    } else {
        ALS_CHECK_ARG(false, "unknown mode");
    }
    int path = a->path;
    // FACTOR mode (graph sweep) stores the Cholesky factor from the tile solver: FFMA path only
    if (path != B200ALS_PATH_F64 && (path == B200ALS_PATH_AUTO || a->mode == B200ALS_STEP_FACTOR))
        path = (a->mode == B200ALS_STEP_SOLVE && umma_preferred(a->k, a->ldf)) ? B200ALS_PATH_UMMA : B200ALS_PATH_FFMA;
    g_last_path = path;
    size_t need = 256;
    if (path == B200ALS_PATH_UMMA) {
        ALS_CHECK_ARG(a->n_f_rows > 0, "n_f_rows must be set for the tcgen05 path");
        need = umma_workspace_bytes((int)a->ldf, a->n_f_rows);
    }
    ALS_CHECK_ARG(a->n_peers >= 0 && a->n_peers <= B200ALS_MAX_PEERS, "n_peers out of range");
    ALS_CHECK_ARG(a->n_peers == 0 || a->mode == B200ALS_STEP_SOLVE, "peer replicas only in SOLVE mode");
    if (!a->workspace || a->workspace_bytes < need) {
        set_error("b200als_row_step: workspace %zu < required %zu", a->workspace_bytes, need);
        return B200ALS_E_WORKSPACE;
    }
    if (a->n_list == 0) return B200ALS_OK;
    cudaStream_t s = as_stream(stream);
    RowStepParams p;
    p.k = a->k; p.mode = a->mode; p.n_list = a->n_list;
    p.indptr = a->indptr; p.indices = a->indices; p.vals = a->vals; p.rows = a->rows;
    p.F = a->F; p.ldf = a->ldf; p.bias_other = a->bias_other; p.mu = a->mu;
    p.reg = a->reg; p.reg_row = a->reg_row; p.lam_bias = a->lam_bias; p.lam_bias_row = a->lam_bias_row;
    p.X = a->X; p.ldx = a->ldx; p.bias_own = a->bias_own;
    p.chol_out = a->chol_out; p.aux_out = a->aux_out; p.status = a->status;
    p.stats = (a->mode == B200ALS_STEP_SOLVE) ? a->stats_acc : nullptr;
    p.norms = (a->mode == B200ALS_STEP_SOLVE) ? a->norm_acc : nullptr;
    p.cursor = reinterpret_cast<unsigned long long*>(a->workspace);
    p.items = nullptr; p.split_hdr = nullptr; p.slot_np = nullptr; p.slot_ctr = nullptr; p.slot_acc = nullptr;
    p.row_ids = a->row_ids;
    p.n_peers = a->n_peers;
    for (int j = 0; j < B200ALS_MAX_PEERS; ++j) {
        p.peer_X[j] = j < a->n_peers ? a->peer_X[j] : nullptr;
        p.peer_bias[j] = j < a->n_peers ? a->peer_bias[j] : nullptr;
        ALS_CHECK_ARG(j >= a->n_peers || (p.peer_X[j] && p.peer_bias[j]), "null peer replica");
    }
    ALS_CUDA(cudaMemsetAsync(a->workspace, 0, 256, s));

    if (path == B200ALS_PATH_UMMA) {
        if (!umma_supported(a->k, a->ldf) || a->ldx < a->ldf) {
            set_error("b200als_row_step: tcgen05 path needs ldf in {64,128}, k <= ldf, zero padding in F and ldx >= ldf "
                      "(got k=%d ldf=%lld ldx=%lld)", a->k, (long long)a->ldf, (long long)a->ldx);
            return B200ALS_E_UNSUPPORTED;
        }
        return launch_umma(p, a->n_f_rows, a->split, static_cast<char*>(a->workspace), a->workspace_bytes, s);
    }
    if (path == B200ALS_PATH_F64) return launch_f64(p, s);
    return dispatch_ffma(p, s);
}

#ifdef B200ALS_TIMING
int b200als_debug_floats(float* out64_host) {
    return cudaMemcpyFromSymbol(out64_host, b200als::g_dbg, 64 * sizeof(float)) == cudaSuccess ? 0 : -2;
}
int b200als_debug_timing(unsigned long long* out16_host) {
    return cudaMemcpyFromSymbol(out16_host, b200als::g_timing, 16 * sizeof(unsigned long long)) == cudaSuccess ? 0 : -2;
}
#endif

int b200als_graph_level_solve(int32_t k, const int32_t* level_rows, int64_t n_level, const float* chol,
                              const float* aux, const int64_t* sptr, const int32_t* sidx, const float* sval,
                              float alpha, float lam_bias, const float* lam_bias_row, float* V, int64_t ldv,
                              float* b_i, b200als_stream_t stream) {
    ALS_CHECK_ARG(k >= 1 && k <= B200ALS_MAX_K, "k out of range");
    ALS_CHECK_ARG(level_rows && chol && aux && sptr && sidx && sval && V && b_i, "null pointer");
    if (n_level <= 0) return B200ALS_OK;
    cudaStream_t s = as_stream(stream);
    graph_level_kernel<B200ALS_MAX_K><<<(unsigned)ceil_div64(n_level, 4), 128, 0, s>>>(
        k, level_rows, n_level, chol, aux, sptr, sidx, sval, alpha, lam_bias, lam_bias_row, V, ldv, b_i);
    ALS_LAUNCH_CHECK();
    return B200ALS_OK;
}

int b200als_graph_sweep(int32_t k, const int64_t* level_ptr, int32_t n_levels, int64_t max_level_rows,
                        const int32_t* level_rows, const float* chol, const float* aux, const int64_t* sptr,
                        const int32_t* sidx, const float* sval, float alpha, float lam_bias, const float* lam_bias_row,
                        float* V, int64_t ldv, float* b_i, b200als_stream_t stream) {
    g_launch_count = 0;
    ALS_CHECK_ARG(k >= 1 && k <= B200ALS_MAX_K, "k out of range");
    ALS_CHECK_ARG(level_ptr && level_rows && chol && aux && sptr && sidx && sval && V && b_i, "null pointer");
    if (n_levels <= 0) return B200ALS_OK;
    cudaStream_t s = as_stream(stream);
    const int nch = (k + 31) / 32;
#define SWEEP_CASE(N) case N: return launch_graph_sweep<N>(k, level_ptr, n_levels, max_level_rows, level_rows, chol, aux, sptr, \
                                                           sidx, sval, alpha, lam_bias, lam_bias_row, V, ldv, b_i, s)
    switch (nch) {
        SWEEP_CASE(1); SWEEP_CASE(2); SWEEP_CASE(3); SWEEP_CASE(4); SWEEP_CASE(5);
    }
#undef SWEEP_CASE
    set_error("graph_sweep: k out of range");
    return B200ALS_E_UNSUPPORTED;
}

}  // extern "C"

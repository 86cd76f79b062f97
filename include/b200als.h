/*
 * b200als.h — C-ABI of libb200als.so: the B200 (sm_100a) ALS block-update path.
 *
 * Drop-in boundary for the hot path of zhukovanadezhda/collaborative-filtering
 * `scripts/als.py` (ALS.fit loop, als.py:408-523, and ALS.predict, als.py:532-574).
 * The reference is pure Python/numpy, so its "FFI for this path" is the set of
 * numpy expressions listed per entry point below; each function names the
 * reference lines it replaces.  INTEGRATION.md shows the ctypes binding.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer unless its name ends in `_host`;
 *  - the caller owns all memory; the library allocates nothing persistent.
 *    Entry points that need scratch take (workspace, workspace_bytes); the
 *    matching *_workspace_bytes call says how much;
 *  - every call is asynchronous on `stream` (a cudaStream_t) unless stated;
 *  - return value: 0 = ok, <0 = error (B200ALS_E_*); b200als_last_error() gives
 *    the thread-local message.  No C++ exception crosses the boundary;
 *  - factor matrices are row-major float32 with a row stride (`ld*`, in
 *    floats, multiple of 4 so rows are 16-byte aligned); index arrays are
 *    int32, row pointers int64 (nnz may exceed 2^31 across a job);
 *  - a device `int32 status[B200ALS_STATUS_WORDS]` word array collects
 *    numerical failures (non-SPD pivot → row id) without a host sync.
 */
#ifndef B200ALS_H_
#define B200ALS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200ALS_ABI_VERSION 8

#define B200ALS_OK            0
#define B200ALS_E_INVALID    -1   /* bad argument */
#define B200ALS_E_CUDA       -2   /* CUDA runtime error (message has the string) */
#define B200ALS_E_WORKSPACE  -3   /* workspace too small */
#define B200ALS_E_UNSUPPORTED -4  /* shape outside what the kernels cover */

#define B200ALS_MAX_K        160  /* tune_params.py:99 N_FACTORS_MAX = 150 */
#define B200ALS_MAX_FEATURES 8
#define B200ALS_STATUS_WORDS 4    /* [0]=#non-SPD rows, [1]=first bad row, [2]=phase, [3]=reserved */
#define B200ALS_MAX_PEERS    7    /* other ranks of an 8-GPU box whose replicas a half-step stores into */

typedef void* b200als_stream_t;   /* cudaStream_t */

int         b200als_abi_version(void);
const char* b200als_last_error(void);
/* 1 if a usable sm_100 device is current, 0 otherwise (sync; never throws). */
int         b200als_device_ok(void);
/* Let kernels on the current device store into memory of `peer_device` (NVLink peer access; idempotent).  Needed
 * before b200als_step_args.peer_X / peer_bias point at another GPU's replicas. */
int         b200als_peer_enable(int32_t peer_device);
/* Replicas of the other ranks of a box for the fused exchange (peer_X / peer_bias of b200als_step_args); no counterpart
 * in the reference, which is one process.  export: the 64-byte CUDA IPC handle of the cudaMalloc block that holds
 * dev_ptr and dev_ptr's offset inside it (the owner sends both to the other ranks by any means).  import: the address
 * of that memory in THIS process, mapped with the CURRENT device's context (cudaIpcMemLazyEnablePeerAccess) so that
 * kernels of the current device can store into it; opens are cached and reference counted per (device, handle).
 * release: drops one reference (the mapping is closed with the last one).  The owner must keep the memory allocated
 * until every importer has released it. */
int         b200als_ipc_export(const void* dev_ptr, void* handle64_host, int64_t* offset_host);
int         b200als_ipc_import(const void* handle64_host, int64_t offset, void** ptr_out_host);
int         b200als_ipc_release(const void* handle64_host);

/* ------------------------------------------------------------------------
 * A1 — CSR / CSC construction (als.py:335-340), index-bit-exact:
 *   CSR  indices[indptr[u] : indptr[u+1]] == np.flatnonzero(mask[u])      (ascending)
 *   CSC cindices[cindptr[i]:cindptr[i+1]] == np.flatnonzero(mask[:, i])   (ascending)
 *   COO order of (rows_u, rows_i) = np.where(mask)  == CSR order
 *   perm[p_csc] = p_csr  (optional, may be NULL)
 * ---------------------------------------------------------------------- */

/* Dense input: R is (m, n) float64 row-major, NaN = missing (als.py:335).
 * Step 1 writes indptr[m+1] and cindptr[n+1]; nnz = indptr[m] (caller reads it
 * back, allocates, then calls step 2). */
int b200als_dense_indptr(const double* R, int64_t m, int64_t n,
                         int64_t* indptr, int64_t* cindptr,
                         b200als_stream_t stream);
int b200als_dense_fill(const double* R, int64_t m, int64_t n,
                       const int64_t* indptr, const int64_t* cindptr,
                       int32_t* indices, float* vals,
                       int32_t* cindices, float* cvals, int64_t* perm,
                       b200als_stream_t stream);

/* COO input: flat_idx[e] = u*n + i (create_folds.py:73-84 convention), unique,
 * any order unless `sorted` != 0 (then ascending, as np.flatnonzero and
 * np.setdiff1d give it, create_folds.py:200).  Replaces the dense
 * matrix_from_indices round trip (create_folds.py:152-174 + als.py:335-340).
 * The input is validated on the device before anything is built from it (one pass): *status3 (device, 3
 * int64, may be NULL) receives [0] the number of indices outside [0, m*n) — the reference's
 * matrix_from_indices raises IndexError on those (create_folds.py:169-174) —, [1] the number of adjacent
 * pairs of the (sorted) sequence that are not strictly increasing, i.e. duplicates, or a `sorted` input that
 * is not, [2] the number of non-finite ratings (NaN means "missing" in the reference's dense form and has no
 * meaning in a list of observed entries).  Out-of-range indices are clamped while building, so nothing is
 * written outside the outputs; the caller reads the status and rejects the arrays if any count is non-zero. */
int b200als_coo_workspace_bytes(int64_t nnz, int64_t m, int64_t n, size_t* bytes_host);
int b200als_coo_to_csr_csc(const int64_t* flat_idx, const float* vals_in,
                           int64_t nnz, int64_t m, int64_t n, int sorted,
                           int64_t* indptr, int32_t* indices, float* vals,
                           int64_t* cindptr, int32_t* cindices, float* cvals,
                           int64_t* perm,
                           void* workspace, size_t workspace_bytes,
                           int64_t* status3,
                           b200als_stream_t stream);

/* Row schedule: order[] = row ids sorted by descending rating count (longest
 * first, ties by ascending id); empty rows are placed last and *n_nonempty
 * (device int64) receives how many rows have >=1 rating (als.py:418-419,
 * 440-441 skip the others). */
int b200als_row_order_workspace_bytes(int64_t n_rows, size_t* bytes_host);
int b200als_row_order(const int64_t* indptr, int64_t n_rows,
                      int32_t* order, int64_t* n_nonempty,
                      void* workspace, size_t workspace_bytes,
                      b200als_stream_t stream);

/* b200als_row_order without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same outputs as b200als_row_order; the scratch (sort keys, row ids, radix-sort temporary storage) is a
 * file-scope `static __device__` array inside the library: 7 MiB, held by every GPU that loads the library until
 * the process exits.  Limit: n_rows <= B200ALS_ROW_ORDER_STATIC_MAX_ROWS (262 144), larger inputs are rejected with
 * B200ALS_E_UNSUPPORTED (use b200als_row_order with a caller-owned workspace).  The call itself performs no
 * cudaMalloc / cudaFree / allocating library call.  Every call shares the array (it keeps the last call's contents;
 * nothing in it needs to start at zero, *n_nonempty is cleared by the call): calls that overlap on the GPU — two
 * streams, two host threads — RACE on it and must be serialised by the caller. */
#define B200ALS_ROW_ORDER_STATIC_MAX_ROWS (1 << 18)
int64_t b200als_row_order_static_max_rows(void);
int b200als_row_order_static(const int64_t* indptr, int64_t n_rows,
                             int32_t* order, int64_t* n_nonempty,
                             b200als_stream_t stream);

/* ------------------------------------------------------------------------
 * A5 — Z = V + sum_f X_f W_f  (als.py:262-281, 411, 504, 568-572)
 * X_f is (n, d_f) float32 row-major, W_f is (d_f, k) float32 row-major with
 * row stride ldw.  `skip` >= 0 leaves feature `skip` out (W step, als.py:485).
 * ---------------------------------------------------------------------- */
int b200als_compose_z(const float* V, int64_t ldv, int64_t n, int32_t k,
                      int32_t n_features,
                      const float* const* X_host_array_of_dev_ptrs,
                      const int32_t* d_host,
                      const float* const* W_host_array_of_dev_ptrs,
                      int64_t ldw, int32_t skip,
                      float* Z, int64_t ldz,
                      b200als_stream_t stream);

/* ------------------------------------------------------------------------
 * A6 / A7 / A8 / A11 — one ALS half-step over a set of rows.
 *
 * For every row r in `rows` (or 0..n_rows-1 when rows == NULL) with cnt > 0:
 *   idx  = indices[indptr[r] : indptr[r+1]]
 *   q_j  = vals_j - (mu + bias_other[idx_j])
 *   A    = F[idx]^T F[idx] + reg_r I                       (als.py:426 / 455)
 *   b    = F[idx]^T (q - bias_own[r])  (+ rhs_extra[r])    (als.py:427 / 456-458)
 *   x    = cholesky_solve(A, b)                            (helpers.py:19-20)
 *   bias_own[r] = sum_j (q_j - F[idx_j]·x) / (cnt + lam_bias_r)   (als.py:431-433 / 464-466)
 * with reg_r = reg_row ? reg_row[r] : reg   (EPS, alpha*D_i already folded in
 * by the caller in float64, als.py:426,450-454) and likewise lam_bias.
 * User step: F = Z, bias_other = b_i, X = U.  Item step: F = U, bias_other =
 * b_u, X = V.  Rows with no ratings are left untouched.
 *
 * mode B200ALS_STEP_SOLVE   : as above.
 * mode B200ALS_STEP_FACTOR  : (graph sweep, A8; always runs on the FFMA path) do not solve; store the lower
 *     Cholesky factor of A in chol_out[r] (k*k floats, row-major, ld = k) and
 *     c_r = F^T q  (k), s_r = F^T 1 (k), t_r = sum q, in aux_out[r] (2k+2
 *     floats: c, s, t, cnt) for b200als_graph_level_solve.
 * ---------------------------------------------------------------------- */
#define B200ALS_STEP_SOLVE  0
#define B200ALS_STEP_FACTOR 1

#define B200ALS_PATH_AUTO   0   /* tcgen05 when ldf in {64,128} and ldf/2 < k <= ldf (zero-padded rows), else FFMA */
#define B200ALS_PATH_FFMA   1
#define B200ALS_PATH_UMMA   2
#define B200ALS_PATH_F64    3   /* normal matrix accumulated, factored and solved in float64 (any k <= 160, both modes):
                                   for rows whose ridge is too small for float32 — fewer ratings than factors and
                                   lambda below ~0.1, a quarter of the tuner's range (tune_params.py:102-103) — and
                                   for callers who want the reference's arithmetic; never chosen by AUTO */

/* ------------------------------------------------------------------------
 * Long rows (tcgen05 path).  One SM consumes a row's ratings 32 per pipeline stage, so a very popular
 * item (als.py:436-441 visits it like any other) would bound the whole half-step.  A split plan cuts
 * rows with more than split_rows ratings into pieces of about split_piece ratings that different SMs
 * accumulate: the partial normal matrices [A | c s] are added into a per-row slot in global memory and
 * the SM adding the last piece factors and solves — same sums, different summation order.
 *   split_rows: 0 = default (rows of more than 6144 ratings, pieces of 4096), < 0 = never split
 *   split_piece: ratings per piece, rounded up to 32; 0 = split_rows / 2
 * The defaults are ABSOLUTE lengths on purpose.  The tensor core adds into its float32 accumulator with
 * truncation: a chain of n accumulation steps (16 ratings each) biases the sums low by ~n 2^-24 relative — measured
 * on B200 at the Netflix shape against the float64 oracle: 1.4e-3 ... 2.1e-3 relative error of the solved rows of
 * 3.4e5 ... 4.8e5-rating items accumulated in one chain, 2.5e-5 with 8192-rating pieces, 6e-6 with 2048
 * (profiles/r02_v1_scale_parity_c4*.json).  A fixed piece bounds the chain and makes the summation tree a function
 * of the row alone, not of how many ranks share the launch.
 *   nnz: ratings held by the listed rows (bounds the number of slots)
 * plan_mem is caller-owned device memory of b200als_split_plan_bytes(); *plan_host describes it.
 * ---------------------------------------------------------------------- */
typedef struct b200als_split_plan {
    const void*    items;        /* int4 {row, first rating (offset in the row), ratings, slot or -1} */
    const int32_t* hdr;          /* device: [0] items, [1] split rows */
    const int32_t* slot_np;      /* pieces per split row */
    int32_t*       slot_ctr;     /* pieces accumulated so far (zeroed by every step) */
    float*         slot_acc;     /* per slot: (K + 2) x (K + 4) floats [A (stored transposed: it is symmetric); c; s], then t, qq, -, -
                                    (K = padded width 64 / 128; zeroed by every step) */
    int64_t        n_list;       /* list entries the plan was built for */
    int32_t        slot_floats, thresh, piece, reserved;
} b200als_split_plan;
/* n_long: how many listed rows hold more than split_rows ratings (sizes the accumulator slots exactly), or -1 for
 * the bound nnz / split_rows.  b200als_split_default(0 | 1) returns the default split_rows | split_piece. */
int b200als_split_default(int32_t which);
int b200als_split_plan_bytes(int32_t k, int64_t n_list, int64_t nnz, int32_t split_rows, int32_t split_piece,
                             int64_t n_long, size_t* bytes_host);
/* returns 0 and fills *plan_host; plan_host->items == NULL when no row can exceed the threshold */
int b200als_split_plan_build(const int64_t* indptr, const int32_t* rows, int64_t n_list, int64_t nnz, int32_t k,
                             int32_t split_rows, int32_t split_piece, int64_t n_long, void* plan_mem, size_t plan_bytes,
                             b200als_split_plan* plan_host, b200als_stream_t stream);

typedef struct b200als_step_args {
    uint32_t struct_bytes;       /* = sizeof(b200als_step_args) */
    int32_t  k;
    int32_t  mode;
    int32_t  path;
    int64_t  n_rows;             /* rows of the sparse matrix (bounds for rows[]) */
    const int64_t* indptr;
    const int32_t* indices;
    const float*   vals;
    const int32_t* rows;         /* optional row list (e.g. b200als_row_order output) */
    int64_t  n_list;             /* entries of rows[] to process (or n_rows if rows==NULL) */
    const float* F;  int64_t ldf;
    int64_t  n_f_rows;           /* rows of F (every index < n_f_rows); sizes the tcgen05 path's bf16 copy */
    const float* bias_other;
    const float* mu;             /* device scalar */
    float  reg;        const float* reg_row;
    float  lam_bias;   const float* lam_bias_row;
    float* X;  int64_t ldx;
    float* bias_own;
    float* chol_out;             /* FACTOR mode: n_rows * k * k */
    float* aux_out;              /* FACTOR mode: n_rows * (2k+2) */
    int32_t* status;             /* B200ALS_STATUS_WORDS ints, may be NULL */
    double*  stats_acc;          /* optional, SOLVE mode: A10 fused into the step (als.py:503-511 when Z == F-side
                                    factors are final): [0] += sum_e d_e, [1] += sum_e d_e^2 over the rows solved,
                                    d_e = R_e - (mu + bias_other + bias_own_new + F[idx_e].x_new).  Caller zeroes. */
    double*  norm_acc;           /* optional, SOLVE mode: the Frobenius norms of als.py:513-516 fused into the step:
                                    [0] += sum over the rows solved of |x_new|^2, [2] += sum of bias_own_new^2
                                    (stride 2: acc6 + 2 for the user step, acc6 + 3 for the item step fill the
                                    layout b200als_finalize_iter reads).  Rows not in the list keep their values:
                                    the caller adds their (constant) share. */
    const struct b200als_split_plan* split;   /* host pointer, optional (tcgen05 path, SOLVE mode): the plan built by
                                    b200als_split_plan_build for exactly this (indptr, rows, n_list) */
    void*  workspace; size_t workspace_bytes;   /* see b200als_step_workspace_bytes */
    const int32_t* row_ids;      /* optional (slab storage): indptr / rows[] are local to the caller's slab of rows and
                                    row_ids[r] is local row r's id in X, bias_own, reg_row and lam_bias_row */
    int32_t n_peers;             /* fused exchange (multi-GPU, SOLVE mode): the solved row and its bias are ALSO stored
                                    into peer_X[j] / peer_bias[j], j < n_peers — the replicas of the other ranks, mapped
                                    into this process (CUDA IPC) — so the half-step's all-gather rides on the solver's
                                    own stores over NVLink instead of a separate collective.  Same layout as X/bias_own. */
    int32_t reserved0;
    float* peer_X[B200ALS_MAX_PEERS];
    float* peer_bias[B200ALS_MAX_PEERS];
} b200als_step_args;

/* workspace for one b200als_row_step call gathering from an F of n_f_rows rows */
int b200als_step_workspace_bytes(int32_t k, int32_t path, int64_t n_f_rows, size_t* bytes_host);
int b200als_row_step(const b200als_step_args* args_host, b200als_stream_t stream);
/* Number of kernel launches the last b200als_row_step on this thread issued. */
int b200als_last_launch_count(void);
/* Gram path (B200ALS_PATH_FFMA / B200ALS_PATH_UMMA) the last b200als_row_step on this thread took. */
int b200als_last_path(void);

/* tcgen05 path and k: the kernel exists for K = 64 and K = 128 and covers every k <= K by zero padding — the
 * caller keeps F (and X) with a row stride of exactly K floats and zeros in columns [k, K).  The padded block of
 * the normal matrix is reg * I with a zero right-hand side and decouples: the k leading entries of the solution
 * are those of the k x k system (als.py:426-428), the padded entries come out as exact zeros (and are written).
 * k in 129..160 (tune_params.py:98-99 samples up to 150) stays on the FFMA path: a K = 160 accumulator needs
 * 176 TMEM columns (2 matrices in flight instead of 3) and an 80 KB column store per matrix. */

/* A8 — one level of the Gauss–Seidel Laplacian sweep (als.py:453-461).
 * For every item i in level_rows[0..n_level):
 *   b = c_i - b_i[i]*s_i + alpha * sum_j S[i,j] V[j]      (live V)
 *   V[i] = L_i^-T L_i^-1 b ;  b_i[i] = (t_i - s_i·V[i]) / (cnt_i + lam_bias_i)
 * S is CSR (sptr int64, sidx int32, sval float32). */
int b200als_graph_level_solve(int32_t k, const int32_t* level_rows, int64_t n_level,
                              const float* chol, const float* aux,
                              const int64_t* sptr, const int32_t* sidx, const float* sval,
                              float alpha,
                              float lam_bias, const float* lam_bias_row,
                              float* V, int64_t ldv, float* b_i,
                              b200als_stream_t stream);

/* The whole sweep in ONE cooperative launch: level_ptr (device int64 [n_levels + 1]) delimits the levels inside
 * level_rows; a grid barrier separates them.  max_level_rows = rows of the widest level (sizes the grid).  Same
 * arithmetic as n_levels calls of b200als_graph_level_solve in order. */
int b200als_graph_sweep(int32_t k, const int64_t* level_ptr, int32_t n_levels, int64_t max_level_rows,
                        const int32_t* level_rows, const float* chol, const float* aux,
                        const int64_t* sptr, const int32_t* sidx, const float* sval,
                        float alpha, float lam_bias, const float* lam_bias_row,
                        float* V, int64_t ldv, float* b_i, b200als_stream_t stream);

/* ------------------------------------------------------------------------
 * A10 — residual statistics over all observed entries (als.py:503-517):
 *   d_e = R_e - (U[u]·Z[i] + b_u[u] + b_i[i]);  acc[0] += sum d, acc[1] += sum d^2
 * in float64.  The caller zeroes acc.  b200als_sumsq adds sum x^2 (Frobenius
 * norms, als.py:514-517) over an (rows, cols, ld) matrix into *acc.
 * b200als_finalize_iter turns them into mu / rmse / norms on the device:
 *   mu = acc[0]/nnz (als.py:507)  → *mu_f32 (read by the next half-steps)
 *   hist[0..5] = {rmse, |U|, |V|, |b_u|, |b_i|, mu} float64 (als.py:510-517)
 * ---------------------------------------------------------------------- */
int b200als_residual_stats(int64_t m, const int64_t* indptr, const int32_t* indices,
                           const float* vals, int32_t k,
                           const float* U, int64_t ldu, const float* Z, int64_t ldz,
                           const float* b_u, const float* b_i,
                           double* acc, b200als_stream_t stream);
int b200als_sumsq(const float* x, int64_t rows, int64_t cols, int64_t ld,
                  double* acc, b200als_stream_t stream);
int b200als_finalize_iter(const double* acc6, int64_t nnz_total,
                          float* mu_f32, double* hist6, b200als_stream_t stream);
/* Same, when acc6[0..1] came from b200als_step_args.stats_acc of the item step, i.e. are sums of
 * d'_e = d_e - mu_old (the step subtracts the mu it ran with):  mu = mu_old + acc[0]/nnz,
 * rmse^2 = acc[1]/nnz - (acc[0]/nnz)^2.  *mu_f32 is read (mu_old) and overwritten.
 * reset6 (optional): acc6 <- reset6 afterwards, the start values of the next iteration's
 * accumulators (zeros, plus the norm share of the rows no step solves). */
int b200als_finalize_iter_fused(double* acc6, const double* reset6, int64_t nnz_total,
                                float* mu_f32, double* hist6, b200als_stream_t stream);

/* ------------------------------------------------------------------------
 * A9 — W_f update (als.py:468-501) in Kronecker form, float64:
 *   G_i = sum_{u in Omega_i} U_u U_u^T,  c_i = sum_u (R_ui - mu - b_u - b_i) U_u
 *   h_i = c_i - G_i (Z_i - X_f[i] W_f)            (own contribution added back)
 *   A   = sum_i (x_i x_i^T) (x) G_i + lam I,  b = sum_i x_i (x) h_i,  index a*k+c
 *   vec(W_f) = cholesky_solve(A, b)
 * b200als_item_gram_f64 fills G (n,k,k) and c (n,k);
 * b200als_wstep_system builds A (dk x dk, lower+upper) and b (dk);
 * b200als_cholesky_solve_f64 solves in place (x overwrites b).
 * ---------------------------------------------------------------------- */
int b200als_item_gram_f64(int64_t n, const int64_t* cindptr, const int32_t* cindices,
                          const float* cvals, int32_t k,
                          const float* U, int64_t ldu,
                          const float* b_u, const float* b_i, const float* mu,
                          double* G, double* c, b200als_stream_t stream);
int b200als_wstep_system(int64_t n, int32_t k, int32_t d,
                         const float* X, const double* G, const double* c,
                         const float* Zmf, int64_t ldz, double lam,
                         double* A, double* b, b200als_stream_t stream);
int b200als_cholesky_solve_f64(double* A, double* b, int32_t N, int32_t* status,
                               b200als_stream_t stream);
/* W_f (float32, ld = ldw) <- vec (float64, d*k) ; also the reverse for upload */
int b200als_f64_to_f32(const double* src, float* dst, int64_t count, b200als_stream_t stream);

/* ------------------------------------------------------------------------
 * A13 — predictions (als.py:574): out[u, i] = U[u]·Z[i] + mu + b_u[u] + b_i[i]
 * dense (m, n) float64, or at (rows, cols) pairs (tune_params.py:165-167).
 * ---------------------------------------------------------------------- */
int b200als_predict_dense(const float* U, int64_t ldu, const float* Z, int64_t ldz,
                          const float* b_u, const float* b_i, const float* mu,
                          int64_t m, int64_t n, int32_t k,
                          double* out, b200als_stream_t stream);
int b200als_predict_pairs(const float* U, int64_t ldu, const float* Z, int64_t ldz,
                          const float* b_u, const float* b_i, const float* mu,
                          const int32_t* rows, const int32_t* cols, int64_t count, int32_t k,
                          float* out, b200als_stream_t stream);

/* ------------------------------------------------------------------------
 * N2 — `make_train_valid_split` (create_folds.py:177-208) on the resident observed
 * set, no dense matrices.  flat_sorted: the ascending flat indices u*n+i of every
 * observed entry (`flatnonzero(~isnan(R))`, create_folds.py:73-76), vals their ratings.
 * train_flat / train_vals (capacity nnz) receive, in ascending order, the entries
 * not listed in val_idx — the sorted `setdiff1d` of create_folds.py:200 —, *n_train
 * (device) their number; val_vals[j] = rating at val_idx[j] (NaN when val_idx[j] is
 * not an observed entry).  status (device, 2 ints): [0] validation indices that are
 * not observed entries, [1] validation indices listed more than once.
 * ---------------------------------------------------------------------- */
int b200als_fold_split_workspace_bytes(int64_t nnz, size_t* bytes_host);
int b200als_fold_split(const int64_t* flat_sorted, const float* vals, int64_t nnz,
                       const int64_t* val_idx, int64_t n_val,
                       int64_t* train_flat, float* train_vals, float* val_vals,
                       int64_t* n_train, int32_t* status,
                       void* workspace, size_t workspace_bytes, b200als_stream_t stream);

/* ------------------------------------------------------------------------
 * N1 — `_rmse_on_indices` (tune_params.py:147-167) and the popularity-bin
 * split of it (evaluate_models.py:173-191, 263-274) without the dense (m, n)
 * prediction: for every flat index f = u*n + i (create_folds.py:73-84),
 *   d = y_true - (U[u]·Z[i] + mu + b_u[u] + b_i[i]),
 *   sse[b] += d^2, cnt[b] += 1,  b = item_bin[i]  (b = 0 when item_bin is NULL).
 * sse / cnt (n_bins entries, n_bins <= 32) are accumulated into: zero them first.
 * ---------------------------------------------------------------------- */
int b200als_sse_on_flat(const float* U, int64_t ldu, const float* Z, int64_t ldz,
                        const float* b_u, const float* b_i, const float* mu,
                        const int64_t* flat_idx, const float* y_true, int64_t count,
                        int64_t n, int32_t k, const int32_t* item_bin, int32_t n_bins,
                        double* sse, int64_t* cnt, b200als_stream_t stream);

/* ------------------------------------------------------------------------
 * N4 — `normalize_feature` (prepare_features.py:131-191) in place on X (n, d) float64 row-major.
 *   impute: 0 none | 1 col_median (NaN / +-Inf -> np.nanmedian of the column, 0 for an all-NaN column; :83-94)
 *   method: 0 none | 1 row_l1 | 2 row_l2 | 3 col_zscore | 4 col_minmax          (:97-128, eps as there)
 * *n_nonfinite (device int64) = NaN / Inf entries of the INPUT (with impute 0 the caller raises the reference's
 * ValueError on a non-zero count).  med_work: d + 1 doubles; workspace from b200als_normalize_workspace_bytes (both
 * only needed with impute 1).
 * ---------------------------------------------------------------------- */
int b200als_normalize_workspace_bytes(int64_t n, size_t* bytes_host);
int b200als_normalize_feature(double* X, int64_t n, int32_t d, int32_t method, int32_t impute, double eps,
                              int64_t* n_nonfinite, double* med_work, void* workspace, size_t workspace_bytes,
                              b200als_stream_t stream);

/* ------------------------------------------------------------------------
 * N3 — cosine similarity + per-row top-k of als.py:224-236 with a DETERMINISTIC tie rule (opt-in; the default
 * graph build stays numpy's on the host because np.argpartition breaks ties in an implementation-defined way).
 *   Xn = X / (||X||_row + eps);  s_ij = Xn_i . Xn_j, s_ii = 0;
 *   row i keeps its topk largest s_ij over all j (the zeroed diagonal takes part, as in the reference);
 *   among entries EQUAL to the topk-th largest value the smaller column index wins.
 * Output: nbr_idx / nbr_val (n x topk), the kept columns of every row in ascending column order.  The caller forms
 * S = max(S_kept, S_kept^T) (als.py:239) from the lists.  Xn_work: n x d floats; scratch: scratch_rows x n floats
 * (one row per resident CTA; 148 * 4 is plenty).  Requires 1 <= topk < n.
 * ---------------------------------------------------------------------- */
int b200als_similarity_topk_rows(const float* X, int64_t n, int32_t d, int32_t topk, float eps, float* Xn_work,
                                 float* scratch, int64_t scratch_rows, int32_t* nbr_idx, float* nbr_val,
                                 b200als_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* B200ALS_H_ */

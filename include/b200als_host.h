/* b200als_host.h — host-side (CPU, no CUDA) helper of the ALS block-update path: libb200als_host.so.
 *
 * b200als_host_normal_fill replaces the reference's factor initialisation draw (scripts/als.py:329, 367-376:
 * `rng = np.random.default_rng(random_state)`, then `rng.normal(scale=0.1, size=...)` for U, V and every W_f,
 * in that order from the one PCG64 stream) with a multi-threaded generator of the SAME values, bit for bit.
 * It links numpy's own `random_standard_normal` (numpy/random/lib/libnpyrandom.a) and only changes how the
 * stream is walked; see collaborative-filtering_b200/csrc/host_rng.c.
 */
#ifndef B200ALS_HOST_H
#define B200ALS_HOST_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* state_* / inc_*: the 128-bit `state` and `inc` of numpy's PCG64
 * (`np.random.default_rng(seed).bit_generator.state["state"]`), as 64-bit halves.
 * Writes counts[s] values `0.0 + scale * standard_normal` to outs[s], s = 0..n_seg-1, consecutively from the one
 * stream; outs[s] is float32 when as_f32[s] != 0 (the float64 value rounded, as `.astype(np.float32)` does),
 * else float64.  n_threads <= 1 walks the stream serially.  raw_words (may be NULL): stream words consumed,
 * reported by the serial walk only (0 otherwise).
 * Returns 0, or -1 when a scratch allocation fails. */
int b200als_host_normal_fill(uint64_t state_hi, uint64_t state_lo, uint64_t inc_hi, uint64_t inc_lo,
                             int32_t n_seg, const int64_t* counts, void* const* outs, const int32_t* as_f32,
                             double scale, int32_t n_threads, uint64_t* raw_words);

/* ---- item-similarity graph set-up (scripts/als.py:220-240, 357; csrc/host_graph.c) ----
 * The per-row top-k stays numpy's argpartition (tie parity, see graph.py); these are the array passes around it.
 * elem_bytes: 4 (float32) or 8 (float64) — S has the feature matrix's dtype.  All return 0, or -1 on bad arguments. */

/* S <- np.maximum(S, S.T) in place, S row-major [n, n]  (scripts/als.py:240). */
int b200als_host_symmetrize_max(void* S, int64_t n, int32_t elem_bytes, int32_t n_threads);

/* row_nnz[r] = count of S[r, :] != 0  (first pass of `np.nonzero(S)`). */
int b200als_host_dense_row_nnz(const void* S, int64_t rows, int64_t cols, int32_t elem_bytes, int64_t* row_nnz,
                               int32_t n_threads);

/* CSR of the nonzeros of S, columns ascending per row (row-major `np.nonzero` order); sptr[rows + 1] is the
 * exclusive prefix sum of row_nnz; sval is float32 (what the device consumes). */
int b200als_host_dense_to_csr(const void* S, int64_t rows, int64_t cols, int32_t elem_bytes, const int64_t* sptr,
                              int32_t* sidx, float* sval, int32_t n_threads);

/* level[i] of every rated row i in the dependency order of the Gauss-Seidel item sweep (scripts/als.py:453-461
 * reads the freshly written V_j of neighbours j < i): 0 if no rated neighbour has a smaller index, else 1 + the
 * deepest such neighbour's level; -1 for unrated rows (rated[i] == 0, skipped by scripts/als.py:440-441). */
int b200als_host_graph_levels(const int64_t* sptr, const int32_t* sidx, const uint8_t* rated, int64_t n,
                              int64_t* level);

/* dst[r, c] = (double)src[r * ld_src + c]: the float64 public copies of the learned float32 arrays
 * (scripts/als.py:404-407 leave U, V, b_u, b_i as float64).  dst is contiguous [rows, cols]. */
int b200als_host_widen_f32(const float* src, int64_t ld_src, double* dst, int64_t rows, int64_t cols,
                           int32_t n_threads);

#ifdef __cplusplus
}
#endif
#endif

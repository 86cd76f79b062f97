// host_graph.c — A4 / A8 host set-up of the item-similarity graph (scripts/als.py:220-240, 357): the parts of it that
// are plain array passes, on all host cores.  The per-row top-k itself stays numpy's `argpartition` (graph.py: its
// choice among tied similarities is implementation-defined, so only numpy's own call gives the reference's graph).
//   b200als_host_symmetrize_max   S <- maximum(S, S^T)   (als.py:240)  in place, tile pairs, exact
//   b200als_host_dense_row_nnz / b200als_host_dense_to_csr   nonzeros of S in row-major order -> CSR for the device
// Plain C + OpenMP; element type float32 or float64 (S follows the feature matrix's dtype).
#include <stdint.h>
#include <stddef.h>

#include "b200als_host.h"

#define TILE 64

#define DEFINE_SYM(NAME, T)                                                                                     \
    static void NAME(T* S, int64_t n, int n_threads) {                                                          \
        const int64_t nt = (n + TILE - 1) / TILE;                                                               \
        _Pragma("omp parallel for schedule(dynamic, 1) num_threads(n_threads)")                                 \
        for (int64_t ti = 0; ti < nt; ++ti) {                                                                   \
            const int64_t i0 = ti * TILE, i1 = i0 + TILE < n ? i0 + TILE : n;                                   \
            for (int64_t tj = ti; tj < nt; ++tj) {                                                              \
                const int64_t j0 = tj * TILE, j1 = j0 + TILE < n ? j0 + TILE : n;                               \
                for (int64_t i = i0; i < i1; ++i) {                                                             \
                    for (int64_t j = (tj == ti ? i + 1 : j0); j < j1; ++j) {                                    \
                        const T a = S[i * n + j], b = S[j * n + i];                                             \
                        /* np.maximum: the larger one, NaN if either is NaN */                                  \
                        const T hi = (a >= b || a != a) ? a : b, lo = (b >= a || b != b) ? b : a;               \
                        S[i * n + j] = hi;                                                                      \
                        S[j * n + i] = lo;                                                                      \
                    }                                                                                           \
                }                                                                                               \
            }                                                                                                   \
        }                                                                                                       \
    }
DEFINE_SYM(sym_f32, float)
DEFINE_SYM(sym_f64, double)

int b200als_host_symmetrize_max(void* S, int64_t n, int32_t elem_bytes, int32_t n_threads) {
    if (n < 0 || (n > 0 && !S) || (elem_bytes != 4 && elem_bytes != 8)) return -1;
    if (n_threads < 1) n_threads = 1;
    if (elem_bytes == 4) sym_f32((float*)S, n, n_threads);
    else sym_f64((double*)S, n, n_threads);
    return 0;
}

int b200als_host_dense_row_nnz(const void* S, int64_t rows, int64_t cols, int32_t elem_bytes, int64_t* row_nnz,
                               int32_t n_threads) {
    if (rows < 0 || cols < 0 || (rows > 0 && (!S || !row_nnz)) || (elem_bytes != 4 && elem_bytes != 8)) return -1;
    if (n_threads < 1) n_threads = 1;
#pragma omp parallel for schedule(static) num_threads(n_threads)
    for (int64_t r = 0; r < rows; ++r) {
        int64_t c = 0;
        if (elem_bytes == 4) {
            const float* s = (const float*)S + r * cols;
            for (int64_t j = 0; j < cols; ++j) c += (s[j] != 0.0f);
        } else {
            const double* s = (const double*)S + r * cols;
            for (int64_t j = 0; j < cols; ++j) c += (s[j] != 0.0);
        }
        row_nnz[r] = c;
    }
    return 0;
}

int b200als_host_dense_to_csr(const void* S, int64_t rows, int64_t cols, int32_t elem_bytes, const int64_t* sptr,
                              int32_t* sidx, float* sval, int32_t n_threads) {
    if (rows < 0 || cols < 0 || cols > INT32_MAX || (rows > 0 && (!S || !sptr)) || (elem_bytes != 4 && elem_bytes != 8))
        return -1;
    if (rows > 0 && sptr[rows] > 0 && (!sidx || !sval)) return -1;
    if (n_threads < 1) n_threads = 1;
#pragma omp parallel for schedule(static) num_threads(n_threads)
    for (int64_t r = 0; r < rows; ++r) {
        int64_t o = sptr[r];
        if (elem_bytes == 4) {
            const float* s = (const float*)S + r * cols;
            for (int64_t j = 0; j < cols; ++j)
                if (s[j] != 0.0f) { sidx[o] = (int32_t)j; sval[o] = s[j]; ++o; }
        } else {
            const double* s = (const double*)S + r * cols;
            for (int64_t j = 0; j < cols; ++j)
                if (s[j] != 0.0) { sidx[o] = (int32_t)j; sval[o] = (float)s[j]; ++o; }
        }
    }
    return 0;
}

// Level of every rated row in the lower-triangular dependency DAG of CSR(S) (graph.py: level_schedule): a rated
// row runs one level after its deepest rated neighbour with a smaller index; unrated rows get -1.  Sequential by
// nature (level[i] needs level[j < i]) and a few hundred thousand steps at most, so one thread.
int b200als_host_graph_levels(const int64_t* sptr, const int32_t* sidx, const uint8_t* rated, int64_t n,
                              int64_t* level) {
    if (n < 0 || (n > 0 && (!sptr || !rated || !level)) || (n > 0 && sptr[n] > 0 && !sidx)) return -1;
    for (int64_t i = 0; i < n; ++i) {
        if (!rated[i]) { level[i] = -1; continue; }
        int64_t lv = -1;
        for (int64_t e = sptr[i]; e < sptr[i + 1]; ++e) {
            const int64_t j = sidx[e];
            if (j < 0 || j >= n) return -1;
            if (j < i && level[j] > lv) lv = level[j];
        }
        level[i] = lv + 1;
    }
    return 0;
}

// float32 [rows, cols] (row stride ld_src elements) -> contiguous float64 [rows, cols]: the public attributes are
// float64 (scripts/als.py:404-407) while the device holds float32.  Exact; used by the chunked read-back (als.py).
int b200als_host_widen_f32(const float* src, int64_t ld_src, double* dst, int64_t rows, int64_t cols,
                           int32_t n_threads) {
    if (rows < 0 || cols < 0 || ld_src < cols || (rows > 0 && cols > 0 && (!src || !dst))) return -1;
    if (n_threads < 1) n_threads = 1;
    const int64_t block = 4096 / (cols > 0 ? cols : 1) + 1;                     // rows per task: at least a page of output
    const int64_t n_blocks = (rows + block - 1) / block;
#pragma omp parallel for schedule(static) num_threads(n_threads)
    for (int64_t b = 0; b < n_blocks; ++b) {
        const int64_t r1 = (b + 1) * block < rows ? (b + 1) * block : rows;
        for (int64_t r = b * block; r < r1; ++r) {
            const float* s = src + r * ld_src;
            double* d = dst + r * cols;
            for (int64_t c = 0; c < cols; ++c) d[c] = (double)s[c];
        }
    }
    return 0;
}

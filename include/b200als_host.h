/* b200als_host.h — host-side (CPU, no CUDA) helper of the ALS block-update path: libb200als_host.so.
 *
 * Replaces the reference's factor initialisation draw (scripts/als.py:329, 367-376:
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

#ifdef __cplusplus
}
#endif
#endif

// host_rng.c — A2 (als.py:329, 367-376): the reference draws U, V, W_f ~ 0.1 N(0, 1) from ONE numpy PCG64 stream
// (`default_rng(random_state).normal(scale=0.1, size=...)`, in that order), and seed parity needs exactly those
// values.  numpy produces them on one core (0.74 s for the Netflix shape — half of the end-to-end fit).  This file
// produces the SAME values on all host cores.
//
// numpy's normal is a ziggurat: every attempt takes one 64-bit word of the stream and, rarely (1.2 %), one or more
// extra words (wedge test, tail loop, retry), so the k-th normal does not sit at a known stream position.  But the
// generator has random access (PCG64 is an LCG: `advance` in O(log n)) and a call that STARTS at a given position
// always returns the same value and consumes the same words.  So:
//   1. cut the raw stream into chunks; every thread decodes chunks speculatively from the chunk start, as if a call
//      began there, with numpy's OWN `random_standard_normal` (linked from numpy's libnpyrandom.a — its tables and
//      code, not a restatement) on a counting PCG64; it records where each call started (bitmap) and what it returned;
//   2. a serial stitch walks the true call sequence: if the true position is a speculative call start, the rest of
//      that chunk's speculative calls are the true ones; otherwise (the previous chunk's last call spilled over) it
//      makes single calls from the true position until it lands on a speculative call start — almost always at once;
//   3. the accepted runs are scaled and copied (float32 or float64) into the caller's arrays in parallel.
// Bit-identical to numpy by construction; tests/test_host_rng.py compares against numpy for many sizes, seeds,
// thread counts and segment splits.  Plain C + OpenMP; no CUDA involved.
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <numpy/random/bitgen.h>

#include "b200als_host.h"

double random_standard_normal(bitgen_t* bitgen_state);   // numpy/random/distributions.h (libnpyrandom.a)

typedef unsigned __int128 u128;

// numpy/random/src/pcg64/pcg64.h: 128-bit LCG, XSL-RR output taken AFTER the step
#define PCG_MULT ((((u128)0x2360ED051FC65DA4ULL) << 64) | (u128)0x4385DF649FCCF645ULL)

typedef struct {
    u128 state, inc;
    uint64_t consumed;      // 64-bit words drawn since the generator was positioned
} cpcg_t;

static inline uint64_t cpcg_next64(void* st) {
    cpcg_t* g = (cpcg_t*)st;
    g->state = g->state * PCG_MULT + g->inc;
    g->consumed++;
    const uint64_t hi = (uint64_t)(g->state >> 64), lo = (uint64_t)g->state;
    const uint64_t x = hi ^ lo;
    const unsigned rot = (unsigned)(hi >> 58);
    return (x >> rot) | (x << ((-rot) & 63));
}
static uint32_t cpcg_next32(void* st) { return (uint32_t)(cpcg_next64(st) >> 32); }   // not used by the double normal
static double cpcg_double(void* st) { return (double)(cpcg_next64(st) >> 11) * (1.0 / 9007199254740992.0); }

static void cpcg_seek(cpcg_t* g, u128 state0, u128 inc, uint64_t pos) {   // state after `pos` steps from state0
    u128 acc_mult = 1, acc_plus = 0, cur_mult = PCG_MULT, cur_plus = inc;
    uint64_t d = pos;
    while (d) {
        if (d & 1) {
            acc_mult *= cur_mult;
            acc_plus = acc_plus * cur_mult + cur_plus;
        }
        cur_plus = (cur_mult + 1) * cur_plus;
        cur_mult *= cur_mult;
        d >>= 1;
    }
    g->state = acc_mult * state0 + acc_plus;
    g->inc = inc;
    g->consumed = 0;
}

static void make_bitgen(bitgen_t* bg, cpcg_t* g) {
    bg->state = g;
    bg->next_uint64 = cpcg_next64;
    bg->next_uint32 = cpcg_next32;
    bg->next_double = cpcg_double;
    bg->next_raw = cpcg_next64;
}

#ifndef CHUNK_LOG
#define CHUNK_LOG 15                     // tests rebuild with 6 so that calls spill over nearly every chunk boundary
#endif
#define CHUNK (1u << CHUNK_LOG)          // raw words per chunk

typedef struct { const double* src; int64_t count; int64_t dst; } run_t;

static void store_run(const run_t* r, int n_seg, const int64_t* seg_end, void* const* outs, const int32_t* as_f32,
                      double scale) {
    int s = 0;
    int64_t o = r->dst;
    while (s < n_seg && o >= seg_end[s]) ++s;
    for (int64_t i = 0; i < r->count && s < n_seg; ++i, ++o) {
        while (s < n_seg && o >= seg_end[s]) ++s;
        if (s >= n_seg) break;
        const int64_t local = o - (s ? seg_end[s - 1] : 0);
        const double v = 0.0 + scale * r->src[i];          // numpy: loc + scale * random_standard_normal()
        if (as_f32[s]) ((float*)outs[s])[local] = (float)v;
        else ((double*)outs[s])[local] = v;
    }
}

// state / inc: `default_rng(seed).bit_generator.state["state"]` as (hi, lo) 64-bit halves.
// counts[s] normals go to outs[s] (float32 if as_f32[s] else float64), consecutively from the one stream.
// Returns 0, or -1 on allocation failure.  *raw_words (optional) = stream words consumed, serial path only.
int b200als_host_normal_fill(uint64_t state_hi, uint64_t state_lo, uint64_t inc_hi, uint64_t inc_lo, int32_t n_seg,
                             const int64_t* counts, void* const* outs, const int32_t* as_f32, double scale,
                             int32_t n_threads, uint64_t* raw_words) {
    const u128 state0 = (((u128)state_hi) << 64) | state_lo, inc = (((u128)inc_hi) << 64) | inc_lo;
    int64_t total = 0;
    int64_t* seg_end = (int64_t*)malloc(sizeof(int64_t) * (size_t)(n_seg > 0 ? n_seg : 1));
    if (!seg_end) return -1;
    for (int s = 0; s < n_seg; ++s) { total += counts[s]; seg_end[s] = total; }
    if (total == 0) { free(seg_end); if (raw_words) *raw_words = 0; return 0; }

    // small jobs: one plain serial pass
    if (total < (int64_t)4 * CHUNK || n_threads <= 1) {
        cpcg_t g; bitgen_t bg;
        cpcg_seek(&g, state0, inc, 0);
        make_bitgen(&bg, &g);
        int s = 0;
        for (int64_t o = 0; o < total; ++o) {
            while (o >= seg_end[s]) ++s;
            const int64_t local = o - (s ? seg_end[s - 1] : 0);
            const double v = 0.0 + scale * random_standard_normal(&bg);
            if (as_f32[s]) ((float*)outs[s])[local] = (float)v;
            else ((double*)outs[s])[local] = v;
        }
        if (raw_words) *raw_words = g.consumed;
        free(seg_end);
        return 0;
    }

    // ---- waves of chunks: 1. speculative decode (parallel), 2. stitch (serial), 3. scale/convert/copy (parallel) ----
    const int64_t wave = (int64_t)n_threads * 32;                                 // chunks per wave: the buffers stay cache-sized
    const int64_t max_runs = 4 * wave + 64;
    double* spec = (double*)malloc(sizeof(double) * (size_t)wave * CHUNK);
    uint64_t* starts = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)wave * (CHUNK / 64));   // call-start bitmap
    int32_t* cnt = (int32_t*)malloc(sizeof(int32_t) * (size_t)wave);
    uint64_t* endpos = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)wave);       // position after the chunk's last call
    run_t* runs = (run_t*)malloc(sizeof(run_t) * (size_t)max_runs);
    double* singles = (double*)malloc(sizeof(double) * (size_t)max_runs);
    if (!spec || !starts || !cnt || !endpos || !runs || !singles) {
        free(spec); free(starts); free(cnt); free(endpos); free(runs); free(singles); free(seg_end);
        return -1;
    }
    int64_t produced = 0;
    uint64_t p = 0;                                   // true stream position of the next call
    while (produced < total) {
        const int64_t c0 = (int64_t)(p >> CHUNK_LOG);                            // the wave starts at the chunk holding p
        // chunks this wave can still need (every chunk yields > CHUNK/2 normals)
        int64_t n_wave = (total - produced) / (CHUNK / 2) + 2;
        if (n_wave > wave) n_wave = wave;
#pragma omp parallel for schedule(dynamic, 1) num_threads(n_threads)
        for (int64_t w = 0; w < n_wave; ++w) {
            cpcg_t g; bitgen_t bg;
            cpcg_seek(&g, state0, inc, (uint64_t)(c0 + w) * CHUNK);
            make_bitgen(&bg, &g);
            double* out = spec + (size_t)w * CHUNK;
            uint64_t* bm = starts + (size_t)w * (CHUNK / 64);
            memset(bm, 0, sizeof(uint64_t) * (CHUNK / 64));
            int32_t r = 0;
            while (g.consumed < CHUNK) {
                bm[g.consumed >> 6] |= 1ull << (g.consumed & 63);
                out[r++] = random_standard_normal(&bg);
            }
            cnt[w] = r;
            endpos[w] = (uint64_t)(c0 + w) * CHUNK + g.consumed;
        }

        int64_t n_runs = 0, n_single = 0;
        while (produced < total && n_runs < max_runs) {
            const int64_t w = (int64_t)(p >> CHUNK_LOG) - c0;
            if (w >= n_wave) break;
            const uint32_t off = (uint32_t)(p & (CHUNK - 1));
            const uint64_t* bm = starts + (size_t)w * (CHUNK / 64);
            if ((bm[off >> 6] >> (off & 63)) & 1ull) {
                // p is a speculative call start: everything from here to the end of the chunk is the true sequence
                int32_t rank = 0;
                for (uint32_t i = 0; i < (off >> 6); ++i) rank += __builtin_popcountll(bm[i]);
                if (off & 63) rank += __builtin_popcountll(bm[off >> 6] & ((1ull << (off & 63)) - 1ull));
                int64_t take = cnt[w] - rank;
                if (take > total - produced) take = total - produced;
                runs[n_runs].src = spec + (size_t)w * CHUNK + rank;
                runs[n_runs].count = take;
                runs[n_runs].dst = produced;
                ++n_runs;
                produced += take;
                p = endpos[w];
            } else {
                // the previous call spilled over this position: one true call from p
                cpcg_t g; bitgen_t bg;
                cpcg_seek(&g, state0, inc, p);
                make_bitgen(&bg, &g);
                singles[n_single] = random_standard_normal(&bg);
                runs[n_runs].src = singles + n_single;
                runs[n_runs].count = 1;
                runs[n_runs].dst = produced;
                ++n_runs; ++n_single;
                produced += 1;
                p += g.consumed;
            }
        }
#pragma omp parallel for schedule(dynamic, 1) num_threads(n_threads)
        for (int64_t i = 0; i < n_runs; ++i) store_run(&runs[i], n_seg, seg_end, outs, as_f32, scale);
    }
    if (raw_words) *raw_words = 0;       // the position after a truncated run is not tracked (nobody draws after W_f)
    free(spec); free(starts); free(cnt); free(endpos); free(runs); free(singles); free(seg_end);
    return 0;
}

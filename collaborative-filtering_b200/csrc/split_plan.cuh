// split_plan.cuh — work items of one half-step for the tcgen05 path.
//
// A row's ratings are consumed by ONE SM, 32 per pipeline stage, so a popular item (or a
// very active user) bounds the half-step from below: the Netflix-shaped item step has rows
// of several hundred thousand ratings against ~85 k ratings per SM once eight GPUs share it.
// Rows with more than `thresh` ratings are therefore cut into pieces of <= `piece` ratings
// (a multiple of the 32-rating stage).  Every piece is an ordinary work item of the
// persistent kernel; its partial normal matrix [A | c s] and (sum q, sum q^2) are added to
// the row's accumulator slot in global memory (red.add.f32, L2) and the SM that adds the
// last piece — an arrival counter per slot — factors and solves from the slot.
//
// The item list keeps the schedule's order (longest first), pieces of a row adjacent:
//   item = {row, first rating (offset inside the row), ratings, slot or -1}
// Built on the device once per schedule (b200als_split_plan_build): piece counts -> exclusive scan (cub) ->
// items; every step then only zeroes the slots in use.
#pragma once
#include <cub/device/device_scan.cuh>

#include "common.cuh"

namespace b200als {

#ifndef B200ALS_SPLIT_ROWS_DEFAULT
#define B200ALS_SPLIT_ROWS_DEFAULT 6144
#endif
#ifndef B200ALS_SPLIT_PIECE_DEFAULT
#define B200ALS_SPLIT_PIECE_DEFAULT 4096
#endif

struct SplitLayout {                 // byte offsets inside the step workspace (after the bf16 copy)
    size_t counts, scanned, cub_tmp, cub_bytes, items, slot_np, slot_ctr, slot_acc, total;
    int64_t max_items, max_slots;
    int thresh, piece, slot_floats;
};

// slot: (k + 2) rows of k + 4 floats — the normal matrix, then c, then s — followed by (sum q, sum q^2, -, -)
__host__ __device__ constexpr int split_slot_floats(int k) { return (k + 2) * (k + 4) + 4; }

static inline bool split_layout(int k, int64_t n_list, int64_t nnz, int32_t split_rows, int32_t split_piece,
                                int64_t n_long, size_t base, SplitLayout& L) {
    L = SplitLayout{};
    L.total = base;
    if (split_rows < 0 || nnz <= 0 || n_list <= 0) return false;
    // default: an ABSOLUTE piece length, not a share of the launch's ratings.  The tensor core adds into its
    // float32 accumulator with truncation, so a chain of n accumulation steps (16 ratings each) biases the sums
    // low by ~n 2^-25 relative; with pieces keyed on the launch size the result drifted with the number of
    // ranks sharing the rows (round-1 SCALE record: train RMSE moved 4.4e-5 between 1 and 8 GPUs).  A fixed
    // piece bounds the chain (B200ALS_SPLIT_PIECE_DEFAULT / 16 steps), makes the summation tree a function of the
    // row alone, and still leaves pieces far below one SM's share at every named shape.
    L.thresh = split_rows == 0 ? B200ALS_SPLIT_ROWS_DEFAULT : split_rows;
    if (L.thresh < 32) L.thresh = 32;
    int piece = split_piece > 0 ? split_piece : (split_rows == 0 ? B200ALS_SPLIT_PIECE_DEFAULT : L.thresh / 2);
    piece = ((piece + 31) / 32) * 32;
    if (piece < 32) piece = 32;
    L.piece = piece;
    L.max_slots = nnz / L.thresh;
    if (L.max_slots > n_list) L.max_slots = n_list;
    if (n_long >= 0 && n_long < L.max_slots) L.max_slots = n_long;   // the caller counted the rows above the threshold
    if (L.max_slots == 0) return false;                      // no row can exceed the threshold
    L.max_items = n_list + nnz / piece + L.max_slots;
    L.slot_floats = split_slot_floats(k);
    size_t cub_bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, cub_bytes, (const unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                  (int)(n_list + 1));
    L.cub_bytes = cub_bytes;
    size_t off = base;
    auto take = [&](size_t bytes) { size_t o = off; off += align_up(bytes, 256); return o; };
    L.counts = take((size_t)(n_list + 1) * 8);
    L.scanned = take((size_t)(n_list + 1) * 8);
    L.cub_tmp = take(cub_bytes);
    L.items = take((size_t)L.max_items * 16);
    L.slot_np = take((size_t)L.max_slots * 4);
    L.slot_ctr = take((size_t)L.max_slots * 4);
    L.slot_acc = take((size_t)L.max_slots * L.slot_floats * 4);
    L.total = off;
    return true;
}

// counts[e] = pieces of list entry e (low 32 bits) | split? (bit 32); counts[n_list] = 0
__global__ void __launch_bounds__(256) split_count_kernel(const int64_t* __restrict__ indptr,
                                                          const int32_t* __restrict__ rows, int64_t n_list, int thresh,
                                                          int piece, unsigned long long* __restrict__ counts) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e > n_list) return;
    unsigned long long v = 0;
    if (e < n_list) {
        const int64_t row = rows ? rows[e] : e;
        const int64_t c = indptr[row + 1] - indptr[row];
        v = 1;
        if (c > thresh) v = (unsigned long long)((c + piece - 1) / piece) | (1ull << 32);
    }
    counts[e] = v;
}

__global__ void __launch_bounds__(256) split_write_kernel(const int64_t* __restrict__ indptr,
                                                          const int32_t* __restrict__ rows, int64_t n_list,
                                                          const unsigned long long* __restrict__ counts,
                                                          const unsigned long long* __restrict__ scanned,
                                                          int4* __restrict__ items, int32_t* __restrict__ slot_np,
                                                          int32_t* __restrict__ hdr) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e > n_list) return;
    const unsigned long long sc = scanned[e];
    const uint32_t first = (uint32_t)sc, slot = (uint32_t)(sc >> 32);
    if (e == n_list) {
        hdr[0] = (int32_t)first;                       // items
        hdr[1] = (int32_t)slot;                        // split rows
        return;
    }
    const int64_t row = rows ? rows[e] : e;
    const int64_t c = indptr[row + 1] - indptr[row];
    const uint32_t np = (uint32_t)counts[e];
    if (np == 1) {
        items[first] = make_int4((int)row, 0, (int)c, -1);
        return;
    }
    // equal pieces, each a multiple of 32 ratings but the last
    int64_t len = (c + np - 1) / np;
    len = ((len + 31) / 32) * 32;
    uint32_t made = 0;
    for (int64_t b = 0; b < c; b += len, ++made) {
        const int64_t l = (c - b) < len ? (c - b) : len;
        items[first + made] = make_int4((int)row, (int)b, (int)l, (int)slot);
    }
    slot_np[slot] = (int32_t)made;
    // rounding the piece length up can need fewer pieces than counted: pad with empty items (no stage, no arrival)
    for (; made < np; ++made) items[first + made] = make_int4((int)row, (int)c, 0, (int)slot);
}

// zero the accumulators and arrival counters of the slots in use
__global__ void __launch_bounds__(256) split_zero_kernel(const int32_t* __restrict__ hdr, int slot_floats,
                                                         float* __restrict__ slot_acc, int32_t* __restrict__ slot_ctr) {
    const int64_t n_slots = hdr[1];
    const int64_t total = n_slots * slot_floats;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) slot_acc[i] = 0.0f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots; i += stride) slot_ctr[i] = 0;
}

// Builds the plan into caller memory `mem` (laid out by split_layout with base 256: the first 256 bytes
// hold the device header {items, split rows}) and describes it in *out.
static inline int split_build(const int64_t* indptr, const int32_t* rows, int64_t n_list, const SplitLayout& L,
                              char* mem, b200als_split_plan* out, cudaStream_t st) {
    auto* counts = reinterpret_cast<unsigned long long*>(mem + L.counts);
    auto* scanned = reinterpret_cast<unsigned long long*>(mem + L.scanned);
    int4* items = reinterpret_cast<int4*>(mem + L.items);
    int32_t* hdr = reinterpret_cast<int32_t*>(mem);
    int32_t* slot_np = reinterpret_cast<int32_t*>(mem + L.slot_np);
    const unsigned blocks = (unsigned)ceil_div64(n_list + 1, 256);
    split_count_kernel<<<blocks, 256, 0, st>>>(indptr, rows, n_list, L.thresh, L.piece, counts);
    ALS_LAUNCH_CHECK();
    size_t cb = L.cub_bytes;
    ALS_CUDA(cub::DeviceScan::ExclusiveSum(mem + L.cub_tmp, cb, counts, scanned, (int)(n_list + 1), st));
    split_write_kernel<<<blocks, 256, 0, st>>>(indptr, rows, n_list, counts, scanned, items, slot_np, hdr);
    ALS_LAUNCH_CHECK();
    out->items = items;
    out->hdr = hdr;
    out->slot_np = slot_np;
    out->slot_ctr = reinterpret_cast<int32_t*>(mem + L.slot_ctr);
    out->slot_acc = reinterpret_cast<float*>(mem + L.slot_acc);
    out->n_list = n_list;
    out->slot_floats = L.slot_floats;
    out->thresh = L.thresh;
    out->piece = L.piece;
    return B200ALS_OK;
}

}  // namespace b200als

// row_step_umma.cuh — tcgen05 / TMEM Gram path for k in {64, 128} (placeholder
// until the kernel lands: reports "unsupported" so AUTO falls back to FFMA).
#pragma once
#include "row_step_ffma.cuh"

namespace b200als {

static inline bool umma_supported(int k, int64_t ldf) {
    (void)k; (void)ldf;
    return false;
}

static inline size_t umma_workspace_bytes(int k, int64_t n_list) {
    (void)k; (void)n_list;
    return 256;
}

static inline int launch_umma(const RowStepParams& p, void* ws, size_t ws_bytes, cudaStream_t s) {
    (void)p; (void)ws; (void)ws_bytes; (void)s;
    set_error("tcgen05 path not built");
    return B200ALS_E_UNSUPPORTED;
}

}  // namespace b200als

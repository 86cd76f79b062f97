// common.cuh — shared helpers for libb200als (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/b200als.h"

namespace b200als {

void set_error(const char* fmt, ...);
extern thread_local int g_launch_count;

#define ALS_CHECK_ARG(cond, msg)                                              \
    do {                                                                      \
        if (!(cond)) {                                                        \
            ::b200als::set_error("%s: invalid argument: %s", __func__, msg);  \
            return B200ALS_E_INVALID;                                         \
        }                                                                     \
    } while (0)

#define ALS_CUDA(call)                                                        \
    do {                                                                      \
        cudaError_t e__ = (call);                                             \
        if (e__ != cudaSuccess) {                                             \
            ::b200als::set_error("%s: %s failed: %s", __func__, #call,        \
                                 cudaGetErrorString(e__));                    \
            return B200ALS_E_CUDA;                                            \
        }                                                                     \
    } while (0)

#define ALS_LAUNCH_CHECK()                                                    \
    do {                                                                      \
        ::b200als::g_launch_count++;                                          \
        cudaError_t e__ = cudaGetLastError();                                 \
        if (e__ != cudaSuccess) {                                             \
            ::b200als::set_error("%s: kernel launch failed: %s", __func__,    \
                                 cudaGetErrorString(e__));                    \
            return B200ALS_E_CUDA;                                            \
        }                                                                     \
    } while (0)

static inline cudaStream_t as_stream(b200als_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

constexpr int kNumSMs = 148;   // B200: 2 dies x 74 SMs

__host__ __device__ constexpr int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }
__host__ __device__ constexpr int round_up(int a, int b) { return (a + b - 1) / b * b; }
static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---- device helpers ------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    unsigned s = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

__device__ __forceinline__ float ld_stream_f32(const float* p) {
    float v;
    asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ int ld_stream_s32(const int* p) {
    int v;
    asm volatile("ld.global.nc.L1::no_allocate.s32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}

// 1: the register-tile updates of the solvers (rank-4 row updates of the LDL^T, Gram tiles of the FFMA path) issue
// packed fma.rn.f32x2 (SASS FFMA2: two IEEE fp32 FMAs per instruction on an aligned register pair, the multiplier
// broadcast) instead of scalar FFMA.  Same operations on every entry in the same order, so the results are
// bit-identical; what changes is the number of instructions on the fma pipe.  Measured on B200 (tools/ffma2_bench.cu,
// profiles/r02_ffma2_microbench.txt): in isolation FFMA2 issues every 2.27 cycles against 1.05 for FFMA (no gain in
// lanes per clock), but inside the LDL^T solver, where the update shares a sub-partition with dependent pivot chains,
// the C4 user half-step went from 34.6 to 32.2 ms (profiles/r02_ffma2_ab.txt).
#ifndef B200ALS_FFMA2
#define B200ALS_FFMA2 1
#endif

__device__ __forceinline__ uint64_t pack_f32x2(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack_f32x2(uint64_t v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t fma_f32x2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}

}  // namespace b200als

// Micro-benchmark: issue rate and latency of scalar FFMA against packed FFMA2 (fma.rn.f32x2) on one SM sub-partition
// and on a full SM.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ffma2_bench.bin tools/ffma2_bench.cu
// Prints cycles per instruction for: 1 warp / 4 warps (one per sub-partition) / 16 warps per SM, 8 independent chains
// per thread (throughput) and 1 chain (latency).
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint64_t pk(float lo, float hi) { uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk(uint64_t v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) { uint64_t d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ float fma1(float a, float b, float c) { float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }

template <int CH, bool PACKED>
__global__ void bench(float* out, long long* cyc, int iters, float w, float v) {
    float a[2 * CH];
    for (int i = 0; i < 2 * CH; ++i) a[i] = (float)(threadIdx.x + i);
    float vv = v + threadIdx.x * 1e-9f, ww = w + threadIdx.x * 1e-9f;
    __syncthreads();
    const long long t0 = clock64();
    if (PACKED) {
        uint64_t x[CH];
        for (int i = 0; i < CH; ++i) x[i] = pk(a[2 * i], a[2 * i + 1]);
        const uint64_t W = pk(ww, ww), V = pk(vv, vv * 1.5f);
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < CH; ++i) x[i] = fma2(W, V, x[i]);
        }
        for (int i = 0; i < CH; ++i) upk(x[i], a[2 * i], a[2 * i + 1]);
    } else {
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 2 * CH; ++i) a[i] = fma1(ww, vv, a[i]);
        }
    }
    const long long t1 = clock64();
    float s = 0.f;
    for (int i = 0; i < 2 * CH; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int CH, bool PACKED>
void run(const char* name, int warps, float* out, long long* cyc) {
    const int iters = 4096;
    bench<CH, PACKED><<<1, 32 * warps>>>(out, cyc, iters, 1.0001f, 0.5f);
    cudaDeviceSynchronize();
    bench<CH, PACKED><<<1, 32 * warps>>>(out, cyc, iters, 1.0001f, 0.5f);
    cudaDeviceSynchronize();
    long long c;
    cudaMemcpy(&c, cyc, sizeof(c), cudaMemcpyDeviceToHost);
    const double instr = (double)iters * (PACKED ? CH : 2 * CH);
    printf("%-8s chains=%d warps=%2d  cycles/instr(warp)=%.3f  fp32 FMA lanes/clk/SM=%.1f\n", name, CH, warps, c / instr,
           (double)iters * 2 * CH * 32 * warps / c);
}

int main() {
    float* out; long long* cyc;
    cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 8);
    for (int warps : {1, 4, 8, 16}) {
        if (warps == 1) { run<8, false>("FFMA", 1, out, cyc); run<8, true>("FFMA2", 1, out, cyc); run<1, false>("FFMA", 1, out, cyc); run<1, true>("FFMA2", 1, out, cyc); }
        else if (warps == 4) { run<8, false>("FFMA", 4, out, cyc); run<8, true>("FFMA2", 4, out, cyc); }
        else if (warps == 8) { run<8, false>("FFMA", 8, out, cyc); run<8, true>("FFMA2", 8, out, cyc); }
        else { run<8, false>("FFMA", 16, out, cyc); run<8, true>("FFMA2", 16, out, cyc); }
    }
    cudaError_t e = cudaGetLastError();
    printf("status %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}

// Micro-benchmark: issue cost of packed FP32 (FFMA2) against scalar FFMA on sm_100a, alone and mixed with ALU-pipe work.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2_bench ffma2_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ float lo(u64 v) { float l, h; asm("mov.b64 {%0,%1}, %2;" : "=f"(l), "=f"(h) : "l"(v)); return l + h; }

template <int MODE> __global__ void __launch_bounds__(256) k(float* out, int n, float a, float b, unsigned q) {
    float s[16]; u64 p[8]; unsigned z[8];
    for (int i = 0; i < 8; ++i) z[i] = q + i;
    for (int i = 8; i < 16; ++i) s[i] = b + i;
    for (int i = 0; i < 8; ++i) { s[i] = a + i; p[i] = pk(a + i, b + i); }
    const u64 m = pk(a, a), c = pk(b, b);
    for (int it = 0; it < n; ++it) {
        if (MODE == 0) {          // 8 scalar FFMA
#pragma unroll
            for (int i = 0; i < 8; ++i) s[i] = fmaf(s[i], a, b);
        } else if (MODE == 1) {   // 8 FFMA2 (16 FMA)
#pragma unroll
            for (int i = 0; i < 8; ++i) p[i] = ffma2(p[i], m, c);
        } else if (MODE == 2) {   // 16 scalar FFMA + 8 independent ALU ops
#pragma unroll
            for (int i = 0; i < 8; ++i) { s[i] = fmaf(s[i], a, b); s[i + 8] = fmaf(s[i + 8], a, b); z[i] = __funnelshift_l(z[i], z[i], 3) ^ q; }
        } else if (MODE == 3) {   // 8 FFMA2 (16 FMA) + 8 independent ALU ops
#pragma unroll
            for (int i = 0; i < 8; ++i) { p[i] = ffma2(p[i], m, c); z[i] = __funnelshift_l(z[i], z[i], 3) ^ q; }
        } else if (MODE == 4) {   // 8 independent ALU ops alone
#pragma unroll
            for (int i = 0; i < 8; ++i) z[i] = __funnelshift_l(z[i], z[i], 3) ^ q;
        }
    }
    float r = 0; unsigned zz = 0; for (int i = 0; i < 8; ++i) { r += s[i] + s[i + 8] + lo(p[i]); zz ^= z[i]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = r + (float)zz;
}
template <int MODE> void run(const char* name, double fma_per_iter, float* d, int sms) {
    const int n = 20000, blocks = sms * 8;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<blocks, 256>>>(d, 100, 1.0001f, 0.5f, 7u);
    cudaEventRecord(e0); k<MODE><<<blocks, 256>>>(d, n, 1.0001f, 0.5f, 7u); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double fma = fma_per_iter * n * blocks * 256.0;
    printf("%-34s %8.3f ms  %7.2f TFLOP/s (FMA=2)  %6.2f cycles/iter/warp-scheduler-slot\n", name, ms, 2 * fma / ms / 1e9,
           ms * 1e-3 * 1.965e9 / n / (8.0 * 8 / 4));
}
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    float* d; cudaMalloc(&d, (size_t)p.multiProcessorCount * 8 * 256 * 4);
    run<0>("8 FFMA", 8, d, p.multiProcessorCount);
    run<1>("8 FFMA2", 16, d, p.multiProcessorCount);
    run<2>("16 FFMA + 8 SHF + 8 LOP", 16, d, p.multiProcessorCount);
    run<3>("8 FFMA2 + 8 SHF + 8 LOP", 16, d, p.multiProcessorCount);
    run<4>("8 SHF + 8 LOP", 0, d, p.multiProcessorCount);
    return 0;
}

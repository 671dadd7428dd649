// Microbenchmark: issue rate of packed FP32 (FFMA2, fma.rn.f32x2) against scalar FFMA on sm_100a, alone and mixed
// with integer ALU work.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -o ffma2 ffma2.cu ; run: ./ffma2
#include <cstdio>
#include <cuda_runtime.h>

typedef unsigned long long u64;
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ float fma1(float a, float b, float c) { float r; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
__device__ __forceinline__ unsigned iadd(unsigned a, unsigned b) { unsigned r; asm volatile("add.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }

constexpr int ITERS = 4096;

// MODE 0: 8 scalar FFMA per iteration; 1: 8 FFMA2 (=16 flop pairs); 2: 8 FFMA + 8 IADD; 3: 4 FFMA2 + 8 IADD (same flops as 2)
template <int MODE>
__global__ void bench(float* out, float a, float b) {
    float x[8]; u64 y[8]; unsigned z[8];
    for (int i = 0; i < 8; ++i) { x[i] = threadIdx.x + i; y[i] = ((u64)__float_as_uint(x[i]) << 32) | __float_as_uint(x[i] + 1.f); z[i] = threadIdx.x + i; }
    const u64 aa = ((u64)__float_as_uint(a) << 32) | __float_as_uint(a), bb = ((u64)__float_as_uint(b) << 32) | __float_as_uint(b);
    for (int it = 0; it < ITERS; ++it) {
        if (MODE == 0 || MODE == 2) {
#pragma unroll
            for (int i = 0; i < 8; ++i) x[i] = fma1(x[i], a, b);
        }
        if (MODE == 1) {
#pragma unroll
            for (int i = 0; i < 8; ++i) y[i] = fma2(y[i], aa, bb);
        }
        if (MODE == 3) {
#pragma unroll
            for (int i = 0; i < 4; ++i) y[i] = fma2(y[i], aa, bb);
        }
        if (MODE >= 2) {
#pragma unroll
            for (int i = 0; i < 8; ++i) z[i] = iadd(z[i], (unsigned)it);
        }
    }
    float s = 0; unsigned zz = 0;
    for (int i = 0; i < 8; ++i) { s += x[i] + __uint_as_float((unsigned)y[i]) + __uint_as_float((unsigned)(y[i] >> 32)); zz += z[i]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + zz;
}

template <int MODE>
void run(const char* name, float* out, int sms, double inst_per_iter, double flop_per_iter) {
    const int blocks = sms * 4, threads = 512;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    bench<MODE><<<blocks, threads>>>(out, 1.0001f, 0.5f);
    cudaEventRecord(e0);
    bench<MODE><<<blocks, threads>>>(out, 1.0001f, 0.5f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double warps = (double)blocks * threads / 32;
    const double winst = warps * ITERS * inst_per_iter;
    printf("%-28s %8.3f ms  %7.2f warp-inst/clk/SM (at 1.965 GHz)  %7.2f TFLOP/s\n", name, ms, winst / (ms * 1e-3) / 1.965e9 / sms,
           warps * 32 * ITERS * flop_per_iter / (ms * 1e-3) / 1e12);
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    float* out; cudaMalloc(&out, (size_t)p.multiProcessorCount * 4 * 512 * 4);
    printf("%s, %d SMs\n", p.name, p.multiProcessorCount);
    run<0>("8 FFMA", out, p.multiProcessorCount, 8, 16);
    run<1>("8 FFMA2", out, p.multiProcessorCount, 8, 32);
    run<2>("8 FFMA + 8 IADD", out, p.multiProcessorCount, 16, 16);
    run<3>("4 FFMA2 + 8 IADD", out, p.multiProcessorCount, 12, 16);
    return 0;
}

// Microbenchmark: issue rate of FFMA2 (fma.rn.f32x2) against FFMA on sm_100a, alone and mixed with ALU work.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench/ffma2 tools/ubench/ffma2.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void upk(u64 v, float& a, float& b) { asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ float fma1(float a, float b, float c) { float r; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }

template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, int iters, float s) {
    // 8 independent chains per thread
    float a[8], b[8];
    u64 p[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { a[i] = threadIdx.x * 1e-3f + i; b[i] = a[i] + 0.5f; p[i] = pk(a[i], b[i]); }
    unsigned x = threadIdx.x;
    const u64 ss = pk(s, s);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) { a[i] = fma1(a[i], s, 1e-6f); }                                  // 8 FFMA
            if (MODE == 1) { p[i] = fma2(p[i], ss, ss); }                                     // 8 FFMA2
            if (MODE == 2) { a[i] = fma1(a[i], s, 1e-6f); b[i] = fma1(b[i], s, 1e-6f); }      // 16 FFMA
            if (MODE == 3) { p[i] = fma2(p[i], ss, ss); asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x) : "r"(it), "r"(i)); }   // FFMA2 + ALU
            if (MODE == 4) { a[i] = fma1(a[i], s, 1e-6f); b[i] = fma1(b[i], s, 1e-6f); asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x) : "r"(it), "r"(i)); }  // 2 FFMA + ALU
            if (MODE == 5) { p[i] = fma2(p[i], ss, ss); asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x) : "r"(it), "r"(i)); asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x) : "r"(i), "r"(it)); }  // FFMA2 + 2 ALU
        }
    }
    float r = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { float u, v; upk(p[i], u, v); r += a[i] + b[i] + u + v; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = r + (float)x;
}

template <int MODE> void run(const char* name, int fma_per_iter, int instr_per_iter) {
    int dev_sms = 148; cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0); dev_sms = pr.multiProcessorCount;
    const int blocks = dev_sms * 8, iters = 20000;
    float* out; cudaMalloc(&out, (size_t)blocks * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<blocks, 256>>>(out, 100, 0.999f);
    cudaEventRecord(e0);
    k<MODE><<<blocks, 256>>>(out, iters, 0.999f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const double warps = (double)blocks * 8, wi = warps * iters * instr_per_iter;
    const double fl = (double)blocks * 256 * iters * fma_per_iter * 2;
    printf("%-22s %8.3f ms  %7.2f TFLOP/s  %6.3f warp-instr/clk/SM (at %d MHz nominal)\n", name, ms, fl / ms * 1e-9,
           wi / (ms * 1e-3) / (clk * 1e3) / dev_sms, clk / 1000);
    cudaFree(out);
}
int main() {
    run<0>("8 FFMA", 8, 8);
    run<2>("16 FFMA", 16, 16);
    run<1>("8 FFMA2", 16, 8);
    run<4>("16 FFMA + 8 LOP3", 16, 24);
    run<3>("8 FFMA2 + 8 LOP3", 16, 16);
    run<5>("8 FFMA2 + 16 LOP3", 16, 24);
    return 0;
}

// Microbenchmark: issue rate of scalar FFMA / FMUL / FADD vs the packed f32x2 forms on sm_100a.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o f32x2_bench f32x2_bench.cu ; run on the GPU box.
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 4096
#define CHAINS 8
template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, float s, float t) {
    float a[CHAINS]; unsigned long long p[CHAINS];
    unsigned long long S, T;
    asm("mov.b64 %0, {%1, %1};" : "=l"(S) : "f"(s));
    asm("mov.b64 %0, {%1, %1};" : "=l"(T) : "f"(t));
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) { a[c] = threadIdx.x * 1e-3f + c; asm("mov.b64 %0, {%1, %1};" : "=l"(p[c]) : "f"(a[c])); }
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) {
            if (MODE == 0) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[c]) : "f"(s), "f"(t));
            if (MODE == 1) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[c]) : "l"(S), "l"(T));
            if (MODE == 2) asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(a[c]) : "f"(s));
            if (MODE == 3) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(p[c]) : "l"(S));
            if (MODE == 4) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(a[c]) : "f"(t));
            if (MODE == 5) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p[c]) : "l"(T));
            if (MODE == 6) { asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[c]) : "l"(S), "l"(T)); asm volatile("lop3.b32 %0, %0, 0x7fffffff, 0, 0xc0;" : "+f"(a[c])); }   // packed fma + an ALU op
            if (MODE == 7) { asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[c]) : "f"(s), "f"(t)); asm volatile("lop3.b32 %0, %0, 0x7fffffff, 0, 0xc0;" : "+f"(a[c])); }
        }
    }
    float r = 0.0f;
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) { float lo, hi; asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p[c])); r += a[c] + lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int MODE> void run(const char* name, float* d, int per_iter_flops_per_thread) {
    const int blocks = 148 * 8;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<blocks, 256>>>(d, 1.0001f, 1e-6f);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<MODE><<<blocks, 256>>>(d, 1.0001f, 1e-6f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double inst = (double)blocks * 8 /*warps*/ * ITERS * CHAINS;
    printf("%-28s %8.3f ms  %7.2f G warp-inst/s  (%.1f per clk per SM at 1.965 GHz)  %7.2f Tlane-op/s\n", name, ms, inst / ms * 1e-6,
           inst / ms * 1e-6 / 148 / 1.965, inst * 32 * per_iter_flops_per_thread / ms * 1e-9);
}
int main() {
    float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    run<0>("fma.f32", d, 1); run<1>("fma.f32x2", d, 2); run<2>("mul.f32", d, 1); run<3>("mul.f32x2", d, 2);
    run<4>("add.f32", d, 1); run<5>("add.f32x2", d, 2); run<6>("fma.f32x2 + lop3 (pairs)", d, 2); run<7>("fma.f32 + lop3 (pairs)", d, 1);
    return 0;
}

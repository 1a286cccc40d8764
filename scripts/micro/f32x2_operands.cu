// Microbenchmark: does the cost of a packed f32x2 instruction depend on how many distinct register operands it reads?
// Each thread keeps N independent values and combines them in a rotating pattern so that no operand is reused.
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 2048
#define N 12
template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, float s) {
    unsigned long long p[N]; float a[N];
#pragma unroll
    for (int c = 0; c < N; ++c) { a[c] = 1.0f + (threadIdx.x + c) * 1e-6f * s; asm("mov.b64 %0, {%1, %1};" : "=l"(p[c]) : "f"(a[c])); }
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int c = 0; c < N; ++c) {
            const int c1 = (c + 4) % N, c2 = (c + 8) % N;
            if (MODE == 0) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(p[c]) : "l"(p[c1]), "l"(p[c2]));       // 3 distinct register pairs
            if (MODE == 1) asm volatile("fma.rn.f32x2 %0, %1, %1, %0;" : "+l"(p[c]) : "l"(p[c1]));                   // 2 distinct
            if (MODE == 2) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(p[c]) : "l"(p[c1]));                       // 2 distinct
            if (MODE == 3) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p[c]) : "l"(p[c1]));
            if (MODE == 4) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(a[c]) : "f"(a[c1]), "f"(a[c2]));         // scalar, 3 distinct
            if (MODE == 5) asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(a[c]) : "f"(a[c1]));
            if (MODE == 6) asm volatile("fma.rn.f32x2 %0, %0, %0, %0;" : "+l"(p[c]));                                // 1 distinct
        }
    }
    float r = 0.0f;
#pragma unroll
    for (int c = 0; c < N; ++c) { float lo, hi; asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p[c])); r += a[c] + lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int MODE> void run(const char* name, float* d) {
    const int blocks = 148 * 8;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<blocks, 256>>>(d, 1.0f);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<MODE><<<blocks, 256>>>(d, 1.0f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double inst = (double)blocks * 8 * ITERS * N;
    printf("%-44s %8.3f ms  %5.2f SMSP-cycles per warp instruction\n", name, ms, ms * 1e-3 * 1.965e9 * 148 * 4 / inst);
}
int main() {
    float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    run<0>("fma.f32x2, 3 distinct operand pairs", d); run<1>("fma.f32x2, 2 distinct", d); run<6>("fma.f32x2, 1 distinct", d);
    run<2>("mul.f32x2, 2 distinct", d); run<3>("add.f32x2, 2 distinct", d); run<4>("fma.f32 scalar, 3 distinct", d); run<5>("mul.f32 scalar, 2 distinct", d);
    return 0;
}

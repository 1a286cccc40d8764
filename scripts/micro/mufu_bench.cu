// Microbenchmark: MUFU.RSQ / MUFU.RCP issue rate on sm_100a, alone and next to FP32 work.
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 4096
#define CHAINS 8
template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, float s, float t) {
    float a[CHAINS], b[CHAINS];
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) { a[c] = threadIdx.x * 1e-3f + c + 1.0f; b[c] = a[c] * 0.5f; }
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) {
            if (MODE == 0) asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(a[c]));
            if (MODE == 1) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(a[c]));
            if (MODE == 2) { asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(a[c])); asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(b[c]) : "f"(s), "f"(t)); }
            if (MODE == 3) { asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(a[c]));
                             asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(b[c]) : "f"(s), "f"(t)); asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(b[c]) : "f"(s), "f"(t));
                             asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(b[c]) : "f"(s), "f"(t)); asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(b[c]) : "f"(s), "f"(t)); }
            if (MODE == 4) { asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(a[c]));
#pragma unroll
                             for (int j = 0; j < 10; ++j) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(b[c]) : "f"(s), "f"(t)); }
        }
    }
    float r = 0.0f;
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) r += a[c] + b[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int MODE> void run(const char* name, float* d, int mufu, int fma) {
    const int blocks = 148 * 8;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<blocks, 256>>>(d, 1.0001f, 1e-6f);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<MODE><<<blocks, 256>>>(d, 1.0001f, 1e-6f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double groups = (double)blocks * 8 * ITERS * CHAINS;     // warp-level groups of (mufu MUFU + fma FMA)
    const double cyc = ms * 1e-3 * 1.965e9 * 148 * 4 / groups;     // SMSP cycles per group
    printf("%-34s %8.3f ms  %6.2f SMSP-cycles per group (%d MUFU + %d FMA)\n", name, ms, cyc, mufu, fma);
}
int main() {
    float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    run<0>("rsqrt", d, 1, 0); run<1>("rcp", d, 1, 0); run<2>("rsqrt + 1 fma", d, 1, 1); run<3>("rsqrt + 4 fma", d, 1, 4); run<4>("rsqrt + 10 fma", d, 1, 10);
    return 0;
}

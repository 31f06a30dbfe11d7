// Issue rates of the packed FP32 instructions one by one: FADD2, FMUL2, FFMA2 (and scalar FADD / FFMA for scale).
// 8 independent chains per thread, 64 warps per SM; lane-ops counted as 2 per packed instruction (an FMA counts as one).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 --fmad=false -o variants/f32x2_rates tools/probes/f32x2_rates.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
template <int OP>
__global__ void __launch_bounds__(256) k(float* out, int iters, float c1, float c2) {
    u64 p[8], C1, C2;
    float a[8];
    asm("mov.b64 %0, {%1, %1};" : "=l"(C1) : "f"(c1));
    asm("mov.b64 %0, {%1, %1};" : "=l"(C2) : "f"(c2));
#pragma unroll
    for (int i = 0; i < 8; i++) { a[i] = threadIdx.x * 0.001f + i; asm("mov.b64 %0, {%1, %1};" : "=l"(p[i]) : "f"(a[i])); }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 4; u++) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (OP == 0) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(C1));
                if (OP == 1) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(C2));
                if (OP == 2) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(C2), "l"(C1));
                if (OP == 3) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(c1));
                if (OP == 4) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(c2), "f"(c1));
            }
        }
    }
    float s = 0;
    for (int i = 0; i < 8; i++) { float lo, hi; asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p[i])); s += lo + hi + a[i]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int OP> void run(float* d, const char* name, double per_instr) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 20000;
    float best = 1e9f;
    for (int r = 0; r < 4; r++) {
        cudaEventRecord(e0);
        k<OP><<<148 * 8, 256>>>(d, iters, 1.0000001f, 0.9999999f);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double instr = 148.0 * 8 * 8 * 32.0 * iters;   // warp instructions
    printf("%-8s %7.3f ms  %6.2f G warp-instr/s = %.2f per clock and scheduler at 1.965 GHz;  %6.2f T results/s\n", name, best,
           instr / best / 1e6, instr / best / 1e6 / (148 * 4 * 1.965), instr * 32 * per_instr / best / 1e9);
}
int main() {
    float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    run<0>(d, "FADD2", 2); run<1>(d, "FMUL2", 2); run<2>(d, "FFMA2", 2); run<3>(d, "FADD", 1); run<4>(d, "FFMA", 1);
    return 0;
}

// A probe that was FOOLED, kept as the demonstration of a ptxas hazard.  It was meant to compare scalar add / mul chains
// with the same chains on packed pairs (add / mul .rn.f32x2 -> FADD2 / FMUL2) and printed 36.7 against 69.5 Tlane-op/s.
// The SASS of its packed loop holds 400 FFMA2 and only 64 FADD2 / FMUL2: ptxas contracted mul.rn.f32x2 followed (one
// iteration later) by add.rn.f32x2 into FFMA2 — explicit .rn modifiers, `asm volatile` and --fmad=false notwithstanding —
// so the "doubled rate" is the ordinary FMA rate counted as two operations, and the bits are not those of the separately
// rounded operations.  The honest measurement is f32x2_rates.cu: FADD2, FMUL2 and FFMA2 each issue at 0.5 per clock and
// scheduler, i.e. the SAME results per second as the scalar instructions, in half the issue slots.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 --fmad=false -o variants/f32x2_probe tools/probes/f32x2_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 r; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 r; asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ float addv(float a, float b) { float r; asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float mulv(float a, float b) { float r; asm volatile("mul.rn.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }

template <bool PACKED, bool INTMIX>
__global__ void __launch_bounds__(256) k(float* out, int iters, float c1, float c2) {
    float a[16];
    unsigned n[16];
#pragma unroll
    for (int i = 0; i < 16; i++) { a[i] = threadIdx.x * 0.001f + i; n[i] = threadIdx.x + i; }
    u64 p[8];
    const float2 cc = make_float2(c1, c1), dd = make_float2(c2, c2);
    const u64 C1 = *reinterpret_cast<const u64*>(&cc), C2 = *reinterpret_cast<const u64*>(&dd);
#pragma unroll
    for (int i = 0; i < 8; i++) { float2 t = make_float2(a[2 * i], a[2 * i + 1]); p[i] = *reinterpret_cast<u64*>(&t); }
    for (int it = 0; it < iters; it++) {
        if (PACKED) {
#pragma unroll
            for (int i = 0; i < 8; i++) p[i] = mul2(add2(p[i], C1), C2);
        } else {
#pragma unroll
            for (int i = 0; i < 16; i++) a[i] = mulv(addv(a[i], c1), c2);
        }
        if (INTMIX) {
#pragma unroll
            for (int i = 0; i < 16; i++) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(n[i]) : "r"(n[(i + 1) & 15]), "r"(it));
        }
    }
    float s = 0;
    if (PACKED) { for (int i = 0; i < 8; i++) { float2 t = *reinterpret_cast<float2*>(&p[i]); s += t.x + t.y; } }
    else for (int i = 0; i < 16; i++) s += a[i];
    unsigned m = 0;
    for (int i = 0; i < 16; i++) m ^= n[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + (float)(m & 1);
}
template <bool P, bool I> float run(float* d, int iters) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e9f;
    for (int r = 0; r < 4; r++) {
        cudaEventRecord(e0);
        k<P, I><<<148 * 8, 256>>>(d, iters, 1.0000001f, 0.9999999f);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return best;
}
int main() {
    float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    const int iters = 20000;
    const double ops = 148.0 * 8 * 256 * 32.0 * iters;   // FP32 operations (2 per accumulator per iteration)
    float t;
    t = run<false, false>(d, iters); printf("scalar FADD/FMUL          : %7.3f ms  %6.2f Tlane-op/s\n", t, ops / t / 1e9);
    t = run<true, false>(d, iters);  printf("packed FADD2/FMUL2        : %7.3f ms  %6.2f Tlane-op/s\n", t, ops / t / 1e9);
    t = run<false, true>(d, iters);  printf("scalar + 16 LOP3 per iter : %7.3f ms  %6.2f Tlane-op/s (FP32 part)\n", t, ops / t / 1e9);
    t = run<true, true>(d, iters);   printf("packed + 16 LOP3 per iter : %7.3f ms  %6.2f Tlane-op/s (FP32 part)\n", t, ops / t / 1e9);
    return 0;
}

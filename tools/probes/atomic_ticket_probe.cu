// How fast does ONE global counter hand out tickets?  444 CTAs x 8 warps (the strip kernel's grid), lane 0 of every warp
// takes tickets with atomicAdd until `total` are gone; either back to back (dependent) or with `work` FMAs in between.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o variants/atomic_ticket_probe tools/probes/atomic_ticket_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(int* counter, int total, int work, float* sink, int ncounters) {
    const int lane = threadIdx.x & 31;
    int* c = counter + 32 * ((blockIdx.x * 8 + (threadIdx.x >> 5)) % ncounters);
    float acc = threadIdx.x;
    int t = 0;
    for (;;) {
        if (lane == 0) t = atomicAdd(c, 1);
        t = __shfl_sync(0xffffffffu, t, 0);
        if (t >= total / ncounters) break;
        for (int i = 0; i < work; i++) acc = acc * 1.0001f + 0.5f;
    }
    if (acc == 123.456f) *sink = acc;
}
int main() {
    int* c; float* s;
    cudaMalloc(&c, 4096 * 4); cudaMalloc(&s, 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int totals[] = {65280, 130560};
    for (int nc : {1, 8, 64})
    for (int total : totals)
        for (int work : {0, 64, 256, 1024}) {
            float best = 1e9f;
            for (int rep = 0; rep < 5; rep++) {
                cudaMemset(c, 0, 4096 * 4);
                cudaEventRecord(e0);
                k<<<444, 256>>>(c, total, work, s, nc);
                cudaEventRecord(e1); cudaEventSynchronize(e1);
                float ms; cudaEventElapsedTime(&ms, e0, e1);
                if (ms < best) best = ms;
            }
            printf("counters %2d tickets %6d work %4d FMAs/ticket: %8.1f us  (%.2f ns per ticket)\n", nc, total, work, best * 1e3, best * 1e6 / total);
        }
    return 0;
}

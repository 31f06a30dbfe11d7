// rtce_measure.cu — measurement support: the FP32-issue roofline denominator.
// MEASURED_PEAKS.json carries HBM and bf16-tensor peaks only; the pass is bound by FP32 CUDA-core
// instruction issue, so the sustained rate of separately rounded FMUL/FADD is measured here.
#include <cuda_runtime.h>

#include "../../include/rtce.h"

namespace {

constexpr int kChains = 8;

// 8 independent chains per thread, each iteration = 1 FMUL + 1 FADD per chain (never contracted):
// 16 FP32 instructions per iteration, register operands only.
__global__ void __launch_bounds__(256) k_fp32_issue(float* sink, int iters, float a, float b) {
    float v[kChains];
#pragma unroll
    for (int i = 0; i < kChains; i++) v[i] = (float)(threadIdx.x + i) * 1e-3f;
    for (int it = 0; it < iters; it++) {
#pragma unroll 4
        for (int u = 0; u < 4; u++) {
#pragma unroll
            for (int i = 0; i < kChains; i++) v[i] = __fadd_rn(__fmul_rn(v[i], a), b);
        }
    }
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < kChains; i++) s += v[i];
    if (s == 12345.678f) sink[0] = s;   // keep the chains alive
}

}  // namespace

extern "C" int rtce_measure_fp32_issue(int device, int iters, double* lane_ops_per_s, double* ms_out) {
    if (!lane_ops_per_s || iters < 1) return RTCE_ERR_INVALID;
    int prev = 0, sms = 0;
    if (cudaGetDevice(&prev) != cudaSuccess) return RTCE_ERR_CUDA;
    if (cudaSetDevice(device) != cudaSuccess) return RTCE_ERR_CUDA;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    float* sink = nullptr;
    cudaEvent_t e0, e1;
    int rc = RTCE_OK;
    if (cudaMalloc(&sink, 4) != cudaSuccess) return RTCE_ERR_CUDA;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int blocks = sms * 8, threads = 256;   // 2048 threads / SM: full occupancy
    k_fp32_issue<<<blocks, threads>>>(sink, 64, 0.999f, 1e-3f);   // warm-up
    float best = 1e30f;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(e0);
        k_fp32_issue<<<blocks, threads>>>(sink, iters, 0.999f, 1e-3f);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { rc = RTCE_ERR_CUDA; break; }
        float ms = 0.0f;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    if (rc == RTCE_OK) {
        const double ops = (double)blocks * threads * (double)iters * 4.0 * kChains * 2.0;
        *lane_ops_per_s = ops / (best * 1e-3);
        if (ms_out) *ms_out = best;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    cudaSetDevice(prev);
    return rc;
}

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

// ------------------------------------------------------------------------------------------------
// rtce_selftest_math — exhaustive proof-by-enumeration that rtce::sqrt_fast / rtce::rcp_fast return
// the bits of sqrt.rn / rcp.rn on their stated domains (all 2^32 float bit patterns are visited).
// ------------------------------------------------------------------------------------------------
#include "rtce_device.cuh"

namespace {

__global__ void __launch_bounds__(256) k_selftest_math(unsigned long long* out /* [4] */) {
    unsigned long long bad_sqrt = 0, bad_rcp = 0, n_sqrt = 0, n_rcp = 0;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long b = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; b < (1ull << 32); b += stride) {
        const float x = __uint_as_float((unsigned)b);
        const float ax = fabsf(x);
        // sqrt domain: +0 and finite x >= 2^-101 (bit pattern 0x0d000000)
        if (b == 0ull || (b >= 0x0d000000ull && b < 0x7f800000ull)) {
            n_sqrt++;
            if (__float_as_uint(rtce::sqrt_fast(x)) != __float_as_uint(__fsqrt_rn(x))) bad_sqrt++;
        }
        // rcp domain: 2^-126 <= |x| <= 2^126
        if (ax >= 1.17549435e-38f && ax <= 8.50705917e37f) {
            n_rcp++;
            if (__float_as_uint(rtce::rcp_fast(x)) != __float_as_uint(__frcp_rn(x))) bad_rcp++;
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        bad_sqrt += __shfl_down_sync(0xffffffffu, bad_sqrt, o);
        bad_rcp += __shfl_down_sync(0xffffffffu, bad_rcp, o);
        n_sqrt += __shfl_down_sync(0xffffffffu, n_sqrt, o);
        n_rcp += __shfl_down_sync(0xffffffffu, n_rcp, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(out + 0, bad_sqrt);
        atomicAdd(out + 1, bad_rcp);
        atomicAdd(out + 2, n_sqrt);
        atomicAdd(out + 3, n_rcp);
    }
}

}  // namespace

extern "C" int rtce_selftest_math(int device, uint64_t out4[4]) {
    if (!out4) return RTCE_ERR_INVALID;
    int prev = 0;
    if (cudaGetDevice(&prev) != cudaSuccess) return RTCE_ERR_CUDA;
    if (cudaSetDevice(device) != cudaSuccess) return RTCE_ERR_CUDA;
    unsigned long long* d = nullptr;
    int rc = RTCE_OK;
    if (cudaMalloc(&d, 32) != cudaSuccess) return RTCE_ERR_CUDA;
    cudaMemset(d, 0, 32);
    k_selftest_math<<<148 * 8, 256>>>(d);
    unsigned long long h[4] = {0, 0, 0, 0};
    if (cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost) != cudaSuccess) rc = RTCE_ERR_CUDA;
    for (int i = 0; i < 4; i++) out4[i] = h[i];
    cudaFree(d);
    cudaSetDevice(prev);
    return rc;
}

// ------------------------------------------------------------------------------------------------
// rtce_selftest_repeat — utilities.glsl:32-34 `repeat` (unused by the reference's shaders) on the device.
// ------------------------------------------------------------------------------------------------
namespace {

__global__ void __launch_bounds__(256) k_selftest_repeat(const float* __restrict__ p, const float* __restrict__ r,
                                                         float* __restrict__ out, size_t n) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = rtce::repeat1(p[i], r[i]);
}

}  // namespace

extern "C" int rtce_selftest_repeat(int device, const float* p, const float* r, float* out, size_t n) {
    if (!p || !r || !out) return RTCE_ERR_INVALID;
    if (!n) return RTCE_OK;
    int prev = 0;
    if (cudaGetDevice(&prev) != cudaSuccess) return RTCE_ERR_CUDA;
    if (cudaSetDevice(device) != cudaSuccess) return RTCE_ERR_CUDA;
    float *dp = nullptr, *dr = nullptr, *dout = nullptr;
    int rc = RTCE_OK;
    const size_t bytes = n * sizeof(float);
    if (cudaMalloc(&dp, bytes) != cudaSuccess || cudaMalloc(&dr, bytes) != cudaSuccess || cudaMalloc(&dout, bytes) != cudaSuccess) {
        rc = RTCE_ERR_NOMEM;
    } else if (cudaMemcpy(dp, p, bytes, cudaMemcpyHostToDevice) != cudaSuccess || cudaMemcpy(dr, r, bytes, cudaMemcpyHostToDevice) != cudaSuccess) {
        rc = RTCE_ERR_CUDA;
    } else {
        k_selftest_repeat<<<148 * 4, 256>>>(dp, dr, dout, n);
        if (cudaMemcpy(out, dout, bytes, cudaMemcpyDeviceToHost) != cudaSuccess) rc = RTCE_ERR_CUDA;
    }
    cudaGetLastError();
    cudaFree(dp);
    cudaFree(dr);
    cudaFree(dout);
    cudaSetDevice(prev);
    return rc;
}

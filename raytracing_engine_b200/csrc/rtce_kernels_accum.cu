// rtce_kernels_accum.cu — the accumulation extension (SURVEY.md §8d "progressive accumulation"): running FP32
// sum of the final pyramid levels of consecutive frames, added in frame order, divided by the frame count at
// the end.  No reference counterpart (the reference has no accumulation); HBM-bound, one pass per chunk.
#include <cuda_runtime.h>

#include "rtce_kernels.cuh"

namespace rtce {

__global__ void __launch_bounds__(256) k_accumulate(float* __restrict__ acc, const float* __restrict__ img, size_t slot_stride,
                                                    int frames, size_t n, int first) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float a = __fadd_rn(first ? 0.0f : acc[i], img[i]);
        for (int f = 1; f < frames; f++) a = __fadd_rn(a, img[(size_t)f * slot_stride + i]);   // frame order
        acc[i] = a;
    }
}

__global__ void __launch_bounds__(256) k_scale(float* __restrict__ acc, size_t n, float divisor) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) acc[i] = __fdiv_rn(acc[i], divisor);
}

cudaError_t launch_accumulate(float* acc, const float* img, size_t slot_stride, int frames, size_t n, bool first,
                              cudaStream_t st) {
    if (!n || frames <= 0) return cudaSuccess;
    k_accumulate<<<148 * 8, 256, 0, st>>>(acc, img, slot_stride, frames, n, first ? 1 : 0);
    return cudaGetLastError();
}

cudaError_t launch_scale(float* acc, size_t n, float divisor, cudaStream_t st) {
    if (!n) return cudaSuccess;
    k_scale<<<148 * 8, 256, 0, st>>>(acc, n, divisor);
    return cudaGetLastError();
}

}  // namespace rtce

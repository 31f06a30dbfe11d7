// rtce_kernels_large.cu — objCount > 8 (extension of utilities.glsl:3 MAX_OBJECTS, BASELINE.json
// configs[3]): the same tracing algorithm 3 (compute.glsl:34-68), re-organised so that its
// per-ray x per-object cache of lower bounds (`float distances[MAX_OBJECTS]`, compute.glsl:35)
// fits on chip: ONE CTA marches ONE ray at a time and the objects are spread over its threads.
//
//   * distances[] lives in shared memory as groups of four (one 128-bit access per group), each
//     thread owns groups tid, tid+T, ...
//   * per march step every thread updates its objects (one subtract and one min each, one compare
//     per group; the exact SDF only when a bound falls inside the cone radius) and the block takes
//     min over all of them.
//     min is exactly associative and commutative on non-NaN floats, so the reduction order
//     cannot change a bit.
//   * the scalar ray state (len, last, position) is computed redundantly by every thread with
//     identical inputs -> identical bits, no broadcast needed.
//   * spheres are read as one 128-bit load per object (float4 {pos.xyz, size}, read-only path).
#include "rtce_device.cuh"
#include "rtce_kernels.cuh"

namespace rtce {

constexpr int kLargeThreads = 256;

RTCE_DEV float block_min(float v, float* red /* [2][8] */, int parity) {
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* r = red + parity * 8;
    if (lane == 0) r[warp] = v;
    __syncthreads();
    float m = r[0];
#pragma unroll
    for (int i = 1; i < kLargeThreads / 32; i++) m = fminf(m, r[i]);
    return m;
}

template <bool COUNT>
__global__ void __launch_bounds__(kLargeThreads)
    k_trace_level_large(const __grid_constant__ LevelParamsLarge P, long long total_rays) {
    extern __shared__ float4 s_dist4[];   // ceil(n / 4) groups of four lower bounds
    __shared__ float s_red[16];
    const int n = P.n_spheres;
    const int n4 = (n + 3) >> 2;
    const int rw = P.x1 - P.x0, rh = P.y1 - P.y0;
    const long long per_frame = (long long)rw * rh;
    unsigned long long cE = 0, cS = 0, cR = 0;
    int parity = 0;   // runs across rays: consecutive block_min calls must alternate buffers

    for (long long ray = blockIdx.x; ray < total_rays; ray += gridDim.x) {
        const int f = (int)(ray / per_frame);
        const long long rem = ray - (long long)f * per_frame;
        const int y = P.y0 + (int)(rem / rw);
        const int x = P.x0 + (int)(rem % rw);
        const Cam& cam = P.cam[f];
        // compute.glsl:71-77 (uniform across the CTA)
        const float ncx = fmul(fsub(fmul((float)(2u * (unsigned)x + 1u), P.isx), 1.0f), P.ratx);
        const float ncy = fmul(fsub(fmul((float)(2u * (unsigned)y + 1u), P.isy), 1.0f), P.raty);
        const Vec3 dir = normalize3(rotate_q(cam.rot[0], cam.rot[1], cam.rot[2], cam.rot[3], ncx, 1.0f, ncy));
        float start = 1.0f;   // :79-82
        if (P.iter > 0)
            start = __ldg(P.parent + (size_t)f * P.slot_stride + (size_t)(y >> 1) * (size_t)P.pw + (size_t)(x >> 1));
        const float ox = fadd(cam.pos[0], fmul(dir.x, start));
        const float oy = fadd(cam.pos[1], fmul(dir.y, start));
        const float oz = fadd(cam.pos[2], fmul(dir.z, start));

        // traceCone :37-39; distances[] is kept as float4 groups (slots past n hold +inf: never refreshed,
        // never the minimum)
        for (int i4 = threadIdx.x; i4 < n4; i4 += kLargeThreads) {
            float v[4];
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const int i = 4 * i4 + c;
                v[c] = __int_as_float(0x7F800000);
                if (i < n) {
                    const float4 s = __ldg(P.spheres + i);
                    v[c] = sphere_sdf(ox, oy, oz, s.x, s.y, s.z, s.w);
                }
            }
            s_dist4[i4] = make_float4(v[0], v[1], v[2], v[3]);
        }
        if (COUNT && threadIdx.x == 0) cE += (unsigned long long)n;
        float len = 0.0f, last = 0.0f;
        while (len < P.render_dist) {   // :44
            const float px = fadd(ox, fmul(dir.x, len));
            const float py = fadd(oy, fmul(dir.y, len));
            const float pz = fadd(oz, fmul(dir.z, len));
            const float radius = fmul(fadd(len, 1.0f), P.threshold);
            float m = P.render_dist;   // :49
            // :51-57, four objects per shared-memory access (each thread touches only its own groups): per object
            // one subtract and one min, plus one compare per group; the exact SDF only inside the cone radius
            for (int i4 = threadIdx.x; i4 < n4; i4 += kLargeThreads) {
                float4 d = s_dist4[i4];
                d.x = fsub(d.x, last);
                d.y = fsub(d.y, last);
                d.z = fsub(d.z, last);
                d.w = fsub(d.w, last);
                float m4 = fminf(fminf(d.x, d.y), fminf(d.z, d.w));
                if (m4 <= radius) {
                    float v[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
                    for (int c = 0; c < 4; c++) {
                        if (v[c] <= radius) {   // never true for the +inf padding
                            const float4 s = __ldg(P.spheres + 4 * i4 + c);
                            v[c] = sphere_sdf(px, py, pz, s.x, s.y, s.z, s.w);
                            if (COUNT) cE++;
                        }
                    }
                    d = make_float4(v[0], v[1], v[2], v[3]);
                    m4 = fminf(fminf(d.x, d.y), fminf(d.z, d.w));
                }
                s_dist4[i4] = d;
                m = fminf(m, m4);
            }
            const float dist = block_min(m, s_red, parity);
            parity ^= 1;
            if (COUNT && threadIdx.x == 0) cS++;
            last = glsl_max(dist, 0.0f);   // :59
            len = fadd(len, last);
            if (dist <= radius) {   // :62-65
                len = fsub(len, radius);
                break;
            }
        }
        if (threadIdx.x == 0) {
            P.out[(size_t)f * P.slot_stride + (size_t)y * (size_t)P.w + (size_t)x] = glsl_max(fadd(start, len), 0.0f);
            if (COUNT) cR++;
        }
        // s_dist slots are private to their owner thread, s_red is double-buffered with a barrier
        // in between: no extra barrier needed before the next ray.
    }
    if (COUNT) {
        for (int o = 16; o > 0; o >>= 1) cE += __shfl_down_sync(0xffffffffu, cE, o);
        if ((threadIdx.x & 31) == 0 && cE) atomicAdd(P.counters + 1, cE);
        if (threadIdx.x == 0) {
            atomicAdd(P.counters + 0, cR);
            atomicAdd(P.counters + 2, cS);
        }
    }
}

cudaError_t launch_trace_level_large(const LevelParamsLarge& p, int frames, bool count, cudaStream_t st) {
    if (p.x1 <= p.x0 || p.y1 <= p.y0 || frames <= 0) return cudaSuccess;
    const long long total = (long long)(p.x1 - p.x0) * (p.y1 - p.y0) * frames;
    const size_t smem = sizeof(float4) * (size_t)((p.n_spheres + 3) / 4);
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // persistent CTAs: a multiple of the SM count, as many as fit by shared memory / threads
    int per_sm = (int)((200u * 1024u) / (smem + 1024u));
    if (per_sm > 8) per_sm = 8;
    if (per_sm < 1) per_sm = 1;
    long long grid = (long long)sms * per_sm;
    if (grid > total) grid = total;
    cudaError_t e;
    if (count) {
        e = cudaFuncSetAttribute(k_trace_level_large<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        k_trace_level_large<true><<<(unsigned)grid, kLargeThreads, smem, st>>>(p, total);
    } else {
        e = cudaFuncSetAttribute(k_trace_level_large<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        k_trace_level_large<false><<<(unsigned)grid, kLargeThreads, smem, st>>>(p, total);
    }
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// accumulation extension: running FP32 sum of final levels, scaled at the end
// ------------------------------------------------------------------------------------------
__global__ void k_accumulate(float* __restrict__ acc, const float* __restrict__ img, size_t slot_stride, int frames,
                             size_t n, int first) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float a = __fadd_rn(first ? 0.0f : acc[i], img[i]);
        for (int f = 1; f < frames; f++) a = __fadd_rn(a, img[(size_t)f * slot_stride + i]);   // frame order
        acc[i] = a;
    }
}
__global__ void k_scale(float* __restrict__ acc, size_t n, float divisor) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        acc[i] = __fdiv_rn(acc[i], divisor);
    }
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

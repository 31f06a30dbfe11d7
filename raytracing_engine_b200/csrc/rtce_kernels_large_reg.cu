// rtce_kernels_large_reg.cu — objCount > 8, second version (BASELINE.json configs[3]): tracing algorithm 3
// (compute.glsl:34-68) with the per-ray x per-object cache of lower bounds (`float distances[MAX_OBJECTS]`,
// compute.glsl:35) held in REGISTERS.
//
// Same organisation as k_trace_level_large (one CTA marches one ray at a time, the objects are spread over
// its threads, the scalar ray state is computed redundantly by every thread from identical inputs), but a
// thread owns 8 * G8 objects as a fully unrolled register array instead of float4 groups in shared memory.
// Why: per march trip the algorithm does one subtract, one min and a share of one compare per object, and the
// shared-memory version paid one LDS.128 + one STS.128 per four objects on top (the shared-memory pipe, not
// instruction issue, bounded it) plus ~8 instructions of loop control per group.  Here a trip is
//     8 FADD + 7 FMNMX + FSETP + BRA + FMNMX   per group of eight objects   (2.25 instructions per object,
//                                                                            algorithmic count 3 N S: 3 per object)
// and the block-wide exact minimum is one CREDUX.MIN.F32 (sm_100a warp reduction to a uniform register) per
// warp, one barrier and eight shared-memory words.  min is exactly associative and commutative on non-NaN
// floats, and fsub never yields -0, so neither the reduction order nor the object-to-thread mapping can change
// a bit.  Object j * 256 + tid is register j of thread tid: the 128-bit sphere loads of the
// initialisation are coalesced, and 8 * G8 of them are in flight per thread.
#include "rtce_device.cuh"
#include "rtce_kernels.cuh"

namespace rtce {

constexpr int kRegThreads = 256;

RTCE_DEV float warp_min_f32(float v) {
    float r;
    asm volatile("redux.sync.min.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
    return r;
}

// utilities.glsl:36-38 with the short correctly rounded sqrt (rtce_device.cuh: same bits as sqrt.rn for
// +0 and every finite x >= 2^-101, checked exhaustively on the GPU).
RTCE_DEV float sphere_sdf_fast(float px, float py, float pz, const float4 s) {
    const float dx = fsub(px, s.x), dy = fsub(py, s.y), dz = fsub(pz, s.z);
    const float d2 = fadd(fadd(fmul(dx, dx), fmul(dy, dy)), fmul(dz, dz));
    return fsub(sqrt_fast(d2), s.w);
}

RTCE_DEV float min8(const float* v) {
    return fminf(fminf(fminf(v[0], v[1]), fminf(v[2], v[3])), fminf(fminf(v[4], v[5]), fminf(v[6], v[7])));
}

// LISTED: the rays are the entries of P.ray_list (the event-driven kernel's fallback rays) and their number is read
// from device memory.  A template parameter, not a run-time test: with the test in the loop ptxas kept the list
// pointers live across the march loop and the G8 = 5 kernel went from 20 to 104 bytes of spills (0.87 -> 1.09 s per
// configs[3] frame).
template <int G8, bool COUNT, bool LISTED>
__global__ void __launch_bounds__(kRegThreads, (G8 <= 5 ? 3 : 2))
    k_trace_level_large_reg(const __grid_constant__ LevelParamsLarge P, long long total_rays) {
    constexpr int NJ = 8 * G8;
    __shared__ float s_red[2][kRegThreads / 32];
    const int n = P.n_spheres;
    const int tid = threadIdx.x;
    const int rw = P.x1 - P.x0, rh = P.y1 - P.y0;
    const long long per_frame = (long long)rw * rh;
    const float4* __restrict__ sph = P.spheres + tid;   // object j * kRegThreads + tid
    const float inf = __int_as_float(0x7F800000);
    unsigned long long cE = 0, cS = 0, cR = 0;
    int parity = 0;   // runs across rays: consecutive block minima alternate buffers (one barrier per trip)

    if (LISTED) total_rays = (long long)*P.ray_count;
    for (long long idx = blockIdx.x; idx < total_rays; idx += gridDim.x) {
        const long long ray = LISTED ? (long long)P.ray_list[idx] : idx;
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

        // traceCone :37-39 (registers past n hold +inf: never refreshed, never the minimum)
        float d[NJ];
#pragma unroll
        for (int j = 0; j < NJ; j++) {
            d[j] = inf;
            if (j * kRegThreads + tid < n) d[j] = sphere_sdf_fast(ox, oy, oz, __ldg(sph + j * kRegThreads));
        }
        if (COUNT && tid == 0) cE += (unsigned long long)n;

        float len = 0.0f, last = 0.0f;
        while (len < P.render_dist) {   // :44
            const float radius = fmul(fadd(len, 1.0f), P.threshold);
            float m = P.render_dist;   // :49
            const float px = fadd(ox, fmul(dir.x, len));   // :46, once per trip (most trips enter several groups)
            const float py = fadd(oy, fmul(dir.y, len));
            const float pz = fadd(oz, fmul(dir.z, len));
#pragma unroll
            for (int g = 0; g < G8; g++) {   // :51-57
                float* v = d + 8 * g;
#pragma unroll
                for (int c = 0; c < 8; c++) v[c] = fsub(v[c], last);
                float m8 = min8(v);
                if (m8 <= radius) {   // some bound of the group fell inside the cone radius
#pragma unroll
                    for (int c = 0; c < 8; c++) {
                        if (v[c] <= radius) {
                            v[c] = sphere_sdf_fast(px, py, pz, __ldg(sph + (8 * g + c) * kRegThreads));
                            if (COUNT) cE++;
                        }
                    }
                    m8 = min8(v);
                }
                m = fminf(m, m8);
            }
            // block-wide exact minimum
            const float wm = warp_min_f32(m);
            if ((tid & 31) == 0) s_red[parity][tid >> 5] = wm;
            __syncthreads();
            const float4 r0 = *reinterpret_cast<const float4*>(&s_red[parity][0]);
            const float4 r1 = *reinterpret_cast<const float4*>(&s_red[parity][4]);
            parity ^= 1;
            const float dist = fminf(fminf(fminf(r0.x, r0.y), fminf(r0.z, r0.w)), fminf(fminf(r1.x, r1.y), fminf(r1.z, r1.w)));
            if (COUNT && tid == 0) cS++;
            last = glsl_max(dist, 0.0f);   // :59
            len = fadd(len, last);
            if (dist <= radius) {   // :62-65
                len = fsub(len, radius);
                break;
            }
        }
        if (tid == 0) {
            P.out[(size_t)f * P.slot_stride + (size_t)y * (size_t)P.w + (size_t)x] = glsl_max(fadd(start, len), 0.0f);
            if (COUNT) cR++;
        }
    }
    if (COUNT) {
        for (int o = 16; o > 0; o >>= 1) cE += __shfl_down_sync(0xffffffffu, cE, o);
        if ((tid & 31) == 0 && cE) atomicAdd(P.counters + 1, cE);
        if (tid == 0) {
            atomicAdd(P.counters + 0, cR);
            atomicAdd(P.counters + 2, cS);
        }
    }
}

template <int G8>
static cudaError_t launch_g8(const LevelParamsLarge& p, long long total, bool count, int sms, cudaStream_t st) {
    const int per_sm = G8 <= 5 ? 3 : 2;
    long long grid = (long long)sms * per_sm;   // persistent CTAs: one resident wave
    if (grid > total) grid = total;
    const bool listed = p.ray_list != nullptr;
    if (count && listed)
        k_trace_level_large_reg<G8, true, true><<<(unsigned)grid, kRegThreads, 0, st>>>(p, total);
    else if (count)
        k_trace_level_large_reg<G8, true, false><<<(unsigned)grid, kRegThreads, 0, st>>>(p, total);
    else if (listed)
        k_trace_level_large_reg<G8, false, true><<<(unsigned)grid, kRegThreads, 0, st>>>(p, total);
    else
        k_trace_level_large_reg<G8, false, false><<<(unsigned)grid, kRegThreads, 0, st>>>(p, total);
    return cudaGetLastError();
}

int large_reg_capacity() { return 8 * 8 * kRegThreads; }

cudaError_t launch_trace_level_large_reg(const LevelParamsLarge& p, int frames, bool count, cudaStream_t st) {
    if (p.x1 <= p.x0 || p.y1 <= p.y0 || frames <= 0) return cudaSuccess;
    if (p.n_spheres > large_reg_capacity()) return cudaErrorInvalidValue;
    const long long total = (long long)(p.x1 - p.x0) * (p.y1 - p.y0) * frames;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int g8 = (p.n_spheres + 8 * kRegThreads - 1) / (8 * kRegThreads);
    switch (g8) {
        case 0:
        case 1: return launch_g8<1>(p, total, count, sms, st);
        case 2: return launch_g8<2>(p, total, count, sms, st);
        case 3: return launch_g8<3>(p, total, count, sms, st);
        case 4: return launch_g8<4>(p, total, count, sms, st);
        case 5: return launch_g8<5>(p, total, count, sms, st);
        case 6: return launch_g8<6>(p, total, count, sms, st);
        default: return launch_g8<8>(p, total, count, sms, st);
    }
}

}  // namespace rtce

// rtce_kernels.cu — sm_100a kernels of the cone-march pass: per-level kernels.
//
//   k_trace_level_small<N>   compute.glsl:main for objCount = N <= 8, one thread per pixel,
//                            all frames of a batch in one launch (grid.z)
//   k_trace_level_large      objCount > 8 (extension): one CTA per ray, objects spread over
//                            the CTA's threads, block-wide min per march step
//
// Arithmetic: rtce_device.cuh (canonical, non-contracted).  Built with --fmad=false as a
// second line of defence; no -use_fast_math.
#include "rtce_march.cuh"

namespace rtce {

// ------------------------------------------------------------------------------------------
// compute.glsl:34-68  traceCone — tracing algorithm 3, distances[] in registers (N static)
// ------------------------------------------------------------------------------------------
template <int N, bool COUNT>
RTCE_DEV float trace_cone_small(float ox, float oy, float oz, float sx, float sy, float sz, float threshold,
                                float render_dist, const float4* __restrict__ sph, unsigned& cntE, unsigned& cntS) {
    float d[N > 0 ? N : 1];
#pragma unroll
    for (int i = 0; i < N; i++) {   // :37-39
        d[i] = sphere_sdf(ox, oy, oz, sph[i].x, sph[i].y, sph[i].z, sph[i].w);
    }
    if (COUNT) cntE += N;
    float len = 0.0f;    // :41
    float last = 0.0f;   // :42
    while (len < render_dist) {   // :44
        const float px = fadd(ox, fmul(sx, len));   // :45
        const float py = fadd(oy, fmul(sy, len));
        const float pz = fadd(oz, fmul(sz, len));
        float dist = render_dist;                              // :49
        const float radius = fmul(fadd(len, 1.0f), threshold); // :50
#pragma unroll
        for (int i = 0; i < N; i++) {   // :51-57
            d[i] = fsub(d[i], last);
            if (d[i] <= radius) {
                d[i] = sphere_sdf(px, py, pz, sph[i].x, sph[i].y, sph[i].z, sph[i].w);
                if (COUNT) cntE++;
            }
            dist = glsl_min(dist, d[i]);
        }
        if (COUNT) cntS++;
        last = glsl_max(dist, 0.0f);   // :59
        len = fadd(len, last);         // :60
        if (dist <= radius) {          // :62-65
            len = fsub(len, radius);
            break;
        }
    }
    return len;   // :67
}

template <bool COUNT>
RTCE_DEV void flush_counters(unsigned long long* counters, unsigned r, unsigned e, unsigned s) {
    if (!COUNT) return;
    // warp-aggregate, then one atomic per warp and counter
    for (int o = 16; o > 0; o >>= 1) {
        r += __shfl_down_sync(0xffffffffu, r, o);
        e += __shfl_down_sync(0xffffffffu, e, o);
        s += __shfl_down_sync(0xffffffffu, s, o);
    }
    if ((threadIdx.x + threadIdx.y * blockDim.x) % 32 == 0) {
        atomicAdd(counters + 0, (unsigned long long)r);
        atomicAdd(counters + 1, (unsigned long long)e);
        atomicAdd(counters + 2, (unsigned long long)s);
    }
}

// ------------------------------------------------------------------------------------------
// compute.glsl:70-87  main() — one thread per pixel of level P.iter, frame = blockIdx.z
// ------------------------------------------------------------------------------------------
constexpr int kTileX = 16, kTileY = 16;

template <int N, bool COUNT>
__global__ void __launch_bounds__(kTileX* kTileY)
    k_trace_level_small(const __grid_constant__ LevelParams P) {
    const int x = P.x0 + blockIdx.x * kTileX + threadIdx.x;
    const int y = P.y0 + blockIdx.y * kTileY + threadIdx.y;
    const int f = blockIdx.z;
    unsigned cR = 0, cE = 0, cS = 0;
    if (x < P.x1 && y < P.y1) {
        const Cam& cam = P.cam[f];
        // :71-72  normCoord = ((gid*2+1) * imageSize - 1.0) * ratio
        const float ncx = fmul(fsub(fmul((float)(2u * (unsigned)x + 1u), P.isx), 1.0f), P.ratx);
        const float ncy = fmul(fsub(fmul((float)(2u * (unsigned)y + 1u), P.isy), 1.0f), P.raty);
        // :77  step = normalize(rotate(rot, vec3(nc.x, 1.0, nc.y)))
        const Vec3 dir = normalize3(rotate_q(cam.rot[0], cam.rot[1], cam.rot[2], cam.rot[3], ncx, 1.0f, ncy));
        // :79-82  start distance: 1.0 at level 0, else the parent texel ivec2(gid*0.5) of level iter-1
        float len = 1.0f;
        if (P.iter > 0) {
            len = __ldg(P.parent + (size_t)f * P.slot_stride + (size_t)(y >> 1) * (size_t)P.pw + (size_t)(x >> 1));
        }
        // :84  len += traceCone(pos + step*len, step, threshold)
        const float ox = fadd(cam.pos[0], fmul(dir.x, len));
        const float oy = fadd(cam.pos[1], fmul(dir.y, len));
        const float oz = fadd(cam.pos[2], fmul(dir.z, len));
        len = fadd(len, trace_cone_small<N, COUNT>(ox, oy, oz, dir.x, dir.y, dir.z, P.threshold, P.render_dist,
                                                   P.sph, cE, cS));
        // :86  imageStore(max(len, 0))
        P.out[(size_t)f * P.slot_stride + (size_t)y * (size_t)P.w + (size_t)x] = glsl_max(len, 0.0f);
        cR = 1;
    }
    flush_counters<COUNT>(P.counters, cR, cE, cS);
}

template <int N>
static cudaError_t launch_small_n(const LevelParams& p, int frames, bool count, cudaStream_t st) {
    dim3 block(kTileX, kTileY, 1);
    dim3 grid((unsigned)((p.x1 - p.x0 + kTileX - 1) / kTileX), (unsigned)((p.y1 - p.y0 + kTileY - 1) / kTileY),
              (unsigned)frames);
    if (count)
        k_trace_level_small<N, true><<<grid, block, 0, st>>>(p);
    else
        k_trace_level_small<N, false><<<grid, block, 0, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_trace_level_small(const LevelParams& p, int frames, bool count, cudaStream_t st) {
    if (p.x1 <= p.x0 || p.y1 <= p.y0 || frames <= 0) return cudaSuccess;
    switch (p.n_spheres) {
        case 0: return launch_small_n<0>(p, frames, count, st);
        case 1: return launch_small_n<1>(p, frames, count, st);
        case 2: return launch_small_n<2>(p, frames, count, st);
        case 3: return launch_small_n<3>(p, frames, count, st);
        case 4: return launch_small_n<4>(p, frames, count, st);
        case 5: return launch_small_n<5>(p, frames, count, st);
        case 6: return launch_small_n<6>(p, frames, count, st);
        case 7: return launch_small_n<7>(p, frames, count, st);
        case 8: return launch_small_n<8>(p, frames, count, st);
        default: return cudaErrorInvalidValue;
    }
}

// ------------------------------------------------------------------------------------------
// The author's alternative marchers (shaders/tracing_algorithms.txt:2-13 "algorithm 1", :16-37 "algorithm 2") dropped
// into the traceCone frame of compute.glsl:34-68, selectable with RTCE_OPT_MARCHER.  The reference holds only
// pseudo-code for them: the checker is a restatement (oracle/rtce_oracle.c), parity UNPINNED by the reference.
// One thread per pixel per level.  kAltMaxTrips: the text has no trip limit; a ray still marching after that many
// trips is cut off, in the checker and here alike.
// ------------------------------------------------------------------------------------------
constexpr unsigned kAltMaxTrips = 65536u;

// distances[closest] for a dynamic index over a register array (N <= 8: a select chain, no local memory)
template <int N>
RTCE_DEV float pick(const float* d, unsigned k) {
    float v = d[0];
#pragma unroll
    for (int i = 1; i < N; i++) v = (k == (unsigned)i) ? d[i] : v;
    return v;
}

template <int ALG, int N, bool COUNT>
RTCE_DEV float trace_cone_alt(float ox, float oy, float oz, float sx, float sy, float sz, float threshold, float render_dist,
                              const float4* __restrict__ sph, unsigned& cntE, unsigned& cntS) {
    float len = 0.0f;
    unsigned trips = 0;
    if (ALG == 1) {
        while (len < render_dist && trips++ < kAltMaxTrips) {
            const float px = fadd(ox, fmul(sx, len)), py = fadd(oy, fmul(sy, len)), pz = fadd(oz, fmul(sz, len));   // frame :45
            float dist = sphere_sdf(px, py, pz, sph[0].x, sph[0].y, sph[0].z, sph[0].w);                         // txt:2
#pragma unroll
            for (int i = 1; i < N; i++) dist = glsl_min(dist, sphere_sdf(px, py, pz, sph[i].x, sph[i].y, sph[i].z, sph[i].w));   // txt:3-5
            if (COUNT) { cntE += N; cntS++; }
            len = fadd(len, dist);                                        // txt:7
            const float radius = fmul(fadd(len, 1.0f), threshold);        // txt:8
            if (dist <= radius) {                                         // txt:9-12
                len = fsub(len, radius);
                break;
            }
        }
    } else {
        float d[N > 0 ? N : 1];
#pragma unroll
        for (int i = 0; i < N; i++) d[i] = sphere_sdf(ox, oy, oz, sph[i].x, sph[i].y, sph[i].z, sph[i].w);   // frame :37-39
        if (COUNT) cntE += N;
        unsigned closest = 0;    // txt:19 (outside the loop)
        float nearest = 0.0f;    // txt:20
        while (len < render_dist && trips++ < kAltMaxTrips) {
            const float px = fadd(ox, fmul(sx, len)), py = fadd(oy, fmul(sy, len)), pz = fadd(oz, fmul(sz, len));   // frame :45
#pragma unroll
            for (int i = 0; i < N; i++) {   // txt:22-27, in this order: distances[closest] may not have been decremented yet
                d[i] = fsub(d[i], nearest);
                if (d[i] < pick<N>(d, closest)) closest = (unsigned)i;
            }
            nearest = pick<N>(d, closest);                                // txt:29
            len = fadd(len, nearest);                                     // txt:30
            float4 sc = sph[0];
#pragma unroll
            for (int i = 1; i < N; i++) sc = (closest == (unsigned)i) ? sph[i] : sc;
            const float fresh = sphere_sdf(px, py, pz, sc.x, sc.y, sc.z, sc.w);   // txt:31
#pragma unroll
            for (int i = 0; i < N; i++) d[i] = (closest == (unsigned)i) ? fresh : d[i];
            if (COUNT) { cntE++; cntS++; }
            const float radius = fmul(fadd(len, 1.0f), threshold);        // txt:33
            if (fresh <= radius) {                                        // txt:34-37
                len = fadd(len, fsub(fresh, radius));
                break;
            }
        }
    }
    return len;
}

template <int ALG, int N, bool COUNT>
__global__ void __launch_bounds__(kTileX* kTileY)
    k_trace_level_alt(const __grid_constant__ LevelParams P) {
    const int x = P.x0 + blockIdx.x * kTileX + threadIdx.x;
    const int y = P.y0 + blockIdx.y * kTileY + threadIdx.y;
    const int f = blockIdx.z;
    unsigned cR = 0, cE = 0, cS = 0;
    if (x < P.x1 && y < P.y1) {
        const Cam& cam = P.cam[f];
        const float ncx = fmul(fsub(fmul((float)(2u * (unsigned)x + 1u), P.isx), 1.0f), P.ratx);   // compute.glsl:71-72
        const float ncy = fmul(fsub(fmul((float)(2u * (unsigned)y + 1u), P.isy), 1.0f), P.raty);
        const Vec3 dir = normalize3(rotate_q(cam.rot[0], cam.rot[1], cam.rot[2], cam.rot[3], ncx, 1.0f, ncy));   // :77
        float len = 1.0f;   // :79-82
        if (P.iter > 0) len = __ldg(P.parent + (size_t)f * P.slot_stride + (size_t)(y >> 1) * (size_t)P.pw + (size_t)(x >> 1));
        const float ox = fadd(cam.pos[0], fmul(dir.x, len));
        const float oy = fadd(cam.pos[1], fmul(dir.y, len));
        const float oz = fadd(cam.pos[2], fmul(dir.z, len));
        len = fadd(len, trace_cone_alt<ALG, N, COUNT>(ox, oy, oz, dir.x, dir.y, dir.z, P.threshold, P.render_dist, P.sph, cE, cS));   // :84
        P.out[(size_t)f * P.slot_stride + (size_t)y * (size_t)P.w + (size_t)x] = glsl_max(len, 0.0f);   // :86
        cR = 1;
    }
    flush_counters<COUNT>(P.counters, cR, cE, cS);
}

template <int ALG, int N>
static cudaError_t launch_alt_n(const LevelParams& p, int frames, bool count, cudaStream_t st) {
    dim3 block(kTileX, kTileY, 1);
    dim3 grid((unsigned)((p.x1 - p.x0 + kTileX - 1) / kTileX), (unsigned)((p.y1 - p.y0 + kTileY - 1) / kTileY), (unsigned)frames);
    if (count)
        k_trace_level_alt<ALG, N, true><<<grid, block, 0, st>>>(p);
    else
        k_trace_level_alt<ALG, N, false><<<grid, block, 0, st>>>(p);
    return cudaGetLastError();
}

template <int ALG>
static cudaError_t launch_alt(const LevelParams& p, int frames, bool count, cudaStream_t st) {
    switch (p.n_spheres) {
        case 1: return launch_alt_n<ALG, 1>(p, frames, count, st);
        case 2: return launch_alt_n<ALG, 2>(p, frames, count, st);
        case 3: return launch_alt_n<ALG, 3>(p, frames, count, st);
        case 4: return launch_alt_n<ALG, 4>(p, frames, count, st);
        case 5: return launch_alt_n<ALG, 5>(p, frames, count, st);
        case 6: return launch_alt_n<ALG, 6>(p, frames, count, st);
        case 7: return launch_alt_n<ALG, 7>(p, frames, count, st);
        case 8: return launch_alt_n<ALG, 8>(p, frames, count, st);
        default: return cudaErrorInvalidValue;   // algorithm 1 reads objs[0] unconditionally: objCount >= 1
    }
}

cudaError_t launch_trace_level_alt(const LevelParams& p, int frames, bool count, int marcher, cudaStream_t st) {
    if (p.x1 <= p.x0 || p.y1 <= p.y0 || frames <= 0) return cudaSuccess;
    if (marcher == 1) return launch_alt<1>(p, frames, count, st);
    if (marcher == 2) return launch_alt<2>(p, frames, count, st);
    return cudaErrorInvalidValue;
}

// ------------------------------------------------------------------------------------------
// Coarse levels (a few thousand rays per frame, but up to ~40 dependent march trips per ray): one
// thread per pixel, 8x4 pixels per warp, marched to the end in place with the branch-free trip
// (march_block) — these launches are bound by the latency of the longest ray, not by issue.
// Requires RENDER_DIST > 0.
// ------------------------------------------------------------------------------------------
template <int N, bool COUNT>
__global__ void __launch_bounds__(256)
    k_trace_level_pix(const __grid_constant__ LevelParams P) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int x = P.x0 + blockIdx.x * 32 + (warp & 3) * 8 + (lane & 7);
    const int y = P.y0 + blockIdx.y * 8 + (warp >> 2) * 4 + (lane >> 3);
    const int f = blockIdx.z;
    unsigned cR = 0, cE = 0, cS = 0, cF = 0;
    if (x < P.x1 && y < P.y1) {
        Cam cam;
        {
            const float4 q = __ldg(reinterpret_cast<const float4*>(P.cams + f));
            const float4 t = __ldg(reinterpret_cast<const float4*>(P.cams + f) + 1);
            cam.rot[0] = q.x; cam.rot[1] = q.y; cam.rot[2] = q.z; cam.rot[3] = q.w;
            cam.pos[0] = t.x; cam.pos[1] = t.y; cam.pos[2] = t.z; cam.pad = t.w;   // pad = far-miss distance
        }
        float start = 1.0f;   // compute.glsl:79
        if (P.iter > 0) start = __ldg(P.parent + (size_t)f * P.slot_stride + (size_t)(y >> 1) * (size_t)P.pw + (size_t)(x >> 1));
        float len = P.render_dist;
        if (P.iter > 0 && start >= cam.pad) {
            // beyond every sphere by more than RENDER_DIST: one trip with dist = RENDER_DIST, no hit (see strip kernel)
            if (COUNT) { cE += (unsigned)N; cS++; cF = 1; }
        } else {
            const float4 col = col_terms(cam, norm_coord(x, P.isx, P.ratx));
            const float4 row = row_terms(cam, norm_coord(y, P.isy, P.raty));
            Ray<N> ray;
            ray_setup<N>(ray, col, row, start, cam, P.sph);
            bool go = first_trip<N, COUNT>(ray, len, P.threshold, P.render_dist, cE, cS);
            float last = len;
            while (go) go = march_block<N, COUNT, RTCE_PRED_BLOCK>(ray, len, last, P.threshold, P.render_dist, P.sph, cE, cS);
        }
        P.out[(size_t)f * P.slot_stride + (size_t)y * (size_t)P.w + (size_t)x] = glsl_max(fadd(start, len), 0.0f);   // :84, :86
        cR = 1;
    }
    if (COUNT) {
        for (int o = 16; o > 0; o >>= 1) {
            cR += __shfl_down_sync(0xffffffffu, cR, o);
            cE += __shfl_down_sync(0xffffffffu, cE, o);
            cS += __shfl_down_sync(0xffffffffu, cS, o);
            cF += __shfl_down_sync(0xffffffffu, cF, o);
        }
        if (lane == 0) {
            atomicAdd(P.counters + 0, (unsigned long long)cR);
            atomicAdd(P.counters + 1, (unsigned long long)cE);
            atomicAdd(P.counters + 2, (unsigned long long)cS);
            if (cF) atomicAdd(P.counters + 3, (unsigned long long)cF);
        }
    }
}

template <int N>
static cudaError_t launch_pix_n(const LevelParams& p, int frames, bool count, cudaStream_t st) {
    dim3 grid((unsigned)((p.x1 - p.x0 + 31) / 32), (unsigned)((p.y1 - p.y0 + 7) / 8), (unsigned)frames);
    if (count)
        k_trace_level_pix<N, true><<<grid, 256, 0, st>>>(p);
    else
        k_trace_level_pix<N, false><<<grid, 256, 0, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_trace_level_pix(const LevelParams& p, int frames, bool count, cudaStream_t st) {
    if (p.x1 <= p.x0 || p.y1 <= p.y0 || frames <= 0) return cudaSuccess;
    if (!(p.render_dist > 0.0f)) return cudaErrorInvalidValue;
    switch (p.n_spheres) {
        case 0: return launch_pix_n<0>(p, frames, count, st);
        case 1: return launch_pix_n<1>(p, frames, count, st);
        case 2: return launch_pix_n<2>(p, frames, count, st);
        case 3: return launch_pix_n<3>(p, frames, count, st);
        case 4: return launch_pix_n<4>(p, frames, count, st);
        case 5: return launch_pix_n<5>(p, frames, count, st);
        case 6: return launch_pix_n<6>(p, frames, count, st);
        case 7: return launch_pix_n<7>(p, frames, count, st);
        case 8: return launch_pix_n<8>(p, frames, count, st);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace rtce

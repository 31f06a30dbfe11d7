// rtce_kernels_coarse.cu — the coarse pyramid levels of a frame batch in ONE launch.
//
// The first levels of the pyramid (main.rs:300-316, iter = 0 .. k-1) hold a few thousand rays per
// frame, but every level needs the one before it (compute.glsl:80-82) and a grazing cone marches
// 20 .. 50 dependent trips, so a launch per level costs ~10 us of launch + drain latency each for
// microseconds of work.  Here ONE thread-block cluster owns one frame: its CTAs (up to 8 SMs)
// trace level after level with a hardware cluster barrier in between, the parent level read back
// from L2.  No tile is recomputed and no level leaves the launch: per level the cost is the
// dependent chain of its longest ray.
//
// Pixel arithmetic: the shared body of rtce_march.cuh (bit-identical to the per-level kernels).
#include <cooperative_groups.h>

#include "rtce_device.cuh"
#include "rtce_kernels.cuh"
#include "rtce_march.cuh"

namespace cg = cooperative_groups;

namespace rtce {

namespace {

template <int N, bool COUNT>
__global__ void __launch_bounds__(kCoarseThreads)
    k_trace_coarse_cluster(const __grid_constant__ FusedParams P) {
    cg::cluster_group cluster = cg::this_cluster();
    const unsigned csize = cluster.num_blocks();
    const int f = blockIdx.x / csize;                      // one cluster per frame
    const unsigned tid = cluster.block_rank() * blockDim.x + threadIdx.x;
    const unsigned nthr = csize * blockDim.x;
    const int lane = threadIdx.x & 31;
    float* const slot = P.base + (size_t)f * P.slot_stride;
    Cam cam;
    {
        const float4 q = __ldg(reinterpret_cast<const float4*>(P.cams + f));
        const float4 t = __ldg(reinterpret_cast<const float4*>(P.cams + f) + 1);
        cam.rot[0] = q.x; cam.rot[1] = q.y; cam.rot[2] = q.z; cam.rot[3] = q.w;
        cam.pos[0] = t.x; cam.pos[1] = t.y; cam.pos[2] = t.z; cam.pad = 0.0f;
    }
    for (int i = 0; i < P.levels; i++) {
        const FusedLevel& L = P.lv[i];
        const int rw = L.rx1 - L.rx0, rh = L.ry1 - L.ry0;
        // 8x4-pixel warp tiles (a ray's trip count follows the silhouettes: compact tiles diverge least)
        const int tw = (rw + 7) >> 3, th = (rh + 3) >> 2;
        const unsigned npx = (unsigned)(tw * th) * 32u;
        float* const out = slot + L.off;
        const float* const par = i > 0 ? slot + P.lv[i - 1].off : nullptr;
        const int pw = i > 0 ? P.lv[i - 1].w : 0;
        unsigned cR = 0, cE = 0, cS = 0;
        for (unsigned p = tid; p < npx; p += nthr) {
            const int tile = (int)(p >> 5);
            const int x = L.rx0 + (tile % tw) * 8 + (lane & 7);
            const int y = L.ry0 + (tile / tw) * 4 + (lane >> 3);
            if (x < L.rx1 && y < L.ry1) {
                const float4 col = col_terms(cam, norm_coord(x, L.isx, P.ratx));
                const float4 row = row_terms(cam, norm_coord(y, L.isy, P.raty));
                float start = 1.0f;   // compute.glsl:79
                if (i > 0) start = __ldcg(par + (size_t)(y >> 1) * (size_t)pw + (size_t)(x >> 1));   // written in this launch: L2
                Ray<N> ray;
                ray_setup<N>(ray, col, row, start, cam, P.sph);
                float len = 0.0f;
                bool go = first_step<N, COUNT>(ray, len, L.threshold, P.render_dist, cE, cS);
                float last = len;
                while (go) go = march_block_pred<N, COUNT, RTCE_PRED_BLOCK>(ray, len, last, L.threshold, P.render_dist, P.sph, cE, cS);
                __stcg(out + (size_t)y * (size_t)L.w + (size_t)x, glsl_max(fadd(start, len), 0.0f));   // :84, :86
                cR++;
            }
        }
        if (COUNT) {
            for (int o = 16; o > 0; o >>= 1) {
                cR += __shfl_down_sync(0xffffffffu, cR, o);
                cE += __shfl_down_sync(0xffffffffu, cE, o);
                cS += __shfl_down_sync(0xffffffffu, cS, o);
            }
            if (lane == 0 && cR) {
                unsigned long long* c = P.counters + 4 * (i - (P.levels - 1));   // P.counters is the finest level's block
                atomicAdd(c + 0, (unsigned long long)cR);
                atomicAdd(c + 1, (unsigned long long)cE);
                atomicAdd(c + 2, (unsigned long long)cS);
            }
        }
        if (i + 1 < P.levels) {
            __threadfence();   // the level's texels are in L2 before any CTA of the cluster reads them
            cluster.sync();
        }
    }
}

template <int N>
cudaError_t launch_coarse_n(const FusedParams& p, int frames, int csize, bool count, cudaStream_t st) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(frames * csize));
    cfg.blockDim = dim3(kCoarseThreads);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned)csize;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    if (count) return cudaLaunchKernelEx(&cfg, k_trace_coarse_cluster<N, true>, p);
    return cudaLaunchKernelEx(&cfg, k_trace_coarse_cluster<N, false>, p);
}

}  // namespace

cudaError_t launch_trace_coarse(const FusedParams& p, int frames, bool count, cudaStream_t st) {
    if (frames <= 0 || p.levels <= 0) return cudaSuccess;
    if (!(p.render_dist > 0.0f) || !p.cams) return cudaErrorInvalidValue;
    // cluster size: enough threads for one pixel each on the finest level of the launch, at most 8 CTAs
    // (the portable limit); small views take fewer SMs per frame
    const FusedLevel& L = p.lv[p.levels - 1];
    const long long px = (long long)(L.rx1 - L.rx0) * (L.ry1 - L.ry0);
    int csize = 1;
    while (csize < 8 && (long long)csize * kCoarseThreads < px) csize <<= 1;
    switch (p.n_spheres) {
        case 0: return launch_coarse_n<0>(p, frames, csize, count, st);
        case 1: return launch_coarse_n<1>(p, frames, csize, count, st);
        case 2: return launch_coarse_n<2>(p, frames, csize, count, st);
        case 3: return launch_coarse_n<3>(p, frames, csize, count, st);
        case 4: return launch_coarse_n<4>(p, frames, csize, count, st);
        case 5: return launch_coarse_n<5>(p, frames, csize, count, st);
        case 6: return launch_coarse_n<6>(p, frames, csize, count, st);
        case 7: return launch_coarse_n<7>(p, frames, csize, count, st);
        case 8: return launch_coarse_n<8>(p, frames, csize, count, st);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace rtce

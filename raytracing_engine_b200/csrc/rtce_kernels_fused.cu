// rtce_kernels_fused.cu — the whole pass (src/main.rs:300-316: one dispatch per pyramid level) as
// ONE kernel launch per frame batch, for objCount <= 8.
//
// Work decomposition (DESIGN.md "Fused pyramid kernel"):
//   CTA (tx, ty, frame) owns the 64x64 final-level tile at (gx0 + 64 tx, gy0 + 64 ty) and walks the
//   pyramid coarse -> fine.  At level l (shift s = levels-1-l) its region is the tile footprint
//   >> s (at least one pixel), clipped to the level's region of interest; the last tile column / row
//   extend to the level border so that the "orphan" texels the reference also computes (level sizes
//   are rounded up to 8, main.rs:211) are covered.  Each level's results go to global memory (the
//   level image, binding 0) and, as the next level's parents, to shared memory.
//   A warp takes 32 consecutive x of one region row: coalesced stores, and all terms of the ray
//   direction that depend on x only / y only are precomputed once per region column / row.
//
// Arithmetic: identical rounding sequence to compute.glsl:70-87 as pinned in rtce_device.cuh.  The
// separable terms below are sub-expressions of rotate() hoisted out of the pixel, not re-associated.
#include "rtce_march.cuh"

namespace rtce {

// Region of a tile at one level, computed once per CTA (thread l computes level l).
struct Region {
    int rx0, ry0, rw, rh;     // first texel and size (may be <= 0: nothing to do)
    int own_x, own_y;         // a texel is written / counted by this tile iff x >= own_x && y >= own_y
    int pad0, pad1;
};

RTCE_DEV Region tile_region(const FusedParams& P, int l) {
    const int s = P.levels - 1 - l;
    const FusedLevel& L = P.lv[l];
    const int X0 = P.gx0 + (int)blockIdx.x * kFusedTile, Y0 = P.gy0 + (int)blockIdx.y * kFusedTile;
    const bool last_col = (int)blockIdx.x == P.ntx - 1, last_row = (int)blockIdx.y == P.nty - 1;
    // tile footprint at this level (>= 1 texel); the last tile column / row extend to the level border
    const int fx0 = X0 >> s, fy0 = Y0 >> s;
    const int fx1 = last_col ? 0x7fffffff : max((X0 + kFusedTile) >> s, fx0 + 1);
    const int fy1 = last_row ? 0x7fffffff : max((Y0 + kFusedTile) >> s, fy0 + 1);
    Region r;
    r.rx0 = max(fx0, L.rx0);
    r.ry0 = max(fy0, L.ry0);
    r.rw = min(fx1, L.rx1) - r.rx0;
    r.rh = min(fy1, L.ry1) - r.ry0;
    // a texel shared with the previous tile column / row (only possible when 64 >> s == 0) belongs to
    // the first tile that maps to it
    const bool own_x0 = blockIdx.x == 0 || ((X0 - kFusedTile) >> s) != fx0;
    const bool own_y0 = blockIdx.y == 0 || ((Y0 - kFusedTile) >> s) != fy0;
    r.own_x = own_x0 ? fx0 : fx0 + 1;
    r.own_y = own_y0 ? fy0 : fy0 + 1;
    r.pad0 = r.pad1 = 0;
    return r;
}

// Per-warp compaction queue (shared memory, structure of arrays, 64 entries):
// word 0-5 origin / step, 6 len, 7 start, 8 packed texel (xx | yy << 8 | owned << 30), 9.. distances[].
constexpr int kQueueCap = 64;
#define RTCE_QUEUE_WORDS(N) (9 + (N))

template <int N>
RTCE_DEV void queue_push(float* q, int idx, const Ray<N>& r, float len, float start, int pix) {
    q[0 * kQueueCap + idx] = r.ox;
    q[1 * kQueueCap + idx] = r.oy;
    q[2 * kQueueCap + idx] = r.oz;
    q[3 * kQueueCap + idx] = r.sx;
    q[4 * kQueueCap + idx] = r.sy;
    q[5 * kQueueCap + idx] = r.sz;
    q[6 * kQueueCap + idx] = len;
    q[7 * kQueueCap + idx] = start;
    q[8 * kQueueCap + idx] = __int_as_float(pix);
#pragma unroll
    for (int i = 0; i < N; i++) q[(9 + i) * kQueueCap + idx] = r.d[i];
}

template <int N>
RTCE_DEV void queue_pop(const float* q, int idx, Ray<N>& r, float& len, float& start, int& pix) {
    r.ox = q[0 * kQueueCap + idx];
    r.oy = q[1 * kQueueCap + idx];
    r.oz = q[2 * kQueueCap + idx];
    r.sx = q[3 * kQueueCap + idx];
    r.sy = q[4 * kQueueCap + idx];
    r.sz = q[5 * kQueueCap + idx];
    len = q[6 * kQueueCap + idx];
    start = q[7 * kQueueCap + idx];
    pix = __float_as_int(q[8 * kQueueCap + idx]);
#pragma unroll
    for (int i = 0; i < N; i++) r.d[i] = q[(9 + i) * kQueueCap + idx];
}

template <int N, bool COUNT>
__global__ void __launch_bounds__(kFusedThreads)
    k_trace_fused(const __grid_constant__ FusedParams P) {
    __shared__ Region s_reg[kMaxLevels];
    __shared__ float4 s_col[2][kFusedMaxSide];
    __shared__ float4 s_row[2][kFusedMaxSide];
    __shared__ float s_odd[kFusedOddPitch * kFusedOddPitch];
    __shared__ float s_even[kFusedEvenPitch * kFusedEvenPitch];
    extern __shared__ float s_queue[];   // kFusedThreads/32 warps x RTCE_QUEUE_WORDS(N) x kQueueCap

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const Cam& cam = P.cam[blockIdx.z];
    float* const frame = P.base + (size_t)blockIdx.z * P.slot_stride;
    const float rd = P.render_dist;
    unsigned cR = 0, cE = 0, cS = 0;

    if (tid < P.levels) s_reg[tid] = tile_region(P, tid);
    __syncthreads();

    // ---- leading levels with <= 32 texels: warp 0 alone, one texel per lane, no block barriers ----
    int nsmall = 0;
    while (nsmall < P.levels && max(s_reg[nsmall].rw, 0) * max(s_reg[nsmall].rh, 0) <= 32) nsmall++;
    if (warp == 0) {
        for (int l = 0; l < nsmall; l++) {
            const int s = P.levels - 1 - l;
            const Region R = s_reg[l];
            const FusedLevel& L = P.lv[l];
            float* cbuf = (s & 1) ? s_odd : s_even;
            const int cpitch = (s & 1) ? kFusedOddPitch : kFusedEvenPitch;
            if (R.rw > 0 && R.rh > 0 && lane < R.rw * R.rh) {
                const int yy = lane / R.rw, xx = lane - yy * R.rw;
                const int x = R.rx0 + xx, y = R.ry0 + yy;
                const float4 col = col_terms(cam, norm_coord(x, L.isx, P.ratx));
                const float4 row = row_terms(cam, norm_coord(y, L.isy, P.raty));
                float start = 1.0f;   // compute.glsl:79
                if (l > 0) {          // :80-82
                    const Region Q = s_reg[l - 1];
                    const float* pbuf = ((s + 1) & 1) ? s_odd : s_even;
                    const int ppitch = ((s + 1) & 1) ? kFusedOddPitch : kFusedEvenPitch;
                    start = pbuf[((y >> 1) - Q.ry0) * ppitch + ((x >> 1) - Q.rx0)];
                }
                unsigned e = 0, st = 0;
                const float v = shade_pixel<N, COUNT>(col, row, start, cam, L.threshold, rd, P.sph, e, st);
                if (x >= R.own_x && y >= R.own_y) {
                    frame[L.off + (size_t)y * (size_t)L.w + (size_t)x] = v;   // :86
                    if (COUNT) { cR++; cE += e; cS += st; }
                }
                if (s > 0) cbuf[yy * cpitch + xx] = v;
            }
            __syncwarp();
        }
    }

    // ---- remaining levels: all warps; rays that leave the first march step unfinished are
    //      compacted through the warp's queue so that the march loop runs with full warps ----
    float* const q = s_queue + warp * (RTCE_QUEUE_WORDS(N) * kQueueCap);
    for (int l = nsmall; l < P.levels; l++) {
        const int s = P.levels - 1 - l;
        const Region R = s_reg[l];
        const FusedLevel& L = P.lv[l];
        const int tb = l & 1;
        if (tid < R.rw) {
            s_col[tb][tid] = col_terms(cam, norm_coord(R.rx0 + tid, L.isx, P.ratx));
        } else if (tid >= 128 && tid - 128 < R.rh) {
            s_row[tb][tid - 128] = row_terms(cam, norm_coord(R.ry0 + tid - 128, L.isy, P.raty));
        }
        __syncthreads();   // tables ready; the previous level's parent buffer is complete

        float* const cbuf = (s & 1) ? s_odd : s_even;
        const int cpitch = (s & 1) ? kFusedOddPitch : kFusedEvenPitch;
        const float* pbuf = ((s + 1) & 1) ? s_odd : s_even;
        const int ppitch = ((s + 1) & 1) ? kFusedOddPitch : kFusedEvenPitch;
        int prx0 = 0, pry0 = 0;
        if (l > 0) {
            prx0 = s_reg[l - 1].rx0;
            pry0 = s_reg[l - 1].ry0;
        }
        float* const img = frame + L.off;
        const float thr = L.threshold;

        // lane -> texel mapping: lx = min(32, pow2ceil(rw)) lanes along x, 32/lx rows per warp pass
        int lxs = 5;
        if (R.rw <= 16) lxs = 4;
        if (R.rw <= 8) lxs = 3;
        if (R.rw <= 4) lxs = 2;
        if (R.rw <= 2) lxs = 1;
        if (R.rw <= 1) lxs = 0;
        const int lx = 1 << lxs, rows_pw = 32 >> lxs;
        const int xx_l = lane & (lx - 1), yy_l = lane >> lxs;
        int qn = 0;   // entries in this warp's queue (warp-uniform)

        // finish a popped ray and store it
        auto finish = [&](Ray<N>& r, float len, float start, int pix) {
            unsigned e = 0, st = 0;
            len = march_rest<N, COUNT>(r, len, thr, rd, P.sph, e, st);
            const float v = glsl_max(fadd(start, len), 0.0f);   // :84, :86
            const int xx = pix & 0xff, yy = (pix >> 8) & 0xff;
            if (pix & (1 << 30)) {
                img[(R.ry0 + yy) * L.w + (R.rx0 + xx)] = v;
                if (COUNT) { cE += e; cS += st; }
            }
            if (s > 0) cbuf[yy * cpitch + xx] = v;
        };

        for (int xs = 0; xs < R.rw; xs += lx) {
            const int xx = xs + xx_l;
            const bool xok = xx < R.rw;
            float4 col = make_float4(0.f, 0.f, 0.f, 0.f);
            if (xok) col = s_col[tb][xx];
            const int x = R.rx0 + xx;
            for (int yb = warp * rows_pw; yb < R.rh; yb += (kFusedThreads / 32) * rows_pw) {   // warp-uniform trip count
                const int yy = yb + yy_l;
                bool cont = false;
                Ray<N> r;
                float len = 0.0f, start = 1.0f;   // compute.glsl:79
                int pix = 0;
                if (xok && yy < R.rh) {
                    const float4 row = s_row[tb][yy];
                    const int y = R.ry0 + yy;
                    if (l > 0) start = pbuf[((y >> 1) - pry0) * ppitch + ((x >> 1) - prx0)];   // :80-82
                    ray_setup<N>(r, col, row, start, cam, P.sph);
                    unsigned e = 0, st = 0;
                    cont = first_step<N, COUNT>(r, len, thr, rd, e, st);
                    const bool owned = x >= R.own_x && y >= R.own_y;
                    if (COUNT && owned) { cR++; cE += e; cS += st; }
                    if (!cont) {
                        const float v = glsl_max(fadd(start, len), 0.0f);   // :84, :86
                        if (owned) img[y * L.w + x] = v;
                        if (s > 0) cbuf[yy * cpitch + xx] = v;
                    }
                    pix = xx | (yy << 8) | (owned ? (1 << 30) : 0);
                }
                const unsigned m = __ballot_sync(0xffffffffu, cont);
                if (m) {
                    if (cont) queue_push<N>(q, qn + __popc(m & ((1u << lane) - 1u)), r, len, start, pix);
                    qn += __popc(m);
                    __syncwarp();
                    if (qn >= 32) {   // a full warp of unfinished rays: march them together
                        qn -= 32;
                        queue_pop<N>(q, qn + lane, r, len, start, pix);
                        finish(r, len, start, pix);
                        __syncwarp();
                    }
                }
            }
        }
        if (qn > 0) {   // flush the partial warp left at the end of the level
            if (lane < qn) {
                Ray<N> r;
                float len, start;
                int pix;
                queue_pop<N>(q, lane, r, len, start, pix);
                finish(r, len, start, pix);
            }
            __syncwarp();
        }
    }
    if (COUNT) {
        for (int o = 16; o > 0; o >>= 1) {
            cR += __shfl_down_sync(0xffffffffu, cR, o);
            cE += __shfl_down_sync(0xffffffffu, cE, o);
            cS += __shfl_down_sync(0xffffffffu, cS, o);
        }
        if (lane == 0) {
            atomicAdd(P.counters + 0, (unsigned long long)cR);
            atomicAdd(P.counters + 1, (unsigned long long)cE);
            atomicAdd(P.counters + 2, (unsigned long long)cS);
        }
    }
}

// Host-side mirror of the kernel's region arithmetic: every region must fit the shared-memory
// tables (<= kFusedMaxSide, and <= 128 because 128 threads fill each table) and, for shift >= 1,
// its parent buffer.
bool fused_geometry_ok(const FusedParams& p) {
    if (p.levels < 1 || p.levels > kMaxLevels || p.ntx < 1 || p.nty < 1) return false;
    for (int l = 0; l < p.levels; l++) {
        const int s = p.levels - 1 - l;
        const FusedLevel& L = p.lv[l];
        const int cap = s == 0 ? kFusedMaxSide : ((s & 1) ? kFusedOddPitch : kFusedEvenPitch);
        for (int axis = 0; axis < 2; axis++) {
            const int g0 = axis ? p.gy0 : p.gx0, nt = axis ? p.nty : p.ntx;
            const int r0 = axis ? L.ry0 : L.rx0, r1 = axis ? L.ry1 : L.rx1;
            for (int t = 0; t < nt; t++) {
                const int X0 = g0 + t * kFusedTile;
                const int f0 = X0 >> s;
                const int f1 = t == nt - 1 ? 0x7fffffff : (((X0 + kFusedTile) >> s) > f0 + 1 ? ((X0 + kFusedTile) >> s) : f0 + 1);
                const int a = f0 > r0 ? f0 : r0, b = f1 < r1 ? f1 : r1;
                if (b - a > cap) return false;
            }
        }
    }
    return true;
}

template <int N, bool COUNT>
static cudaError_t launch_fused_nc(const FusedParams& p, int frames, cudaStream_t st) {
    const int smem = (kFusedThreads / 32) * RTCE_QUEUE_WORDS(N) * kQueueCap * (int)sizeof(float);
    static bool configured = false;   // per instantiation; the attribute is per function and per device context
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(k_trace_fused<N, COUNT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(k_trace_fused<N, COUNT>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    dim3 grid((unsigned)p.ntx, (unsigned)p.nty, (unsigned)frames);
    k_trace_fused<N, COUNT><<<grid, kFusedThreads, smem, st>>>(p);
    return cudaGetLastError();
}

template <int N>
static cudaError_t launch_fused_n(const FusedParams& p, int frames, bool count, cudaStream_t st) {
    return count ? launch_fused_nc<N, true>(p, frames, st) : launch_fused_nc<N, false>(p, frames, st);
}

cudaError_t launch_trace_fused(const FusedParams& p, int frames, bool count, cudaStream_t st) {
    if (frames <= 0) return cudaSuccess;
    switch (p.n_spheres) {
        case 0: return launch_fused_n<0>(p, frames, count, st);
        case 1: return launch_fused_n<1>(p, frames, count, st);
        case 2: return launch_fused_n<2>(p, frames, count, st);
        case 3: return launch_fused_n<3>(p, frames, count, st);
        case 4: return launch_fused_n<4>(p, frames, count, st);
        case 5: return launch_fused_n<5>(p, frames, count, st);
        case 6: return launch_fused_n<6>(p, frames, count, st);
        case 7: return launch_fused_n<7>(p, frames, count, st);
        case 8: return launch_fused_n<8>(p, frames, count, st);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace rtce

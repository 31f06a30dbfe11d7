// rtce_kernels_shade.cu — the presentation shading pass.
//
// Replaces the draw of main.rs:318-338 (fullscreen triangle vertex.glsl:3-10, one fragment per
// window pixel) with fragment.glsl:123-187 as a CUDA kernel: final pyramid level (binding 0) +
// MutableData (binding 1) + ConstantBuffer (binding 2) + push constants rot / pos -> an RGBA32F
// colour image in device memory (the headless stand-in for the swapchain image), and a second
// kernel that packs it to RGBA8 UNORM.  Arithmetic: rtce_shade_math.cuh (canonical, pow pinned);
// results are bit-identical to the reference's fragment shader compiled as C++ (oracle/_ref).
//
// One thread per pixel, 8x4 pixels per warp (hit pixels are spatially coherent: 22 % of the default
// view); background pixels leave after one load and one store.
#include "rtce_device.cuh"
#include "rtce_kernels.cuh"
#include "rtce_shade_math.cuh"

#ifndef RTCE_SHADE_VOTE
// refresh of fragment.glsl:107-109: 0 = per-lane branch (default), 1 = warp vote per object + select, 2 = the same with one
// activemask per trip.  Measured on the bench batch (16 x 1080p): 853 / 906 / 875 us; identical bits.
#define RTCE_SHADE_VOTE 0
#endif

namespace rtce {

namespace {

struct V3 {
    float x, y, z;
};
RTCE_DEV V3 v_sub(V3 a, V3 b) { return {fsub(a.x, b.x), fsub(a.y, b.y), fsub(a.z, b.z)}; }
RTCE_DEV V3 v_neg(V3 a) { return {-a.x, -a.y, -a.z}; }
RTCE_DEV float v_dot(V3 a, V3 b) { return fadd(fadd(fmul(a.x, b.x), fmul(a.y, b.y)), fmul(a.z, b.z)); }
RTCE_DEV float v_length(V3 a) { return sqrt_fast(v_dot(a, a)); }
RTCE_DEV V3 v_normalize(V3 a) {   // v * (1.0 / length(v)); rcp_fast is exact for 2^-126 <= length <= 2^126
    const float inv = rcp_fast(v_length(a));
    return {fmul(a.x, inv), fmul(a.y, inv), fmul(a.z, inv)};
}
RTCE_DEV V3 xyz(float4 v) { return {v.x, v.y, v.z}; }

// utilities.glsl:36-38
RTCE_DEV float sdf(V3 p, float4 s) { return fsub(v_length(v_sub(p, xyz(s))), s.w); }

// fragment.glsl:89-121.
// min / max inside shadow_ray: the specification's ternaries (lit_min / lit_max) differ from FMNMX only on a (+0, -0)
// pair or a NaN.  Here a zero `dist` returns before its sign can matter (0 <= RAY_RADIUS), `last` and `nearest` are
// only updated with dist > RAY_RADIUS > 0, and finite inputs produce no NaN: FMNMX yields the same bits in one
// instruction instead of two.
#ifndef RTCE_SHADE_FMNMX
#define RTCE_SHADE_FMNMX 1
#endif
#if RTCE_SHADE_FMNMX
#define SR_MIN(a, b) fminf(a, b)
#define SR_MAX(a, b) fmaxf(a, b)
#else
#define SR_MIN(a, b) lit_min(a, b)
#define SR_MAX(a, b) lit_max(a, b)
#endif
#ifndef RTCE_SHADE_CTAS
#define RTCE_SHADE_CTAS 4
#endif

template <int N, bool COUNT>
RTCE_DEV float shadow_ray(const ShadeParams& P, V3 origin, V3 step, float end, unsigned& cD) {
    const float RAY_RADIUS = 0.01f;
    float d[N];
#pragma unroll
    for (int i = 0; i < N; i++) d[i] = sdf(origin, P.obj[i]);   // :92-94
    if (COUNT) cD += (unsigned)N;
    float last = 0.0f, nearest = 1.0f;
    for (float len = 0.0f; len < end; len = fadd(len, fadd(last, RAY_RADIUS))) {   // :99
        const V3 pos = {fadd(origin.x, fmul(step.x, len)), fadd(origin.y, fmul(step.y, len)), fadd(origin.z, fmul(step.z, len))};
        float dist = end;   // :104
#if RTCE_SHADE_VOTE == 2
        // the lanes marching together do not change inside a trip (the votes below are warp-uniform among them, the
        // only exit is at the end of the trip): one activemask per trip instead of one per object
        const unsigned marching = __activemask();
#endif
#pragma unroll
        for (int i = 0; i < N; i++) {   // :105-111
            const float bound = fsub(d[i], last);
            const bool refresh = bound <= nearest;
            if (COUNT) cD += refresh ? 1u : 0u;
            d[i] = bound;
#if RTCE_SHADE_VOTE == 0
            if (refresh) d[i] = sdf(pos, P.obj[i]);
#else
            // skipped when no lane marching with this one needs it (the far spheres, most trips); a lane
            // that does need it always sees its own vote
#if RTCE_SHADE_VOTE == 2
            if (__any_sync(marching, refresh)) {
#else
            if (__any_sync(__activemask(), refresh)) {
#endif
                const float exact = sdf(pos, P.obj[i]);
                d[i] = refresh ? exact : bound;
            }
#endif
            dist = SR_MIN(dist, d[i]);
        }
        if (dist <= RAY_RADIUS) return 0.0f;   // :113-115
        last = SR_MAX(dist, 0.0f);
        nearest = SR_MIN(nearest, dist);
    }
    return nearest;
}

template <int N, bool COUNT>
__global__ void __launch_bounds__(256, RTCE_SHADE_CTAS)
    k_shade(const __grid_constant__ ShadeParams P) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int x = P.x0 + blockIdx.x * 32 + (warp & 3) * 8 + (lane & 7);
    const int y = P.y0 + blockIdx.y * 8 + (warp >> 2) * 4 + (lane >> 3);
    const int f = blockIdx.z;
    unsigned cP = 0, cH = 0, cD = 0, cL = 0;
    if (x < P.x1 && y < P.y1) {
        cP = 1;
        const float CAM_FALL_OFF = 0.01f, LIGHT_FALL_OFF = 0.01f;
        const size_t px = (size_t)y * (size_t)P.w + (size_t)x;
        const float total = __ldg(P.depth + (size_t)f * P.depth_stride + px);   // :132
        float r = 0.0f, g = 0.0f, b = 0.0f;
        if (!(total >= P.render_dist)) {   // :134-137 (a background fragment needs no view direction)
            const float4 q = __ldg(reinterpret_cast<const float4*>(P.cams + f));
            const float4 t = __ldg(reinterpret_cast<const float4*>(P.cams + f) + 1);
            const V3 cam = {t.x, t.y, t.z};
            // gl_FragCoord.xy = pixel centre; :126-130
            const float fx = fadd((float)x, 0.5f), fy = fadd((float)y, 0.5f);
            const float ncx = fmul(fsub(__fdiv_rn(fmul(fx, 2.0f), P.vw), 1.0f), P.ratx);
            const float ncy = fmul(fsub(__fdiv_rn(fmul(fy, 2.0f), P.vh), 1.0f), P.raty);
            const Vec3 rv = rotate_q(q.x, q.y, q.z, q.w, ncx, 1.0f, ncy);
            const V3 step = v_normalize({rv.x, rv.y, rv.z});
            cH = 1;
            const V3 position = {fadd(cam.x, fmul(step.x, total)), fadd(cam.y, fmul(step.y, total)), fadd(cam.z, fmul(step.z, total))};
            // nearest object and its material, :141-153
            int best = 0;
            float dist = sdf(position, P.obj[0]);
#pragma unroll
            for (int i = 1; i < N; i++) {
                const float nd = sdf(position, P.obj[i]);
                if (nd < dist) {
                    dist = nd;
                    best = i;
                }
            }
            if (COUNT) cD += (unsigned)N + 1u;
            const float4 object = P.obj[best];
            const float4 matc = P.matc[best];   // color.xyz, shine
            const float ambient = P.amb[best];
            const float cam_dist = v_length(v_sub(position, cam));   // :159
            const float cam_fall = lit_max(fmul(CAM_FALL_OFF, fadd(fmul(cam_dist, cam_dist), 1.0f)), 1.0f);   // :160
            const V3 normal = v_normalize(v_sub(position, xyz(object)));   // :163
            const V3 cam_dir = v_neg(step);
            const float normal_fall = lit_max(v_dot(normal, cam_dir), 0.0f);   // :164
            for (int i = 0; i < P.n_light; i++) {   // :167-186
                const V3 lpos = xyz(P.lpos[i]);
                const V3 lcol = xyz(P.lcol[i]);
                const V3 ldir = v_normalize(v_sub(lpos, position));   // :170
                const float ldist = v_length(v_sub(position, lpos));   // :171
                if (COUNT) { cD++; cL++; }
                const V3 so = {fadd(position.x, ldir.x), fadd(position.y, ldir.y), fadd(position.z, ldir.z)};
                const float soft = lit_min(shadow_ray<N, COUNT>(P, so, ldir, ldist, cD), 1.0f);   // :173
                const float lfall = lit_max(fmul(fmul(LIGHT_FALL_OFF, ldist), ldist), 1.0f);   // :175
                const float diff = lit_max(v_dot(normal, ldir), 0.0f);   // :177
                const V3 inc = v_neg(ldir);   // :178, reflect(-lightDir, normal)
                const float k = fmul(2.0f, v_dot(normal, inc));
                const V3 refl = {fsub(inc.x, fmul(k, normal.x)), fsub(inc.y, fmul(k, normal.y)), fsub(inc.z, fmul(k, normal.z))};
                const float spec = lit_max(fmul(diff, pinned_pow(v_dot(refl, cam_dir), matc.w)), 0.0f);
                const float m = lit_max(fadd(diff, spec), 0.0f);   // :180
                const float dx = fmul(__fdiv_rn(fmul(m, lcol.x), lfall), soft);
                const float dy = fmul(__fdiv_rn(fmul(m, lcol.y), lfall), soft);
                const float dz = fmul(__fdiv_rn(fmul(m, lcol.z), lfall), soft);
                // :182
                r = fadd(r, fmul(fmul(__fdiv_rn(fadd(ambient, dx), cam_fall), normal_fall), matc.x));
                g = fadd(g, fmul(fmul(__fdiv_rn(fadd(ambient, dy), cam_fall), normal_fall), matc.y));
                b = fadd(b, fmul(fmul(__fdiv_rn(fadd(ambient, dz), cam_fall), normal_fall), matc.z));
            }
        }
        P.out[(size_t)f * P.out_stride + (size_t)y * (size_t)P.out_w + (size_t)x] = make_float4(r, g, b, 1.0f);   // fragColor.a is never written: reported as 1
    }
    if (COUNT) {
        for (int o = 16; o > 0; o >>= 1) {
            cP += __shfl_down_sync(0xffffffffu, cP, o);
            cH += __shfl_down_sync(0xffffffffu, cH, o);
            cD += __shfl_down_sync(0xffffffffu, cD, o);
            cL += __shfl_down_sync(0xffffffffu, cL, o);
        }
        if (lane == 0) {
            atomicAdd(P.counters + 0, (unsigned long long)cP);
            atomicAdd(P.counters + 1, (unsigned long long)cH);
            atomicAdd(P.counters + 2, (unsigned long long)cD);
            atomicAdd(P.counters + 3, (unsigned long long)cL);
        }
    }
}

__global__ void k_rgba8(const float4* __restrict__ in, uchar4* __restrict__ out, size_t n) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float4 c = in[i];
        out[i] = make_uchar4((unsigned char)unorm8(c.x), (unsigned char)unorm8(c.y), (unsigned char)unorm8(c.z),
                             (unsigned char)unorm8(c.w));
    }
}

}  // namespace

template <int N>
static cudaError_t launch_shade_n(const ShadeParams& p, dim3 grid, bool count, cudaStream_t st) {
    if (count)
        k_shade<N, true><<<grid, 256, 0, st>>>(p);
    else
        k_shade<N, false><<<grid, 256, 0, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_shade(const ShadeParams& p, int frames, bool count, cudaStream_t st) {
    if (p.x1 <= p.x0 || p.y1 <= p.y0 || frames <= 0) return cudaSuccess;
    if (p.n_obj < 1 || p.n_obj > kSmallSceneMax || p.n_light < 0 || p.n_light > kSmallSceneMax || !p.cams) return cudaErrorInvalidValue;
    dim3 grid((unsigned)((p.x1 - p.x0 + 31) / 32), (unsigned)((p.y1 - p.y0 + 7) / 8), (unsigned)frames);
    switch (p.n_obj) {
        case 1: return launch_shade_n<1>(p, grid, count, st);
        case 2: return launch_shade_n<2>(p, grid, count, st);
        case 3: return launch_shade_n<3>(p, grid, count, st);
        case 4: return launch_shade_n<4>(p, grid, count, st);
        case 5: return launch_shade_n<5>(p, grid, count, st);
        case 6: return launch_shade_n<6>(p, grid, count, st);
        case 7: return launch_shade_n<7>(p, grid, count, st);
        case 8: return launch_shade_n<8>(p, grid, count, st);
        default: return cudaErrorInvalidValue;
    }
}

cudaError_t launch_rgba8(const float4* in, uchar4* out, size_t n, cudaStream_t st) {
    if (!n) return cudaSuccess;
    k_rgba8<<<148 * 8, 256, 0, st>>>(in, out, n);
    return cudaGetLastError();
}

}  // namespace rtce

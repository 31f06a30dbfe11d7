// rtce_kernels_large_ev.cu — objCount > 8 (BASELINE.json configs[3]), event-driven form of tracing algorithm 3
// (compute.glsl:34-68), bit-exact with it.
//
// The reference keeps, per ray, a cache of lower bounds `distances[objCount]` and on every march trip subtracts the
// last step from ALL of them, refreshes the ones that fell inside the cone radius and takes the minimum
// (compute.glsl:51-57): 3 N operations per trip, of which only the ~60 refreshed objects and the minimum matter.
// The register kernel (rtce_kernels_large_reg.cu) does exactly that, eagerly.  Here a trip touches only the objects
// that CAN be due or be the minimum:
//
//   * ONE RAY PER WARP.  Every object of the ray is in exactly one of two places:
//       W   the working set (shared memory, <= kEvWCap entries): objects whose bound is near — their EXACT current
//           bound is kept and updated on every trip, as the reference does (c = c - last; refresh when c <= radius);
//       far an unordered array in global memory of {value at epoch e, e, id}: objects whose bound is far.  Nothing is
//           done for them per trip.  Their exact bound at trip j is the chain ((v - last[e+1]) - last[e+2]) ... - last[j]
//           of separately rounded subtractions; it differs from  key - S_j  (key = v + S_e, S = running sum of the steps,
//           kept in double) by at most (j - e) half-ulps of |v|, bounded by the slack G_j below.
//   * kmin = the smallest key in `far` is tracked exactly.  While  kmin - S_j - G_j > max(radius, best bound in W)
//     no far object can be due (its bound is > radius) nor be the minimum, so the trip is decided by W alone.
//     Otherwise `far` is scanned once (coalesced), the objects with key < kmin + span move to W — their exact bound is
//     REPLAYED from the per-ray history of steps (the same rounded subtractions the reference made, so the same bits) —
//     and the test repeats.  W entries whose bound grew far (a refreshed object off to the side) are evicted to `far`.
//   * A ray whose working set or history would overflow is handed to the register kernel (fallback list): the
//     result is exact in every case, the event-driven form is only the fast path.
//
// min is exactly associative and commutative on non-NaN floats and `c - last` is applied to every W entry on every
// trip, so neither the order of W nor the moment an object enters W can change a bit; the work counters (E = SDF
// evaluations: N per ray at the start + every refresh; S = trips) equal the reference's because the set of objects
// refreshed on a trip is the same set.
#include "rtce_march.cuh"

namespace rtce {

constexpr int kEvThreads = 128;            // 4 rays per CTA
constexpr int kEvWarps = kEvThreads / 32;
#ifndef RTCE_EV_WCAP
#define RTCE_EV_WCAP 1024
#endif
#ifndef RTCE_EV_SPAN
#define RTCE_EV_SPAN 8.0f
#endif
constexpr int kEvWCap = RTCE_EV_WCAP;      // working-set entries per ray
constexpr int kEvHist = 4096;              // march trips of history per ray (longer rays: fallback)
constexpr float kEvSpan = RTCE_EV_SPAN;    // keys loaded ahead of the smallest far key per scan

struct EvFar {            // 8 bytes
    float val;            // exact bound after `epoch` trips (the initial SDF: 0 trips; evicted during trip j: j + 1 trips)
    unsigned epid;        // epoch | id << 16
};

// Scratch layout (device memory, owned by the context): header | fallback list | per-warp slots
struct EvHeader {
    unsigned ticket;      // next ray
    unsigned fb_count;    // rays handed to the register kernel
    unsigned pad[2];
};

RTCE_DEV float warp_min_f(float v) {
    float r;
    asm volatile("redux.sync.min.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
    return r;
}
RTCE_DEV float warp_max_f(float v) {
    float r;
    asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
    return r;
}
RTCE_DEV unsigned warp_min_u(unsigned v) { return __reduce_min_sync(0xffffffffu, v); }

size_t large_ev_slot_bytes(int n_spheres) {
    size_t far = ((size_t)n_spheres * sizeof(EvFar) + 255) & ~(size_t)255;
    return far + (size_t)(kEvHist + 8) * (sizeof(float) + sizeof(double));
}

template <bool COUNT>
__global__ void __launch_bounds__(kEvThreads, 6)
    k_trace_level_large_ev(const __grid_constant__ LevelParamsLarge P, long long total_rays, EvHeader* __restrict__ hdr,
                           int* __restrict__ fb_list, unsigned char* __restrict__ slots, size_t slot_bytes) {
    __shared__ float s_val[kEvWarps][kEvWCap];
    __shared__ unsigned short s_id[kEvWarps][kEvWCap];
    __shared__ unsigned short s_ep[kEvWarps][kEvWCap];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    float* const Wv = s_val[warp];
    unsigned short* const Wi = s_id[warp];
    unsigned short* const We = s_ep[warp];
    unsigned char* const slot = slots + (size_t)(blockIdx.x * kEvWarps + warp) * slot_bytes;
    EvFar* const far = reinterpret_cast<EvFar*>(slot);
    const size_t far_bytes = ((size_t)P.n_spheres * sizeof(EvFar) + 255) & ~(size_t)255;
    double* const hS = reinterpret_cast<double*>(slot + far_bytes);                        // hS[t] = sum of the steps subtracted on trips < t
    float* const hL = reinterpret_cast<float*>(slot + far_bytes + (size_t)(kEvHist + 8) * sizeof(double));   // step subtracted on trip t
    const int n = P.n_spheres;
    const int rw = P.x1 - P.x0, rh = P.y1 - P.y0;
    const long long per_frame = (long long)rw * rh;
    const float4* __restrict__ sph = P.spheres;
    const float inf = __int_as_float(0x7F800000);
    const float rd = P.render_dist, thr = P.threshold;
    unsigned long long tE = 0, tS = 0, tR = 0;

    for (;;) {
        unsigned tk = 0;
        if (lane == 0) tk = atomicAdd(&hdr->ticket, 1u);
        tk = __shfl_sync(0xffffffffu, tk, 0);
        if ((long long)tk >= total_rays) break;
        const long long ray = tk;
        const int f = (int)(ray / per_frame);
        const long long rem = ray - (long long)f * per_frame;
        const int y = P.y0 + (int)(rem / rw);
        const int x = P.x0 + (int)(rem % rw);
        const Cam& cam = P.cam[f];
        // compute.glsl:71-77 (uniform across the warp)
        const float ncx = fmul(fsub(fmul((float)(2u * (unsigned)x + 1u), P.isx), 1.0f), P.ratx);
        const float ncy = fmul(fsub(fmul((float)(2u * (unsigned)y + 1u), P.isy), 1.0f), P.raty);
        const Vec3 dir = normalize3(rotate_q(cam.rot[0], cam.rot[1], cam.rot[2], cam.rot[3], ncx, 1.0f, ncy));
        float start = 1.0f;   // :79-82
        if (P.iter > 0)
            start = __ldg(P.parent + (size_t)f * P.slot_stride + (size_t)(y >> 1) * (size_t)P.pw + (size_t)(x >> 1));
        const float ox = fadd(cam.pos[0], fmul(dir.x, start));
        const float oy = fadd(cam.pos[1], fmul(dir.y, start));
        const float oz = fadd(cam.pos[2], fmul(dir.z, start));

        // traceCone :37-39 — every object starts in `far` with its initial SDF, epoch 0
        float kmin = inf, vmax = 0.0f;
        __syncwarp();
        for (int k = lane; k < n; k += 32) {
            const float d = sphere_sdf_fast(ox, oy, oz, __ldg(sph + k));
            far[k] = EvFar{d, (unsigned)k << 16};
            kmin = fminf(kmin, d);
            vmax = fmaxf(vmax, fabsf(d));
        }
        kmin = warp_min_f(kmin);
        vmax = warp_max_f(vmax);           // bounds |value| and |key| of every record for the slack G
        unsigned long long cE = (unsigned long long)n, cS = 0;
        int n_far = n, n_w = 0;
        float len = 0.0f, last = 0.0f;     // :41-42
        double S = 0.0;
        float t_hi = -inf;                 // keys below t_hi have been moved to W at some point
        int j = 0;
        bool bail = false;
        if (lane == 0) hS[0] = 0.0;
        while (len < rd) {   // :44
            if (j >= kEvHist) { bail = true; break; }
            const float px = fadd(ox, fmul(dir.x, len));   // :45
            const float py = fadd(oy, fmul(dir.y, len));
            const float pz = fadd(oz, fmul(dir.z, len));
            const float radius = fmul(fadd(len, 1.0f), thr);   // :50
            S += (double)last;
            if (lane == 0) {
                hL[j] = last;
                hS[j + 1] = S;
            }
            __syncwarp();
            const float Sf = (float)S;
            // ---- W: the reference's loop body (:51-57) on the working set, compacting as it goes ----
            const float evict_above = fmaxf(fsub(t_hi, Sf), 0.0f) + kEvSpan;
            float m = rd;   // :49
            float kmin_l = inf;
            int out = 0;
            for (int base = 0; base < n_w; base += 32) {
                const int k = base + lane;
                const bool valid = k < n_w;
                float c = valid ? Wv[k] : inf;
                const unsigned id = valid ? Wi[k] : 0u;
                __syncwarp();
                bool evict = false;
                if (valid) {
                    c = fsub(c, last);                       // :52
                    if (c <= radius) {                       // :53-55
                        c = sphere_sdf_fast(px, py, pz, __ldg(sph + id));
                        if (COUNT) cE++;
                    }
                    m = fminf(m, c);                         // :56
                    evict = c > evict_above;
                }
                const unsigned mk = __ballot_sync(0xffffffffu, valid && !evict);
                const unsigned me = __ballot_sync(0xffffffffu, evict);
                if (valid && !evict) {
                    const int o = out + __popc(mk & lt);
                    Wv[o] = c;
                    Wi[o] = (unsigned short)id;
                }
                if (evict) {
                    far[n_far + __popc(me & lt)] = EvFar{c, (unsigned)(j + 1) | (id << 16)};   // this trip is applied
                    kmin_l = fminf(kmin_l, (float)((double)c + S));
                    vmax = fmaxf(vmax, fabsf(c) + Sf);
                }
                out += __popc(mk);
                n_far += __popc(me);
            }
            n_w = out;
            if (__any_sync(0xffffffffu, kmin_l < inf)) {
                kmin = fminf(kmin, warp_min_f(kmin_l));
                vmax = warp_max_f(vmax);
            }
            float best = warp_min_f(m);
            // ---- far: can any of its objects be due or be the minimum on this trip? ----
            // exact bound of a far record >= key - S - G, G = (j + 4) ulps-ish of the largest magnitude around
            const float G = (float)(j + 4) * 1.1920929e-7f * fadd(vmax, Sf);
            while (n_far > 0 && fsub(fsub(kmin, Sf), G) <= fmaxf(radius, best)) {
                const float t_new = fmaxf(kmin, fadd(fadd(Sf, radius), G)) + kEvSpan;
                const int nw0 = n_w;
                int fo = 0;
                float kmin_n = inf;
                __syncwarp();
                for (int base = 0; base < n_far; base += 32) {
                    const int k = base + lane;
                    const bool valid = k < n_far;
                    EvFar rec = valid ? far[k] : EvFar{0.0f, 0u};
                    const unsigned ep = rec.epid & 0xffffu;
                    const float key = valid ? (float)((double)rec.val + hS[ep]) : inf;
                    const bool sel = valid && key < t_new;
                    const unsigned ms = __ballot_sync(0xffffffffu, sel);
                    const unsigned mkp = __ballot_sync(0xffffffffu, valid && !sel);
                    if (sel) {
                        const int o = n_w + __popc(ms & lt);
                        if (o < kEvWCap) {
                            Wv[o] = rec.val;
                            Wi[o] = (unsigned short)(rec.epid >> 16);
                            We[o] = (unsigned short)ep;
                        }
                    }
                    if (valid && !sel) {
                        far[fo + __popc(mkp & lt)] = rec;
                        kmin_n = fminf(kmin_n, key);
                    }
                    n_w += __popc(ms);
                    fo += __popc(mkp);
                    __syncwarp();
                }
                n_far = fo;
                kmin = warp_min_f(kmin_n);
                t_hi = fmaxf(t_hi, t_new);
                if (n_w > kEvWCap) { bail = true; break; }
                __syncwarp();
                // the objects that came in: replay their bound to this trip, then the loop body
                for (int base = nw0; base < n_w; base += 32) {
                    const int k = base + lane;
                    const bool valid = k < n_w;
                    float c = valid ? Wv[k] : inf;
                    const unsigned ep = valid ? (unsigned)We[k] : (unsigned)(j + 1);
                    const unsigned e0 = warp_min_u(ep);
                    for (unsigned t = e0; t <= (unsigned)j; t++) {
                        const float l = hL[t];
                        if (t >= ep) c = fsub(c, l);         // :52 of trip t
                    }
                    if (valid) {
                        if (ep <= (unsigned)j && c <= radius) {   // :53-55 of this trip (not again for an object evicted on this trip)
                            c = sphere_sdf_fast(px, py, pz, __ldg(sph + Wi[k]));
                            if (COUNT) cE++;
                        }
                        Wv[k] = c;
                        m = fminf(m, c);
                    }
                }
                best = warp_min_f(m);
            }
            if (bail) break;
            if (COUNT) cS++;
            const float dist = best;
            last = glsl_max(dist, 0.0f);   // :59
            len = fadd(len, last);         // :60
            j++;
            if (dist <= radius) {          // :62-65
                len = fsub(len, radius);
                break;
            }
        }
        if (bail) {
            if (lane == 0) fb_list[atomicAdd(&hdr->fb_count, 1u)] = (int)tk;
            continue;
        }
        if (lane == 0) P.out[(size_t)f * P.slot_stride + (size_t)y * (size_t)P.w + (size_t)x] = glsl_max(fadd(start, len), 0.0f);
        if (COUNT) {
            const unsigned long long e_lanes = cE - (unsigned long long)n;   // per-lane refresh counts; n was added on every lane
            unsigned long long e = e_lanes;
            for (int o = 16; o > 0; o >>= 1) e += __shfl_down_sync(0xffffffffu, e, o);
            if (lane == 0) {
                tE += e + (unsigned long long)n;
                tS += cS;
                tR++;
            }
        }
    }
    if (COUNT && lane == 0) {
        atomicAdd(P.counters + 0, tR);
        atomicAdd(P.counters + 1, tE);
        atomicAdd(P.counters + 2, tS);
    }
}

size_t large_ev_scratch_bytes(int n_spheres, long long max_rays_per_launch, int* warps_out) {
    const int warps = device_sm_count() * 6 * kEvWarps;
    if (warps_out) *warps_out = warps;
    return 256 + (((size_t)max_rays_per_launch * sizeof(int) + 255) & ~(size_t)255) + (size_t)warps * large_ev_slot_bytes(n_spheres);
}

cudaError_t launch_trace_level_large_ev(LevelParamsLarge p, int frames, bool count, void* scratch, long long max_rays,
                                        cudaStream_t st) {
    if (p.x1 <= p.x0 || p.y1 <= p.y0 || frames <= 0) return cudaSuccess;
    const long long total = (long long)(p.x1 - p.x0) * (p.y1 - p.y0) * frames;
    if (total > max_rays || total >= (1ll << 31)) return cudaErrorInvalidValue;
    unsigned char* base = static_cast<unsigned char*>(scratch);
    EvHeader* hdr = reinterpret_cast<EvHeader*>(base);
    int* fb = reinterpret_cast<int*>(base + 256);
    unsigned char* slots = base + 256 + (((size_t)max_rays * sizeof(int) + 255) & ~(size_t)255);
    cudaError_t e = cudaMemsetAsync(hdr, 0, sizeof(EvHeader), st);
    if (e != cudaSuccess) return e;
    const size_t slot_bytes = large_ev_slot_bytes(p.n_spheres);
    long long ctas = (long long)device_sm_count() * 6;
    const long long need = (total + kEvWarps - 1) / kEvWarps;
    if (ctas > need) ctas = need;
    if (count)
        k_trace_level_large_ev<true><<<(unsigned)ctas, kEvThreads, 0, st>>>(p, total, hdr, fb, slots, slot_bytes);
    else
        k_trace_level_large_ev<false><<<(unsigned)ctas, kEvThreads, 0, st>>>(p, total, hdr, fb, slots, slot_bytes);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    // the rays the fast path handed back (working set or history overflow): the register kernel, from the list
    p.ray_list = fb;
    p.ray_count = &hdr->fb_count;
    return launch_trace_level_large_reg(p, frames, count, st);
}

}  // namespace rtce

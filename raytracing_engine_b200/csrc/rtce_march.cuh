// rtce_march.cuh — the per-pixel body of compute.glsl:70-87, split where rays diverge, shared by
// the fused-pyramid kernel (rtce_kernels_fused.cu) and the strip kernels (rtce_kernels_strip.cu).
//
// Arithmetic: identical rounding sequence to compute.glsl:70-87 as pinned in rtce_device.cuh.  The
// separable column / row terms are sub-expressions of rotate() hoisted out of the pixel, not
// re-associated.
#pragma once
#include "rtce_device.cuh"
#include "rtce_kernels.cuh"

namespace rtce {

// Column terms (depend on gid.x only) and row terms (gid.y only) of
//   v = (nx, 1, ny);  t = cross(q.xyz, v) + q.w * v           (utilities.glsl:27)
//   t.x = (q.y*ny - 1*q.z) + q.w*nx = A(y) + B(x)
//   t.y = (q.z*nx - ny*q.x) + q.w*1 = (C(x) - G(y)) + q.w
//   t.z = (q.x*1 - nx*q.y) + q.w*ny = E(x) + F(y)
// (1*q.z, q.x*1, q.w*1 are exact.)
RTCE_DEV float norm_coord(int g, float image_size, float ratio) {   // compute.glsl:71-72
    return fmul(fsub(fmul((float)(2u * (unsigned)g + 1u), image_size), 1.0f), ratio);
}
RTCE_DEV float4 col_terms(const Cam& c, float nx) {   // {B, C, E, nx}
    return make_float4(fmul(c.rot[3], nx), fmul(c.rot[2], nx), fsub(c.rot[0], fmul(nx, c.rot[1])), nx);
}
RTCE_DEV float4 row_terms(const Cam& c, float ny) {   // {A, G, F, ny}
    return make_float4(fsub(fmul(c.rot[1], ny), c.rot[2]), fmul(ny, c.rot[0]), fmul(c.rot[3], ny), ny);
}

// compute.glsl:70-87 for one pixel whose column / row terms are known.  Returns max(len, 0).
template <int N, bool COUNT>
RTCE_DEV float shade_pixel(const float4 col, const float4 row, const float start, const Cam& c, const float threshold,
                           const float rd, const float4* __restrict__ sph, unsigned& cE, unsigned& cS) {
    const float qx = c.rot[0], qy = c.rot[1], qz = c.rot[2], qw = c.rot[3];
    const float tx = fadd(row.x, col.x);
    const float ty = fadd(fsub(col.y, row.y), qw);
    const float tz = fadd(col.z, row.z);
    // cross(q.xyz, t)  (utilities.glsl:28)
    const float c2x = fsub(fmul(qy, tz), fmul(ty, qz));
    const float c2y = fsub(fmul(qz, tx), fmul(tz, qx));
    const float c2z = fsub(fmul(qx, ty), fmul(tx, qy));
    // v + 2.0 * c2: the product by 2 is exact, so one FMA rounds exactly like FMUL then FADD
    const float rx = __fmaf_rn(2.0f, c2x, col.w);
    const float ry = __fmaf_rn(2.0f, c2y, 1.0f);
    const float rz = __fmaf_rn(2.0f, c2z, row.w);
    // normalize (compute.glsl:77): v * (1 / sqrt(dot(v, v)))
    const float l2 = fadd(fadd(fmul(rx, rx), fmul(ry, ry)), fmul(rz, rz));
    const float inv = rcp_fast(sqrt_fast(l2));
    const float sx = fmul(rx, inv), sy = fmul(ry, inv), sz = fmul(rz, inv);
    // origin = pos + step * len  (compute.glsl:84)
    const float ox = fadd(c.pos[0], fmul(sx, start));
    const float oy = fadd(c.pos[1], fmul(sy, start));
    const float oz = fadd(c.pos[2], fmul(sz, start));

    // traceCone (compute.glsl:34-68)
    float d[N > 0 ? N : 1];
#pragma unroll
    for (int i = 0; i < N; i++) {   // :37-39
        const float dx = fsub(ox, sph[i].x), dy = fsub(oy, sph[i].y), dz = fsub(oz, sph[i].z);
        d[i] = fsub(sqrt_fast(fadd(fadd(fmul(dx, dx), fmul(dy, dy)), fmul(dz, dz))), sph[i].w);
    }
    if (COUNT) cE += N;
    float len = 0.0f;
    if (0.0f < rd) {   // :44, first trip
        // First iteration peeled: len = 0 and last = 0, so position == origin, radius == threshold,
        // `distances[i] -= last` changes nothing and a refresh re-evaluates the SDF at the origin,
        // i.e. returns the value distances[i] already holds.  Nothing to recompute.
        float dist = rd;
#pragma unroll
        for (int i = 0; i < N; i++) {
            if (COUNT) cE += (d[i] <= threshold) ? 1u : 0u;
            dist = glsl_min(dist, d[i]);
        }
        if (COUNT) cS++;
        float last = glsl_max(dist, 0.0f);
        len = last;   // 0 + last
        if (dist <= threshold) {
            len = fsub(len, threshold);
        } else {
            while (len < rd) {   // :44
                const float px = fadd(ox, fmul(sx, len));   // :45
                const float py = fadd(oy, fmul(sy, len));
                const float pz = fadd(oz, fmul(sz, len));
                dist = rd;                                                 // :49
                const float radius = fmul(fadd(len, 1.0f), threshold);    // :50
#pragma unroll
                for (int i = 0; i < N; i++) {   // :51-57
                    d[i] = fsub(d[i], last);
                    if (d[i] <= radius) {
                        const float dx = fsub(px, sph[i].x), dy = fsub(py, sph[i].y), dz = fsub(pz, sph[i].z);
                        d[i] = fsub(sqrt_fast(fadd(fadd(fmul(dx, dx), fmul(dy, dy)), fmul(dz, dz))), sph[i].w);
                        if (COUNT) cE++;
                    }
                    dist = glsl_min(dist, d[i]);
                }
                if (COUNT) cS++;
                last = glsl_max(dist, 0.0f);   // :59
                len = fadd(len, last);         // :60
                if (dist <= radius) {          // :62-65
                    len = fsub(len, radius);
                    break;
                }
            }
        }
    }
    return glsl_max(fadd(start, len), 0.0f);   // :84, :86
}

// ---- the same pixel, split at the point where rays diverge ------------------------------------
// ray_setup   : compute.glsl:71-84 up to and including traceCone's init loop (:37-39) — straight line
// first_step  : the first trip of the march loop (:44-66), which needs no SDF evaluation (see above)
// march_rest  : trips 2.. of the loop — only ~1/4 of the rays get here, with varying trip counts
template <int N>
struct Ray {
    float ox, oy, oz, sx, sy, sz;
    float d[N > 0 ? N : 1];
};

template <int N>
RTCE_DEV void ray_setup(Ray<N>& r, const float4 col, const float4 row, const float start, const Cam& c,
                        const float4* __restrict__ sph) {
    const float qx = c.rot[0], qy = c.rot[1], qz = c.rot[2], qw = c.rot[3];
    const float tx = fadd(row.x, col.x);
    const float ty = fadd(fsub(col.y, row.y), qw);
    const float tz = fadd(col.z, row.z);
    const float c2x = fsub(fmul(qy, tz), fmul(ty, qz));
    const float c2y = fsub(fmul(qz, tx), fmul(tz, qx));
    const float c2z = fsub(fmul(qx, ty), fmul(tx, qy));
    const float rx = __fmaf_rn(2.0f, c2x, col.w);
    const float ry = __fmaf_rn(2.0f, c2y, 1.0f);
    const float rz = __fmaf_rn(2.0f, c2z, row.w);
    const float l2 = fadd(fadd(fmul(rx, rx), fmul(ry, ry)), fmul(rz, rz));
    const float inv = rcp_fast(sqrt_fast(l2));
    r.sx = fmul(rx, inv);
    r.sy = fmul(ry, inv);
    r.sz = fmul(rz, inv);
    r.ox = fadd(c.pos[0], fmul(r.sx, start));
    r.oy = fadd(c.pos[1], fmul(r.sy, start));
    r.oz = fadd(c.pos[2], fmul(r.sz, start));
#pragma unroll
    for (int i = 0; i < N; i++) {
        const float dx = fsub(r.ox, sph[i].x), dy = fsub(r.oy, sph[i].y), dz = fsub(r.oz, sph[i].z);
        r.d[i] = fsub(sqrt_fast(fadd(fadd(fmul(dx, dx), fmul(dy, dy)), fmul(dz, dz))), sph[i].w);
    }
}

// Returns true when the ray must go on marching; `len` is traceCone's local length so far
// (and equals `last`, because len = 0 + last).
template <int N, bool COUNT>
RTCE_DEV bool first_step(const Ray<N>& r, float& len, const float threshold, const float rd, unsigned& cE, unsigned& cS) {
    len = 0.0f;
    if (!(0.0f < rd)) return false;   // :44 never entered
    float dist = rd;
#pragma unroll
    for (int i = 0; i < N; i++) {
        if (COUNT) cE += (r.d[i] <= threshold) ? 1u : 0u;
        dist = glsl_min(dist, r.d[i]);
    }
    if (COUNT) {
        cE += N;
        cS++;
    }
    len = glsl_max(dist, 0.0f);
    if (dist <= threshold) {
        len = fsub(len, threshold);
        return false;
    }
    return len < rd;
}

template <int N, bool COUNT>
RTCE_DEV float march_rest(Ray<N>& r, float len, const float threshold, const float rd, const float4* __restrict__ sph,
                          unsigned& cE, unsigned& cS) {
    float last = len;
    while (len < rd) {   // :44
        const float px = fadd(r.ox, fmul(r.sx, len));   // :45
        const float py = fadd(r.oy, fmul(r.sy, len));
        const float pz = fadd(r.oz, fmul(r.sz, len));
        float dist = rd;                                           // :49
        const float radius = fmul(fadd(len, 1.0f), threshold);    // :50
#pragma unroll
        for (int i = 0; i < N; i++) {   // :51-57
            r.d[i] = fsub(r.d[i], last);
            if (r.d[i] <= radius) {
                const float dx = fsub(px, sph[i].x), dy = fsub(py, sph[i].y), dz = fsub(pz, sph[i].z);
                r.d[i] = fsub(sqrt_fast(fadd(fadd(fmul(dx, dx), fmul(dy, dy)), fmul(dz, dz))), sph[i].w);
                if (COUNT) cE++;
            }
            dist = glsl_min(dist, r.d[i]);
        }
        if (COUNT) cS++;
        last = glsl_max(dist, 0.0f);   // :59
        len = fadd(len, last);         // :60
        if (dist <= radius) {          // :62-65
            len = fsub(len, radius);
            break;
        }
    }
    return len;
}

// ONE trip of the march loop (compute.glsl:44-66) for a ray whose loop test `len < rd` has passed.
// `last` is the previous trip's max(dist, 0).  Returns true when the ray goes on (no hit and the
// next loop test passes); on a hit `len` already has the cone radius subtracted (:63).
template <int N, bool COUNT>
RTCE_DEV bool march_step(Ray<N>& r, float& len, float& last, const float threshold, const float rd,
                         const float4* __restrict__ sph, unsigned& cE, unsigned& cS) {
    const float px = fadd(r.ox, fmul(r.sx, len));   // :45
    const float py = fadd(r.oy, fmul(r.sy, len));
    const float pz = fadd(r.oz, fmul(r.sz, len));
    float dist = rd;                                           // :49
    const float radius = fmul(fadd(len, 1.0f), threshold);    // :50
#pragma unroll
    for (int i = 0; i < N; i++) {   // :51-57
        r.d[i] = fsub(r.d[i], last);
        if (r.d[i] <= radius) {
            const float dx = fsub(px, sph[i].x), dy = fsub(py, sph[i].y), dz = fsub(pz, sph[i].z);
            r.d[i] = fsub(sqrt_fast(fadd(fadd(fmul(dx, dx), fmul(dy, dy)), fmul(dz, dz))), sph[i].w);
            if (COUNT) cE++;
        }
        dist = glsl_min(dist, r.d[i]);
    }
    if (COUNT) cS++;
    last = glsl_max(dist, 0.0f);   // :59
    len = fadd(len, last);         // :60
    if (dist <= radius) {          // :62-65
        len = fsub(len, radius);
        return false;
    }
    return len < rd;               // :44
}

// march_step with a branch-free refresh: the exact SDF of EVERY object is evaluated (N independent
// chains that depend only on `len`) and selected where the cached bound fell inside the cone.  Same
// values as march_step (a refresh that is not selected has no effect), ~2x the instructions, but
// the dependent chain of one trip drops from N serial branch-and-refresh blocks to one SDF chain.
// Used where marching is latency-bound (long rays, coarse levels), not in the issue-bound kernels.
template <int N, bool COUNT>
RTCE_DEV bool march_step_pred(Ray<N>& r, float& len, float& last, const float threshold, const float rd,
                              const float4* __restrict__ sph, unsigned& cE, unsigned& cS) {
    const float px = fadd(r.ox, fmul(r.sx, len));   // :45
    const float py = fadd(r.oy, fmul(r.sy, len));
    const float pz = fadd(r.oz, fmul(r.sz, len));
    float dist = rd;                                           // :49
    const float radius = fmul(fadd(len, 1.0f), threshold);    // :50
#pragma unroll
    for (int i = 0; i < N; i++) {   // :51-57
        const float dx = fsub(px, sph[i].x), dy = fsub(py, sph[i].y), dz = fsub(pz, sph[i].z);
        const float exact = fsub(sqrt_fast(fadd(fadd(fmul(dx, dx), fmul(dy, dy)), fmul(dz, dz))), sph[i].w);
        const float bound = fsub(r.d[i], last);
        const bool refresh = bound <= radius;
        if (COUNT) cE += refresh ? 1u : 0u;
        r.d[i] = refresh ? exact : bound;
        dist = glsl_min(dist, r.d[i]);
    }
    if (COUNT) cS++;
    last = glsl_max(dist, 0.0f);   // :59
    len = fadd(len, last);         // :60
    if (dist <= radius) {          // :62-65
        len = fsub(len, radius);
        return false;
    }
    return len < rd;               // :44
}

// K branch-free trips in a row for the latency-bound marchers.  A single warp per scheduler pays
// ~150 cycles of loop control (votes, convergence barriers, branches) around a ~200-cycle trip;
// here the control is paid once per K trips.  A lane whose ray ends inside the block keeps its
// final state (every later trip of the block is computed and discarded), so the values are those
// of march_step.  Returns true while the ray is still going.
#ifndef RTCE_PRED_BLOCK
#define RTCE_PRED_BLOCK 4
#endif
template <int N, bool COUNT, int K>
RTCE_DEV bool march_block_pred(Ray<N>& r, float& len, float& last, const float threshold, const float rd,
                               const float4* __restrict__ sph, unsigned& cE, unsigned& cS) {
    bool go = true;
#pragma unroll
    for (int k = 0; k < K; k++) {
        const float px = fadd(r.ox, fmul(r.sx, len));   // :45
        const float py = fadd(r.oy, fmul(r.sy, len));
        const float pz = fadd(r.oz, fmul(r.sz, len));
        float dist = rd;                                           // :49
        const float radius = fmul(fadd(len, 1.0f), threshold);    // :50
        float dn[N > 0 ? N : 1];
#pragma unroll
        for (int i = 0; i < N; i++) {   // :51-57
            const float dx = fsub(px, sph[i].x), dy = fsub(py, sph[i].y), dz = fsub(pz, sph[i].z);
            const float exact = fsub(sqrt_fast(fadd(fadd(fmul(dx, dx), fmul(dy, dy)), fmul(dz, dz))), sph[i].w);
            const float bound = fsub(r.d[i], last);
            const bool refresh = bound <= radius;
            if (COUNT) cE += (go && refresh) ? 1u : 0u;
            dn[i] = refresh ? exact : bound;
            dist = glsl_min(dist, dn[i]);
        }
        if (COUNT) cS += go ? 1u : 0u;
        const float last_n = glsl_max(dist, 0.0f);   // :59
        const float len_a = fadd(len, last_n);       // :60
        const bool hit = dist <= radius;             // :62-65
        const float len_n = hit ? fsub(len_a, radius) : len_a;
#pragma unroll
        for (int i = 0; i < N; i++) r.d[i] = go ? dn[i] : r.d[i];
        last = go ? last_n : last;
        len = go ? len_n : len;
        go = go && !hit && len_n < rd;               // :44
    }
    return go;
}

}  // namespace rtce

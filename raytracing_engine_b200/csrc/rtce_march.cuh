// rtce_march.cuh — the per-pixel body of compute.glsl:70-87, split where rays diverge, shared by
// every kernel that traces objCount <= 8 (rtce_kernels.cu, rtce_kernels_strip.cu).
//
// Arithmetic: identical rounding sequence to compute.glsl:70-87 as pinned in rtce_device.cuh.  The
// separable column / row terms are sub-expressions of rotate() hoisted out of the pixel, not
// re-associated.  There is ONE restatement of the march-loop trip (march_trip); the branchy and
// the branch-free forms are two instantiations of it.
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

// Per-ray state of traceCone (compute.glsl:34-68): origin, step, and the per-object cache of lower
// bounds `float distances[MAX_OBJECTS]` (:35) in registers (N is a template parameter).
template <int N>
struct Ray {
    float ox, oy, oz, sx, sy, sz;
    float d[N > 0 ? N : 1];
};

// utilities.glsl:36-38 with the short correctly rounded sqrt (same bits as sqrt.rn, rtce_device.cuh)
RTCE_DEV float sphere_sdf_fast(float px, float py, float pz, const float4 s) {
    const float dx = fsub(px, s.x), dy = fsub(py, s.y), dz = fsub(pz, s.z);
    return fsub(sqrt_fast(fadd(fadd(fmul(dx, dx), fmul(dy, dy)), fmul(dz, dz))), s.w);
}

// compute.glsl:71-84 up to and including traceCone's init loop (:37-39) — straight-line code.
template <int N>
RTCE_DEV void ray_setup(Ray<N>& r, const float4 col, const float4 row, const float start, const Cam& c,
                        const float4* __restrict__ sph) {
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
    r.sx = fmul(rx, inv);
    r.sy = fmul(ry, inv);
    r.sz = fmul(rz, inv);
    // origin = pos + step * len  (compute.glsl:84)
    r.ox = fadd(c.pos[0], fmul(r.sx, start));
    r.oy = fadd(c.pos[1], fmul(r.sy, start));
    r.oz = fadd(c.pos[2], fmul(r.sz, start));
#pragma unroll
    for (int i = 0; i < N; i++) r.d[i] = sphere_sdf_fast(r.ox, r.oy, r.oz, sph[i]);   // :37-39
}

// The FIRST trip of the march loop (compute.glsl:44-66) for RENDER_DIST > 0, branch-free.  len = 0 and
// last = 0, so position == origin, radius == threshold, `distances[i] -= last` changes nothing and a
// refresh re-evaluates the SDF at the origin, i.e. returns the value distances[i] already holds: no SDF
// needs evaluating.  On return len = traceCone's local length so far; when the ray goes on (return
// value true) it equals `last`, because len = 0 + last.
template <int N, bool COUNT>
RTCE_DEV bool first_trip(const Ray<N>& r, float& len, const float thr, const float rd, unsigned& cE, unsigned& cS) {
    float dist = rd;   // compute.glsl:49
#pragma unroll
    for (int i = 0; i < N; i++) {
        if (COUNT) cE += (r.d[i] <= thr) ? 1u : 0u;
        dist = glsl_min(dist, r.d[i]);
    }
    if (COUNT) {
        cE += N;   // the init loop :37-39
        cS++;
    }
    const float last = glsl_max(dist, 0.0f);   // :59; len = 0 + last
    const bool hit = dist <= thr;              // :62 (radius = (0 + 1) * threshold)
    len = hit ? fsub(last, thr) : last;        // :63
    return !hit && last < rd;                  // :44
}

// ONE trip of the march loop (compute.glsl:44-66) for a ray whose loop test `len < rd` has passed.
// `last` is the previous trip's max(dist, 0).  Returns true when the ray goes on (no hit and the next
// loop test passes); on a hit `len` already has the cone radius subtracted (:63).
//   PRED = false  the refresh (:53-55) is a per-lane branch: fewest instructions; for the issue-bound kernels.
//   PRED = true   branch-free: the exact SDF of EVERY object is evaluated (N independent chains that depend
//                 only on `len`) and selected where the cached bound fell inside the cone — same values
//                 (a refresh that is not selected has no effect), ~2x the instructions, but the dependent
//                 chain of a trip drops from N serial branch-and-refresh blocks to one SDF chain; for the
//                 latency-bound marchers.  `live` = false freezes the lane's state (a ray that ended inside
//                 a block of trips keeps its final values).
template <int N, bool COUNT, bool PRED>
RTCE_DEV bool march_trip(Ray<N>& r, float& len, float& last, const float thr, const float rd,
                         const float4* __restrict__ sph, unsigned& cE, unsigned& cS, const bool live = true) {
    const float px = fadd(r.ox, fmul(r.sx, len));   // :45
    const float py = fadd(r.oy, fmul(r.sy, len));
    const float pz = fadd(r.oz, fmul(r.sz, len));
    float dist = rd;                                    // :49
    const float radius = fmul(fadd(len, 1.0f), thr);    // :50
    float dn[N > 0 ? N : 1];
#pragma unroll
    for (int i = 0; i < N; i++) {   // :51-57
        if (PRED) {
            const float exact = sphere_sdf_fast(px, py, pz, sph[i]);
            const float bound = fsub(r.d[i], last);
            const bool refresh = bound <= radius;
            if (COUNT) cE += (live && refresh) ? 1u : 0u;
            dn[i] = refresh ? exact : bound;
        } else {
            dn[i] = fsub(r.d[i], last);
            if (dn[i] <= radius) {
                dn[i] = sphere_sdf_fast(px, py, pz, sph[i]);
                if (COUNT) cE++;
            }
        }
        dist = glsl_min(dist, dn[i]);
    }
    if (COUNT) cS += live ? 1u : 0u;
    const float last_n = glsl_max(dist, 0.0f);   // :59
    const float len_a = fadd(len, last_n);       // :60
    const bool hit = dist <= radius;             // :62-65
    const float len_n = hit ? fsub(len_a, radius) : len_a;
    if (PRED) {
#pragma unroll
        for (int i = 0; i < N; i++) r.d[i] = live ? dn[i] : r.d[i];
        last = live ? last_n : last;
        len = live ? len_n : len;
    } else {
#pragma unroll
        for (int i = 0; i < N; i++) r.d[i] = dn[i];
        last = last_n;
        len = len_n;
    }
    return live && !hit && len_n < rd;           // :44
}

// K branch-free trips in a row for the latency-bound marchers.  A single warp per scheduler pays
// ~150 cycles of loop control (votes, convergence barriers, branches) around a ~200-cycle trip;
// here the control is paid once per K trips.  Returns true while the ray is still going.
#ifndef RTCE_PRED_BLOCK
#define RTCE_PRED_BLOCK 4
#endif
template <int N, bool COUNT, int K>
RTCE_DEV bool march_block(Ray<N>& r, float& len, float& last, const float thr, const float rd,
                          const float4* __restrict__ sph, unsigned& cE, unsigned& cS) {
    bool go = true;
#pragma unroll
    for (int k = 0; k < K; k++) go = march_trip<N, COUNT, true>(r, len, last, thr, rd, sph, cE, cS, go);
    return go;
}

}  // namespace rtce

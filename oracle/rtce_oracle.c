/* TEST INFRASTRUCTURE ONLY — not part of the product path.
 *
 * Plain-C restatement ("port") of the reference's per-pixel compute pass.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library; it is the checker, never the
 * thing measured as the product and never a fallback for the CUDA path.
 *
 * Parity pin: the reference ships no tests, golden images or known answers
 * (SURVEY.md §4), so this port is pinned against the reference ITSELF run in
 * the build container: oracle/_ref/librtce_ref.so is the reference's own
 * shaders/compute.glsl + shaders/utilities.glsl compiled as C++ (oracle/Makefile,
 * oracle/glsl_shim.hpp, oracle/ref_driver.cpp).  tests/test_oracle.py checks
 * port == _ref bit-for-bit, and both against the committed tests/golden/
 * fixtures (made by tests/golden/make_golden.py from _ref).
 *
 * Canonical arithmetic (glsl_shim.hpp header): every operation
 * below is ONE binary32 IEEE-754 operation, never contracted
 * (-ffp-contract=off), sqrtf and / correctly rounded.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <omp.h>

#include "rtce_host_math.h"

#define RTCE_EXPORT __attribute__((visibility("default")))
#define RTCE_ORACLE_MAX_LEVELS 32

typedef struct { float x, y, z; } v3;

/* GLSL min/max (spec 8.3): min(x,y) = y < x ? y : x; max(x,y) = x < y ? y : x. */
static inline float f_min(float x, float y) { return (y < x) ? y : x; }
static inline float f_max(float x, float y) { return (x < y) ? y : x; }

/* GLSL cross (spec 8.5). */
static inline v3 v_cross(v3 a, v3 b) {
    v3 r;
    r.x = a.y * b.z - b.y * a.z;
    r.y = a.z * b.x - b.z * a.x;
    r.z = a.x * b.y - b.x * a.y;
    return r;
}

/* utilities.glsl:26-29  rotate(q, v): t = cross(q.xyz, v) + q.w*v; v + 2.0*cross(q.xyz, t) */
static inline v3 rotate_q(const float q[4], v3 v) {
    v3 qv = { q[0], q[1], q[2] };
    v3 c = v_cross(qv, v);
    v3 t = { c.x + q[3] * v.x, c.y + q[3] * v.y, c.z + q[3] * v.z };
    v3 c2 = v_cross(qv, t);
    v3 r = { v.x + 2.0f * c2.x, v.y + 2.0f * c2.y, v.z + 2.0f * c2.z };
    return r;
}

/* utilities.glsl:36-38  sphereSDF(p, s) = distance(p, s.pos) - s.size */
static inline float sphere_sdf(v3 p, const float* s /* x,y,z,r */) {
    float dx = p.x - s[0], dy = p.y - s[1], dz = p.z - s[2];
    float d2 = ((dx * dx) + (dy * dy)) + (dz * dz);
    return sqrtf(d2) - s[3];
}

typedef struct {
    unsigned long long E, S;
} counters_t;

/* compute.glsl:34-68  traceCone(origin, step, threshold) — tracing algorithm 3 */
static float trace_cone(v3 origin, v3 step, float threshold, const float* spheres, uint32_t n,
                        float render_dist, float* distances, counters_t* cnt) {
    for (uint32_t i = 0; i < n; i++) {                       /* :37-39 */
        distances[i] = sphere_sdf(origin, spheres + 4 * i);
    }
    cnt->E += n;
    float len = 0.0f;                                        /* :41 */
    float last = 0.0f;                                       /* :42 */
    while (len < render_dist) {                              /* :44 */
        v3 position = { origin.x + step.x * len, origin.y + step.y * len, origin.z + step.z * len }; /* :45 */
        float dist = render_dist;                            /* :49 */
        float radius = (len + 1.0f) * threshold;             /* :50 */
        for (uint32_t i = 0; i < n; i++) {                   /* :51-57 */
            distances[i] -= last;
            if (distances[i] <= radius) {
                distances[i] = sphere_sdf(position, spheres + 4 * i);
                cnt->E++;
            }
            dist = f_min(dist, distances[i]);
        }
        cnt->S++;
        last = f_max(dist, 0.0f);                            /* :59 */
        len += last;                                         /* :60 */
        if (dist <= radius) {                                /* :62-65 */
            len -= radius;
            break;
        }
    }
    return len;                                              /* :67 */
}

/* ---- the author's alternative marchers (shaders/tracing_algorithms.txt) -----------------------------------------
 * The reference holds only PSEUDO-CODE for algorithms 1 and 2 (the shader ships algorithm 3): these are restatements of
 * that text dropped into the traceCone frame of compute.glsl:34-68 (`len` is the text's `totalDist`, `threshold` its
 * `thresholdDist`, `position` the frame's :45).  Parity of anything checked against them is UNPINNED by the reference.
 * RTCE_ALT_MAX_TRIPS: the text has no trip limit; a ray still marching after that many trips is cut off here and in
 * the kernels alike (it would hang the shader). */
#define RTCE_ALT_MAX_TRIPS 65536u

/* tracing_algorithms.txt:2-13  "algorithm 1: checks real distance for each object every update" */
static float trace_cone_alg1(v3 origin, v3 step, float threshold, const float* spheres, uint32_t n,
                             float render_dist, counters_t* cnt) {
    float len = 0.0f;                                        /* frame :41 */
    uint32_t trips = 0;
    while (len < render_dist && trips++ < RTCE_ALT_MAX_TRIPS) {   /* frame :44 */
        v3 position = { origin.x + step.x * len, origin.y + step.y * len, origin.z + step.z * len }; /* frame :45 */
        float dist = sphere_sdf(position, spheres);          /* txt:2 */
        for (uint32_t i = 1; i < n; i++) {                   /* txt:3-5 */
            dist = f_min(dist, sphere_sdf(position, spheres + 4 * i));
        }
        cnt->E += n;
        cnt->S++;
        len += dist;                                         /* txt:7 */
        float radius = (len + 1.0f) * threshold;             /* txt:8 */
        if (dist <= radius) {                                /* txt:9-12 */
            len -= radius;
            break;
        }
    }
    return len;
}

/* tracing_algorithms.txt:16-37  "algorithm 2: only checks real distance when necessary" */
static float trace_cone_alg2(v3 origin, v3 step, float threshold, const float* spheres, uint32_t n,
                             float render_dist, float* distances, counters_t* cnt) {
    for (uint32_t i = 0; i < n; i++) {                       /* frame :37-39 */
        distances[i] = sphere_sdf(origin, spheres + 4 * i);
    }
    cnt->E += n;
    float len = 0.0f;
    uint32_t closest = 0;                                    /* txt:19 OUTSIDE LOOP */
    float nearest = 0.0f;                                    /* txt:20 */
    uint32_t trips = 0;
    while (len < render_dist && trips++ < RTCE_ALT_MAX_TRIPS) {
        v3 position = { origin.x + step.x * len, origin.y + step.y * len, origin.z + step.z * len }; /* frame :45 */
        for (uint32_t i = 0; i < n; i++) {                   /* txt:22-27, in this order */
            distances[i] -= nearest;
            if (distances[i] < distances[closest]) {
                closest = i;
            }
        }
        nearest = distances[closest];                        /* txt:29 */
        len += nearest;                                      /* txt:30 */
        distances[closest] = sphere_sdf(position, spheres + 4 * closest);   /* txt:31 */
        cnt->E++;
        cnt->S++;
        float radius = (len + 1.0f) * threshold;             /* txt:33 */
        if (distances[closest] <= radius) {                  /* txt:34-37 */
            len += distances[closest] - radius;
            break;
        }
    }
    return len;
}

static int g_marcher = 3;   /* set by rtce_oracle_trace_marcher around one frame (callers are single-threaded tests) */

/* compute.glsl:70-87  main() for one invocation (gx, gy) of level `iter`. */
static float shade_pixel(uint32_t gx, uint32_t gy, uint32_t iter, const float image_size[2],
                         const float ratio[2], const float rot[4], const float pos[3],
                         const float* parent, int parent_w, const float* spheres, uint32_t n,
                         float render_dist, float* distances, counters_t* cnt) {
    /* :71  (gid*2+1) * imageSize - 1.0   (uint -> float exact) */
    float ncx = (float)(gx * 2u + 1u) * image_size[0] - 1.0f;
    float ncy = (float)(gy * 2u + 1u) * image_size[1] - 1.0f;
    ncx = ncx * ratio[0];                                    /* :72 */
    ncy = ncy * ratio[1];
    /* :75  1.4142135 * gl_WorkGroupSize.x * imageSize.x, left to right */
    float threshold = (1.4142135f * 8.0f) * image_size[0];
    /* :77  normalize(rotate(rot, vec3(nc.x, 1.0, nc.y))) ; normalize(v) = v * (1/length(v)) */
    v3 v = { ncx, 1.0f, ncy };
    v3 r = rotate_q(rot, v);
    float l2 = ((r.x * r.x) + (r.y * r.y)) + (r.z * r.z);
    float inv = 1.0f / sqrtf(l2);
    v3 step = { r.x * inv, r.y * inv, r.z * inv };
    float len = 1.0f;                                        /* :79 */
    if (iter > 0u) {                                         /* :80-82  ivec2(gid * 0.5) = gid >> 1 */
        len = parent[(size_t)(gy >> 1) * (size_t)parent_w + (size_t)(gx >> 1)];
    }
    v3 origin = { pos[0] + step.x * len, pos[1] + step.y * len, pos[2] + step.z * len };
    if (g_marcher == 1) len += trace_cone_alg1(origin, step, threshold, spheres, n, render_dist, cnt);
    else if (g_marcher == 2) len += trace_cone_alg2(origin, step, threshold, spheres, n, render_dist, distances, cnt);
    else len += trace_cone(origin, step, threshold, spheres, n, render_dist, distances, cnt);   /* :84 */
    return f_max(len, 0.0f);                                 /* :86 */
}

/* utilities.glsl:32-34  repeat(p, r) = mod(p + 0.5*r, r) - 0.5*r, componentwise; GLSL mod(x, y) = x - y * floor(x / y)
 * (spec 8.3).  Unused by the reference's shaders. */
RTCE_EXPORT void rtce_oracle_repeat(const float* p3, const float* r3, float* out3, size_t n) {
    for (size_t i = 0; i < 3 * n; i++) {
        const float h = 0.5f * r3[i];
        const float a = p3[i] + h;
        const float q = a / r3[i];
        const float m = a - r3[i] * floorf(q);
        out3[i] = m - h;
    }
}

RTCE_EXPORT int rtce_oracle_max_objects(void) { return 1 << 24; }
RTCE_EXPORT int rtce_oracle_max_levels(void) { return RTCE_ORACLE_MAX_LEVELS; }
RTCE_EXPORT float rtce_oracle_render_dist(void) { return 0.0f; /* any */ }
RTCE_EXPORT int rtce_oracle_level_count(float view_w, int resize_rule) { return rtce_hm_level_count(view_w, resize_rule); }
RTCE_EXPORT void rtce_oracle_level_dims(int levels, float vw, float vh, int i, int* w, int* h) {
    rtce_hm_level_dims(levels, vw, vh, i, w, h);
}
RTCE_EXPORT void rtce_oracle_image_size(int levels, int i, float vw, float vh, float* out2) {
    rtce_hm_image_size(levels, i, vw, vh, out2);
}

/* One frame of the pass: the loop of main.rs:300-316 over compute.glsl:main.
 * Same contract as rtce_ref_trace in ref_driver.cpp. */
RTCE_EXPORT int rtce_oracle_trace(const float* spheres, uint32_t n, const float* view2, const float* ratio2,
                                  const float* rot4, const float* pos3, float render_dist, int levels,
                                  const int* rect4, int threads, float* const* level_out,
                                  unsigned long long* counters3) {
    if (levels < 1 || levels > RTCE_ORACLE_MAX_LEVELS) return -4;
    if (threads < 1) threads = omp_get_max_threads();
    int lw[RTCE_ORACLE_MAX_LEVELS], lh[RTCE_ORACLE_MAX_LEVELS];
    for (int i = 0; i < levels; i++) rtce_hm_level_dims(levels, view2[0], view2[1], i, &lw[i], &lh[i]);

    unsigned long long R = 0, E = 0, S = 0;
    int fail = 0;
    int rc[4] = { 0, 0, 0, 0 };
    float isz[2] = { 0.0f, 0.0f };
    /* One thread team per frame; the barrier that ends each level's loop is the image barrier
     * between dispatches (level i reads image i-1, compute.glsl:81). */
#pragma omp parallel num_threads(threads) reduction(+ : R, E, S) reduction(| : fail)
    {
        float* distances = (float*)malloc(sizeof(float) * (n ? n : 1));
        counters_t cnt = { 0, 0 };
        if (!distances) fail = 1;
        for (int i = 0; i < levels; i++) {
#pragma omp single
            {
                rtce_hm_image_size(levels, i, view2[0], view2[1], isz);
                rtce_hm_level_rect(levels, i, rect4, lw[i], lh[i], rc);
            } /* implicit barrier */
            const float* parent = i > 0 ? level_out[i - 1] : NULL;
            const int pw = i > 0 ? lw[i - 1] : 0;
            float* out = level_out[i];
            const int w = lw[i];
#pragma omp for schedule(dynamic, 1)
            for (int y = rc[1]; y < rc[3]; y++) {
                if (!distances) continue;
                for (int x = rc[0]; x < rc[2]; x++) {
                    out[(size_t)y * (size_t)w + (size_t)x] =
                        shade_pixel((uint32_t)x, (uint32_t)y, (uint32_t)i, isz, ratio2, rot4, pos3, parent, pw,
                                    spheres, n, render_dist, distances, &cnt);
                }
                R += (unsigned long long)(rc[2] - rc[0]);
            } /* implicit barrier */
        }
        E += cnt.E;
        S += cnt.S;
        free(distances);
    }
    if (fail) return -5;
    if (counters3) {
        counters3[0] = R;
        counters3[1] = E;
        counters3[2] = S;
    }
    return 0;
}

/* The same frame with the author's alternative marchers (1, 2; 3 = the shipped algorithm).  objCount >= 1 for 1 and 2:
 * algorithm 1 reads objs[0] unconditionally (tracing_algorithms.txt:2). */
RTCE_EXPORT int rtce_oracle_trace_marcher(const float* spheres, uint32_t n, const float* view2, const float* ratio2,
                                          const float* rot4, const float* pos3, float render_dist, int levels,
                                          const int* rect4, int threads, float* const* level_out,
                                          unsigned long long* counters3, int marcher) {
    if (marcher < 1 || marcher > 3) return -6;
    if (marcher != 3 && n < 1) return -6;
    g_marcher = marcher;
    const int r = rtce_oracle_trace(spheres, n, view2, ratio2, rot4, pos3, render_dist, levels, rect4, threads, level_out, counters3);
    g_marcher = 3;
    return r;
}

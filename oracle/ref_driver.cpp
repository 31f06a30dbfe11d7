// TEST INFRASTRUCTURE ONLY — not part of the product path.
//
// Driver around the REFERENCE'S OWN shader text.  The build recipe
// (oracle/Makefile) points RTCE_REF_COMPUTE_GLSL at a temporary copy of
// /root/reference/shaders/compute.glsl whose `#version` line is dropped (the
// only edit; the `_ext` variant additionally raises MAX_OBJECTS and the imgs[]
// array length, SURVEY.md finding 4) and compiles it as C++ behind
// glsl_shim.hpp.  Nothing of the reference is copied into the repository; the
// output is oracle/_ref/*.so (git-ignored).
//
// What this file restates itself is only the host side of the pass:
//   * the pass loop                       src/main.rs:291-316
//   * per-level push constants            src/main.rs:301-307
//   * pyramid level sizes                 src/main.rs:203-213   (rtce_host_math.h)
// The per-pixel arithmetic is compute.glsl:70-87 / utilities.glsl:26-38 verbatim.
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <vector>
#include <omp.h>

#include "rtce_host_math.h"
#include "glsl_shim.hpp"

thread_local uvec3 gl_GlobalInvocationID;
#ifdef RTCE_COUNT
thread_local unsigned long long rtce_cnt_distance = 0;
thread_local unsigned long long rtce_cnt_max = 0;
#endif

// ---- the reference shader, as it lies under /root/reference ----------------
#include RTCE_REF_COMPUTE_GLSL
// -------------------------------------------------------------------------------
#undef layout
#undef uniform
#undef readonly
#undef in
#undef main

#ifndef RTCE_REF_PREFIX
#define RTCE_REF_PREFIX rtce_ref
#endif
#define RTCE_CAT2(a, b) a##b
#define RTCE_CAT(a, b) RTCE_CAT2(a, b)
#define RTCE_FN(name) RTCE_CAT(RTCE_REF_PREFIX, name)

#define RTCE_EXPORT __attribute__((visibility("default")))

extern "C" {

RTCE_EXPORT int RTCE_FN(_max_objects)(void) { return (int)MAX_OBJECTS; }
RTCE_EXPORT int RTCE_FN(_max_levels)(void) { return (int)(sizeof(imgs) / sizeof(imgs[0])); }
RTCE_EXPORT float RTCE_FN(_render_dist)(void) { return RENDER_DIST; }
RTCE_EXPORT int RTCE_FN(_level_count)(float view_w, int resize_rule) { return rtce_hm_level_count(view_w, resize_rule); }
RTCE_EXPORT void RTCE_FN(_level_dims)(int levels, float vw, float vh, int i, int* w, int* h) {
    rtce_hm_level_dims(levels, vw, vh, i, w, h);
}
RTCE_EXPORT void RTCE_FN(_image_size)(int levels, int i, float vw, float vh, float* out2) {
    rtce_hm_image_size(levels, i, vw, vh, out2);
}

// utilities.glsl:32-34 `repeat` — defined by the reference, called by none of its shaders; exposed so that the
// device function can be compared with the reference's own text.  p3 / r3 / out3: n triples.
RTCE_EXPORT void RTCE_FN(_repeat)(const float* p3, const float* r3, float* out3, size_t n) {
    for (size_t i = 0; i < n; i++) {
        const vec3 v = repeat(vec3(p3[3 * i], p3[3 * i + 1], p3[3 * i + 2]), vec3(r3[3 * i], r3[3 * i + 1], r3[3 * i + 2]));
        out3[3 * i] = v.x;
        out3[3 * i + 1] = v.y;
        out3[3 * i + 2] = v.z;
    }
}

// One frame of the pass.  spheres = n x {x,y,z,r}; level_out[i] = w_i*h_i floats
// (row-major).  rect4 (final-level x0,y0,x1,y1) limits the work to that
// rectangle's ancestor chain; NULL / all-zero = whole image.  counters3 (may be
// NULL) receives {R rays, E sphereSDF evaluations, S march steps} when the
// library was built with RTCE_COUNT.  Returns 0, or a negative code.
RTCE_EXPORT int RTCE_FN(_trace)(const float* spheres, uint32_t n, const float* view2, const float* ratio2,
                    const float* rot4, const float* pos3, float render_dist, int levels,
                    const int* rect4, int threads, float* const* level_out,
                    unsigned long long* counters3) {
    if (n > MAX_OBJECTS) return -3;
    if (levels < 1 || levels > RTCE_FN(_max_levels)()) return -4;
    if (render_dist != RENDER_DIST) return -2;   // specialization constant 0 (compute.glsl:32) is fixed here
    if (threads < 1) threads = omp_get_max_threads();

    // binding 1: MutableData.  The pass reads objCount and objs only (compute.glsl:37-38,51,54).
    memset(&buf, 0, sizeof(buf));
    buf.objCount = n;
    for (uint32_t i = 0; i < n; i++) {
        buf.objs[i].pos = vec3(spheres[4 * i + 0], spheres[4 * i + 1], spheres[4 * i + 2]);
        buf.objs[i].size = spheres[4 * i + 3];
    }
    // binding 2: ConstantBuffer (main.rs:608-611)
    cs.view = vec2(view2[0], view2[1]);
    cs.ratio = vec2(ratio2[0], ratio2[1]);
    // push constants (main.rs:624-629, 770-773)
    pc.rot = vec4(rot4[0], rot4[1], rot4[2], rot4[3]);
    pc.pos = vec3(pos3[0], pos3[1], pos3[2]);

    // binding 0: the pyramid (main.rs:203-234)
    std::vector<int> lw(levels), lh(levels);
    for (int i = 0; i < levels; i++) {
        rtce_hm_level_dims(levels, view2[0], view2[1], i, &lw[i], &lh[i]);
        imgs[i].data = level_out[i];
        imgs[i].w = lw[i];
        imgs[i].h = lh[i];
    }

    unsigned long long R = 0, E = 0, M = 0;
    int rc[4] = {0, 0, 0, 0};
    // the pass loop, main.rs:300-316 (the per-level clear at :312 is dropped: every texel in range is
    // overwritten).  One thread team for the whole frame; the barrier after each level is the image
    // barrier vulkano inserts between dispatches (level i reads image i-1, compute.glsl:81).
#pragma omp parallel num_threads(threads) reduction(+ : R, E, M)
    {
#ifdef RTCE_COUNT
        rtce_cnt_distance = 0;
        rtce_cnt_max = 0;
#endif
        for (int i = 0; i < levels; i++) {
#pragma omp single
            {
                float isz[2];
                rtce_hm_image_size(levels, i, view2[0], view2[1], isz);
                pc.iter = (uint)i;
                pc.imageSize = vec2(isz[0], isz[1]);
                rtce_hm_level_rect(levels, i, rect4, lw[i], lh[i], rc);
            }   // implicit barrier
            const int x0 = rc[0], y0 = rc[1], x1 = rc[2], y1 = rc[3];
#pragma omp for schedule(dynamic, 1)
            for (int y = y0; y < y1; y++) {
                for (int x = x0; x < x1; x++) {
                    gl_GlobalInvocationID = uvec3((uint)x, (uint)y, 0u);   // dispatch([w/8,h/8,1]) x local 8x8
                    rtce_ref_shader_main();
                }
                R += (unsigned long long)(x1 - x0);
            }   // implicit barrier
        }
#ifdef RTCE_COUNT
        E += rtce_cnt_distance;
        M += rtce_cnt_max;
#endif
    }
    if (counters3) {
        counters3[0] = R;
        counters3[1] = E;
        counters3[2] = M >= R ? M - R : 0;
    }
    return 0;
}

}  // extern "C"

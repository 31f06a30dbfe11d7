// TEST INFRASTRUCTURE ONLY — not part of the product path.
//
// Driver around the REFERENCE'S OWN presentation shader (shaders/fragment.glsl +
// shaders/utilities.glsl, read from /root/reference at build time, `#version` line dropped,
// never copied into this repository), compiled as C++ behind glsl_shim.hpp.  Output:
// oracle/_ref/librtce_ref_frag.so (git-ignored).
//
// What this file restates itself is only the raster side of the draw:
//   * one fragment per window pixel, gl_FragCoord.xy = (x + 0.5, y + 0.5): the fullscreen
//     triangle of vertex.glsl:3-10 drawn by main.rs:318-338 covers every pixel once;
//   * the descriptor set of main.rs:644-667: binding 0 = the LAST pyramid level, binding 1 =
//     MutableData (std140, 656 bytes, main.rs:592-605), binding 2 = ConstantBuffer;
//   * push constants rot / pos (main.rs:770-773).
// The per-pixel arithmetic is fragment.glsl:123-187 verbatim, with pow() pinned as
// rtce_pinned_math.h says.  fragColor.a is never written by the shader; it is reported as 1.
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <omp.h>

#include "rtce_pinned_math.h"
#define RTCE_SHADER_MAIN rtce_ref_fragment_main
#include "glsl_shim.hpp"

thread_local uvec3 gl_GlobalInvocationID;
thread_local vec4 gl_FragCoord;
#ifdef RTCE_COUNT
thread_local unsigned long long rtce_cnt_distance = 0;
thread_local unsigned long long rtce_cnt_max = 0;
#endif

// GLSL lets a local variable shadow a function of the same name inside its own initialiser
// (fragment.glsl:180-181: `float diffuse = diffuse(...)`); C++ does not.  Function-like macros
// rename the functions and leave the variables alone.
#define diffuse(...) rtce_fn_diffuse(__VA_ARGS__)
#define specular(...) rtce_fn_specular(__VA_ARGS__)
#define pow rtce_pinned_pow
#define out thread_local

// ---- the reference shader, as it lies under /root/reference ----------------
#include RTCE_REF_FRAGMENT_GLSL
// -------------------------------------------------------------------------------
#undef out
#undef pow
#undef diffuse
#undef specular
#undef layout
#undef uniform
#undef readonly
#undef in
#undef main

#define RTCE_EXPORT __attribute__((visibility("default")))

static float f32_at(const unsigned char* b, size_t off) {
    float f;
    memcpy(&f, b + off, 4);
    return f;
}

extern "C" {

RTCE_EXPORT float rtce_ref_frag_render_dist(void) { return RENDER_DIST; }

// Shade one frame.  md656 = std140 MutableData; depth = the final pyramid level (w x h, row-major: the
// level image, whose size main.rs:209-213 rounds up from the window); the fragments are the WINDOW's
// pixels, win_w x win_h = cs.view; rgba_out = win_w*win_h*4 floats.  counters4 (may be NULL) = {pixels, hit pixels, sphereSDF evaluations,
// light evaluations}.  Returns 0 or a negative code.
RTCE_EXPORT int rtce_ref_shade(const void* md656, const float* view2, const float* ratio2, const float* rot4,
                               const float* pos3, float render_dist, const float* depth, int w, int h, int threads,
                               float* rgba_out, unsigned long long* counters4) {
    if (!md656 || !depth || !rgba_out || w < 1 || h < 1) return -1;
    const int win_w = (int)view2[0], win_h = (int)view2[1];
    if (win_w < 1 || win_h < 1 || win_w > w || win_h > h) return -6;
    if (render_dist != RENDER_DIST) return -2;
    const unsigned char* b = (const unsigned char*)md656;
    memset(&buf, 0, sizeof(buf));
    memcpy(&buf.matCount, b + 0, 4);
    memcpy(&buf.objCount, b + 4, 4);
    memcpy(&buf.lightCount, b + 8, 4);
    if (buf.objCount > MAX_OBJECTS || buf.lightCount > MAX_LIGHTS || buf.objCount < 1) return -3;
    for (uint i = 0; i < MAX_MATERIALS; i++) {   // std140: Material = 32 bytes at 16 + 32 i
        const size_t o = 16 + 32 * (size_t)i;
        buf.mats[i].color = vec3(f32_at(b, o), f32_at(b, o + 4), f32_at(b, o + 8));
        buf.mats[i].diffuse = f32_at(b, o + 12);
        buf.mats[i].specular = f32_at(b, o + 16);
        buf.mats[i].shine = f32_at(b, o + 20);
        buf.mats[i].ambient = f32_at(b, o + 24);
    }
    for (uint i = 0; i < MAX_OBJECTS; i++) {     // Object = 16 bytes at 272 + 16 i
        const size_t o = 272 + 16 * (size_t)i;
        buf.objs[i].pos = vec3(f32_at(b, o), f32_at(b, o + 4), f32_at(b, o + 8));
        buf.objs[i].size = f32_at(b, o + 12);
    }
    for (uint i = 0; i < MAX_LIGHTS; i++) {      // Light = 32 bytes at 400 + 32 i
        const size_t o = 400 + 32 * (size_t)i;
        buf.lights[i].pos = vec3(f32_at(b, o), f32_at(b, o + 4), f32_at(b, o + 8));
        buf.lights[i].color = vec3(f32_at(b, o + 16), f32_at(b, o + 20), f32_at(b, o + 24));
    }
    cs.view = vec2(view2[0], view2[1]);
    cs.ratio = vec2(ratio2[0], ratio2[1]);
    pc.rot = vec4(rot4[0], rot4[1], rot4[2], rot4[3]);
    pc.pos = vec3(pos3[0], pos3[1], pos3[2]);
    pc.iter = 0;
    pc.imageSize = vec2(view2[0], view2[1]);
    img.data = const_cast<float*>(depth);
    img.w = w;
    img.h = h;
    if (threads < 1) threads = omp_get_max_threads();
    unsigned long long E = 0, H = 0;
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads) reduction(+ : E, H)
    for (int y = 0; y < win_h; y++) {
#ifdef RTCE_COUNT
        rtce_cnt_distance = 0;
#endif
        for (int x = 0; x < win_w; x++) {
            gl_FragCoord = vec4((float)x + 0.5f, (float)y + 0.5f, 0.0f, 1.0f);
            fragColor = vec4(0.0f, 0.0f, 0.0f, 1.0f);
            rtce_ref_fragment_main();
            float* o = rgba_out + 4 * ((size_t)y * (size_t)win_w + (size_t)x);
            o[0] = fragColor.x;
            o[1] = fragColor.y;
            o[2] = fragColor.z;
            o[3] = 1.0f;
            if (depth[(size_t)y * (size_t)w + (size_t)x] < RENDER_DIST) H++;
        }
#ifdef RTCE_COUNT
        E += rtce_cnt_distance;
#endif
    }
    if (counters4) {
        counters4[0] = (unsigned long long)win_w * (unsigned long long)win_h;
        counters4[1] = H;
        counters4[2] = E;
        counters4[3] = H * (unsigned long long)buf.lightCount;
    }
    return 0;
}

}  // extern "C"

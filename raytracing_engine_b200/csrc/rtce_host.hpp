// rtce_host.hpp — host-side arithmetic of the pass (product code; the oracle has its own copy).
// Restates, in binary32 with one rounding per operation as glam/Rust do:
//   level count      src/main.rs:639 (startup, floor) and :845 (after resize, ceil)
//   level sizes      src/main.rs:209-213 (get_compute_images)
//   imageSize        src/main.rs:301-307 (per-level push constants)
//   threshold        shaders/compute.glsl:75 (level-uniform, hoisted to the host)
#pragma once
#include <cmath>
#include <cstdint>

namespace rtce {

inline int host_level_count(float view_w, bool resize_rule) {
    float l = std::log2(view_w / 8.0f);           // f32::log2
    if (resize_rule) l = std::ceil(l);
    if (!(l > 0.0f)) return 1;                    // `as usize` saturates negatives / NaN to 0
    return (int)l + 1;
}

inline void host_level_dims(int levels, float view_w, float view_h, int i, int* w, int* h) {
    const float denom = (float)((int64_t)4 << levels);     // (4 << image_count) as f32
    const float rx = view_w / denom, ry = view_h / denom;  // Vec2 / f32
    const float s = (float)((int64_t)1 << i);              // (1 << i) as f32
    const float fx = std::ceil(s * rx), fy = std::ceil(s * ry);
    const uint32_t ux = fx > 0.0f ? (uint32_t)fx : 0u, uy = fy > 0.0f ? (uint32_t)fy : 0u;   // as_uvec2
    *w = (int)(ux * 8u);
    *h = (int)(uy * 8u);
}

inline void host_image_size(int levels, int i, float view_w, float view_h, float out[2]) {
    const float s = (float)((int64_t)1 << (levels - 1 - i));
    out[0] = s / view_w;                                    // f32 / Vec2
    out[1] = s / view_h;
}

// compute.glsl:75: 1.4142135 * gl_WorkGroupSize.x * pc.imageSize.x, evaluated left to right in binary32
inline float host_threshold(float image_size_x) {
    const float a = 1.4142135f * 8.0f;
    return a * image_size_x;
}

// A final-level rectangle [x0,x1) x [y0,y1) needs, at level i, the ancestors of its pixels:
// parent = ivec2(gid * 0.5) = gid >> 1 (compute.glsl:81) => shift by (levels-1-i).
inline void host_level_rect(int levels, int i, const int rect[4], int lw, int lh, int out[4]) {
    const int s = levels - 1 - i;
    if (rect[0] == 0 && rect[1] == 0 && rect[2] == 0 && rect[3] == 0) {
        out[0] = 0; out[1] = 0; out[2] = lw; out[3] = lh;
        return;
    }
    out[0] = rect[0] >> s;
    out[1] = rect[1] >> s;
    out[2] = ((rect[2] - 1) >> s) + 1;
    out[3] = ((rect[3] - 1) >> s) + 1;
    if (out[2] > lw) out[2] = lw;
    if (out[3] > lh) out[3] = lh;
}

}  // namespace rtce

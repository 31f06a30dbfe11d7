/* TEST INFRASTRUCTURE ONLY — not part of the product path.
 *
 * Plain-C restatement of the HOST-side arithmetic of the reference's compute
 * pass (src/main.rs), shared by the C-port oracle (rtce_oracle.c) and by the
 * driver that runs the reference's own shader text (ref_driver.cpp).
 * Each function cites the reference lines it follows.  All arithmetic is
 * binary32, one rounding per operation, as in the Rust/glam source.
 */
#ifndef RTCE_HOST_MATH_H
#define RTCE_HOST_MATH_H
#include <math.h>
#include <stdint.h>

/* main.rs:639 (startup): (W / 8.0).log2() as usize + 1   -- `as usize` saturates at 0.
 * main.rs:845 (after a resize): (W / 8.0).log2().ceil() as usize + 1. */
static inline int rtce_hm_level_count(float view_w, int resize_rule) {
    float l = log2f(view_w / 8.0f);
    if (resize_rule) l = ceilf(l);
    if (!(l > 0.0f)) return 1;
    return (int)l + 1;
}

/* main.rs:209-213: ratio = max_resolution / (4 << n) as f32;
 * dims(i) = ((1 << i) as f32 * ratio).ceil().as_uvec2() * 8. */
static inline void rtce_hm_level_dims(int levels, float view_w, float view_h, int i, int* w, int* h) {
    float denom = (float)((int64_t)4 << levels);
    float rx = view_w / denom, ry = view_h / denom;
    float s = (float)((int64_t)1 << i);
    float fx = ceilf(s * rx), fy = ceilf(s * ry);
    uint32_t ux = fx > 0.0f ? (uint32_t)fx : 0u, uy = fy > 0.0f ? (uint32_t)fy : 0u;
    *w = (int)(ux * 8u);
    *h = (int)(uy * 8u);
}

/* main.rs:301-307: imageSize = (1 << (n - 1 - i)) as f32 / Vec2::from_array(view). */
static inline void rtce_hm_image_size(int levels, int i, float view_w, float view_h, float out[2]) {
    float s = (float)((int64_t)1 << (levels - 1 - i));
    out[0] = s / view_w;
    out[1] = s / view_h;
}

/* main.rs:608-611: ratio = [FOV, FOV * H / W], FOV = 1.0 (main.rs:364). */
static inline void rtce_hm_view_ratio(float fov, float view_w, float view_h, float out[2]) {
    out[0] = fov;
    out[1] = fov * view_h / view_w;
}

/* Sub-rectangle support (multi-GPU row bands, windowed CPU timing): the final-
 * level rectangle [x0,x1) x [y0,y1) needs, at level i, its ancestors
 * [x0>>s, ((x1-1)>>s)+1) x [y0>>s, ((y1-1)>>s)+1) with s = levels-1-i, because a
 * pixel's parent is ivec2(gid * 0.5) = gid >> 1 (compute.glsl:81). */
static inline void rtce_hm_level_rect(int levels, int i, const int rect[4], int lw, int lh, int out[4]) {
    int s = levels - 1 - i;
    if (!rect || (rect[0] == 0 && rect[1] == 0 && rect[2] == 0 && rect[3] == 0)) {
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

#endif

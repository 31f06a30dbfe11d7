/* TEST INFRASTRUCTURE ONLY — not part of the product path.
 *
 * Plain-C restatement ("port") of the reference's presentation shading pass
 * (shaders/fragment.glsl:39-187 drawn over the fullscreen triangle of vertex.glsl:3-10 by
 * main.rs:318-338), SURVEY.md §8(f) rank 1.  Pinned against the reference ITSELF:
 * oracle/_ref/librtce_ref_frag.so is fragment.glsl + utilities.glsl compiled as C++
 * (oracle/ref_fragment_driver.cpp); tests/test_oracle.py checks port == _ref bit for bit.
 *
 * Canonical arithmetic: glsl_shim.hpp's definitions (one binary32 operation each, never
 * contracted, dot left to right, normalize(v) = v * (1 / length(v)), reflect(i, n) =
 * i - (2 * dot(n, i)) * n); pow() as pinned in rtce_pinned_math.h.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <omp.h>

#include "rtce_pinned_math.h"

#define RTCE_EXPORT __attribute__((visibility("default")))
#define MAX_N 8   /* MAX_MATERIALS = MAX_OBJECTS = MAX_LIGHTS = 8, utilities.glsl:1-3 */

typedef struct { float x, y, z; } v3;

static inline float f_min(float x, float y) { return (y < x) ? y : x; }
static inline float f_max(float x, float y) { return (x < y) ? y : x; }
static inline v3 v_sub(v3 a, v3 b) { v3 r = { a.x - b.x, a.y - b.y, a.z - b.z }; return r; }
static inline v3 v_neg(v3 a) { v3 r = { -a.x, -a.y, -a.z }; return r; }
static inline float v_dot(v3 a, v3 b) { return ((a.x * b.x) + (a.y * b.y)) + (a.z * b.z); }
static inline float v_length(v3 a) { return sqrtf(v_dot(a, a)); }
static inline v3 v_normalize(v3 a) {
    const float inv = 1.0f / v_length(a);
    v3 r = { a.x * inv, a.y * inv, a.z * inv };
    return r;
}
static inline v3 v_cross(v3 a, v3 b) {
    v3 r = { a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y };
    return r;
}
/* utilities.glsl:26-29 */
static inline v3 rotate_q(const float q[4], v3 v) {
    v3 qv = { q[0], q[1], q[2] };
    v3 c = v_cross(qv, v);
    v3 t = { c.x + q[3] * v.x, c.y + q[3] * v.y, c.z + q[3] * v.z };
    v3 c2 = v_cross(qv, t);
    v3 r = { v.x + 2.0f * c2.x, v.y + 2.0f * c2.y, v.z + 2.0f * c2.z };
    return r;
}

typedef struct { v3 color; float diffuse, specular, shine, ambient; } material_t;
typedef struct { v3 pos; float size; } object_t;
typedef struct { v3 pos, color; } light_t;
typedef struct {
    uint32_t mat_count, obj_count, light_count;
    material_t mats[MAX_N];
    object_t objs[MAX_N];
    light_t lights[MAX_N];
} scene_t;

static float f32_at(const unsigned char* b, size_t off) {
    float f;
    memcpy(&f, b + off, 4);
    return f;
}

/* std140 image of MutableData (fragment.glsl:18-25, main.rs:592-605): 656 bytes */
static int parse_mutable_data(const unsigned char* b, scene_t* s) {
    memset(s, 0, sizeof(*s));
    memcpy(&s->mat_count, b + 0, 4);
    memcpy(&s->obj_count, b + 4, 4);
    memcpy(&s->light_count, b + 8, 4);
    if (s->obj_count < 1 || s->obj_count > MAX_N || s->light_count > MAX_N) return -3;
    for (int i = 0; i < MAX_N; i++) {
        const size_t m = 16 + 32 * (size_t)i, o = 272 + 16 * (size_t)i, l = 400 + 32 * (size_t)i;
        s->mats[i].color.x = f32_at(b, m); s->mats[i].color.y = f32_at(b, m + 4); s->mats[i].color.z = f32_at(b, m + 8);
        s->mats[i].diffuse = f32_at(b, m + 12); s->mats[i].specular = f32_at(b, m + 16);
        s->mats[i].shine = f32_at(b, m + 20); s->mats[i].ambient = f32_at(b, m + 24);
        s->objs[i].pos.x = f32_at(b, o); s->objs[i].pos.y = f32_at(b, o + 4); s->objs[i].pos.z = f32_at(b, o + 8);
        s->objs[i].size = f32_at(b, o + 12);
        s->lights[i].pos.x = f32_at(b, l); s->lights[i].pos.y = f32_at(b, l + 4); s->lights[i].pos.z = f32_at(b, l + 8);
        s->lights[i].color.x = f32_at(b, l + 16); s->lights[i].color.y = f32_at(b, l + 20); s->lights[i].color.z = f32_at(b, l + 24);
    }
    return 0;
}

typedef struct { unsigned long long E; } cnt_t;

/* utilities.glsl:36-38 */
static inline float sphere_sdf(v3 p, const object_t* s, cnt_t* c) {
    c->E++;
    return v_length(v_sub(p, s->pos)) - s->size;
}

/* fragment.glsl:89-121 */
static float shadow_ray(v3 origin, v3 step, float end, const scene_t* s, cnt_t* c) {
    const float RAY_RADIUS = 0.01f;
    float distances[MAX_N];
    for (uint32_t i = 0; i < s->obj_count; i++) distances[i] = sphere_sdf(origin, &s->objs[i], c);   /* :92-94 */
    float last = 0.0f, nearest = 1.0f;
    for (float len = 0.0f; len < end; len += last + RAY_RADIUS) {   /* :99 */
        v3 position = { origin.x + step.x * len, origin.y + step.y * len, origin.z + step.z * len };
        float dist = end;   /* :104 */
        for (uint32_t i = 0; i < s->obj_count; i++) {   /* :105-111 */
            distances[i] -= last;
            if (distances[i] <= nearest) distances[i] = sphere_sdf(position, &s->objs[i], c);
            dist = f_min(dist, distances[i]);
        }
        if (dist <= RAY_RADIUS) return 0.0f;   /* :113-115 */
        last = f_max(dist, 0.0f);
        nearest = f_min(nearest, dist);
    }
    return nearest;
}

/* fragment.glsl:123-187, one fragment */
static void shade_fragment(int x, int y, const scene_t* s, const float view[2], const float ratio[2], const float rot[4],
                           const float pos3[3], float render_dist, const float* depth, int w, float out[4], cnt_t* c) {
    const float CAM_FALL_OFF = 0.01f, LIGHT_FALL_OFF = 0.01f;
    const float fx = (float)x + 0.5f, fy = (float)y + 0.5f;   /* gl_FragCoord.xy */
    float ncx = (fx * 2.0f) / view[0] - 1.0f;   /* :126 */
    float ncy = (fy * 2.0f) / view[1] - 1.0f;
    ncx = ncx * ratio[0];   /* :128 */
    ncy = ncy * ratio[1];
    v3 dir0 = { ncx, 1.0f, ncy };
    const v3 step = v_normalize(rotate_q(rot, dir0));   /* :130 */
    const float total = depth[(size_t)y * (size_t)w + (size_t)x];   /* :132 */
    out[0] = out[1] = out[2] = 0.0f;
    out[3] = 1.0f;   /* never written by the shader; reported as 1 */
    if (total >= render_dist) return;   /* :134-137 */
    const v3 pos = { pos3[0], pos3[1], pos3[2] };
    const v3 position = { pos.x + step.x * total, pos.y + step.y * total, pos.z + step.z * total };   /* :139 */
    material_t mat = s->mats[0];
    object_t object = s->objs[0];
    float dist = sphere_sdf(position, &object, c);
    for (uint32_t i = 1; i < s->obj_count; i++) {   /* :145-153 */
        const float nd = sphere_sdf(position, &s->objs[i], c);
        if (nd < dist) {
            object = s->objs[i];
            dist = nd;
            mat = s->mats[i];
        }
    }
    c->E++;   /* distance() of :159 */
    const float cam_dist = v_length(v_sub(position, pos));
    const float cam_fall = f_max(CAM_FALL_OFF * (cam_dist * cam_dist + 1.0f), 1.0f);   /* :160 */
    const v3 normal = v_normalize(v_sub(position, object.pos));   /* :163 */
    const float normal_fall = f_max(v_dot(normal, v_neg(step)), 0.0f);   /* :164 */
    float r = 0.0f, g = 0.0f, b = 0.0f;
    for (uint32_t i = 0; i < s->light_count; i++) {   /* :167-186 */
        const light_t* L = &s->lights[i];
        const v3 ldir = v_normalize(v_sub(L->pos, position));   /* :170 */
        c->E++;   /* distance() of :171 */
        const float ldist = v_length(v_sub(position, L->pos));
        v3 so = { position.x + ldir.x, position.y + ldir.y, position.z + ldir.z };
        const float soft = f_min(shadow_ray(so, ldir, ldist, s, c), 1.0f);   /* :173 */
        const float lfall = f_max(LIGHT_FALL_OFF * ldist * ldist, 1.0f);   /* :175 */
        const float diff = f_max(v_dot(normal, ldir), 0.0f);   /* :177, :41-43 */
        /* :178, :45-48: reflect(-lightDir, normal) = i - (2 * dot(n, i)) * n */
        const v3 inc = v_neg(ldir);
        const float k = 2.0f * v_dot(normal, inc);
        const v3 refl = { inc.x - k * normal.x, inc.y - k * normal.y, inc.z - k * normal.z };
        const float spec = f_max(diff * rtce_pinned_pow(v_dot(refl, v_neg(step)), mat.shine), 0.0f);
        const float m = f_max(diff + spec, 0.0f);   /* :180 */
        v3 direct = { ((m * L->color.x) / lfall) * soft, ((m * L->color.y) / lfall) * soft, ((m * L->color.z) / lfall) * soft };
        /* :182 */
        r = r + (((mat.ambient + direct.x) / cam_fall) * normal_fall) * mat.color.x;
        g = g + (((mat.ambient + direct.y) / cam_fall) * normal_fall) * mat.color.y;
        b = b + (((mat.ambient + direct.z) / cam_fall) * normal_fall) * mat.color.z;
    }
    out[0] = r;
    out[1] = g;
    out[2] = b;
}

/* Same contract as rtce_ref_shade in ref_fragment_driver.cpp. */
RTCE_EXPORT int rtce_oracle_shade(const void* md656, const float* view2, const float* ratio2, const float* rot4,
                                  const float* pos3, float render_dist, const float* depth, int w, int h, int threads,
                                  float* rgba_out, unsigned long long* counters4) {
    if (!md656 || !depth || !rgba_out || w < 1 || h < 1) return -1;
    /* the fragments are the window's pixels (cs.view); the level image may be larger (main.rs:209-213) */
    const int win_w = (int)view2[0], win_h = (int)view2[1];
    if (win_w < 1 || win_h < 1 || win_w > w || win_h > h) return -6;
    scene_t s;
    int r = parse_mutable_data((const unsigned char*)md656, &s);
    if (r) return r;
    if (threads < 1) threads = omp_get_max_threads();
    unsigned long long E = 0, H = 0;
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads) reduction(+ : E, H)
    for (int y = 0; y < win_h; y++) {
        cnt_t c = { 0 };
        for (int x = 0; x < win_w; x++) {
            shade_fragment(x, y, &s, view2, ratio2, rot4, pos3, render_dist, depth, w,
                           rgba_out + 4 * ((size_t)y * (size_t)win_w + (size_t)x), &c);
            if (depth[(size_t)y * (size_t)w + (size_t)x] < render_dist) H++;
        }
        E += c.E;
    }
    if (counters4) {
        counters4[0] = (unsigned long long)win_w * (unsigned long long)win_h;
        counters4[1] = H;
        counters4[2] = E;
        counters4[3] = H * (unsigned long long)s.light_count;
    }
    return 0;
}

/* RGBA32F -> RGBA8 UNORM (the headless stand-in for the swapchain image, main.rs:909-914). */
RTCE_EXPORT void rtce_oracle_rgba8(const float* rgba, size_t n_pixels, unsigned char* out) {
    for (size_t i = 0; i < 4 * n_pixels; i++) out[i] = rtce_unorm8(rgba[i]);
}

RTCE_EXPORT float rtce_oracle_pinned_pow(float x, float y) { return rtce_pinned_pow(x, y); }

// TEST INFRASTRUCTURE ONLY — not part of the product path.
//
// GLSL-4.60 -> C++17 compatibility shim that lets the reference's own compute
// shader text (shaders/compute.glsl + shaders/utilities.glsl, read from
// /root/reference at build time, never copied into this repo) compile as a C++
// translation unit.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load what is built from this.
//
// This header is also where the ONE canonical arithmetic of the build is pinned
// (SURVEY.md §7 H1): GLSL leaves contraction and the precision of normalize /
// distance / sqrt implementation-defined, so "the reference output" is only
// defined up to that noise.  The canonical choice, reproduced bit-for-bit by
// the CUDA kernels, is:
//   * every + - * / sqrt is one separately rounded IEEE-754 binary32 operation
//     (build flags: -ffp-contract=off -fsingle-precision-constant, SSE2 math);
//   * dot(a,b)    = ((a.x*b.x) + (a.y*b.y)) + (a.z*b.z)      (left to right);
//   * cross(a,b)  = (a.y*b.z - b.y*a.z, a.z*b.x - b.z*a.x, a.x*b.y - b.x*a.y)
//                   (the GLSL specification's own formula);
//   * length(v)   = sqrt(dot(v,v));  distance(a,b) = length(a-b);
//   * normalize(v)= v * (1.0f / length(v));
//   * min(x,y)    = (y < x) ? y : x;   max(x,y) = (x < y) ? y : x   (GLSL spec);
//   * uint -> float conversions are exact (all indices < 2^24).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstddef>

typedef unsigned int uint;

#ifdef RTCE_COUNT
// Work counters: E = sphereSDF evaluations (one distance() call
// each, utilities.glsl:37), maxcalls = S + R (compute.glsl:59 once per march
// step, compute.glsl:86 once per ray).
extern thread_local unsigned long long rtce_cnt_distance;
extern thread_local unsigned long long rtce_cnt_max;
#define RTCE_CNT(x) (++(x))
#else
#define RTCE_CNT(x) ((void)0)
#endif

struct uvec2;
struct vec2 {
    float x, y;
    vec2() = default;
    vec2(float a, float b) : x(a), y(b) {}
    explicit vec2(float a) : x(a), y(a) {}
    inline vec2(const uvec2& u);
    vec2& operator*=(const vec2& o) { x = x * o.x; y = y * o.y; return *this; }
};
struct uvec2 {
    uint x, y;
    uvec2() = default;
    uvec2(uint a, uint b) : x(a), y(b) {}
};
inline vec2::vec2(const uvec2& u) : x((float)u.x), y((float)u.y) {}
struct ivec2 {
    int x, y;
    ivec2() = default;
    ivec2(int a, int b) : x(a), y(b) {}
    explicit ivec2(const vec2& v) : x((int)v.x), y((int)v.y) {}   // truncation, as GLSL
    explicit ivec2(const uvec2& v) : x((int)v.x), y((int)v.y) {}
};
// swizzle views must be constructor-free to live in an anonymous aggregate
struct uvec2_view { uint x, y; operator uvec2() const { return uvec2(x, y); } };
struct uvec3 {
    union {
        struct { uint x, y, z; };
        struct { uvec2_view xy; uint z_; };
    };
    uvec3() = default;
    uvec3(uint a, uint b, uint c) : x(a), y(b), z(c) {}
};
struct vec3 {
    float x, y, z;
    vec3() = default;
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
    explicit vec3(float a) : x(a), y(a), z(a) {}
};
struct vec3_view {
    float x, y, z;
    operator vec3() const { return vec3(x, y, z); }
    vec3_view& operator=(const vec3& v) { x = v.x; y = v.y; z = v.z; return *this; }           // fragColor.rgb = ...
    vec3_view& operator+=(const vec3& v) { x = x + v.x; y = y + v.y; z = z + v.z; return *this; }   // fragColor.rgb += ...
};
struct vec2_view { float x, y; operator vec2() const { return vec2(x, y); } };
struct vec4 {
    union {
        struct { float x, y, z, w; };
        struct { float r, g, b, a; };
        struct { vec3_view xyz; float w_; };
        struct { vec3_view rgb; float a_; };
        struct { vec2_view xy; float z_, w2_; };
    };
    vec4() = default;
    vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
    explicit vec4(float s) : x(s), y(s), z(s), w(s) {}
};

// ---- operators actually used by compute.glsl / utilities.glsl --------------
inline uvec2 operator*(const uvec2& a, int s) { return uvec2(a.x * (uint)s, a.y * (uint)s); }
inline uvec2 operator+(const uvec2& a, int s) { return uvec2(a.x + (uint)s, a.y + (uint)s); }
inline vec2 operator*(const uvec2& a, const vec2& b) { return vec2((float)a.x * b.x, (float)a.y * b.y); }
inline vec2 operator*(const uvec2& a, float s) { return vec2((float)a.x * s, (float)a.y * s); }
inline vec2 operator-(const vec2& a, float s) { return vec2(a.x - s, a.y - s); }
inline vec2 operator*(const vec2& a, const vec2& b) { return vec2(a.x * b.x, a.y * b.y); }
inline vec2 operator*(const vec2& a, int s) { return vec2(a.x * (float)s, a.y * (float)s); }   // fragment.glsl:126
inline vec2 operator/(const vec2& a, const vec2& b) { return vec2(a.x / b.x, a.y / b.y); }

inline vec3 operator+(const vec3& a, const vec3& b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(const vec3& a, const vec3& b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator-(const vec3& a) { return vec3(-a.x, -a.y, -a.z); }
inline vec3 operator*(const vec3& a, const vec3& b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator*(const vec3& a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, const vec3& a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator/(const vec3& a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline vec3 operator+(const vec3& a, float s) { return vec3(a.x + s, a.y + s, a.z + s); }
inline vec3 operator+(float s, const vec3& a) { return vec3(s + a.x, s + a.y, s + a.z); }   // fragment.glsl:185
inline vec3 operator-(const vec3& a, float s) { return vec3(a.x - s, a.y - s, a.z - s); }

// ---- built-ins (canonical definitions, see header comment) ------------------
inline float min(float x, float y) { return (y < x) ? y : x; }
inline float max(float x, float y) { RTCE_CNT(rtce_cnt_max); return (x < y) ? y : x; }
inline float dot(const vec3& a, const vec3& b) { return ((a.x * b.x) + (a.y * b.y)) + (a.z * b.z); }
inline vec3 cross(const vec3& a, const vec3& b) {
    return vec3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
inline float length(const vec3& v) { return sqrtf(dot(v, v)); }
inline float distance(const vec3& a, const vec3& b) { RTCE_CNT(rtce_cnt_distance); return length(a - b); }
inline vec3 normalize(const vec3& v) { float inv = 1.0f / length(v); return v * inv; }
inline float mod(float x, float y) { return x - y * floorf(x / y); }
inline vec3 mod(const vec3& a, const vec3& b) { return vec3(mod(a.x, b.x), mod(a.y, b.y), mod(a.z, b.z)); }
inline vec3 min(const vec3& a, const vec3& b) { return vec3(min(a.x, b.x), min(a.y, b.y), min(a.z, b.z)); }
inline vec3 max(const vec3& a, const vec3& b) { return vec3(max(a.x, b.x), max(a.y, b.y), max(a.z, b.z)); }
inline float clamp(float x, float lo, float hi) { return min(max(x, lo), hi); }
inline vec3 reflect(const vec3& i, const vec3& n) { return i - (2.0f * dot(n, i)) * n; }

// ---- storage images (binding 0): r32f, row-major, one float per texel --------
struct image2D {
    float* data;
    int w, h;
};
inline vec4 imageLoad(const image2D& img, const ivec2& p) {
    // Vulkan returns 0 for out-of-bounds storage-image loads (robust access);
    // never hit on the pass (2*ceil(a) >= ceil(2a) keeps every parent in range).
    if (p.x < 0 || p.y < 0 || p.x >= img.w || p.y >= img.h) return vec4(0.0f, 0.0f, 0.0f, 1.0f);
    return vec4(img.data[(size_t)p.y * (size_t)img.w + (size_t)p.x], 0.0f, 0.0f, 1.0f);
}
inline void imageStore(const image2D& img, const ivec2& p, const vec4& v) {
    if (p.x < 0 || p.y < 0 || p.x >= img.w || p.y >= img.h) return;
    img.data[(size_t)p.y * (size_t)img.w + (size_t)p.x] = v.x;
}

// ---- compute-stage built-in variables ------------------------------------------
extern thread_local uvec3 gl_GlobalInvocationID;
static const uvec3 gl_WorkGroupSize(8u, 8u, 1u);   // compute.glsl:5

// ---- declaration syntax ------------------------------------------------------------
#define layout(...)
#define uniform struct
#define readonly
#define in
#ifndef RTCE_SHADER_MAIN
#define RTCE_SHADER_MAIN rtce_ref_shader_main
#endif
#define main RTCE_SHADER_MAIN

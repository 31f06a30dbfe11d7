// rtce_shade_math.cuh — device arithmetic of the presentation shading pass (fragment.glsl).
//
// Canonical arithmetic as in rtce_device.cuh (one separately rounded binary32 operation per GLSL
// operation), plus the PINNED pow(): GLSL leaves pow's precision implementation-defined and its
// value undefined for x <= 0, so the build fixes one definition and states it operation for
// operation: pow(x, y) = 0 when !(x > 0) (what max(NaN, 0) yields on the hardware path: no
// highlight), else exp2_p(y * log2_p(x)) with
//   log2_p: x = m 2^e, m in [sqrt 1/2, sqrt 2); s = (m - 1) / (m + 1); z = s s;
//           e + s (c1 + z (c3 + z (c5 + z (c7 + z c9)))),  c_k = 2 / (k ln 2), Horner in fma;
//   exp2_p: n = floor(t + 0.5), f = t - n; degree-7 Taylor of e^(f ln 2), Horner in fma;
//           scaled by 2^n as two exponent-constructed factors.
// Relative error < 2e-5 for shine <= 256 (the error of log2_p times y).  The CPU checker evaluates
// the same sequence, so both sides agree bit for bit.
#pragma once
#include "rtce_device.cuh"

namespace rtce {

// GLSL min / max exactly as the specification words them (the shading pass can meet +-0: a
// surface point exactly on a sphere): min(x, y) = y < x ? y : x;  max(x, y) = x < y ? y : x.
RTCE_DEV float lit_min(float x, float y) { return (y < x) ? y : x; }
RTCE_DEV float lit_max(float x, float y) { return (x < y) ? y : x; }

RTCE_DEV float log2_p(float x) {   // x > 0, finite
    int e = 0;
    if (x < 1.17549435e-38f) {
        x = fmul(x, 16777216.0f);
        e = -24;
    }
    unsigned u = __float_as_uint(x);
    e += (int)(u >> 23) - 127;
    u = (u & 0x007FFFFFu) | 0x3F800000u;
    float m = __uint_as_float(u);
    if (m >= 1.41421354f) {
        m = fmul(m, 0.5f);
        e += 1;
    }
    const float s = __fdiv_rn(fsub(m, 1.0f), fadd(m, 1.0f));
    const float z = fmul(s, s);
    float p = 0.320598892f;
    p = __fmaf_rn(p, z, 0.412198573f);
    p = __fmaf_rn(p, z, 0.577078009f);
    p = __fmaf_rn(p, z, 0.961796701f);
    p = __fmaf_rn(p, z, 2.88539004f);
    return __fmaf_rn(s, p, (float)e);
}

RTCE_DEV float exp2_p(float t) {
    if (!(t > -150.0f)) return 0.0f;
    if (t >= 128.0f) return __int_as_float(0x7F800000);
    const float n = floorf(fadd(t, 0.5f));
    const float f = fsub(t, n);
    float p = 1.52527338e-5f;
    p = __fmaf_rn(p, f, 1.54035304e-4f);
    p = __fmaf_rn(p, f, 1.33335581e-3f);
    p = __fmaf_rn(p, f, 9.61812911e-3f);
    p = __fmaf_rn(p, f, 5.55041087e-2f);
    p = __fmaf_rn(p, f, 2.40226507e-1f);
    p = __fmaf_rn(p, f, 6.93147182e-1f);
    p = __fmaf_rn(p, f, 1.0f);
    const int ni = (int)n;
    const int a = ni / 2, b = ni - a;
    const float fa = __uint_as_float((unsigned)(a + 127) << 23);
    const float fb = __uint_as_float((unsigned)(b + 127) << 23);
    return fmul(fmul(p, fa), fb);
}

RTCE_DEV float pinned_pow(float x, float y) {
    if (!(x > 0.0f)) return 0.0f;
    if (x == __int_as_float(0x7F800000)) return y > 0.0f ? x : (y < 0.0f ? 0.0f : 1.0f);
    return exp2_p(fmul(y, log2_p(x)));
}

// Float -> UNORM8 (Vulkan "conversion from floating point to normalized fixed-point": clamp, scale,
// round to nearest even; NaN -> 0).
RTCE_DEV unsigned unorm8(float c) {
    if (!(c > 0.0f)) return 0u;
    if (c >= 1.0f) return 255u;
    return (unsigned)__float2int_rn(fmul(c, 255.0f));
}

}  // namespace rtce

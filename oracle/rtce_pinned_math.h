/* TEST INFRASTRUCTURE ONLY — not part of the product path.
 *
 * Pinned definition of GLSL pow() for the presentation shading pass (fragment.glsl:47).
 *
 * GLSL leaves pow(x, y) implementation-defined in precision and undefined for x < 0 (and for
 * x = 0, y <= 0).  GPUs evaluate it as exp2(y * log2(x)) with approximate lg2 / ex2 units, whose
 * bits no CPU can reproduce.  "The reference output" of the shading pass is therefore only defined
 * up to that noise; the canonical choice of this build — restated operation for operation by the
 * CUDA kernel (csrc/rtce_shade_math.cuh) — is:
 *
 *   pow(x, y) = 0                                   if !(x > 0)      (what max(NaN, 0) gives on the
 *                                                                     hardware path: no highlight)
 *             = exp2_p(y * log2_p(x))               otherwise
 *
 *   log2_p(x): x = m * 2^e with m in [sqrt(1/2), sqrt(2)); s = (m - 1) / (m + 1); z = s * s;
 *              log2(m) = s * (c1 + z * (c3 + z * (c5 + z * (c7 + z * c9)))), Horner with fmaf,
 *              c_k = 2 / (k ln 2); result = (float)e + that.            (abs error < 2e-7)
 *   exp2_p(t): t <= -150 -> 0, t >= 128 -> +inf; n = floor(t + 0.5); f = t - n in [-0.5, 0.5];
 *              2^f = Horner of degree 7 in f with fmaf (Taylor of e^(f ln 2)); times 2^n by
 *              exponent construction (two factors, so that subnormal results round once more
 *              the same way everywhere).                                 (rel error < 3e-7)
 *
 * Every operation is one IEEE-754 binary32 operation (fmaf is the correctly rounded fused
 * multiply-add on both sides), so CPU and GPU agree bit for bit.
 */
#ifndef RTCE_PINNED_MATH_H
#define RTCE_PINNED_MATH_H
#include <math.h>
#include <stdint.h>
#include <string.h>

static inline float rtce_bits_to_float(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint32_t rtce_float_to_bits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

static inline float rtce_log2_p(float x) {   /* x > 0, finite */
    int e = 0;
    if (x < 1.17549435e-38f) {   /* subnormal: scale into the normal range */
        x = x * 16777216.0f;
        e = -24;
    }
    uint32_t u = rtce_float_to_bits(x);
    e += (int)(u >> 23) - 127;
    u = (u & 0x007FFFFFu) | 0x3F800000u;   /* m in [1, 2) */
    float m = rtce_bits_to_float(u);
    if (m >= 1.41421354f) {   /* [sqrt 2, 2) -> [sqrt 1/2, 1) */
        m = m * 0.5f;
        e += 1;
    }
    const float s = (m - 1.0f) / (m + 1.0f);
    const float z = s * s;
    float p = 0.320598892f;              /* 2 / (9 ln 2) */
    p = fmaf(p, z, 0.412198573f);        /* 2 / (7 ln 2) */
    p = fmaf(p, z, 0.577078009f);        /* 2 / (5 ln 2) */
    p = fmaf(p, z, 0.961796701f);        /* 2 / (3 ln 2) */
    p = fmaf(p, z, 2.88539004f);         /* 2 / ln 2 */
    return fmaf(s, p, (float)e);
}

static inline float rtce_exp2_p(float t) {
    if (!(t > -150.0f)) return 0.0f;     /* also NaN -> 0 */
    if (t >= 128.0f) return INFINITY;
    const float n = floorf(t + 0.5f);
    const float f = t - n;               /* exact: |f| <= 0.5 */
    float p = 1.52527338e-5f;            /* (ln 2)^7 / 7! */
    p = fmaf(p, f, 1.54035304e-4f);      /* (ln 2)^6 / 6! */
    p = fmaf(p, f, 1.33335581e-3f);      /* (ln 2)^5 / 5! */
    p = fmaf(p, f, 9.61812911e-3f);      /* (ln 2)^4 / 4! */
    p = fmaf(p, f, 5.55041087e-2f);      /* (ln 2)^3 / 3! */
    p = fmaf(p, f, 2.40226507e-1f);      /* (ln 2)^2 / 2! */
    p = fmaf(p, f, 6.93147182e-1f);      /* ln 2 */
    p = fmaf(p, f, 1.0f);
    /* 2^n = 2^a * 2^b with a = n / 2 rounded toward zero: both factors are normal for n in [-150, 128) */
    const int ni = (int)n;
    const int a = ni / 2, b = ni - a;
    const float fa = rtce_bits_to_float((uint32_t)(a + 127) << 23);
    const float fb = rtce_bits_to_float((uint32_t)(b + 127) << 23);
    return (p * fa) * fb;
}

static inline float rtce_pinned_pow(float x, float y) {
    if (!(x > 0.0f)) return 0.0f;
    if (x == INFINITY) return y > 0.0f ? INFINITY : (y < 0.0f ? 0.0f : 1.0f);
    return rtce_exp2_p(y * rtce_log2_p(x));
}

/* UNORM8 conversion of a colour channel for the headless RGBA8 framebuffer (Vulkan spec, "Conversion from
 * floating point to normalized fixed-point": clamp to [0, 1], scale by 255, round to nearest even; NaN -> 0). */
static inline unsigned char rtce_unorm8(float c) {
    if (!(c > 0.0f)) return 0;
    if (c >= 1.0f) return 255;
    return (unsigned char)(int)rintf(c * 255.0f);
}

#endif

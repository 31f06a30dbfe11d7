// rtce_camera.cu — the step before the pass: camera state -> PushConstantData.
// Host code only.  Restates, in binary32 with one rounding per operation (built with -ffp-contract=off):
//   Data::rotation   src/main.rs:402-404   Quat::from_rotation_z(-yaw) * Quat::from_rotation_x(pitch)
//   Data::position   src/main.rs:406-414   x * right + y * forward + z * up, axes rotated by Quat::mul_vec3
//   the frame update src/main.rs:770-773   pitch clamp, rot = rotation().to_array(), pos += position()
// glam 0.21.3 (Cargo.lock:465-466) is not vendored; the formulas are its published ones:
//   from_rotation_z(a) = (0, 0, sin(a/2), cos(a/2)), from_rotation_x(a) = (sin(a/2), 0, 0, cos(a/2))  [x, y, z, w]
//   q * v = v (w^2 - b.b) + b (2 (v.b)) + (b x v) (2 w),  b = q.xyz                                  (Quat::mul_vec3)
// sin / cos are this machine's libm sinf / cosf (glam: f32::sin_cos): parity with glam on the author's machine is
// UNPINNED (the Rust host cannot be built or run anywhere but there, Cargo.toml:9).  What IS exact: the product of
// the two axis rotations has one non-zero term per component, so its bits do not depend on glam's summation order
// (scalar or SSE2 path), only on sinf / cosf.
#include <cmath>

#include "../../include/rtce.h"

namespace {

struct V3 { float x, y, z; };

inline float dot3(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline V3 cross3(V3 a, V3 b) { return V3{a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
inline V3 scale3(V3 a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
inline V3 add3(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }

// glam Quat::mul_vec3
inline V3 quat_mul_vec3(const float q[4], V3 v) {
    const float w = q[3];
    const V3 b{q[0], q[1], q[2]};
    const float b2 = dot3(b, b);
    return add3(add3(scale3(v, w * w - b2), scale3(b, dot3(v, b) * 2.0f)), scale3(cross3(b, v), w * 2.0f));
}

void rotation(float yaw, float pitch, float q[4]) {
    // from_rotation_z(-yaw) = (0, 0, sz, cz); from_rotation_x(pitch) = (sx, 0, 0, cx)
    const float hz = -yaw * 0.5f, hx = pitch * 0.5f;
    const float sz = sinf(hz), cz = cosf(hz), sx = sinf(hx), cx = cosf(hx);
    // Hamilton product (x0, y0, z0, w0) * (x1, y1, z1, w1) with x0 = y0 = 0, y1 = z1 = 0: every other term is a zero;
    // + 0.0f normalises a -0 product the way any summation with the +0 terms does
    q[0] = cz * sx + 0.0f;   // w0 x1
    q[1] = sz * sx + 0.0f;   // z0 x1
    q[2] = sz * cx + 0.0f;   // z0 w1
    q[3] = cz * cx + 0.0f;   // w0 w1
}

}  // namespace

extern "C" int rtce_camera_rotation(float yaw, float pitch, float rot_xyzw[4]) {
    if (!rot_xyzw) return RTCE_ERR_INVALID;
    rotation(yaw, pitch, rot_xyzw);
    return RTCE_OK;
}

extern "C" int rtce_camera_pose(float yaw, float pitch, const float local_delta[3], rtce_push_constants* inout) {
    if (!inout) return RTCE_ERR_INVALID;
    const float half_pi = 0.5f * 3.14159265358979323846f;   // 0.5 * std::f32::consts::PI
    if (pitch < -half_pi) pitch = -half_pi;                   // f32::clamp (main.rs:770)
    if (pitch > half_pi) pitch = half_pi;
    float q[4];
    rotation(yaw, pitch, q);
    for (int i = 0; i < 4; i++) inout->rot[i] = q[i];          // main.rs:771
    if (local_delta) {                                         // main.rs:406-414, 772
        const V3 right = quat_mul_vec3(q, V3{1.0f, 0.0f, 0.0f});
        const V3 forward = quat_mul_vec3(q, V3{0.0f, 1.0f, 0.0f});
        const V3 up = quat_mul_vec3(q, V3{0.0f, 0.0f, 1.0f});
        const V3 d = add3(add3(scale3(right, local_delta[0]), scale3(forward, local_delta[1])), scale3(up, local_delta[2]));
        inout->pos[0] = inout->pos[0] + d.x;
        inout->pos[1] = inout->pos[1] + d.y;
        inout->pos[2] = inout->pos[2] + d.z;
    }
    return RTCE_OK;
}

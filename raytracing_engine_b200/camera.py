"""Camera -> push constants, the step before the pass (src/main.rs:343-357, 402-414, 770-773).

rotation = Quat::from_rotation_z(-yaw) * Quat::from_rotation_x(pitch)  (main.rs:402-404), stored
as glam's Quat::to_array() = [x, y, z, w] in PushConstantData.rot; forward = +Y, up = +Z, right = +X
(main.rs:351-357).  glam 0.21 computes sin/cos in f32 via libm; to keep libm out of the parity
question the pose generator here evaluates in float64 and rounds once to float32 — oracle and
kernels always receive the SAME quaternion bytes.
"""
from __future__ import annotations

import math

import numpy as np

UP = (0.0, 0.0, 1.0)
FORWARD = (0.0, 1.0, 0.0)
RIGHT = (1.0, 0.0, 0.0)
IDENTITY_ROT = (0.0, 0.0, 0.0, 1.0)


def quat_mul(a, b):
    """Hamilton product of [x,y,z,w] quaternions (glam Quat * Quat)."""
    ax, ay, az, aw = a
    bx, by, bz, bw = b
    return (aw * bx + ax * bw + ay * bz - az * by,
            aw * by - ax * bz + ay * bw + az * bx,
            aw * bz + ax * by - ay * bx + az * bw,
            aw * bw - ax * bx - ay * by - az * bz)


def rotation(yaw: float, pitch: float) -> np.ndarray:
    """main.rs:402-404: from_rotation_z(-yaw) * from_rotation_x(pitch) as float32 [x,y,z,w]."""
    qz = (0.0, 0.0, math.sin(-yaw * 0.5), math.cos(-yaw * 0.5))
    qx = (math.sin(pitch * 0.5), 0.0, 0.0, math.cos(pitch * 0.5))
    return np.array(quat_mul(qz, qx), dtype=np.float32)


def quat_rotate(q, v):
    """glam Quat::mul_vec3 in float64 (used for the position update main.rs:406-414)."""
    x, y, z, w = [float(c) for c in q]
    qv = np.array([x, y, z])
    v = np.asarray(v, dtype=np.float64)
    t = 2.0 * np.cross(qv, v)
    return v + w * t + np.cross(qv, t)


def advance(pos, yaw: float, pitch: float, local_delta) -> np.ndarray:
    """main.rs:406-414 + 772: pos += dx*right + dy*forward + dz*up, axes rotated by the camera."""
    q = rotation(yaw, pitch)
    d = (local_delta[0] * quat_rotate(q, RIGHT) + local_delta[1] * quat_rotate(q, FORWARD)
         + local_delta[2] * quat_rotate(q, UP))
    return (np.asarray(pos, dtype=np.float64) + d).astype(np.float32)


def bench_camera_path(frames: int):
    """SURVEY.md §8d config 1 pose k: yaw = 0.01k, pitch = 0.005k, pos = (0.05k, 0, 0)."""
    rots = np.stack([rotation(0.01 * k, 0.005 * k) for k in range(frames)])
    poss = np.array([[np.float32(0.05 * k), 0.0, 0.0] for k in range(frames)], dtype=np.float32)
    return rots, poss


def static_path(frames: int, rot=IDENTITY_ROT, pos=(0.0, 0.0, 0.0)):
    rots = np.tile(np.asarray(rot, dtype=np.float32), (frames, 1))
    poss = np.tile(np.asarray(pos, dtype=np.float32), (frames, 1))
    return rots, poss

"""Case table shared by make_golden.py (generator, needs oracle/_ref) and the tests.

Every case is fully determined by these literals, so the GPU box can rebuild identical inputs
without /root/reference.  `kind` names the checker that minted the fixture: "ref" = the
reference's own shader text, unmodified limits; "refx" = same text with MAX_OBJECTS / imgs[] raised.
"""
import numpy as np

from raytracing_engine_b200 import camera, scene

DEFAULT = scene.spheres_of(scene.default_mutable_data())


def _rng_spheres(n, seed, box=8.0, rmax=2.0):
    rs = np.random.RandomState(seed)
    s = np.zeros((n, 4), dtype=np.float32)
    s[:, 0] = rs.uniform(-box, box, n)
    s[:, 1] = rs.uniform(2.0, 2.0 + 2 * box, n)
    s[:, 2] = rs.uniform(-box, box, n)
    s[:, 3] = rs.uniform(0.2, rmax, n)
    return s


def cases():
    ident = np.array(camera.IDENTITY_ROT, dtype=np.float32)
    zero = np.zeros(3, dtype=np.float32)
    rots, poss = camera.bench_camera_path(16)
    c = []
    # (name, kind, spheres, W, H, levels(0 = startup rule, -1 = resize rule), rot, pos, keep_images)
    c.append(("default_800x600", "ref", DEFAULT, 800, 600, 0, ident, zero, False))
    c.append(("default_1080p", "ref", DEFAULT, 1920, 1080, 0, ident, zero, False))
    c.append(("default_1080p_pose5", "ref", DEFAULT, 1920, 1080, 0, rots[5], poss[5], False))
    c.append(("default_1080p_pose15", "ref", DEFAULT, 1920, 1080, 0, rots[15], poss[15], False))
    c.append(("default_4k", "ref", DEFAULT, 3840, 2160, 0, ident, zero, False))
    c.append(("default_8k_10levels", "refx", DEFAULT, 7680, 4320, 0, ident, zero, False))
    c.append(("default_64x48", "ref", DEFAULT, 64, 48, 0, ident, zero, True))
    c.append(("default_250x130_resize_rule", "ref", DEFAULT, 250, 130, -1, rots[3], poss[3], True))
    c.append(("default_800x600_resize_rule", "ref", DEFAULT, 800, 600, -1, ident, zero, False))
    c.append(("empty_scene_160x96", "ref", np.zeros((0, 4), np.float32), 160, 96, 0, ident, zero, True))
    c.append(("one_sphere_160x96", "ref", DEFAULT[:1], 160, 96, 0, rots[7], poss[7], True))
    c.append(("eight_spheres_320x200", "ref", _rng_spheres(8, 1), 320, 200, 0, rots[9], poss[9], False))
    # MAX_OBJECTS = 8 (utilities.glsl:3) at BASELINE's 1080p: the default four spheres plus four more
    c.append(("eight_spheres_1080p", "ref", np.concatenate([DEFAULT, _rng_spheres(4, 11, box=9.0, rmax=2.5)]), 1920, 1080, 0,
              rots[6], poss[6], False))
    # camera inside a sphere (negative SDF at the origin) and looking away (mostly misses)
    c.append(("inside_sphere_200x120", "ref", np.array([[0, 0.5, 0, 3.0], [4, 9, 1, 2.0]], np.float32), 200, 120, 0,
              ident, zero, True))
    c.append(("looking_back_320x200", "ref", DEFAULT, 320, 200, 0, camera.rotation(3.0, -0.4),
              np.array([1.0, -2.0, 0.5], np.float32), False))
    c.append(("hundred_spheres_128x72", "refx", _rng_spheres(100, 2, box=20.0), 128, 72, 0, ident, zero, True))
    c.append(("thousand_spheres_64x40", "refx", _rng_spheres(1000, 3, box=40.0, rmax=1.5), 64, 40, 0, rots[2], poss[2], True))
    return c

"""Cases of the presentation shading pass (fragment.glsl) shared by make_shade_golden.py and the tests."""
import numpy as np

from raytracing_engine_b200 import camera, scene


def random_mutable_data(seed: int) -> bytes:
    """Seeded scene with materials and lights: 1..8 spheres, 0..8 lights, shine 1..200."""
    rs = np.random.RandomState(seed)
    n = int(rs.randint(1, 9))
    nl = int(rs.randint(0, 9))
    objs = [((float(rs.uniform(-8, 8)), float(rs.uniform(4, 20)), float(rs.uniform(-6, 6))), float(rs.uniform(0.3, 4.0)))
            for _ in range(n)]
    mats = [((float(rs.uniform(0, 1)), float(rs.uniform(0, 1)), float(rs.uniform(0, 1))), float(rs.uniform(0, 1)),
             float(rs.uniform(0, 1)), float(rs.uniform(1, 200)), float(rs.uniform(0, 0.3))) for _ in range(n)]
    lights = [((float(rs.uniform(-15, 15)), float(rs.uniform(-5, 25)), float(rs.uniform(-15, 15))),
               (float(rs.uniform(0, 3)), float(rs.uniform(0, 3)), float(rs.uniform(0, 3)))) for _ in range(nl)]
    return scene.pack_mutable_data(materials=mats, objects=objs, lights=lights)


def shade_cases():
    """name -> (mutable_data bytes, W, H, rot, pos)"""
    out = {}
    rots, poss = camera.bench_camera_path(16)
    out["default_800x600"] = (scene.default_mutable_data(), 800, 600, camera.IDENTITY_ROT, np.zeros(3, np.float32))
    out["default_800x600_pose9"] = (scene.default_mutable_data(), 800, 600, rots[9], poss[9])
    out["default_1080p"] = (scene.default_mutable_data(), 1920, 1080, camera.IDENTITY_ROT, np.zeros(3, np.float32))
    out["default_1080p_pose15"] = (scene.default_mutable_data(), 1920, 1080, rots[15], poss[15])
    for seed in (1, 2, 3, 4, 5, 6):
        out[f"random{seed}_333x201"] = (random_mutable_data(seed), 333, 201, camera.rotation(0.1 * seed, -0.05 * seed),
                                        np.array([0.2 * seed, -1.0, 0.1], np.float32))
    return out

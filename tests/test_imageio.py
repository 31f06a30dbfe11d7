"""Headless writers: PFM / PPM / scene file round trips, on checker output."""
import numpy as np

import oracle
from raytracing_engine_b200 import camera, imageio, scene


def test_pfm_ppm_scene_round_trip(tmp_path):
    md = scene.default_mutable_data()
    o = oracle.load("port")
    W, H = 160, 96
    lv, _ = o.trace(scene.spheres_of(md), [W, H], scene.view_ratio(W, H), camera.IDENTITY_ROT, [0, 0, 0], 1000.0, 0)
    rgba, _ = o.shade(md, [W, H], scene.view_ratio(W, H), camera.IDENTITY_ROT, [0, 0, 0], lv[-1])
    imageio.write_pfm(tmp_path / "depth.pfm", lv[-1])
    imageio.write_pfm(tmp_path / "colour.pfm", rgba)
    imageio.write_ppm(tmp_path / "frame.ppm", oracle.rgba8(rgba))
    assert np.array_equal(imageio.read_pfm(tmp_path / "depth.pfm").view(np.uint32), lv[-1].view(np.uint32))
    assert np.array_equal(imageio.read_pfm(tmp_path / "colour.pfm"), rgba[..., :3])
    assert np.array_equal(imageio.read_ppm(tmp_path / "frame.ppm"), oracle.rgba8(rgba)[..., :3])
    assert (tmp_path / "frame.ppm").read_bytes().startswith(b"P6\n160 96\n255\n")
    imageio.write_scene(tmp_path / "a.scn", mutable_data=md)
    assert imageio.read_scene(tmp_path / "a.scn") == ("mutable_data", md)
    sph = scene.synthetic_spheres(100)
    imageio.write_scene(tmp_path / "b.scn", spheres=sph)
    kind, back = imageio.read_scene(tmp_path / "b.scn")
    assert kind == "spheres" and np.array_equal(back, sph)

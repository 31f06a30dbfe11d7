"""Headless render of one frame through the C ABI: compute pass + presentation shading pass -> PFM (depth) and
PPM (RGBA8 colour).  GPU only.   usage: render_frame.py <out_prefix> [W H] [pose index on the bench camera path]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracing_engine_b200 as rt
from raytracing_engine_b200 import camera, imageio, scene

out = sys.argv[1]
W, H = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (800, 600)
k = int(sys.argv[4]) if len(sys.argv) > 4 else 0
rots, poss = camera.bench_camera_path(k + 1)
pcs = rt.make_push_constants(rots[k:k + 1], poss[k:k + 1])
with rt.Engine(0, scene.RENDER_DIST) as e:
    e.set_scene(scene.default_mutable_data())
    e.set_view(W, H)
    e.trace(pcs)
    e.shade(pcs)
    depth = e.readback(-1, 0)
    imageio.write_pfm(out + "_depth.pfm", depth[:H, :W])
    imageio.write_ppm(out + ".ppm", e.readback_color_rgba8(0))
    print(f"wrote {out}.ppm and {out}_depth.pfm ({W}x{H}, pose {k})")

"""Diagnostics: event timeline of the kernel launches of one bench step (GPU only)."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracing_engine_b200 as rt
from raytracing_engine_b200 import camera, scene

groups = int(sys.argv[1]) if len(sys.argv) > 1 else 2
min_rays = int(sys.argv[2]) if len(sys.argv) > 2 else 400000
frames = 16
rots, poss = camera.bench_camera_path(frames)
pcs = rt.make_push_constants(rots, poss)
with rt.Engine(0, 1000.0) as e:
    e.set_option(rt.OPT_FRAME_SLOTS, frames)
    e.set_option(rt.OPT_GROUPS, groups)
    e.set_option(rt.OPT_STRIP_MIN_RAYS, min_rays)
    e.set_scene(scene.default_mutable_data())
    e.set_view(1920, 1080)
    for _ in range(3):
        e.trace(pcs)
    e.sync()
    e.launch_timeline(fetch=False)
    e.trace(pcs)
    t = e.launch_timeline()
    names = {0: "perpixel", 1: "strip", 2: "overflow", 3: "pix", 4: "cluster", 10: "fused"}
    print(f"groups={groups} min_rays={min_rays}: {len(t)} launches, span {t[:, 1].max() * 1e3:.1f} us")
    for r in t:
        print(f"  g{int(r[4]):2d} {names[int(r[2])]:9s} L{int(r[3])}  {r[0] * 1e3:8.1f} -> {r[1] * 1e3:8.1f}  ({(r[1] - r[0]) * 1e3:7.1f} us)")

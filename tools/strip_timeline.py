"""Diagnostics: per-warp timeline of the strip kernel on the bench workload (GPU only)."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracing_engine_b200 as rt
from raytracing_engine_b200 import camera, scene

level = int(sys.argv[1]) if len(sys.argv) > 1 else -1
groups = int(sys.argv[2]) if len(sys.argv) > 2 else 1
frames = 16
rots, poss = camera.bench_camera_path(frames)
pcs = rt.make_push_constants(rots, poss)
with rt.Engine(0, 1000.0) as e:
    e.set_option(rt.OPT_FRAME_SLOTS, frames)
    e.set_option(rt.OPT_GROUPS, groups)
    e.set_scene(scene.default_mutable_data())
    e.set_view(1920, 1080)
    for _ in range(3):
        e.trace(pcs)
    e.sync()
    e.strip_timeline(level, fetch=False)
    e.trace(pcs)
    t = e.strip_timeline(level).astype(np.int64)
    t0 = t[:, 0].min()
    start, tick, end, n = t[:, 0] - t0, t[:, 1] - t0, t[:, 2] - t0, t[:, 3]
    print(f"level {level}: warps {len(t)}, kernel span {end.max() / 1e3:.1f} us")
    q = [0, 10, 50, 90, 99, 100]
    print("  start        us pct", dict(zip(q, np.round(np.percentile(start, q) / 1e3, 1))))
    print("  tickets out  us pct", dict(zip(q, np.round(np.percentile(tick, q) / 1e3, 1))))
    print("  exit         us pct", dict(zip(q, np.round(np.percentile(end, q) / 1e3, 1))))
    print("  drain        us pct", dict(zip(q, np.round(np.percentile(end - tick, q) / 1e3, 1))))
    print("  tasks/warp      pct", dict(zip(q, np.percentile(n, q))))

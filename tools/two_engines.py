"""Experiment: two contexts tracing the bench batch alternately on their own streams (frames in flight)."""
import sys, os, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracing_engine_b200 as rt
from raytracing_engine_b200 import camera, scene

n_eng = int(sys.argv[1]) if len(sys.argv) > 1 else 2
groups = int(sys.argv[2]) if len(sys.argv) > 2 else 2
frames = 16
rots, poss = camera.bench_camera_path(frames)
pcs = rt.make_push_constants(rots, poss)
engs = []
for i in range(n_eng):
    e = rt.Engine(0, 1000.0)
    e.set_option(rt.OPT_FRAME_SLOTS, frames)
    e.set_option(rt.OPT_GROUPS, groups)
    e.set_scene(scene.default_mutable_data())
    e.set_view(1920, 1080)
    engs.append(e)
for _ in range(5):
    for e in engs:
        e.trace(pcs)
for e in engs:
    e.sync()
K = 200
t0 = time.perf_counter()
for _ in range(K):
    for e in engs:
        e.trace(pcs)
for e in engs:
    e.sync()
dt = time.perf_counter() - t0
rays = 44340224 * K * n_eng
print(f"engines={n_eng} groups={groups}: {dt / (K * n_eng) * 1e6:.1f} us per 16-frame batch, {rays / dt / 1e9:.1f} Grays/s")

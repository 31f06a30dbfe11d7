"""Is the crowded configs[1] loop bound by the host?  Time spent ENQUEUEING n batches (host clock, before any sync) against
the time until they have all finished, for the bench workload and for a tiny view whose device work is negligible."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracing_engine_b200 as rt
from raytracing_engine_b200 import camera, scene

def run(W, H, n=3000):
    rots, poss = camera.bench_camera_path(16)
    pcs = rt.make_push_constants(rots, poss)
    with rt.Engine(0, 1000.0) as e:
        e.set_scene(scene.default_mutable_data())
        e.set_view(W, H)
        e.prepare(16)
        for _ in range(50):
            e.trace(pcs)
        e.sync()
        t0 = time.perf_counter()
        for _ in range(n):
            e.trace(pcs)
        t1 = time.perf_counter()
        e.sync()
        t2 = time.perf_counter()
        print(f"{W}x{H}: enqueue {1e6 * (t1 - t0) / n:.1f} us per batch, finished after {1e6 * (t2 - t0) / n:.1f} us per batch")

run(1920, 1080)
run(64, 48)

"""Diagnostics: device time of the shading pass on the bench batch (GPU only)."""
import sys, os
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracing_engine_b200 as rt
from raytracing_engine_b200 import camera, scene

frames = 16
rots, poss = camera.bench_camera_path(frames)
pcs = rt.make_push_constants(rots, poss)
with rt.Engine(0, 1000.0) as e:
    st = torch.cuda.Stream()
    e.set_stream(st.cuda_stream)
    e.set_option(rt.OPT_FRAME_SLOTS, frames)
    e.set_scene(scene.default_mutable_data())
    e.set_view(1920, 1080)
    e.set_option(rt.OPT_COUNTERS, 1)
    e.trace(pcs); e.shade(pcs); c = e.shade_counters()
    e.set_option(rt.OPT_COUNTERS, 0)
    with torch.cuda.stream(st):
        for _ in range(3):
            e.trace(pcs); e.shade(pcs)
        e.sync()
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        K = 50
        e0.record(st)
        for _ in range(K):
            e.trace(pcs)
        e1.record(st)
        for _ in range(K):
            e.shade(pcs)
        e2.record(st)
        e.sync()
    print(f"trace {e0.elapsed_time(e1) / K * 1e3:.1f} us, shade {e1.elapsed_time(e2) / K * 1e3:.1f} us per {frames} frames; counters {c}")

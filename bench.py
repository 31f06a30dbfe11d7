#!/usr/bin/env python
"""bench.py — the reference's headline metric (Mrays/s of the cone-march compute pass) on B200.

    python bench.py --gpus 1 --steps K --warmup W            # this build (CUDA, through the C ABI)
    python bench.py --impl reference --steps K --warmup W     # the reference's own shader on host cores
    torchrun ... bench.py --gpus N ...                        # row-band sharded, NCCL gather to rank 0

A "step" is one pass of the hot path over one batch: workload c1 (BASELINE.json configs[1]) =
default scene, 1920x1080, 8 pyramid levels, 16 frames (the "16 spp" axis: SURVEY.md §8d) on a
seeded camera path.  ray = one invocation of compute.glsl:main at one pyramid level.
One JSON line on stdout (rank 0).  See DESIGN.md "Measurement" for every definition used here.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (W, H, frames, camera path, accumulate, spheres, levels, BASELINE.json config)
    "c1": dict(W=1920, H=1080, frames=16, path="bench", accumulate=False, scene="default", levels=0,
               name="configs[1]: default scene 1920x1080, 8 levels, 16 frames (16 spp axis) on the seeded camera path"),
    "c2": dict(W=3840, H=2160, frames=256, path="static", accumulate=True, scene="default", levels=0,
               name="configs[2]: default scene 3840x2160, 9 levels, 256 frames static pose, FP32 accumulation (extension)"),
    "c3": dict(W=1920, H=1080, frames=1, path="static", accumulate=False, scene="synthetic10k", levels=0,
               name="configs[3]: 10,000 synthetic spheres (SplitMix64 0xB200), 1920x1080, 8 levels, 1 frame"),
    "c4": dict(W=7680, H=4320, frames=1024, path="static", accumulate=True, scene="default", levels=0,
               name="configs[4]: default scene 7680x4320, 10 levels (extension), 1024 frames static pose, FP32 accumulation, row bands + one gather"),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=None)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c1", choices=sorted(WORKLOADS))
    ap.add_argument("--in-flight", type=int, default=None,
                    help="RTCE_OPT_FRAMES_IN_FLIGHT for frame-by-frame workloads (default 6): batches of consecutive steps "
                         "rotate over this many windows of frame slots and overlap on the device; 1 = strict stream order")
    ap.add_argument("--strip-ctas", type=int, default=None, help="RTCE_OPT_STRIP_CTAS_PER_SM")
    ap.add_argument("--large-kernel", type=int, default=None, help="RTCE_OPT_LARGE_KERNEL (objCount > 8): 0 auto, 1 shared-memory kernel, 2 register kernel")
    ap.add_argument("--graph", type=int, default=None, help="RTCE_OPT_GRAPH: 1 = replay the launch chain from a CUDA graph (default), 0 = issue every launch")
    ap.add_argument("--frames", type=int, default=None, help="override frames per step (labelled in config)")
    ap.add_argument("--levels", type=int, default=None, help="override pyramid levels (configs[3] sweep)")
    ap.add_argument("--kernel", type=int, default=0, help="0 auto (= 3), 1 per-pixel kernel per level, 2 fused pyramid kernel, 3 strip kernels + fused coarse levels")
    ap.add_argument("--strip-min-rays", type=int, default=None, help="kernel plan 3: strip-kernel threshold (rays per launch)")
    ap.add_argument("--groups", type=int, default=None, help="kernel plan 3: frame groups per batch (1..8)")
    ap.add_argument("--strip-rows", type=int, default=None, help="kernel plan 3: rows per strip task")
    ap.add_argument("--slots", type=int, default=None, help="RTCE_OPT_FRAME_SLOTS: frames per chunk / launch (default min(frames, 16), at most 32)")
    ap.add_argument("--band-weight", type=float, default=None,
                    help="N > 1, accumulating workloads: cost of a marched texel in background texels for the row-band cut "
                         "(1 = equal bands; default: picked at set-up from timed steps)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--profile", action="store_true", help="short run for ncu: no e2e, no CPU baseline, no clock hold")
    return ap.parse_args()


def workload_inputs(wl):
    from raytracing_engine_b200 import camera, scene
    if wl["scene"] == "default":
        spheres = scene.spheres_of(scene.default_mutable_data())
    else:
        spheres = scene.synthetic_spheres(10000, seed=0xB200)
    if wl["path"] == "bench":
        rots, poss = camera.bench_camera_path(wl["frames"])
    else:
        rots, poss = camera.static_path(wl["frames"])
    return spheres, rots, poss


def algorithmic_ops(cnt, n_spheres):
    """ops = 57 R + 10 E + (12 + 3 N) S  (1 op = 1 separately rounded FP32 operation)."""
    return 57 * cnt["R"] + 10 * cnt["E"] + (12 + 3 * n_spheres) * cnt["S"]


def algorithmic_bytes(rays_all_levels, rays_level0):
    """4 B written per pixel of every level + 4 B read per parent texel (shared by 4 children)."""
    return 4 * rays_all_levels + (rays_all_levels - rays_level0)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as fh:
            d = json.load(fh)
        return float(d.get("hbm_gbs", 6650.0)), float(d.get("sm_max_mhz", 1965.0)), "measured"
    return 6650.0, 1965.0, "fallback"


class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU while the timed region runs (pynvml)."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
               0x80: "hw_power_brake_slowdown"}

    def __init__(self, index, period=0.05):
        # NVML queries take a driver lock that CUDA launches also want: 20 Hz, and only on rank 0 (bench_ours)
        self.period = period
        self.samples, self.reasons, self.max_mhz, self.ok = [], set(), None, False
        self._stop = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception:
            self.ok = False
        self.t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(int(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
                r = int(self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                for bit, name in self.REASONS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(self.period)

    def start(self):
        if self.ok:
            self.t.start()

    def stop(self):
        self._stop.set()
        if self.ok and self.t.is_alive():
            self.t.join(1.0)
        med = int(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def cpu_reference_step(wl, spheres, rots, poss, levels, frames, rect=None):
    """One bounded sample of the workload on the host cores with the reference's own shader text
    (oracle/_ref) when it was built, else the C port.  Returns (seconds, rays, kind, cores, oracle)."""
    import oracle
    from raytracing_engine_b200 import scene
    need_ext = len(spheres) > 8 or levels > 9
    kind = ("refx" if need_ext else "ref")
    if not oracle.available(kind):
        kind = "port"
    o = oracle.load(kind)
    cores = os.cpu_count() or 1
    ratio = scene.view_ratio(wl["W"], wl["H"])
    t0 = time.perf_counter()
    rays = 0
    for k in range(frames):
        _, cnt = o.trace(spheres, [wl["W"], wl["H"]], ratio, rots[k], poss[k], 1000.0, levels, rect=rect, threads=cores)
        rays += cnt["R"]
    return time.perf_counter() - t0, rays, kind, cores, o


def run_reference(args, wl):
    """--impl reference: the reference's CPU-compiled shader over the same workload, host threads only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle
    from raytracing_engine_b200 import engine
    spheres, rots, poss = workload_inputs(wl)
    levels = wl["levels"] or oracle.load("port").level_count(float(wl["W"]))
    steps = args.steps if args.steps is not None else 3
    warmup = args.warmup if args.warmup is not None else 1
    # bounded sample per step: at most 4 frames of the batch (each frame costs the same work +-10 %)
    frames = min(wl["frames"], max(1, int(64e6 / (wl["W"] * wl["H"] * 1.34))))
    rect = None
    if wl["scene"] != "default":
        cx, cy = wl["W"] // 2, wl["H"] // 2
        rect = [cx - 128, cy - 72, cx + 128, cy + 72]   # centred 256x144 window for 10k spheres
    for _ in range(warmup):
        cpu_reference_step(wl, spheres, rots, poss, levels, frames, rect)
    tot_t, tot_r = 0.0, 0
    for _ in range(steps):
        t, r, kind, cores, _o = cpu_reference_step(wl, spheres, rots, poss, levels, frames, rect)
        tot_t += t
        tot_r += r
    value = tot_r / tot_t / 1e6
    sample = f"{frames} of {wl['frames']} frames per step" + (", centred 256x144 window of every level" if rect else ", full image")
    line = {
        "impl": "reference", "metric": "Mrays/s (pixel-invocations of compute.glsl:main over all pyramid levels)",
        "value": value, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": steps, "warmup": warmup,
        "ms_per_step": tot_t / steps * 1e3 * (wl["frames"] / frames), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["name"], "levels": levels, "frames_per_step": wl["frames"]},
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": cores,
                         "kind": "reference" if kind in ("ref", "refx") else "port", "sample": sample},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


class _DevView:
    """Zero-copy torch view of engine-owned device memory (CUDA array interface)."""

    def __init__(self, ptr, shape, strides_bytes):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f4", "data": (int(ptr), False),
                                         "version": 3, "strides": tuple(strides_bytes)}


def run_ours(args, wl):
    import torch
    import torch.distributed as dist
    import raytracing_engine_b200 as rt
    from raytracing_engine_b200 import scene, sharding

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the compute pass has no CPU fallback "
                         "(use --impl reference for the host-core baseline)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        # one slice of the host cores per rank: the launching threads of N ranks do not migrate over each other
        try:
            cores = sorted(os.sched_getaffinity(0))
            per = len(cores) // world
            if per >= 2:
                os.sched_setaffinity(0, set(cores[local_rank * per:(local_rank + 1) * per]))
        except (AttributeError, OSError):
            pass
    # stdout carries ONE JSON line: native libraries (NCCL prints its version there) get stderr meanwhile
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    n_gpus = world

    # N > 1.  Frames are independent units (main.rs:291-316 renders each from its own push constants), so a
    # frame-by-frame workload shards FRAMES: the job is N batches of the workload's shape, rank r traces batch r
    # (the same seeded poses on every rank, so that the work per GPU is exactly fixed) over the whole image
    # into its own HBM, and there is no data-path collective (weak scaling).  An accumulating workload has one
    # result image whose sum is order-defined, so it shards row bands of the final level instead (strong
    # scaling) and gathers the one accumulated image per step.
    weak = world > 1 and not wl["accumulate"]
    frames_rank = wl["frames"]
    spheres, rots, poss = workload_inputs(wl)            # this rank's batch
    if weak:
        wl = dict(wl)
        wl["frames"] = wl["frames"] * world
    frames = wl["frames"]                                # of the whole job
    if not weak:
        frames_rank = frames
    pcs = rt.make_push_constants(rots, poss)
    steps = args.steps if args.steps is not None else (1000 if frames_rank * wl["W"] * wl["H"] < 1e9 and wl["scene"] == "default" else (3 if world == 1 or weak else 10))
    warmup = args.warmup if args.warmup is not None else (20 if steps >= 50 else 3)
    warmup = max(warmup, 3)

    eng = rt.Engine(local_rank, scene.RENDER_DIST)
    stream = torch.cuda.Stream(device=local_rank)
    eng.set_stream(stream.cuda_stream)
    eng.set_option(rt.OPT_KERNEL, args.kernel)
    if args.strip_min_rays is not None:
        eng.set_option(rt.OPT_STRIP_MIN_RAYS, args.strip_min_rays)
    if args.groups is not None:
        eng.set_option(rt.OPT_GROUPS, args.groups)
    if args.strip_rows is not None:
        eng.set_option(rt.OPT_STRIP_ROWS, args.strip_rows)
    if args.graph is not None:
        eng.set_option(rt.OPT_GRAPH, args.graph)
    if args.strip_ctas is not None:
        eng.set_option(rt.OPT_STRIP_CTAS_PER_SM, args.strip_ctas)
    if args.large_kernel is not None:
        eng.set_option(rt.OPT_LARGE_KERNEL, args.large_kernel)
    # accumulating workloads: rtce_trace_accumulate rotates its chunks over max(in_flight, 4) lanes internally; 8 lanes x
    # 32-frame chunks keep the machine busier than 4 x 16 (c4, N = 1: 220.9 -> 210.1 ms per step; with row bands a launch
    # carries 1/N of the rays and it matters more: N = 8, 37.9 -> 33.4 ms)
    in_flight = (args.in_flight or 8) if wl["accumulate"] else (args.in_flight if args.in_flight is not None else 6)
    eng.set_option(rt.OPT_FRAMES_IN_FLIGHT, in_flight)
    slots = min(frames_rank, args.slots or (32 if wl["accumulate"] else 16), 32)
    eng.set_option(rt.OPT_FRAME_SLOTS, slots)
    if len(spheres) <= 8:
        eng.set_scene(scene.pack_mutable_data(objects=[(tuple(s[:3]), s[3]) for s in spheres]))
    else:
        eng.set_spheres(spheres)
    eng.set_view(wl["W"], wl["H"], levels=wl["levels"])
    levels = eng.levels
    Wf, Hf, _ = eng.level_info(-1, 0)
    def dev_final(nslots):
        _, _, p0 = eng.level_info(-1, 0)
        stride = (eng.level_info(-1, 1)[2] - p0) if nslots > 1 else Wf * Hf * 4
        return torch.as_tensor(_DevView(p0, (nslots, Hf, Wf), (stride, Wf * 4, 4)), device=f"cuda:{local_rank}")

    # Row bands (accumulating workloads at N > 1).  Equal bands are badly balanced: the spheres sit in the middle
    # rows and a background texel costs a fraction of a marched one (far-miss shortcut), so the bands are cut to
    # equal ESTIMATED cost from one pilot frame (every rank traces the same pose, hence computes the same cuts
    # without communication), and the one free parameter of the estimate (cost of a marched texel in background
    # texels) is picked at set-up from a few timed steps, max over ranks.  Not in the timed region.
    bands = sharding.all_bands(Hf, world)
    band_info = None
    y0, y1 = bands[rank]
    if world > 1 and not weak:
        with torch.cuda.stream(stream):
            eng.trace(pcs[:1])
            eng.sync()
            pilot = dev_final(1)[0].clone()
            cands = [args.band_weight] if args.band_weight is not None else [1.0, 2.0, 4.0]
            trials = []
            for hw in cands:
                cb = sharding.all_bands(Hf, world) if hw == 1.0 else \
                    sharding.weighted_bands(sharding.row_cost(pilot, scene.RENDER_DIST, hw), world)
                ms = None
                if len(cands) > 1:
                    eng.set_rect(0, cb[rank][0], Wf, cb[rank][1])
                    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    for k in range(2):   # the first pass captures the launch graphs of the new rect
                        dist.barrier()
                        torch.cuda.synchronize()
                        ev0.record(stream)
                        eng.trace_accumulate(pcs)
                        ev1.record(stream)
                        torch.cuda.synchronize()
                    t = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=f"cuda:{local_rank}")
                    dist.all_reduce(t, op=dist.ReduceOp.MAX)
                    ms = float(t.item())
                trials.append((hw, ms, cb))
            hw, ms, bands = min(trials, key=lambda t: (t[1] is None, t[1] or 0.0)) if len(trials) > 1 else trials[0]
            y0, y1 = bands[rank]
            eng.set_rect(0, y0, Wf, y1)
            band_info = {"rows_per_rank": [b1 - b0 for b0, b1 in bands], "marched_texel_weight": hw,
                         "set_up_trials_ms_per_step": {str(t[0]): t[1] for t in trials},
                         "note": "contiguous bands of equal estimated cost (row cost from one pilot frame: width + (weight - 1) x marched texels), "
                                 "weight picked at set-up by timing one step per candidate (max over ranks); weight 1 = equal bands"}

    gather_state = {}

    def finish_gathers():
        """Order the launching stream after every gather still in flight (before the end event of a timed region)."""
        for ev in gather_state.get("done", []):
            stream.wait_event(ev)

    def step():
        """One pass of the hot path over the batch, inputs resident; N>1: + the one gather to rank 0."""
        if wl["accumulate"]:
            eng.trace_accumulate(pcs)
            if world > 1:
                import ctypes as C
                pw, ph, pp = C.c_int(), C.c_int(), C.c_void_p()
                eng._check(eng.lib.rtce_accum_info(eng._ctx, C.byref(pw), C.byref(ph), C.byref(pp)), "rtce_accum_info")
                acc = torch.as_tensor(_DevView(pp.value, (1, Hf, Wf), (Wf * Hf * 4, Wf * 4, 4)), device=f"cuda:{local_rank}")
                # The gather of step k overlaps the tracing of step k + 1: the band is copied out of the accumulator
                # (which the next step overwrites) into one of two buffers and gathered from there on a side stream.
                G = gather_state
                if not G:
                    G["side"] = torch.cuda.Stream(device=local_rank)
                    G["band"] = [torch.empty((1, y1 - y0, Wf), dtype=torch.float32, device=f"cuda:{local_rank}") for _ in range(2)]
                    G["full"] = [torch.empty((1, Hf, Wf), dtype=torch.float32, device=f"cuda:{local_rank}") if rank == 0 else None for _ in range(2)]
                    G["ready"] = [torch.cuda.Event() for _ in range(2)]
                    G["done"] = [torch.cuda.Event() for _ in range(2)]
                    G["k"] = 0
                k = G["k"] % 2
                G["k"] += 1
                stream.wait_event(G["done"][k])          # the gather that last read buffer k has finished
                G["band"][k].copy_(acc[:, y0:y1, :])
                G["ready"][k].record(stream)
                with torch.cuda.stream(G["side"]):
                    G["side"].wait_event(G["ready"][k])
                    full = sharding.gather_bands(G["band"][k], Hf, dst=0, bands=bands, out=G["full"][k])
                    G["done"][k].record(G["side"])
                return full
        else:
            eng.trace(pcs)
        return None

    # ---- work counters of one step (separate counted run; also the parity spot-check input) ----
    with torch.cuda.stream(stream):
        eng.set_option(rt.OPT_COUNTERS, 1)
        if wl["accumulate"]:
            eng.trace_accumulate(pcs[: min(frames_rank, slots)])
        else:
            eng.trace(pcs[: min(frames_rank, slots)])
        cnt = eng.counters()
        cnt_last = eng.level_counters(-1)
        far_all, far_last = eng.far_miss_rays(-1), eng.far_miss_rays(eng.levels - 1)
        eng.set_option(rt.OPT_COUNTERS, 0)
    scale = frames_rank / min(frames_rank, slots)   # static-pose workloads repeat the same frames
    cnt_step = {k: int(cnt[k] * scale) for k in "RES"}
    cnt_last_step = {k: int(cnt_last[k] * scale) for k in "RES"}   # this rank's final-level launches
    far_step, far_last_step = int(far_all * scale) * world, int(far_last * scale)
    # a far-miss ray costs the reference algorithm one trip with objCount evaluations: 57 + 10 N + 12 + 3 N operations
    far_ops_each = 57 + 10 * min(len(spheres), 8) + 12 + 3 * min(len(spheres), 8)
    if world > 1:
        t = torch.tensor([cnt_step["R"], cnt_step["E"], cnt_step["S"]], dtype=torch.int64, device=f"cuda:{local_rank}")
        dist.all_reduce(t)
        cnt_step = dict(zip("RES", [int(v) for v in t.tolist()]))
    dims = rt.host_level_dims(levels, wl["W"], wl["H"])
    rays_full_frame = sum(w * h for w, h in dims)
    units_per_step = rays_full_frame * frames          # rays of the WHOLE job (all ranks)
    ops_step = algorithmic_ops(cnt_step, len(spheres))
    bytes_step = algorithmic_bytes(rays_full_frame, dims[0][0] * dims[0][1]) * frames

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.cuda.stream(stream):
        for _ in range(warmup):
            step()
        barrier()
        eng.set_option(rt.OPT_COUNTERS, 0)
        launches0 = 0
        sampler = ClockSampler(local_rank if "CUDA_VISIBLE_DEVICES" not in os.environ else
                               int(os.environ["CUDA_VISIBLE_DEVICES"].split(",")[local_rank]))
        if rank == 0:
            sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            step()
        eng.flush()   # frames in flight: the end event comes after every batch
        finish_gathers()   # ... and after every gather
        e1.record(stream)
        barrier()
        ms_total = e0.elapsed_time(e1)
        launches_per_step = eng.counters()["launches"]
        # keep the GPU under the same load long enough for the sampler to see clocks (not timed)
        t_end = time.perf_counter() + (0.0 if args.profile else max(0.0, 1.0 - ms_total * 1e-3))
        while time.perf_counter() < t_end:   # time-based, so no collective in here (ranks would diverge)
            if wl["accumulate"]:
                eng.trace_accumulate(pcs)
            else:
                eng.trace(pcs)
            torch.cuda.synchronize()
        clocks = sampler.stop()
    if world > 1:
        t = torch.tensor([ms_total], dtype=torch.float64, device=f"cuda:{local_rank}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ms_step = ms_total / steps
    value = units_per_step / (ms_step * 1e-3) / 1e6

    # ---- the dominant kernel (the launch that traces the final level: 75 % of all rays), timed alone ----
    # CUDA events around each of its launches (rtce_debug_launch_timeline), one frame group so that no
    # other launch shares the machine with it, same workload, same warm state.
    dom = None
    if not wl["accumulate"] and len(spheres) <= 8:
        with torch.cuda.stream(stream):
            eng.set_option(rt.OPT_GROUPS, 1)
            for _ in range(3):
                step()
            barrier()
            eng.launch_timeline(fetch=False)
            k_dom = max(3, min(steps, 20))
            for _ in range(k_dom):
                step()
            tl = eng.launch_timeline(max_launches=4096)
            eng.launch_timeline(max_launches=0)   # disarm
            eng.set_option(rt.OPT_GROUPS, args.groups if args.groups is not None else 0)
            barrier()
        sel = tl[(tl[:, 2] == 1) & (tl[:, 3] == levels - 1)]
        if len(sel):
            dur = sel[:, 1] - sel[:, 0]
            dom = {"kernel": "k_trace_strip (final pyramid level)", "launches_timed": int(len(sel)),
                   "launches_per_step": int(len(sel) // k_dom), "ms_per_launch": float(dur.mean()),
                   "ms_per_launch_min": float(dur.min()), "counters_per_launch": None}

    # ---- the presentation shading pass (fragment.glsl) over the same batch, timed on its own ----
    # Not part of `value`: the metric counts the compute pass.  Reported so that the widened row has a number.
    shade = None
    if not wl["accumulate"] and len(spheres) <= 8 and wl["scene"] == "default" and not args.profile:
        with torch.cuda.stream(stream):
            eng.set_option(rt.OPT_FRAMES_IN_FLIGHT, 1)
            eng.trace(pcs)
            eng.set_option(rt.OPT_COUNTERS, 1)
            eng.shade(pcs)
            sc = eng.shade_counters()
            eng.set_option(rt.OPT_COUNTERS, 0)
            for _ in range(3):
                eng.shade(pcs)
            barrier()
            k_sh = max(3, min(steps, 20))
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s0.record(stream)
            for _ in range(k_sh):
                eng.shade(pcs)
            s1.record(stream)
            barrier()
            ms_sh = s0.elapsed_time(s1) / k_sh
            eng.set_option(rt.OPT_FRAMES_IN_FLIGHT, in_flight)
        # algorithmic FP32 operations: ray direction 45 per pixel; 10 per distance() call (3 sub, 3 mul, 2 add,
        # sqrt, sub); per hit pixel 40 (position, nearest-object select, falloffs, normal); per light evaluation
        # 140 (normalize, reflect, pinned pow = 45, falloffs, colour sum); shadow-march bookkeeping rides on D
        ops_sh = 45 * sc["P"] + 10 * sc["D"] + 40 * sc["H"] + 140 * sc["L"]
        shade = {"kernel": "k_shade (fragment.glsl:123-187)", "ms_per_step": ms_sh, "frames": frames_rank,
                 "mpixels_per_s": sc["P"] / (ms_sh * 1e-3) / 1e6, "counters_per_step": sc,
                 "algorithmic_ops_per_step": ops_sh, "achieved_tlaneop_per_s": ops_sh / (ms_sh * 1e-3) / 1e12,
                 "note": "timed alone on this rank after the main measurement; not included in value"}

    # ---- end to end through the public API with HOST buffers (push constants in, images out) ----
    e2e = None
    if not args.no_e2e:
        band = Hf if weak else y1 - y0
        if wl["accumulate"]:
            host = np.empty((Hf, Wf), dtype=np.float32)
            def e2e_step():
                eng.trace_accumulate(pcs)
                return eng.readback_accum()
            d2h = Wf * band * 4
        else:
            pin = rt.PinnedBuffer((frames_rank, Hf, Wf))
            def e2e_step():
                eng.render_to_host(pcs, pin.array)
            d2h = frames_rank * Wf * band * 4
        k_e2e = max(3, min(steps, int(2.0 / max(ms_step * 1e-3 * 8, 1e-4))))
        with torch.cuda.stream(stream):
            for _ in range(3):
                e2e_step()
            barrier()
            t0 = time.perf_counter()
            for _ in range(k_e2e):
                e2e_step()
            barrier()
            dt = (time.perf_counter() - t0) / k_e2e
        if world > 1:
            t = torch.tensor([dt], dtype=torch.float64, device=f"cuda:{local_rank}")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
            tb = torch.tensor([d2h], dtype=torch.int64, device=f"cuda:{local_rank}")
            dist.all_reduce(tb)
            d2h = int(tb.item())
        e2e = {"value": units_per_step / dt / 1e6, "unit": "Mrays/s", "ms_per_step": dt * 1e3,
               "h2d_bytes_per_step": int(pcs.nbytes) * world, "d2h_bytes_per_step": int(d2h), "steps": k_e2e,
               "note": "host PushConstantData in (staged through a pinned ring into the device camera buffer), every frame's final level "
                       "copied to pinned host memory (accumulate workloads: the one accumulated image); N>1: each rank "
                       "returns its own frames (accumulate workloads: its row band) to its host buffer"}

    # ---- N > 1, row bands: the gathered image on rank 0 against the same job traced whole on rank 0 (not timed) ----
    gather_check = None
    if world > 1 and not weak:
        import ctypes as C
        with torch.cuda.stream(stream):
            full = step()
            finish_gathers()
            barrier()
            if rank == 0:
                got = full.clone()
                eng.set_rect(0, 0, Wf, Hf)
                eng.trace_accumulate(pcs)
                eng.sync()
                pw, ph, pp = C.c_int(), C.c_int(), C.c_void_p()
                eng._check(eng.lib.rtce_accum_info(eng._ctx, C.byref(pw), C.byref(ph), C.byref(pp)), "rtce_accum_info")
                whole = torch.as_tensor(_DevView(pp.value, (1, Hf, Wf), (Wf * Hf * 4, Wf * 4, 4)), device=f"cuda:{local_rank}")
                same = bool(torch.equal(got.view(torch.int32), whole.view(torch.int32)))
                torch.cuda.synchronize()
                gather_check = "bit-identical to the whole image traced on rank 0" if same else "MISMATCH"

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    hbm_peak, sm_max, peak_src = measured_peaks()
    fp32_peak, fp32_ms = rt.measure_fp32_issue(local_rank, 4096)
    if shade is not None:
        shade["frac_of_fp32_issue_peak"] = shade["achieved_tlaneop_per_s"] * 1e12 / fp32_peak
    nominal = 148 * 128 * sm_max * 1e6
    achieved = ops_step / (ms_step * 1e-3)
    step_achieved = ops_step / (ms_step * 1e-3)
    step_roof = {"achieved": step_achieved / 1e12, "frac": step_achieved / (fp32_peak * n_gpus),
                 "far_miss_rays_per_step": far_step,
                 "frac_excluding_far_miss": (ops_step - far_ops_each * far_step) / (ms_step * 1e-3) / (fp32_peak * n_gpus),
                 "frac_of_nominal": step_achieved / (nominal * n_gpus), "ops_per_step": ops_step,
                 "counters_per_step": cnt_step, "scope": "all launches of the step, gaps included",
                 "hbm": {"achieved": bytes_step / (ms_step * 1e-3) / 1e9, "peak": hbm_peak * n_gpus, "unit": "GB/s",
                         "frac": bytes_step / (ms_step * 1e-3) / 1e9 / (hbm_peak * n_gpus),
                         "peak_source": peak_src + " MEASURED_PEAKS.json", "bytes_per_step": bytes_step}}
    traffic = None
    tp = os.path.join(ROOT, "profiles", "dominant_kernel_ncu.json")
    if os.path.exists(tp):
        with open(tp) as fh:
            traffic = json.load(fh).get("dram_bytes_per_launch")
    peak_note = ("measured live: rtce_measure_fp32_issue (non-contracted FMUL/FADD, register operands); "
                 f"nominal 148 SM x 128 lanes x {sm_max:.0f} MHz = {nominal / 1e12:.1f}")
    if dom is not None:
        ops_launch = algorithmic_ops(cnt_last_step, len(spheres)) / dom["launches_per_step"]
        dom["counters_per_launch"] = {k: cnt_last_step[k] // dom["launches_per_step"] for k in "RES"}
        achieved = ops_launch / (dom["ms_per_launch"] * 1e-3)
        far_launch = far_last_step // dom["launches_per_step"]
        roofline = {"bound": "fp32_issue", "achieved": achieved / 1e12, "peak": fp32_peak / 1e12, "unit": "Tlane-op/s",
                    "frac": achieved / fp32_peak, "traffic": traffic, "peak_source": peak_note,
                    "far_miss_rays_per_launch": far_launch,
                    "frac_excluding_far_miss": (ops_launch - far_ops_each * far_launch) / (dom["ms_per_launch"] * 1e-3) / fp32_peak,
                    "far_miss_note": "achieved counts the operations the REFERENCE algorithm performs (device counters R / E / S, equal "
                                     "to the oracle's); rays whose parent texel lies beyond every sphere by more than RENDER_DIST are "
                                     "resolved exactly without evaluating them (RTCE_OPT_FAR_MISS); frac_excluding_far_miss removes "
                                     "their operations from the numerator",
                    "frac_of_nominal": achieved / nominal, "ops_per_launch": ops_launch,
                    "scope": "dominant kernel: algorithmic FP32 operations (SURVEY.md 8d: 57 R + 10 E + (12 + 3 N) S, from "
                             "the device's own work counters of the final level) per launch / its mean launch duration "
                             "(CUDA events on its stream, one frame group, this rank)",
                    "dominant_kernel": dom, "step": step_roof}
    else:
        roofline = {"bound": "fp32_issue", "achieved": step_achieved / 1e12, "peak": fp32_peak * n_gpus / 1e12,
                    "unit": "Tlane-op/s", "frac": step_achieved / (fp32_peak * n_gpus),
                    "traffic": None,   # the committed ncu capture is of the configs[1] strip launch, not of this workload
                    "peak_source": peak_note, "frac_of_nominal": step_achieved / (nominal * n_gpus),
                    "scope": "all launches of the step, gaps included (no single dominant launch was timed)",
                    "step": step_roof}

    cpu_baseline = None
    if not args.no_cpu_baseline and n_gpus == 1:
        rect = None
        f_cpu = min(frames, max(1, int(64e6 / (wl["W"] * wl["H"] * 1.34))))
        if len(spheres) > 8:
            cx, cy = wl["W"] // 2, wl["H"] // 2
            rect = [cx - 128, cy - 72, cx + 128, cy + 72]
        cpu_reference_step(wl, spheres, rots, poss, levels, 1, rect)   # warm the threads
        best, rays_cpu = 1e30, 0
        for _ in range(3):
            t, r, kind, cores, _o = cpu_reference_step(wl, spheres, rots, poss, levels, f_cpu, rect)
            if t < best:
                best, rays_cpu = t, r
        cpu_baseline = {"value": rays_cpu / best / 1e6, "unit": "Mrays/s", "cores": cores,
                        "kind": "reference" if kind in ("ref", "refx") else "port",
                        "sample": f"{f_cpu} of {frames} frames" + (", centred 256x144 window of every level" if rect else ", full image")
                                  + ", best of 3, OpenMP over rows"}

    line = {
        "metric": "Mrays/s (pixel-invocations of compute.glsl:main over all pyramid levels)",
        "value": value, "unit": "Mrays/s", "n_gpus": n_gpus, "steps": steps, "warmup": warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong" if (n_gpus > 1 and not weak) else "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["name"], "levels": levels, "frames_per_step": frames,
                   "rays_per_step": units_per_step, "primary_rays_per_step": Wf * Hf * frames,
                   "ms_per_frame": ms_step / frames,
                   "parallelism": "1 GPU" if n_gpus == 1 else
                                  (f"frames sharded: {n_gpus} batches of {frames_rank} frames, one per rank (the same seeded poses on every rank: "
                                   "fixed work per GPU), whole image per frame, results stay in each rank's HBM; no data-path "
                                   "collective (frames are independent units)" if weak else
                                   f"{n_gpus} row bands of the final level + one NCCL gather of the accumulated image to rank 0 per step"),
                   "bands": band_info, "gather_check": gather_check,
                   "frames_in_flight": in_flight if not wl["accumulate"] else f"{max(in_flight, 4)} chunks of {slots} frames (internal to rtce_trace_accumulate)",
                   "l2": f"no flush: each step rewrites {bytes_step / 1e6:.0f} MB of pyramid (> 126 MB L2 when frames >= 12 at 1080p); parents are meant to be L2 hits",
                   "kernel": {0: "auto (strip kernel + overflow launch per fine level, one per-pixel launch per coarse level)", 1: "per-level", 2: "fused", 3: "strip kernel + overflow launch per fine level, one per-pixel launch per coarse level"}[args.kernel]
                   if len(spheres) <= 8 else {None: "large-scene kernel, auto (register kernel: per-ray bounds in registers, one CTA per ray)", 0: "large-scene kernel, auto (register kernel)", 1: "large-scene kernel, shared-memory version", 2: "large-scene kernel, register version"}[args.large_kernel]},
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches_per_step * steps) * n_gpus,
        "roofline": roofline, "cpu_baseline": cpu_baseline,
        "mprimary_per_s": Wf * Hf * frames / (ms_step * 1e-3) / 1e6,
        "shade_pass": shade,
    }
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    wl = dict(WORKLOADS[args.workload])
    if args.frames:
        wl["frames"] = args.frames
        wl["name"] += f" [frames overridden to {args.frames}]"
    if args.levels:
        wl["levels"] = args.levels
        wl["name"] += f" [levels overridden to {args.levels}]"
    if args.profile:
        args.no_e2e = args.no_cpu_baseline = True
    if args.impl == "reference":
        run_reference(args, wl)
        return
    run_ours(args, wl)


if __name__ == "__main__":
    main()

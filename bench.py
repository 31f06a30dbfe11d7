#!/usr/bin/env python
"""bench.py — the reference's headline metric (Mrays/s of the cone-march compute pass) on B200.

    python bench.py --gpus 1 --steps K --warmup W            # this build (CUDA, through the C ABI)
    python bench.py --impl reference --steps K --warmup W     # the reference's own shader on host cores
    torchrun ... bench.py --gpus N ...                        # N ranks: configs[1] weak + configs[4] / configs[1] strong

Workload of the line (`value`): c1 = BASELINE.json configs[1] = default scene, 1920x1080, 8 pyramid levels, batches
of 16 frames (the "16 spp" axis: SURVEY.md §8d) on a seeded camera path.  ray = one invocation of compute.glsl:main
at one pyramid level.  A "step" is R back-to-back batches, R chosen at set-up so that the K timed steps last >= 0.6 s
whatever K is (a batch is 0.25 ms: K batches alone would be a 5 ms window); `config.batches_per_step` = R,
`frames_per_step` = 16 R.  The same invocation also measures the other BASELINE configs as sub-objects of the line
(N = 1: `c2`, `c3`, `c4`; N > 1: `strong_c4`, `strong_c1`: row bands + the NCCL gather to rank 0 behind the C ABI).
One JSON line on stdout (rank 0).  See DESIGN.md "Measurement" for every definition used here.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import math
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    "c1": dict(W=1920, H=1080, frames=16, path="bench", accumulate=False, scene="default", levels=0,
               name="configs[1]: default scene 1920x1080, 8 levels, 16 frames (16 spp axis) on the seeded camera path"),
    "c2": dict(W=3840, H=2160, frames=256, path="static", accumulate=True, scene="default", levels=0,
               name="configs[2]: default scene 3840x2160, 9 levels, 256 frames static pose, FP32 accumulation (extension)"),
    "c3": dict(W=1920, H=1080, frames=1, path="static", accumulate=False, scene="synthetic10k", levels=0,
               name="configs[3]: 10,000 synthetic spheres (SplitMix64 0xB200), 1920x1080, 8 levels, 1 frame"),
    "c4": dict(W=7680, H=4320, frames=1024, path="static", accumulate=True, scene="default", levels=0,
               name="configs[4]: default scene 7680x4320, 10 levels (extension), 1024 frames static pose, FP32 accumulation, row bands + one gather"),
    "c1x8": dict(W=1920, H=1080, frames=16, path="bench", accumulate=False, scene="eight", levels=0,
                 name="configs[1] with MAX_OBJECTS = 8 spheres (utilities.glsl:3): 1920x1080, 8 levels, 16 frames"),
}
TARGET_TIMED_S = 0.6


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=None)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c1", choices=sorted(WORKLOADS))
    ap.add_argument("--in-flight", type=int, default=None,
                    help="RTCE_OPT_FRAMES_IN_FLIGHT (default: the library's, 8)")
    ap.add_argument("--strip-ctas", type=int, default=None, help="RTCE_OPT_STRIP_CTAS_PER_SM")
    ap.add_argument("--large-kernel", type=int, default=None, help="RTCE_OPT_LARGE_KERNEL (objCount > 8): 0 auto, 2 register kernel, 3 event-driven kernel")
    ap.add_argument("--graph", type=int, default=None, help="RTCE_OPT_GRAPH")
    ap.add_argument("--frames", type=int, default=None, help="override frames per batch (labelled in config)")
    ap.add_argument("--levels", type=int, default=None, help="override pyramid levels (configs[3] sweep)")
    ap.add_argument("--kernel", type=int, default=0, help="0 auto (= 3), 1 per-pixel kernel per level, 3 strip kernels")
    ap.add_argument("--strip-min-rays", type=int, default=None)
    ap.add_argument("--groups", type=int, default=None)
    ap.add_argument("--strip-rows", type=int, default=None)
    ap.add_argument("--marcher", type=int, default=None, help="RTCE_OPT_MARCHER: 3 (default), 1, 2")
    ap.add_argument("--slots", type=int, default=None, help="RTCE_OPT_FRAME_SLOTS")
    ap.add_argument("--batches-per-step", type=int, default=None, help="R (default: sized for a >= 0.6 s timed region)")
    ap.add_argument("--band-weight", type=float, default=None)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="only the workload of the line (no c2 / c3 / c4 / strong sub-objects)")
    ap.add_argument("--profile", action="store_true", help="short run for ncu: one batch per step, no e2e, no CPU baseline, no extras")
    return ap.parse_args()


def workload_inputs(wl):
    from raytracing_engine_b200 import camera, scene
    if wl["scene"] == "default":
        spheres = scene.spheres_of(scene.default_mutable_data())
    elif wl["scene"] == "eight":
        sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
        from cases import cases
        spheres = {c[0]: c for c in cases()}["eight_spheres_1080p"][2]
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
        # NVML queries take a driver lock that CUDA launches also want: 20 Hz, and only on rank 0
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


# --------------------------------------------------------------------------------------------------------------------
# the reference's CPU implementation of the path (oracle/_ref: its shader text compiled as C++), host cores only
# --------------------------------------------------------------------------------------------------------------------
def cpu_reference_step(wl, spheres, rots, poss, levels, frames, rect=None):
    """One bounded sample of the workload on the host cores with the reference's own shader text
    (oracle/_ref) when it was built, else the C port.  Returns (seconds, rays, kind, cores)."""
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
    return time.perf_counter() - t0, rays, kind, cores


def cpu_window(wl, spheres):
    if len(spheres) > 8:   # centred 256x144 window of every level for the 10k-sphere scene
        cx, cy = wl["W"] // 2, wl["H"] // 2
        return [cx - 128, cy - 72, cx + 128, cy + 72]
    return None


def cpu_baseline_of(wl, spheres, rots, poss, levels, budget_frames=None):
    rect = cpu_window(wl, spheres)
    f_cpu = min(wl["frames"], budget_frames or max(1, int(64e6 / (wl["W"] * wl["H"] * 1.34))))
    cpu_reference_step(wl, spheres, rots, poss, levels, 1, rect)   # warm the threads
    best, rays_cpu, kind, cores = 1e30, 0, "port", 1
    for _ in range(3):
        t, r, kind, cores = cpu_reference_step(wl, spheres, rots, poss, levels, f_cpu, rect)
        if t < best:
            best, rays_cpu = t, r
    return {"value": rays_cpu / best / 1e6, "unit": "Mrays/s", "cores": cores,
            "kind": "reference" if kind in ("ref", "refx") else "port",
            "sample": f"{f_cpu} of {wl['frames']} frames of one batch" + (", centred 256x144 window of every level" if rect else ", full image")
                      + ", best of 3, OpenMP over rows"}


def run_reference(args, wl):
    """--impl reference: the reference's CPU-compiled shader over the same workload, host threads only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle
    spheres, rots, poss = workload_inputs(wl)
    levels = wl["levels"] or oracle.load("port").level_count(float(wl["W"]))
    steps = args.steps if args.steps is not None else 3
    warmup = args.warmup if args.warmup is not None else 1
    # bounded sample per step: at most ~4 frames of ONE batch of the step (every batch traces the same poses, each
    # frame costs the same work +-10 %)
    frames = min(wl["frames"], max(1, int(64e6 / (wl["W"] * wl["H"] * 1.34))))
    rect = cpu_window(wl, spheres)
    for _ in range(warmup):
        cpu_reference_step(wl, spheres, rots, poss, levels, frames, rect)
    tot_t, tot_r = 0.0, 0
    kind, cores = "port", 1
    for _ in range(steps):
        t, r, kind, cores = cpu_reference_step(wl, spheres, rots, poss, levels, frames, rect)
        tot_t += t
        tot_r += r
    value = tot_r / tot_t / 1e6
    sample = f"{frames} of {wl['frames']} frames of one batch per step" + (", centred 256x144 window of every level" if rect else ", full image")
    line = {
        "impl": "reference", "metric": "Mrays/s (pixel-invocations of compute.glsl:main over all pyramid levels)",
        "value": value, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": steps, "warmup": warmup,
        "ms_per_step": tot_t / steps * 1e3 * (wl["frames"] / frames), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["name"], "levels": levels, "frames_per_batch": wl["frames"],
                   "note": "ms_per_step is the sample scaled to ONE batch of the workload; throughput (value) does not depend on the step size"},
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": cores,
                         "kind": "reference" if kind in ("ref", "refx") else "port", "sample": sample},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------------------------
# this build
# --------------------------------------------------------------------------------------------------------------------
class _DevView:
    """Zero-copy torch view of engine-owned device memory (CUDA array interface)."""

    def __init__(self, ptr, shape, strides_bytes):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f4", "data": (int(ptr), False),
                                         "version": 3, "strides": tuple(strides_bytes)}


class Ctx:
    """What every measurement below shares: ranks, the launching stream, torch."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.args = torch, dist, args
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device - the compute pass has no CPU fallback "
                             "(use --impl reference for the host-core baseline)")
        torch.cuda.set_device(self.local_rank)
        self.dev = f"cuda:{self.local_rank}"
        if self.world > 1:
            # one slice of the host cores per rank: the launching threads of N ranks do not migrate over each other
            try:
                cores = sorted(os.sched_getaffinity(0))
                per = len(cores) // self.world
                if per >= 2:
                    os.sched_setaffinity(0, set(cores[self.local_rank * per:(self.local_rank + 1) * per]))
            except (AttributeError, OSError):
                pass
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local_rank))
        self.stream = torch.cuda.Stream(device=self.local_rank)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, x: float) -> float:
        if self.world == 1:
            return float(x)
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, xs):
        if self.world == 1:
            return [int(x) for x in xs]
        t = self.torch.tensor(list(xs), dtype=self.torch.int64, device=self.dev)
        self.dist.all_reduce(t)
        return [int(v) for v in t.tolist()]

    def gather_floats(self, x: float):
        if self.world == 1:
            return [float(x)]
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        out = [self.torch.zeros_like(t) for _ in range(self.world)]
        self.dist.all_gather(out, t)
        return [float(o.item()) for o in out]

    def timed(self, fn, steps, finish=None):
        """K calls of fn bracketed by barrier + synchronize and CUDA events on the launching stream -> ms total (this rank)."""
        torch = self.torch
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(self.stream)
        for _ in range(steps):
            fn()
        if finish:
            finish()
        e1.record(self.stream)
        self.barrier()
        return e0.elapsed_time(e1)


def make_engine(cx: Ctx, wl, spheres, in_flight=None, slots=None, levels=None):
    import raytracing_engine_b200 as rt
    from raytracing_engine_b200 import scene
    a = cx.args
    eng = rt.Engine(cx.local_rank, scene.RENDER_DIST)
    eng.set_stream(cx.stream.cuda_stream)
    eng.set_option(rt.OPT_KERNEL, a.kernel)
    for opt, val in ((rt.OPT_STRIP_MIN_RAYS, a.strip_min_rays), (rt.OPT_GROUPS, a.groups), (rt.OPT_STRIP_ROWS, a.strip_rows),
                     (rt.OPT_GRAPH, a.graph), (rt.OPT_STRIP_CTAS_PER_SM, a.strip_ctas), (rt.OPT_LARGE_KERNEL, a.large_kernel),
                     (rt.OPT_MARCHER, a.marcher)):
        if val is not None:
            eng.set_option(opt, val)
    if in_flight is not None:
        eng.set_option(rt.OPT_FRAMES_IN_FLIGHT, in_flight)
    if slots is not None:
        eng.set_option(rt.OPT_FRAME_SLOTS, slots)
    if len(spheres) <= 8:
        eng.set_scene(scene.pack_mutable_data(objects=[(tuple(s[:3]), s[3]) for s in spheres]))
    else:
        eng.set_spheres(spheres)
    eng.set_view(wl["W"], wl["H"], levels=levels if levels is not None else wl["levels"])
    return eng


def counted_pass(eng, wl, pcs, frames_rank, slots):
    """Work counters of one batch (separate counted run; R / E / S equal the oracle's, tests/test_parity_gpu.py)."""
    import raytracing_engine_b200 as rt
    n = min(frames_rank, slots)
    eng.set_option(rt.OPT_COUNTERS, 1)
    (eng.trace_accumulate if wl["accumulate"] else eng.trace)(pcs[:n])
    cnt = eng.counters()
    cnt_last = eng.level_counters(-1)
    far_all, far_last = eng.far_miss_rays(-1), eng.far_miss_rays(eng.levels - 1)
    eng.set_option(rt.OPT_COUNTERS, 0)
    scale = frames_rank / n   # static-pose workloads repeat the same frames
    return ({k: int(cnt[k] * scale) for k in "RES"}, {k: int(cnt_last[k] * scale) for k in "RES"},
            int(far_all * scale), int(far_last * scale))


def roofline_step(ops, far_ops, ms, fp32_peak, n_gpus, nominal, bytes_step, hbm_peak, peak_src, cnt):
    t = ms * 1e-3
    executed = ops - far_ops
    return {"achieved": executed / t / 1e12, "frac": executed / t / (fp32_peak * n_gpus),
            "frac_reference_algorithm": ops / t / (fp32_peak * n_gpus),
            "frac_of_nominal": executed / t / (nominal * n_gpus), "ops_executed": executed, "ops_reference_algorithm": ops,
            "counters": cnt, "scope": "all launches, gaps included",
            "two_term_bound_ms": max(executed / (fp32_peak * n_gpus), bytes_step / (hbm_peak * 1e9 * n_gpus)) * 1e3,
            "hbm": {"achieved": bytes_step / t / 1e9, "peak": hbm_peak * n_gpus, "unit": "GB/s",
                    "frac": bytes_step / t / 1e9 / (hbm_peak * n_gpus), "peak_source": peak_src + " MEASURED_PEAKS.json",
                    "bytes": bytes_step}}


def measure_accumulating(cx: Ctx, key, steps=3, warmup=2, with_cpu=True, with_e2e=True):
    """N = 1 sub-object for an accumulating config (c2 / c4): whole job on this GPU."""
    import raytracing_engine_b200 as rt
    wl = dict(WORKLOADS[key])
    spheres, rots, poss = workload_inputs(wl)
    pcs = rt.make_push_constants(rots, poss)
    slots = 32
    with cx.torch.cuda.stream(cx.stream):
        eng = make_engine(cx, wl, spheres, in_flight=cx.args.in_flight or 8, slots=slots)
        levels = eng.levels
        Wf, Hf, _ = eng.level_info(-1, 0)
        cnt, _, far, _ = counted_pass(eng, wl, pcs, wl["frames"], slots)
        eng.prepare(wl["frames"], accumulate=True)
        for _ in range(warmup):
            eng.trace_accumulate(pcs)
        ms = cx.timed(lambda: eng.trace_accumulate(pcs), steps) / steps
        e2e = None
        if with_e2e:
            eng.readback_accum()
            cx.barrier()
            t0 = time.perf_counter()
            for _ in range(steps):
                eng.trace_accumulate(pcs)
                eng.readback_accum()
            dt = (time.perf_counter() - t0) / steps
            e2e = {"ms_per_step": dt * 1e3, "d2h_bytes_per_step": Wf * Hf * 4, "h2d_bytes_per_step": int(pcs.nbytes)}
        eng.close()
    dims = rt.host_level_dims(levels, wl["W"], wl["H"])
    rays = sum(w * h for w, h in dims) * wl["frames"]
    out = {"workload": wl["name"], "levels": levels, "frames_per_step": wl["frames"], "rays_per_step": rays, "ms_per_step": ms,
           "value": rays / (ms * 1e-3) / 1e6, "unit": "Mrays/s", "steps": steps, "warmup": warmup,
           "lanes": f"{max(cx.args.in_flight or 8, 4)} chunks of {slots} frames in flight (internal to rtce_trace_accumulate)",
           "_cnt": cnt, "_far": far, "_dims": dims, "_spheres": len(spheres)}
    if e2e:
        e2e["value"] = rays / (e2e["ms_per_step"] * 1e-3) / 1e6
        out["e2e"] = e2e
    if with_cpu:
        out["cpu_baseline"] = cpu_baseline_of(wl, spheres, rots, poss, levels, budget_frames=1)
    return out


def measure_c3(cx: Ctx, with_cpu=True):
    """configs[3] sub-object: 10 000 spheres, 1080p, one frame per step; levels 8 (the 1 / 4 sweep: --levels)."""
    import raytracing_engine_b200 as rt
    wl = dict(WORKLOADS["c3"])
    spheres, rots, poss = workload_inputs(wl)
    pcs = rt.make_push_constants(rots, poss)
    out = {"workload": wl["name"], "kernels": {}}
    with cx.torch.cuda.stream(cx.stream):
        for name, mode in (("register", rt.LARGE_REGISTER), ("event_driven", rt.LARGE_EVENT)):
            eng = make_engine(cx, wl, spheres, in_flight=1, slots=1)
            eng.set_option(rt.OPT_LARGE_KERNEL, mode)
            levels = eng.levels
            if "counters" not in out:
                eng.set_option(rt.OPT_COUNTERS, 1)
                eng.trace(pcs)
                c = eng.counters()
                out["counters"] = {k: c[k] for k in "RES"}
                eng.set_option(rt.OPT_COUNTERS, 0)
            eng.trace(pcs)
            steps = 2
            ms = cx.timed(lambda: eng.trace(pcs), steps) / steps
            img = eng.readback(-1, 0)
            out["kernels"][name] = {"ms_per_frame": ms, "sha256_final_level": hashlib.sha256(img.tobytes()).hexdigest()[:16]}
            eng.close()
    dims = rt.host_level_dims(levels, wl["W"], wl["H"])
    rays = sum(w * h for w, h in dims)
    best = min(out["kernels"], key=lambda k: out["kernels"][k]["ms_per_frame"])
    ms = out["kernels"]["register"]["ms_per_frame"]
    out.update({"levels": levels, "rays_per_step": rays, "ms_per_step": ms, "value": rays / (ms * 1e-3) / 1e6, "unit": "Mrays/s",
                "kernel": "register kernel (library default; RTCE_OPT_LARGE_KERNEL = 3 selects the event-driven one)", "fastest": best,
                "same_image": len({k["sha256_final_level"] for k in out["kernels"].values()}) == 1,
                "_cnt": out["counters"], "_far": 0, "_dims": dims, "_spheres": len(spheres)})
    if with_cpu:
        out["cpu_baseline"] = cpu_baseline_of(wl, spheres, rots, poss, levels)
    return out


def finish_sub(sub, fp32_peak, nominal, hbm_peak, peak_src, n_gpus=1):
    """Attach the roofline of a sub-object (needs the FP32 peak measured after the timed regions)."""
    if sub is None:
        return None
    cnt, far, dims, n_sph = sub.pop("_cnt"), sub.pop("_far"), sub.pop("_dims"), sub.pop("_spheres")
    ops = algorithmic_ops(cnt, n_sph)
    far_each = 57 + 10 * min(n_sph, 8) + 12 + 3 * min(n_sph, 8)
    frames = sub.get("frames_per_step", 1)
    bytes_step = algorithmic_bytes(sum(w * h for w, h in dims), dims[0][0] * dims[0][1]) * frames
    sub["roofline"] = roofline_step(ops, far_each * far, sub["ms_per_step"], fp32_peak, n_gpus, nominal, bytes_step, hbm_peak, peak_src, cnt)
    sub["roofline"]["bound"] = "fp32_issue"
    return sub


def strong_c4(cx: Ctx):
    """N > 1: configs[4] (8K, 10 levels, 1024 static frames accumulated), final-level ROW BANDS per rank + ONE gather of the
    accumulated image to rank 0 per step — rtce_set_rect + rtce_trace_accumulate + rtce_gather_accum, all behind the C ABI;
    torch.distributed only distributes the 128-byte NCCL id and reduces the timings."""
    import raytracing_engine_b200 as rt
    from raytracing_engine_b200 import scene, sharding
    torch, dist = cx.torch, cx.dist
    wl = dict(WORKLOADS["c4"])
    spheres, rots, poss = workload_inputs(wl)
    pcs = rt.make_push_constants(rots, poss)
    slots, lanes = 32, 8
    steps, warm = 10, 2
    with torch.cuda.stream(cx.stream):
        eng = make_engine(cx, wl, spheres, in_flight=lanes, slots=slots)
        Wf, Hf, _ = eng.level_info(-1, 0)
        ident = [rt.comm_unique_id() if cx.rank == 0 else None]
        dist.broadcast_object_list(ident, src=0)
        eng.comm_init(ident[0], cx.rank, cx.world)
        # ---- the N = 1 time of the same job, measured here by rank 0 alone (the others wait) ----
        ms_n1, whole = None, None
        if cx.rank == 0:
            eng.prepare(wl["frames"], accumulate=True)
            for _ in range(2):
                eng.trace_accumulate(pcs)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(cx.stream)
            for _ in range(3):
                eng.trace_accumulate(pcs)
            e1.record(cx.stream)
            torch.cuda.synchronize()
            ms_n1 = e0.elapsed_time(e1) / 3
            whole = torch.from_numpy(eng.readback_accum()).to(cx.dev)
        cx.barrier()
        # ---- bands: equal estimated cost from one pilot frame, weight picked from timed trials (not timed) ----
        eng.trace(pcs[:1])
        eng.sync()
        _, _, p0 = eng.level_info(-1, 0)
        pilot = torch.as_tensor(_DevView(p0, (Hf, Wf), (Wf * 4, 4)), device=cx.dev).clone()
        cands = [cx.args.band_weight] if cx.args.band_weight is not None else [1.0, 2.0, 4.0]
        trials = []
        for hw in cands:
            cb = sharding.all_bands(Hf, cx.world) if hw == 1.0 else \
                sharding.weighted_bands(sharding.row_cost(pilot, scene.RENDER_DIST, hw), cx.world)
            ms = None
            if len(cands) > 1:
                eng.set_rect(0, cb[cx.rank][0], Wf, cb[cx.rank][1])
                eng.prepare(wl["frames"], accumulate=True)
                eng.trace_accumulate(pcs)
                ms = cx.max_over_ranks(cx.timed(lambda: eng.trace_accumulate(pcs), 1))
            trials.append((hw, ms, cb))
        # measured-cost balancing on top of the best estimate: time every rank on its band, cut bands of equal predicted
        # time (sharding.rebalanced_bands), keep the result only if the slowest rank got faster
        if len(cands) > 1:
            cb = min(trials, key=lambda t: t[1])[2]
            for it in range(3):
                eng.set_rect(0, cb[cx.rank][0], Wf, cb[cx.rank][1])
                eng.prepare(wl["frames"], accumulate=True)
                eng.trace_accumulate(pcs)
                t_rank = cx.gather_floats(cx.timed(lambda: eng.trace_accumulate(pcs), 2) / 2)
                if it > 0:
                    trials.append((f"measured-{it}", max(t_rank), cb))
                cb = sharding.rebalanced_bands(cb, t_rank, Hf)
            eng.set_rect(0, cb[cx.rank][0], Wf, cb[cx.rank][1])
            eng.prepare(wl["frames"], accumulate=True)
            eng.trace_accumulate(pcs)
            trials.append(("measured-3", cx.max_over_ranks(cx.timed(lambda: eng.trace_accumulate(pcs), 2) / 2), cb))
        hw, trial_ms, bands = min(trials, key=lambda t: (t[1] is None, t[1] or 0.0)) if len(trials) > 1 else trials[0]
        y0, y1 = bands[cx.rank]
        eng.set_rect(0, y0, Wf, y1)
        eng.prepare(wl["frames"], accumulate=True)

        def step():
            eng.trace_accumulate(pcs)
            eng.gather_accum(0, bands)      # overlaps the next step's tracing

        for _ in range(warm):
            step()
        eng.gather_wait()
        ms_rank = cx.timed(step, steps, finish=eng.gather_wait) / steps
        per_rank = cx.gather_floats(ms_rank)
        ms_step = max(per_rank)
        gather_ms = eng.gather_last_ms()
        # tracing alone (no exchange), same bands: what the exchange and its interference cost
        ms_trace_only = cx.max_over_ranks(cx.timed(lambda: eng.trace_accumulate(pcs), 3) / 3)
        # ---- the gathered image against the whole job traced on rank 0 ----
        step()
        check = None
        if cx.rank == 0:
            got = torch.from_numpy(eng.readback_gathered()).to(cx.dev)
            check = "bit-identical to the whole image traced on rank 0" if bool(torch.equal(got.view(torch.int32), whole.view(torch.int32))) else "MISMATCH"
        else:
            eng.gather_wait()
        cx.barrier()
        eng.close()
    if cx.rank != 0:
        return None
    dims = rt.host_level_dims(10, wl["W"], wl["H"])
    rays = sum(w * h for w, h in dims) * wl["frames"]
    return {"workload": wl["name"], "scaling": "strong", "n_gpus": cx.world, "ms_per_step": ms_step, "ms_per_step_n1_same_process": ms_n1,
            "speedup": ms_n1 / ms_step, "target_speedup_at_8": 7.0, "value": rays / (ms_step * 1e-3) / 1e6, "unit": "Mrays/s",
            "steps": steps, "warmup": warm, "per_rank_ms_per_step": per_rank, "ms_per_step_tracing_only": ms_trace_only,
            "speedup_tracing_only": ms_n1 / ms_trace_only, "gather_ms": gather_ms,
            "gather_bytes_into_rank0": int((Hf - (bands[0][1] - bands[0][0])) * Wf * 4), "gather_check": check,
            "gather": "rtce_gather_accum: grouped ncclSend / ncclRecv inside librtce.so, bands received in place, on a communication "
                      "stream out of a staging copy: the gather of step k overlaps the tracing of step k + 1; the timed region ends after the last gather",
            "bands": {"rows_per_rank": [b1 - b0 for b0, b1 in bands], "cut_by": hw if isinstance(hw, str) else f"estimate, marched-texel weight {hw}",
                      "set_up_trials_ms_per_step": {str(t[0]): t[1] for t in trials}},
            "lanes": f"{lanes} chunks of {slots} frames in flight"}


def strong_c1(cx: Ctx, batches=200):
    """N > 1: configs[1] STRONG: every frame's final level is cut in row bands (rtce_set_rect), every rank traces its band of
    every frame of every batch, and the bands of every frame are gathered to rank 0 (rtce_gather_frames: in place into rank
    0's own frame slots).  The per-frame work is small and its coarse levels are latency-bound, so this scales far worse
    than frame sharding (the `value` of the line); it is reported so that the two are not confused."""
    import raytracing_engine_b200 as rt
    torch = cx.torch
    if not hasattr(rt.Engine, "gather_frames"):
        return None
    wl = dict(WORKLOADS["c1"])
    spheres, rots, poss = workload_inputs(wl)
    pcs = rt.make_push_constants(rots, poss)
    with torch.cuda.stream(cx.stream):
        eng = make_engine(cx, wl, spheres, slots=16)
        Wf, Hf, _ = eng.level_info(-1, 0)
        ident = [rt.comm_unique_id() if cx.rank == 0 else None]
        cx.dist.broadcast_object_list(ident, src=0)
        eng.comm_init(ident[0], cx.rank, cx.world)
        ms_n1 = None
        if cx.rank == 0:
            eng.prepare(16)
            for _ in range(20):
                eng.trace(pcs)
            eng.flush()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(cx.stream)
            for _ in range(batches):
                eng.trace(pcs)
            eng.flush()
            e1.record(cx.stream)
            torch.cuda.synchronize()
            ms_n1 = e0.elapsed_time(e1) / batches
            eng.sync()
            want = eng.readback(-1, 15)
        cx.barrier()
        bands = [rt.host_band_rows(Hf, cx.world, r) for r in range(cx.world)]
        y0, y1 = bands[cx.rank]
        eng.set_rect(0, y0, Wf, y1)
        eng.prepare(16)

        def step():
            eng.trace(pcs)
            eng.gather_frames(0, bands, 16)

        for _ in range(20):
            step()
        eng.gather_wait()
        ms = cx.max_over_ranks(cx.timed(step, batches, finish=lambda: (eng.flush(), eng.gather_wait())) / batches)
        check = None
        if cx.rank == 0:
            got = eng.readback(-1, 15)
            check = "bit-identical to the whole frame traced on rank 0" if np.array_equal(got.view(np.uint32), want.view(np.uint32)) else "MISMATCH"
        cx.barrier()
        eng.close()
    if cx.rank != 0:
        return None
    return {"workload": wl["name"], "scaling": "strong", "n_gpus": cx.world, "ms_per_batch": ms, "ms_per_batch_n1_same_process": ms_n1,
            "speedup": ms_n1 / ms, "gather_check": check, "batches_timed": batches,
            "parallelism": f"{cx.world} row bands of every frame + NCCL gather of every frame's bands to rank 0 (rtce_gather_frames)"}


def run_ours(args, wl):
    import raytracing_engine_b200 as rt
    from raytracing_engine_b200 import scene

    # stdout carries ONE JSON line: native libraries (NCCL prints its version there when the process group comes up)
    # get stderr from before torch.distributed starts until the line is ready
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    cx = Ctx(args)
    torch, dist = cx.torch, cx.dist
    world, rank, local_rank = cx.world, cx.rank, cx.local_rank
    n_gpus = world

    # N > 1, the line's `value`.  Frames are independent units (main.rs:291-316 renders each from its own push constants),
    # so a frame-by-frame workload shards FRAMES: the job is N batches of the workload's shape, rank r traces batch r (the
    # same seeded poses on every rank, so that the work per GPU is exactly fixed) over the whole image into its own HBM,
    # and there is no data-path collective (weak scaling).  The row-band + gather forms are the `strong_*` sub-objects.
    accumulate = wl["accumulate"]
    if world > 1 and accumulate:
        raise SystemExit("bench.py: at N > 1 the accumulating configs are measured as the strong_c4 sub-object of the default line")
    spheres, rots, poss = workload_inputs(wl)
    frames_rank = wl["frames"]
    frames = frames_rank * world                         # of the whole job, per batch
    pcs = rt.make_push_constants(rots, poss)
    steps = args.steps if args.steps is not None else 20
    warmup = max(args.warmup if args.warmup is not None else 5, 3)
    in_flight = args.in_flight if args.in_flight is not None else 8   # = the library default (kDefaultInFlight)
    slots = min(frames_rank, args.slots or (32 if accumulate else 16), 32)

    with torch.cuda.stream(cx.stream):
        eng = make_engine(cx, wl, spheres, in_flight=in_flight, slots=slots)
        levels = eng.levels
        Wf, Hf, _ = eng.level_info(-1, 0)
        trace = eng.trace_accumulate if accumulate else eng.trace
        cnt_b, cnt_last_b, far_b, far_last_b = counted_pass(eng, wl, pcs, frames_rank, slots)
        eng.prepare(frames_rank, accumulate=accumulate)   # every lane's windows and graphs exist before anything is timed
        # ---- size the step: R batches so that K steps last >= TARGET_TIMED_S ----
        for _ in range(in_flight + 2):
            trace(pcs)
        eng.flush()
        t_batch = cx.max_over_ranks(cx.timed(lambda: trace(pcs), 12, finish=eng.flush) / 12) * 1e-3
        R = args.batches_per_step or (1 if args.profile else max(1, math.ceil(TARGET_TIMED_S / (steps * t_batch))))

        def step():
            for _ in range(R):
                trace(pcs)

        for _ in range(warmup):
            step()
        eng.flush()
        sampler = ClockSampler(local_rank if "CUDA_VISIBLE_DEVICES" not in os.environ else
                               int(os.environ["CUDA_VISIBLE_DEVICES"].split(",")[local_rank]))
        cx.barrier()
        if rank == 0:
            sampler.start()
        ms_total = cx.max_over_ranks(cx.timed(step, steps, finish=eng.flush))
        clocks = sampler.stop()
        launches_per_batch = eng.counters()["launches"]
        # ---- what was timed is what the reference computes: frame 15 of the last batch against the golden hash ----
        verify = None
        if wl is not None and args.workload == "c1" and not args.frames and not args.levels and args.marcher in (None, 3):
            with open(os.path.join(ROOT, "tests", "golden", "golden.json")) as fh:
                gold = json.load(fh)["default_1080p_pose15"]["sha256"][-1]
            got = hashlib.sha256(eng.readback(-1, 15).tobytes()).hexdigest()
            verify = {"frame": 15, "level": "final", "sha256_equals_golden(default_1080p_pose15)": got == gold}
    ms_step = ms_total / steps
    dims = rt.host_level_dims(levels, wl["W"], wl["H"])
    rays_full_frame = sum(w * h for w, h in dims)
    units_per_step = rays_full_frame * frames * R          # rays of the WHOLE job (all ranks)
    value = units_per_step / (ms_step * 1e-3) / 1e6
    cnt_step = dict(zip("RES", cx.sum_over_ranks([cnt_b[k] * R for k in "RES"])))
    far_step = far_b * R * world
    ops_step = algorithmic_ops(cnt_step, len(spheres))
    far_ops_each = 57 + 10 * min(len(spheres), 8) + 12 + 3 * min(len(spheres), 8)
    bytes_step = algorithmic_bytes(rays_full_frame, dims[0][0] * dims[0][1]) * frames * R

    # ---- the dominant kernel (the launch that traces the final level: 75 % of all rays), timed alone ----
    dom = None
    if not accumulate and len(spheres) <= 8 and args.kernel in (0, 3) and args.marcher in (None, 3):
        with torch.cuda.stream(cx.stream):
            eng.set_option(rt.OPT_GROUPS, 1)
            for _ in range(3):
                trace(pcs)
            cx.barrier()
            eng.launch_timeline(fetch=False)
            k_dom = 20
            for _ in range(k_dom):
                trace(pcs)
            tl = eng.launch_timeline(max_launches=4096)
            eng.launch_timeline(max_launches=0)   # disarm
            eng.set_option(rt.OPT_GROUPS, args.groups if args.groups is not None else 0)
            cx.barrier()
        sel = tl[(tl[:, 2] == 1) & (tl[:, 3] == levels - 1)]
        if len(sel):
            dur = sel[:, 1] - sel[:, 0]
            dom = {"kernel": "k_trace_strip (final pyramid level)", "launches_timed": int(len(sel)),
                   "launches_per_batch": int(len(sel) // k_dom), "ms_per_launch": float(dur.mean()),
                   "ms_per_launch_min": float(dur.min())}

    # ---- the presentation shading pass (fragment.glsl) over the same batch, timed on its own ----
    shade = None
    if not accumulate and len(spheres) <= 8 and wl["scene"] == "default" and not args.profile:
        with torch.cuda.stream(cx.stream):
            eng.set_option(rt.OPT_FRAMES_IN_FLIGHT, 1)
            eng.trace(pcs)
            eng.set_option(rt.OPT_COUNTERS, 1)
            eng.shade(pcs)
            sc = eng.shade_counters()
            eng.set_option(rt.OPT_COUNTERS, 0)
            for _ in range(3):
                eng.shade(pcs)
            k_sh = 20
            ms_sh = cx.timed(lambda: eng.shade(pcs), k_sh) / k_sh
            eng.set_option(rt.OPT_FRAMES_IN_FLIGHT, in_flight)
        # algorithmic FP32 operations: ray direction 45 per pixel; 10 per distance() call (3 sub, 3 mul, 2 add,
        # sqrt, sub); per hit pixel 40 (position, nearest-object select, falloffs, normal); per light evaluation
        # 140 (normalize, reflect, pinned pow = 45, falloffs, colour sum); shadow-march bookkeeping rides on D
        ops_sh = 45 * sc["P"] + 10 * sc["D"] + 40 * sc["H"] + 140 * sc["L"]
        shade = {"kernel": "k_shade (fragment.glsl:123-187)", "ms_per_batch": ms_sh, "frames": frames_rank,
                 "mpixels_per_s": sc["P"] / (ms_sh * 1e-3) / 1e6, "counters_per_batch": sc,
                 "algorithmic_ops_per_batch": ops_sh, "achieved_tlaneop_per_s": ops_sh / (ms_sh * 1e-3) / 1e12,
                 "note": "timed alone on this rank after the main measurement; not included in value"}

    # ---- end to end through the public API with HOST buffers (push constants in, images out) ----
    e2e, e2e_rgba8 = None, None
    if not args.no_e2e:
        with torch.cuda.stream(cx.stream):
            if accumulate:
                def e2e_batch():
                    eng.trace_accumulate(pcs)
                    eng.readback_accum()
                d2h = Wf * Hf * 4
            else:
                # two page-locked buffers, rtce_render_to_host_async: when call k returns, batch k - 1 is complete in its
                # buffer (where a consumer would read it) while the device works on batch k; rtce_sync ends the region
                pins = [rt.PinnedBuffer((frames_rank, Hf, Wf)), rt.PinnedBuffer((frames_rank, Hf, Wf))]
                turn = [0]
                def e2e_batch():
                    eng.render_to_host_async(pcs, pins[turn[0] & 1].array)
                    turn[0] += 1
                d2h = frames_rank * Wf * Hf * 4
            for _ in range(3):
                e2e_batch()
            eng.sync()
            t1 = time.perf_counter()
            e2e_batch()
            eng.sync()
            t_one = time.perf_counter() - t1
            k_e2e = 3
            R_e2e = max(1, min(R, math.ceil(TARGET_TIMED_S / (k_e2e * max(t_one, 1e-5)))))
            cx.barrier()
            t0 = time.perf_counter()
            for _ in range(k_e2e * R_e2e):
                e2e_batch()
            eng.sync()          # the last batch is in host memory
            cx.barrier()
            dt = cx.max_over_ranks((time.perf_counter() - t0) / (k_e2e * R_e2e))   # per batch
        d2h_all = cx.sum_over_ranks([d2h])[0]
        e2e_verify = None
        if verify is not None and not accumulate:   # the frames that arrived in host memory are the reference's
            with open(os.path.join(ROOT, "tests", "golden", "golden.json")) as fh:
                gold = json.load(fh)["default_1080p_pose15"]["sha256"][-1]
            e2e_verify = all(hashlib.sha256(np.ascontiguousarray(p.array[15]).tobytes()).hexdigest() == gold for p in pins)
        e2e = {"value": rays_full_frame * frames / dt / 1e6, "unit": "Mrays/s", "ms_per_step": dt * 1e3 * R,
               "h2d_bytes_per_step": int(pcs.nbytes) * world * R, "d2h_bytes_per_step": int(d2h_all) * R,
               "steps": k_e2e, "batches_per_e2e_step_timed": R_e2e,
               "d2h_gb_per_s": d2h_all / dt / 1e9, "host_frame15_sha256_equals_golden(both buffers)": e2e_verify,
               "note": "every batch: host PushConstantData in (staged through a pinned ring into the device camera buffer), every frame's "
                       "final level copied to pinned host memory (accumulate workloads: the one accumulated image); N>1: each rank "
                       "returns its own frames to its host buffers; frame-by-frame workloads use rtce_render_to_host_async with two buffers "
                       "(call k returns when batch k - 1 is in host memory); timed on the host clock from the first call to the rtce_sync "
                       "after the last; "
                       "ms_per_step and the byte counts are scaled to the R batches of a step"}
        if not accumulate and wl["scene"] == "default" and not args.profile and world == 1:
            # the whole headless frame loop body (main.rs:886-914): trace + shade + RGBA8 "present" to host memory
            with torch.cuda.stream(cx.stream):
                pin8 = [rt.PinnedBuffer((frames_rank, int(wl["H"]), int(wl["W"]))) for _ in range(2)]   # 4 bytes per pixel = one RGBA8 texel
                out8 = [p.array.view(np.uint8).reshape(frames_rank, int(wl["H"]), int(wl["W"]), 4) for p in pin8]
                for k in range(3):
                    eng.render_rgba8_to_host_async(pcs, out8[k & 1])
                eng.sync()
                n8 = 40
                t0 = time.perf_counter()
                for k in range(n8):
                    eng.render_rgba8_to_host_async(pcs, out8[k & 1])
                eng.sync()
                dt8 = (time.perf_counter() - t0) / n8
            e2e_rgba8 = {"value": rays_full_frame * frames_rank / dt8 / 1e6, "unit": "Mrays/s", "ms_per_batch": dt8 * 1e3,
                         "frames_per_s": frames_rank / dt8, "d2h_bytes_per_batch": int(out8[0].nbytes),
                         "note": "rtce_render_rgba8_to_host_async, two buffers: compute pass + presentation shading pass + RGBA8 UNORM images to pinned host memory"}

    # ---- sub-objects: the other BASELINE configs, measured by this same invocation ----
    subs = {}
    extras = not (args.no_extras or args.profile) and args.workload == "c1" and not args.frames and not args.levels
    if extras and world == 1:
        eng.close()
        subs["c2"] = measure_accumulating(cx, "c2", with_cpu=not args.no_cpu_baseline)
        subs["c4"] = measure_accumulating(cx, "c4", steps=3, warmup=2, with_cpu=not args.no_cpu_baseline)
        subs["c3"] = measure_c3(cx, with_cpu=not args.no_cpu_baseline)
    elif extras:
        eng.close()
        subs["strong_c4"] = strong_c4(cx)
        subs["strong_c1"] = strong_c1(cx)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    hbm_peak, sm_max, peak_src = measured_peaks()
    fp32_peak, fp32_ms = rt.measure_fp32_issue(local_rank, 4096)
    if shade is not None:
        shade["frac_of_fp32_issue_peak"] = shade["achieved_tlaneop_per_s"] * 1e12 / fp32_peak
    nominal = 148 * 128 * sm_max * 1e6
    step_roof = roofline_step(ops_step, far_ops_each * far_step, ms_step, fp32_peak, n_gpus, nominal, bytes_step, hbm_peak, peak_src, cnt_step)
    step_roof["far_miss_rays"] = far_step
    traffic = None
    tp = os.path.join(ROOT, "profiles", "dominant_kernel_ncu.json")
    if os.path.exists(tp):
        with open(tp) as fh:
            traffic = json.load(fh).get("dram_bytes_per_launch")
    peak_note = ("measured live: rtce_measure_fp32_issue (non-contracted FMUL/FADD, register operands); "
                 f"nominal 148 SM x 128 lanes x {sm_max:.0f} MHz = {nominal / 1e12:.1f}")
    frac_note = ("frac counts EXECUTED work: the algorithmic FP32 operations (SURVEY.md 8d: 57 R + 10 E + (12 + 3 N) S, device counters "
                 "equal to the oracle's) of the rays that are really marched; rays resolved by the far-miss shortcut (RTCE_OPT_FAR_MISS: "
                 "parent texel beyond every sphere by more than RENDER_DIST, exactly start + RENDER_DIST) are NOT credited. "
                 "frac_reference_algorithm credits them with the operations the reference algorithm spends on them")
    if dom is not None:
        lpb = dom["launches_per_batch"]
        cnt_launch = {k: cnt_last_b[k] // lpb for k in "RES"}
        ops_launch = algorithmic_ops(cnt_launch, len(spheres))
        far_launch = far_last_b // lpb
        dom["counters_per_launch"] = cnt_launch
        t = dom["ms_per_launch"] * 1e-3
        executed = ops_launch - far_ops_each * far_launch
        bytes_launch = 5 * cnt_launch["R"]
        roofline = {"bound": "fp32_issue", "achieved": executed / t / 1e12, "peak": fp32_peak / 1e12, "unit": "Tlane-op/s",
                    "frac": executed / t / fp32_peak, "traffic": traffic, "peak_source": peak_note,
                    "frac_reference_algorithm": ops_launch / t / fp32_peak, "far_miss_rays_per_launch": far_launch,
                    "frac_note": frac_note, "frac_of_nominal": executed / t / nominal,
                    "ops_executed_per_launch": executed, "ops_reference_algorithm_per_launch": ops_launch,
                    "two_term_bound_ms": max(executed / fp32_peak, (traffic or bytes_launch) / (hbm_peak * 1e9)) * 1e3,
                    "scope": "dominant kernel: per launch / its mean launch duration (CUDA events on its stream around every launch, "
                             "one frame group, nothing else on the device, this rank)",
                    "dominant_kernel": dom, "step": step_roof}
    else:
        roofline = dict(step_roof)
        roofline.update({"bound": "fp32_issue", "peak": fp32_peak * n_gpus / 1e12, "unit": "Tlane-op/s", "traffic": None,
                         "peak_source": peak_note, "frac_note": frac_note, "step": step_roof,
                         "scope": "all launches of the step, gaps included (no single dominant launch was timed)"})

    cpu_baseline = None
    if not args.no_cpu_baseline and n_gpus == 1:
        cpu_baseline = cpu_baseline_of(wl, spheres, rots, poss, levels)
    for k in ("c2", "c3", "c4"):
        if k in subs:
            subs[k] = finish_sub(subs[k], fp32_peak, nominal, hbm_peak, peak_src)

    line = {
        "metric": "Mrays/s (pixel-invocations of compute.glsl:main over all pyramid levels)",
        "value": value, "unit": "Mrays/s", "n_gpus": n_gpus, "steps": steps, "warmup": warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["name"], "levels": levels, "batches_per_step": R, "frames_per_batch": frames,
                   "frames_per_step": frames * R, "rays_per_step": units_per_step, "primary_rays_per_step": Wf * Hf * frames * R,
                   "ms_per_batch": ms_step / R, "ms_per_frame": ms_step / (frames * R),
                   "step": f"{R} back-to-back batches of {frames_rank} frames per rank: sized at set-up so that the {steps} timed steps last >= {TARGET_TIMED_S} s",
                   "parallelism": "1 GPU" if n_gpus == 1 else
                                  (f"frames sharded: {n_gpus} batches of {frames_rank} frames per batch slot, one per rank (the same seeded poses on every "
                                   "rank: fixed work per GPU), whole image per frame, results stay in each rank's HBM; no data-path "
                                   "collective (frames are independent units); the row-band + NCCL-gather forms are strong_c4 / strong_c1"),
                   "verify": verify, "frames_in_flight": in_flight,
                   "l2": f"no flush: each batch rewrites {bytes_step / R / 1e6:.0f} MB of pyramid (> 126 MB L2); parents are meant to be L2 hits",
                   "kernel": {0: "auto (strip kernel + overflow launch per fine level, one per-pixel launch per coarse level)", 1: "per-pixel kernel per level",
                              3: "strip kernel + overflow launch per fine level, one per-pixel launch per coarse level"}[args.kernel]
                   if len(spheres) <= 8 else "large-scene kernel"},
        "clocks": clocks, "e2e": e2e, "e2e_rgba8": e2e_rgba8, "gpu_launches": int(launches_per_batch * R * steps) * n_gpus,
        "roofline": roofline, "cpu_baseline": cpu_baseline,
        "mprimary_per_s": Wf * Hf * frames * R / (ms_step * 1e-3) / 1e6,
        "shade_pass": shade,
    }
    line.update(subs)
    if world > 1:
        dist.destroy_process_group()
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    print(json.dumps(line), flush=True)


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

"""ctypes binding of the C ABI in include/rtce.h (librtce.so, built by build.py).

This is the Python host-side mirror of the reference's caller, `get_command_buffer` +
queue submit (src/main.rs:268-316, 909-914): set scene (binding 1), set view (binding 2 +
the pyramid of binding 0), then trace frames from push constants.  There is NO CPU path:
if librtce.so is missing or no CUDA device is present, construction raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librtce.so")

RTCE_OK = 0
OPT_COUNTERS, OPT_FRAME_SLOTS, OPT_KERNEL, OPT_STRIP_MIN_RAYS, OPT_GROUPS, OPT_STRIP_ROWS = 1, 2, 3, 4, 5, 7
OPT_GRAPH = 8
OPT_FRAMES_IN_FLIGHT = 9
OPT_STRIP_CTAS_PER_SM = 10
OPT_FAR_MISS = 11
OPT_LARGE_KERNEL = 12
OPT_MARCHER = 13
KERNEL_AUTO, KERNEL_PER_LEVEL, KERNEL_STRIP = 0, 1, 3
LARGE_AUTO, LARGE_REGISTER, LARGE_EVENT = 0, 2, 3
ABI_VERSION = 2
COMM_ID_BYTES = 128

# compute.glsl:7-12 PushConstantData, 40 bytes
PUSH_CONSTANTS_DTYPE = np.dtype(
    [("rot", "<f4", (4,)), ("pos", "<f4", (3,)), ("iter", "<u4"), ("imageSize", "<f4", (2,))]
)
assert PUSH_CONSTANTS_DTYPE.itemsize == 40

EXPORTS = [
    "rtce_trace_noalloc", "rtce_abi_version", "rtce_create", "rtce_destroy", "rtce_last_error", "rtce_status_string",
    "rtce_set_option", "rtce_set_stream", "rtce_set_scene", "rtce_set_spheres", "rtce_set_view",
    "rtce_set_rect", "rtce_level_count", "rtce_level_info", "rtce_host_level_count",
    "rtce_host_level_dims", "rtce_host_image_size", "rtce_trace", "rtce_trace_batch",
    "rtce_trace_accumulate", "rtce_accum_info", "rtce_prepare", "rtce_readback", "rtce_readback_accum",
    "rtce_render_to_host", "rtce_render_to_host_async", "rtce_render_rgba8_to_host", "rtce_render_rgba8_to_host_async", "rtce_host_alloc", "rtce_host_free", "rtce_sync", "rtce_flush", "rtce_get_counters",
    "rtce_get_level_counters", "rtce_get_far_miss_rays",
    "rtce_shade_batch", "rtce_color_info", "rtce_readback_color", "rtce_readback_color_rgba8", "rtce_get_shade_counters",
    "rtce_measure_fp32_issue", "rtce_selftest_math", "rtce_selftest_repeat", "rtce_debug_strip_timeline",
    "rtce_debug_launch_timeline",
    "rtce_camera_rotation", "rtce_camera_pose",
    "rtce_comm_unique_id", "rtce_comm_init", "rtce_host_band_rows", "rtce_gather_accum", "rtce_gather_frames", "rtce_gather_wait",
    "rtce_gather_info", "rtce_readback_gathered", "rtce_gather_last_ms",
]


class RtceError(RuntimeError):
    pass


class _Counters(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("sdf_evals", C.c_uint64), ("march_steps", C.c_uint64),
                ("kernel_launches", C.c_uint64)]


_lib = None


def load_library() -> C.CDLL:
    """Load librtce.so; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RtceError(f"{LIB_PATH} not built - run `python -m raytracing_engine_b200.build` "
                        "(there is no CPU fallback for the compute pass)")
    lib = C.CDLL(LIB_PATH)
    vp, i32, f32, sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
    sig = {
        "rtce_abi_version": (i32, []),
        "rtce_create": (i32, [i32, f32, C.POINTER(vp)]),
        "rtce_destroy": (None, [vp]),
        "rtce_last_error": (C.c_char_p, [vp]),
        "rtce_status_string": (C.c_char_p, [i32]),
        "rtce_set_option": (i32, [vp, i32, C.c_int64]),
        "rtce_set_stream": (i32, [vp, vp]),
        "rtce_set_scene": (i32, [vp, vp, sz]),
        "rtce_set_spheres": (i32, [vp, vp, C.c_uint32]),
        "rtce_set_view": (i32, [vp, vp, i32]),
        "rtce_set_rect": (i32, [vp, i32, i32, i32, i32]),
        "rtce_level_count": (i32, [vp]),
        "rtce_level_info": (i32, [vp, i32, i32, C.POINTER(i32), C.POINTER(i32), C.POINTER(vp)]),
        "rtce_host_level_count": (i32, [f32, i32]),
        "rtce_host_level_dims": (i32, [i32, f32, f32, i32, C.POINTER(i32), C.POINTER(i32)]),
        "rtce_host_image_size": (i32, [i32, i32, f32, f32, vp]),
        "rtce_trace": (i32, [vp, vp]),
        "rtce_trace_batch": (i32, [vp, vp, i32]),
        "rtce_trace_accumulate": (i32, [vp, vp, i32]),
        "rtce_accum_info": (i32, [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(vp)]),
        "rtce_readback": (i32, [vp, i32, i32, vp, sz]),
        "rtce_readback_accum": (i32, [vp, vp, sz]),
        "rtce_render_to_host": (i32, [vp, vp, i32, vp]),
        "rtce_render_to_host_async": (i32, [vp, vp, i32, vp]),
        "rtce_render_rgba8_to_host": (i32, [vp, vp, i32, vp]),
        "rtce_render_rgba8_to_host_async": (i32, [vp, vp, i32, vp]),
        "rtce_prepare": (i32, [vp, i32, i32]),
        "rtce_selftest_repeat": (i32, [i32, vp, vp, vp, sz]),
        "rtce_camera_rotation": (i32, [f32, f32, vp]),
        "rtce_camera_pose": (i32, [f32, f32, vp, vp]),
        "rtce_comm_unique_id": (i32, [vp]),
        "rtce_comm_init": (i32, [vp, vp, i32, i32]),
        "rtce_host_band_rows": (i32, [i32, i32, i32, C.POINTER(i32), C.POINTER(i32)]),
        "rtce_gather_accum": (i32, [vp, i32, vp]),
        "rtce_gather_frames": (i32, [vp, i32, vp, i32]),
        "rtce_gather_wait": (i32, [vp]),
        "rtce_gather_info": (i32, [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(vp)]),
        "rtce_readback_gathered": (i32, [vp, vp, sz]),
        "rtce_gather_last_ms": (i32, [vp, C.POINTER(f32)]),
        "rtce_host_alloc": (i32, [C.POINTER(vp), sz]),
        "rtce_host_free": (i32, [vp]),
        "rtce_sync": (i32, [vp]),
        "rtce_flush": (i32, [vp]),
        "rtce_get_counters": (i32, [vp, C.POINTER(_Counters)]),
        "rtce_get_level_counters": (i32, [vp, i32, C.POINTER(_Counters)]),
        "rtce_get_far_miss_rays": (i32, [vp, i32, C.POINTER(C.c_uint64)]),
        "rtce_shade_batch": (i32, [vp, vp, i32]),
        "rtce_color_info": (i32, [vp, i32, C.POINTER(i32), C.POINTER(i32), C.POINTER(vp)]),
        "rtce_readback_color": (i32, [vp, i32, vp]),
        "rtce_readback_color_rgba8": (i32, [vp, i32, vp]),
        "rtce_get_shade_counters": (i32, [vp, C.POINTER(C.c_uint64)]),
        "rtce_measure_fp32_issue": (i32, [i32, i32, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
        "rtce_selftest_math": (i32, [i32, C.POINTER(C.c_uint64)]),
        "rtce_debug_strip_timeline": (i32, [vp, i32, vp, i32]),
        "rtce_debug_launch_timeline": (i32, [vp, vp, i32]),
        "rtce_trace_noalloc": (i32, [i32, vp, vp, i32, vp, i32, f32, vp, i32, vp, sz, C.c_char_p, sz]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def host_level_count(view_w: float, resize_rule: bool = False) -> int:
    """main.rs:639 (startup) / main.rs:845 (after a resize)."""
    return int(load_library().rtce_host_level_count(float(view_w), int(resize_rule)))


def host_level_dims(levels: int, view_w: float, view_h: float):
    """main.rs:209-213 -> [(w, h)] per level."""
    lib = load_library()
    out = []
    w, h = C.c_int(), C.c_int()
    for i in range(levels):
        r = lib.rtce_host_level_dims(levels, float(view_w), float(view_h), i, C.byref(w), C.byref(h))
        if r != RTCE_OK:
            raise RtceError(f"rtce_host_level_dims: {r}")
        out.append((w.value, h.value))
    return out


def host_image_size(levels: int, level: int, view_w: float, view_h: float) -> np.ndarray:
    """main.rs:303-305."""
    o = np.zeros(2, dtype=np.float32)
    r = load_library().rtce_host_image_size(levels, level, float(view_w), float(view_h), o.ctypes.data)
    if r != RTCE_OK:
        raise RtceError(f"rtce_host_image_size: {r}")
    return o


def measure_fp32_issue(device: int = 0, iters: int = 4096):
    """Sustained non-contracted FMUL/FADD rate of the device -> (lane-ops per second, kernel ms)."""
    ops, ms = C.c_double(), C.c_double()
    r = load_library().rtce_measure_fp32_issue(int(device), int(iters), C.byref(ops), C.byref(ms))
    if r != RTCE_OK:
        raise RtceError(f"rtce_measure_fp32_issue: {r}")
    return ops.value, ms.value


def selftest_math(device: int = 0) -> dict:
    """Exhaustive GPU check of the short sqrt / rcp sequences against sqrt.rn / rcp.rn."""
    out = (C.c_uint64 * 4)()
    r = load_library().rtce_selftest_math(int(device), out)
    if r != RTCE_OK:
        raise RtceError(f"rtce_selftest_math: {r}")
    return {"sqrt_mismatches": int(out[0]), "rcp_mismatches": int(out[1]),
            "sqrt_checked": int(out[2]), "rcp_checked": int(out[3])}


def selftest_repeat(p, r, device: int = 0) -> np.ndarray:
    """utilities.glsl:32-34 `repeat` evaluated on the GPU, componentwise."""
    p = np.ascontiguousarray(p, dtype=np.float32).ravel()
    r = np.ascontiguousarray(r, dtype=np.float32).ravel()
    assert p.size == r.size
    out = np.empty_like(p)
    rc = load_library().rtce_selftest_repeat(int(device), p.ctypes.data, r.ctypes.data, out.ctypes.data, p.size)
    if rc != RTCE_OK:
        raise RtceError(f"rtce_selftest_repeat: {rc}")
    return out


def camera_rotation(yaw: float, pitch: float) -> np.ndarray:
    """main.rs:402-404 in binary32 (rtce_camera_rotation): [x, y, z, w]."""
    q = np.zeros(4, dtype=np.float32)
    rc = load_library().rtce_camera_rotation(float(yaw), float(pitch), q.ctypes.data)
    if rc != RTCE_OK:
        raise RtceError(f"rtce_camera_rotation: {rc}")
    return q


def camera_pose(yaw: float, pitch: float, local_delta, pc: np.ndarray) -> np.ndarray:
    """main.rs:770-773 (rtce_camera_pose): clamps the pitch, writes pc['rot'], adds the rotated local delta to pc['pos'].
    `pc` is ONE PushConstantData record (0-d or length-1 array of PUSH_CONSTANTS_DTYPE), updated in place."""
    assert pc.dtype == PUSH_CONSTANTS_DTYPE and pc.size == 1 and pc.flags["C_CONTIGUOUS"]
    d = None if local_delta is None else np.ascontiguousarray(local_delta, dtype=np.float32)
    rc = load_library().rtce_camera_pose(float(yaw), float(pitch), d.ctypes.data if d is not None else None, pc.ctypes.data)
    if rc != RTCE_OK:
        raise RtceError(f"rtce_camera_pose: {rc}")
    return pc


def host_band_rows(height: int, world: int, rank: int):
    """Equal split of the final level's rows in whole 8-row dispatch rows (rtce_host_band_rows)."""
    y0, y1 = C.c_int(), C.c_int()
    rc = load_library().rtce_host_band_rows(int(height), int(world), int(rank), C.byref(y0), C.byref(y1))
    if rc != RTCE_OK:
        raise RtceError(f"rtce_host_band_rows: {rc}")
    return y0.value, y1.value


def comm_unique_id() -> bytes:
    """ncclGetUniqueId through the library (rank 0 calls it and distributes the 128 bytes)."""
    buf = C.create_string_buffer(COMM_ID_BYTES)
    rc = load_library().rtce_comm_unique_id(buf)
    if rc != RTCE_OK:
        raise RtceError(f"rtce_comm_unique_id: {rc} (libnccl.so.2 not found?)")
    return buf.raw


def trace_noalloc(device: int, stream_ptr, spheres, view_w: float, view_h: float, pcs, d_pyramid: int, frame_stride: int,
                  fov: float = 1.0, levels: int = 0, render_dist: float = 1000.0, ratio=None) -> None:
    """rtce_trace_noalloc (include/rtce.h; no counterpart in the reference): the strip-kernel pass with its scratch in a
    static device array - no device allocation.  `d_pyramid` is a DEVICE address the caller owns (frames x frame_stride
    floats, levels packed back to back); asynchronous on `stream_ptr`.  Raises RtceError with the library's message."""
    w, h, fov = np.float32(view_w), np.float32(view_h), np.float32(fov)
    if ratio is None:
        ratio = (fov, np.float32(np.float32(fov * h) / w))
    cb = np.array([w, h, ratio[0], ratio[1]], dtype=np.float32)
    sph = np.ascontiguousarray(spheres, dtype=np.float32).reshape(-1, 4)
    pcs = np.ascontiguousarray(pcs, dtype=PUSH_CONSTANTS_DTYPE).reshape(-1)
    err = C.create_string_buffer(256)
    rc = load_library().rtce_trace_noalloc(int(device), C.c_void_p(stream_ptr or 0), sph.ctypes.data, int(sph.shape[0]), cb.ctypes.data,
                                           int(levels), float(render_dist), pcs.ctypes.data, int(pcs.size), C.c_void_p(int(d_pyramid)),
                                           int(frame_stride), err, len(err))
    if rc != RTCE_OK:
        e = RtceError(f"{err.value.decode()} ({rc})")
        e.code = rc
        raise e


def make_push_constants(rots, poss) -> np.ndarray:
    """Pack per-frame (rot[4] = quaternion xyzw, pos[3]) into PushConstantData records."""
    rots = np.asarray(rots, dtype=np.float32).reshape(-1, 4)
    poss = np.asarray(poss, dtype=np.float32).reshape(-1, 3)
    pcs = np.zeros(rots.shape[0], dtype=PUSH_CONSTANTS_DTYPE)
    pcs["rot"] = rots
    pcs["pos"] = poss
    pcs["imageSize"] = 1.0   # main.rs:624-629 initial value; overwritten per level
    return pcs


class PinnedBuffer:
    """Page-locked float32 host buffer (rtce_host_alloc) exposed as a numpy array."""

    def __init__(self, shape):
        self.lib = load_library()
        n = int(np.prod(shape))
        p = C.c_void_p()
        r = self.lib.rtce_host_alloc(C.byref(p), max(n, 1) * 4)
        if r != RTCE_OK:
            raise RtceError("rtce_host_alloc failed")
        self.ptr = p
        self.array = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(max(n, 1),))[:n].reshape(shape)

    def free(self):
        if self.ptr:
            self.array = None
            self.lib.rtce_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Engine:
    """One context = one GPU, one stream; mirrors the reference's per-frame compute recording."""

    def __init__(self, device: int = 0, render_dist: float = 1000.0):
        self.lib = load_library()
        self._ctx = C.c_void_p()
        r = self.lib.rtce_create(int(device), float(render_dist), C.byref(self._ctx))
        if r != RTCE_OK:
            raise RtceError(f"rtce_create(device={device}) failed: {self.lib.rtce_status_string(r).decode()} "
                            "(a CUDA device is required; there is no CPU fallback)")
        self.device = device
        self.view = None
        self.levels = 0

    # -- plumbing --------------------------------------------------------------------------
    def _check(self, r: int, what: str):
        if r != RTCE_OK:
            msg = self.lib.rtce_last_error(self._ctx)
            raise RtceError(f"{what}: {self.lib.rtce_status_string(r).decode()} ({msg.decode() if msg else ''})")

    def close(self):
        if self._ctx:
            self.lib.rtce_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def set_option(self, option: int, value: int):
        self._check(self.lib.rtce_set_option(self._ctx, option, value), "rtce_set_option")

    def set_stream(self, cuda_stream_handle: int):
        self._check(self.lib.rtce_set_stream(self._ctx, C.c_void_p(cuda_stream_handle)), "rtce_set_stream")

    # -- bindings --------------------------------------------------------------------------
    def set_scene(self, mutable_data: bytes):
        """Binding 1: the 656-byte std140 MutableData image (main.rs:592-605)."""
        buf = bytes(mutable_data)
        self._check(self.lib.rtce_set_scene(self._ctx, buf, len(buf)), "rtce_set_scene")

    def set_spheres(self, spheres):
        s = np.ascontiguousarray(np.asarray(spheres, dtype=np.float32).reshape(-1, 4))
        self._check(self.lib.rtce_set_spheres(self._ctx, s.ctypes.data, s.shape[0]), "rtce_set_spheres")

    def set_view(self, view_w: float, view_h: float, fov: float = 1.0, levels: int = 0, ratio=None):
        """Binding 2 (main.rs:608-611: ratio = [FOV, FOV*H/W]) and the pyramid (main.rs:639-648)."""
        w, h, fov = np.float32(view_w), np.float32(view_h), np.float32(fov)
        if ratio is None:
            ratio = (fov, np.float32(np.float32(fov * h) / w))
        cb = np.array([w, h, ratio[0], ratio[1]], dtype=np.float32)
        self._check(self.lib.rtce_set_view(self._ctx, cb.ctypes.data, int(levels)), "rtce_set_view")
        self.view = (float(w), float(h))
        self.ratio = (float(cb[2]), float(cb[3]))
        self.levels = int(self.lib.rtce_level_count(self._ctx))

    def set_rect(self, x0: int, y0: int, x1: int, y1: int):
        self._check(self.lib.rtce_set_rect(self._ctx, x0, y0, x1, y1), "rtce_set_rect")

    def level_info(self, level: int = -1, slot: int = 0):
        w, h, p = C.c_int(), C.c_int(), C.c_void_p()
        self._check(self.lib.rtce_level_info(self._ctx, level, slot, C.byref(w), C.byref(h), C.byref(p)),
                    "rtce_level_info")
        return w.value, h.value, p.value

    # -- the pass --------------------------------------------------------------------------
    def trace(self, pcs: np.ndarray):
        pcs = np.ascontiguousarray(pcs, dtype=PUSH_CONSTANTS_DTYPE).reshape(-1)
        self._check(self.lib.rtce_trace_batch(self._ctx, pcs.ctypes.data, pcs.shape[0]), "rtce_trace_batch")

    def trace_accumulate(self, pcs: np.ndarray):
        pcs = np.ascontiguousarray(pcs, dtype=PUSH_CONSTANTS_DTYPE).reshape(-1)
        self._check(self.lib.rtce_trace_accumulate(self._ctx, pcs.ctypes.data, pcs.shape[0]),
                    "rtce_trace_accumulate")

    def prepare(self, frames: int, accumulate: bool = False):
        """Allocate every lane's windows and capture the launch graphs of a batch of `frames` frames now."""
        self._check(self.lib.rtce_prepare(self._ctx, int(frames), int(bool(accumulate))), "rtce_prepare")

    def accum_info(self):
        w, h, p = C.c_int(), C.c_int(), C.c_void_p()
        self._check(self.lib.rtce_accum_info(self._ctx, C.byref(w), C.byref(h), C.byref(p)), "rtce_accum_info")
        return w.value, h.value, p.value

    # -- several GPUs: the one exchange ---------------------------------------------------------
    def comm_init(self, unique_id: bytes, rank: int, world: int):
        assert len(unique_id) == COMM_ID_BYTES
        self._check(self.lib.rtce_comm_init(self._ctx, unique_id, int(rank), int(world)), "rtce_comm_init")

    def gather_accum(self, dst: int = 0, bands=None):
        """Gather the accumulated image's row bands to rank `dst` (stream-ordered, overlaps later tracing)."""
        tab = None
        if bands is not None:
            tab = np.ascontiguousarray(np.asarray(bands, dtype=np.int32).reshape(-1, 2))
        self._check(self.lib.rtce_gather_accum(self._ctx, int(dst), tab.ctypes.data if tab is not None else None),
                    "rtce_gather_accum")

    def gather_frames(self, dst: int, bands, frames: int):
        """Gather the final-level row bands of slots 0..frames-1 of the latest batch to rank `dst`, in place."""
        tab = None
        if bands is not None:
            tab = np.ascontiguousarray(np.asarray(bands, dtype=np.int32).reshape(-1, 2))
        self._check(self.lib.rtce_gather_frames(self._ctx, int(dst), tab.ctypes.data if tab is not None else None, int(frames)),
                    "rtce_gather_frames")

    def gather_wait(self):
        self._check(self.lib.rtce_gather_wait(self._ctx), "rtce_gather_wait")

    def gather_info(self):
        w, h, p = C.c_int(), C.c_int(), C.c_void_p()
        self._check(self.lib.rtce_gather_info(self._ctx, C.byref(w), C.byref(h), C.byref(p)), "rtce_gather_info")
        return w.value, h.value, p.value

    def readback_gathered(self) -> np.ndarray:
        w, h, _ = self.level_info(-1, 0)
        out = np.empty((h, w), dtype=np.float32)
        self._check(self.lib.rtce_readback_gathered(self._ctx, out.ctypes.data, 0), "rtce_readback_gathered")
        return out

    def gather_last_ms(self) -> float:
        ms = C.c_float()
        self._check(self.lib.rtce_gather_last_ms(self._ctx, C.byref(ms)), "rtce_gather_last_ms")
        return float(ms.value)

    def sync(self):
        self._check(self.lib.rtce_sync(self._ctx), "rtce_sync")

    def flush(self):
        """Frames in flight: order the context's stream after every batch traced so far (no host wait)."""
        self._check(self.lib.rtce_flush(self._ctx), "rtce_flush")

    def readback(self, level: int = -1, slot: int = 0) -> np.ndarray:
        w, h, _ = self.level_info(level, slot)
        out = np.empty((h, w), dtype=np.float32)
        self._check(self.lib.rtce_readback(self._ctx, level, slot, out.ctypes.data, 0), "rtce_readback")
        return out

    def readback_all_levels(self, slot: int = 0):
        return [self.readback(i, slot) for i in range(self.levels)]

    def readback_accum(self) -> np.ndarray:
        w, h, _ = self.level_info(-1, 0)
        out = np.empty((h, w), dtype=np.float32)
        self._check(self.lib.rtce_readback_accum(self._ctx, out.ctypes.data, 0), "rtce_readback_accum")
        return out

    # ---- presentation shading pass (fragment.glsl) ----
    def shade(self, pcs: np.ndarray):
        """Shade the final level of slots 0..len(pcs)-1 of the latest batch with the batch's push constants."""
        pcs = np.ascontiguousarray(pcs, dtype=PUSH_CONSTANTS_DTYPE).reshape(-1)
        self._check(self.lib.rtce_shade_batch(self._ctx, pcs.ctypes.data, pcs.shape[0]), "rtce_shade_batch")

    def color_info(self, slot: int = 0):
        w, h, p = C.c_int(), C.c_int(), C.c_void_p()
        self._check(self.lib.rtce_color_info(self._ctx, slot, C.byref(w), C.byref(h), C.byref(p)), "rtce_color_info")
        return w.value, h.value, p.value

    def readback_color(self, slot: int = 0) -> np.ndarray:
        w, h, _ = self.color_info(slot)
        out = np.empty((h, w, 4), dtype=np.float32)
        self._check(self.lib.rtce_readback_color(self._ctx, slot, out.ctypes.data), "rtce_readback_color")
        return out

    def readback_color_rgba8(self, slot: int = 0) -> np.ndarray:
        w, h, _ = self.color_info(slot)
        out = np.empty((h, w, 4), dtype=np.uint8)
        self._check(self.lib.rtce_readback_color_rgba8(self._ctx, slot, out.ctypes.data), "rtce_readback_color_rgba8")
        return out

    def shade_counters(self) -> dict:
        o = (C.c_uint64 * 4)()
        self._check(self.lib.rtce_get_shade_counters(self._ctx, o), "rtce_get_shade_counters")
        return {"P": int(o[0]), "H": int(o[1]), "D": int(o[2]), "L": int(o[3])}

    def render_to_host(self, pcs: np.ndarray, out: np.ndarray | None = None) -> np.ndarray:
        """Host push constants in, every frame's final level out to host memory (end-to-end call)."""
        pcs = np.ascontiguousarray(pcs, dtype=PUSH_CONSTANTS_DTYPE).reshape(-1)
        w, h, _ = self.level_info(-1, 0)
        if out is None:
            out = np.empty((pcs.shape[0], h, w), dtype=np.float32)
        assert out.dtype == np.float32 and out.size == pcs.shape[0] * h * w and out.flags["C_CONTIGUOUS"]
        self._check(self.lib.rtce_render_to_host(self._ctx, pcs.ctypes.data, pcs.shape[0], out.ctypes.data),
                    "rtce_render_to_host")
        return out

    def render_to_host_async(self, pcs: np.ndarray, out: np.ndarray) -> None:
        """rtce_render_to_host_async: returns when the PREVIOUS asynchronous call's frames are in its buffer; `out`
        (page-locked, e.g. PinnedBuffer.array) is complete after the next call or sync() returns."""
        pcs = np.ascontiguousarray(pcs, dtype=PUSH_CONSTANTS_DTYPE).reshape(-1)
        w, h, _ = self.level_info(-1, 0)
        assert out.dtype == np.float32 and out.size == pcs.shape[0] * h * w and out.flags["C_CONTIGUOUS"]
        self._check(self.lib.rtce_render_to_host_async(self._ctx, pcs.ctypes.data, pcs.shape[0], out.ctypes.data),
                    "rtce_render_to_host_async")

    def render_rgba8_to_host(self, pcs: np.ndarray, out: np.ndarray | None = None) -> np.ndarray:
        """Headless frame loop body (main.rs:886-914): trace + shade + RGBA8 UNORM images to host memory."""
        pcs = np.ascontiguousarray(pcs, dtype=PUSH_CONSTANTS_DTYPE).reshape(-1)
        vw, vh = int(self.view[0]), int(self.view[1])
        if out is None:
            out = np.empty((pcs.shape[0], vh, vw, 4), dtype=np.uint8)
        assert out.dtype == np.uint8 and out.size == pcs.shape[0] * vh * vw * 4 and out.flags["C_CONTIGUOUS"]
        self._check(self.lib.rtce_render_rgba8_to_host(self._ctx, pcs.ctypes.data, pcs.shape[0], out.ctypes.data),
                    "rtce_render_rgba8_to_host")
        return out

    def render_rgba8_to_host_async(self, pcs: np.ndarray, out: np.ndarray) -> None:
        """rtce_render_rgba8_to_host_async: same contract as render_to_host_async, RGBA8 frames."""
        pcs = np.ascontiguousarray(pcs, dtype=PUSH_CONSTANTS_DTYPE).reshape(-1)
        vw, vh = int(self.view[0]), int(self.view[1])
        assert out.dtype == np.uint8 and out.size == pcs.shape[0] * vh * vw * 4 and out.flags["C_CONTIGUOUS"]
        self._check(self.lib.rtce_render_rgba8_to_host_async(self._ctx, pcs.ctypes.data, pcs.shape[0], out.ctypes.data),
                    "rtce_render_rgba8_to_host_async")

    def strip_timeline(self, level: int = -1, max_warps: int = 8192, fetch: bool = True):
        """Diagnostics: arm / fetch the per-warp timeline of the strip launch that traces `level`."""
        out = np.zeros((max_warps, 4), dtype=np.uint64)
        self._check(self.lib.rtce_debug_strip_timeline(self._ctx, level, out.ctypes.data if fetch else None, max_warps),
                    "rtce_debug_strip_timeline")
        return out[out[:, 2] != 0]

    def launch_timeline(self, fetch: bool = True, max_launches: int = 4096):
        """Diagnostics: arm (fetch=False) / fetch the event timeline of the kernel launches since arming."""
        out = np.zeros((max_launches, 6), dtype=np.float64)
        n = self.lib.rtce_debug_launch_timeline(self._ctx, out.ctypes.data if fetch else None, max_launches)
        if n < 0:
            self._check(n, "rtce_debug_launch_timeline")
        return out[:n]

    def level_counters(self, level: int = -1) -> dict:
        c = _Counters()
        self._check(self.lib.rtce_get_level_counters(self._ctx, level, C.byref(c)), "rtce_get_level_counters")
        return {"R": int(c.rays), "E": int(c.sdf_evals), "S": int(c.march_steps)}

    def far_miss_rays(self, level: int = -1) -> int:
        """Rays of the last counted call resolved by the far-miss shortcut (level < 0: every level)."""
        o = C.c_uint64()
        self._check(self.lib.rtce_get_far_miss_rays(self._ctx, level, C.byref(o)), "rtce_get_far_miss_rays")
        return int(o.value)

    def counters(self) -> dict:
        c = _Counters()
        self._check(self.lib.rtce_get_counters(self._ctx, C.byref(c)), "rtce_get_counters")
        return {"R": int(c.rays), "E": int(c.sdf_evals), "S": int(c.march_steps),
                "launches": int(c.kernel_launches)}

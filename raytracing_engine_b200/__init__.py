"""raytracing_engine_b200 — B200-native (sm_100a CUDA) build of ONE path of
IvoteSligte/raytracing_engine: the per-pixel cone-march compute pass
(shaders/compute.glsl + utilities.glsl, driven by src/main.rs:291-316).

The product is the C-ABI library `librtce.so` (include/rtce.h); this package is its Python
host-side mirror.  Nothing here imports `oracle/` and nothing computes the pass on the CPU.
"""
from .engine import (Engine, PinnedBuffer, RtceError, PUSH_CONSTANTS_DTYPE, make_push_constants, load_library,
                     host_level_count, host_level_dims, host_image_size, LIB_PATH, measure_fp32_issue, selftest_math,
                     OPT_COUNTERS, OPT_FRAME_SLOTS, OPT_KERNEL, OPT_STRIP_MIN_RAYS, OPT_GROUPS, OPT_STRIP_ROWS, OPT_GRAPH,
                     OPT_FRAMES_IN_FLIGHT, OPT_STRIP_CTAS_PER_SM, OPT_FAR_MISS, OPT_LARGE_KERNEL, OPT_MARCHER, KERNEL_AUTO,
                     KERNEL_PER_LEVEL, KERNEL_STRIP, LARGE_AUTO, LARGE_REGISTER, LARGE_EVENT, ABI_VERSION, selftest_repeat, trace_noalloc,
                     camera_rotation, camera_pose, host_band_rows, comm_unique_id)
from . import scene, camera

__all__ = ["Engine", "PinnedBuffer", "RtceError", "PUSH_CONSTANTS_DTYPE", "make_push_constants", "load_library",
           "host_level_count", "host_level_dims", "host_image_size", "scene", "camera", "LIB_PATH"]

"""rtce_trace_noalloc (include/rtce.h) — the allocation-free variant of the strip-kernel pass, called through the
C ABI and compared BIT FOR BIT (the bar of every parity test here) with the ordinary entry point
rtce_trace_batch + rtce_readback on the same inputs, on two calls in a row that share the library's static scratch
(the second with other poses, another scene and fewer frames: nothing of the first call may leak into it).

The entry point has no counterpart in the reference.  This file
is named so that pytest collects it last.
"""
import numpy as np
import pytest

import raytracing_engine_b200 as rt
from raytracing_engine_b200 import camera, scene

pytestmark = pytest.mark.gpu

W, H = 1000, 600          # odd-sized view: 7 levels, the last one 1000 x 600 (a multiple of 8 in both)
RENDER_DIST = 1000.0


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _expected(spheres, pcs):
    """The existing entry point: every level of every frame, strip kernel on every level as the variant does."""
    with rt.Engine(0, RENDER_DIST) as e:
        e.set_option(rt.OPT_KERNEL, rt.KERNEL_STRIP)
        e.set_option(rt.OPT_STRIP_MIN_RAYS, 0)
        e.set_option(rt.OPT_FRAME_SLOTS, len(pcs))
        e.set_spheres(spheres)
        e.set_view(W, H)
        e.trace(pcs)
        return e.levels, [e.readback_all_levels(slot=f) for f in range(len(pcs))]


def _call(torch, spheres, pcs, buf, stride, stream):
    rt.trace_noalloc(0, stream.cuda_stream, spheres, W, H, pcs, buf.data_ptr(), stride, render_dist=RENDER_DIST)
    stream.synchronize()
    return buf.cpu().numpy()


def _compare(got, dims, expected, stride):
    for f, levels in enumerate(expected):
        off = f * stride
        for i, (w, h) in enumerate(dims):
            a = got[off:off + w * h].reshape(h, w)
            assert np.array_equal(_bits(a), _bits(levels[i])), (f, i)
            off += w * h


def test_noalloc_matches_trace_batch_on_two_calls_in_a_row():
    import torch
    levels = rt.host_level_count(W)
    dims = rt.host_level_dims(levels, W, H)
    total = sum(w * h for w, h in dims)
    stride = total + 37                      # any stride >= the packed frame
    stream = torch.cuda.Stream(device=0)

    # call 1: the default scene, 16 frames (the maximum) on the bench camera path
    sph1 = scene.spheres_of(scene.default_mutable_data())
    pcs1 = rt.make_push_constants(*camera.bench_camera_path(16))
    # call 2: 8 spheres incl. one around the camera, 3 frames, other poses
    rs = np.random.RandomState(77)
    sph2 = np.zeros((8, 4), np.float32)
    sph2[:, :3] = rs.uniform(-15, 15, (8, 3))
    sph2[:, 3] = rs.uniform(0.5, 4.0, 8)
    sph2[0] = (0.0, 0.0, 0.0, 2.0)
    rots2 = np.stack([camera.rotation(0.3 + 0.2 * k, -0.1 * k) for k in range(3)])
    poss2 = np.array([[0.1 * k, -0.5, 0.25 * k] for k in range(3)], dtype=np.float32)
    pcs2 = rt.make_push_constants(rots2, poss2)

    lv1, exp1 = _expected(sph1, pcs1)
    lv2, exp2 = _expected(sph2, pcs2)
    assert lv1 == lv2 == levels

    # the caller owns the output; poisoned so that an unwritten texel cannot pass
    buf = torch.full((16 * stride,), float("nan"), dtype=torch.float32, device="cuda:0")
    free_before = torch.cuda.mem_get_info(0)[0]
    got1 = _call(torch, sph1, pcs1, buf, stride, stream)
    _compare(got1, dims, exp1, stride)
    buf.fill_(float("nan"))
    got2 = _call(torch, sph2, pcs2, buf, stride, stream)
    _compare(got2, dims, exp2, stride)
    # and the first inputs again: the scratch holds call 2's leftovers now
    buf.fill_(float("nan"))
    got3 = _call(torch, sph1, pcs1, buf, stride, stream)
    _compare(got3, dims, exp1, stride)
    # once the module is loaded (before the first call returned) the calls take no device memory
    free_mid = torch.cuda.mem_get_info(0)[0]
    _call(torch, sph2, pcs2, buf, stride, stream)
    assert torch.cuda.mem_get_info(0)[0] == free_mid
    del free_before


def test_noalloc_rejects_sizes_beyond_the_scratch():
    import torch
    levels = rt.host_level_count(W)
    total = sum(w * h for w, h in rt.host_level_dims(levels, W, H))
    sph = scene.spheres_of(scene.default_mutable_data())
    buf = torch.zeros((total,), dtype=torch.float32, device="cuda:0")
    with pytest.raises(rt.RtceError, match="RTCE_NOALLOC_MAX_FRAMES") as ei:
        rt.trace_noalloc(0, 0, sph, W, H, rt.make_push_constants(*camera.bench_camera_path(17)), buf.data_ptr(), total)
    assert ei.value.code == -4          # RTCE_ERR_LIMIT
    with pytest.raises(rt.RtceError, match="8 spheres") as ei:
        rt.trace_noalloc(0, 0, np.ones((9, 4), np.float32), W, H, rt.make_push_constants(*camera.static_path(1)), buf.data_ptr(), total)
    assert ei.value.code == -4
    with pytest.raises(rt.RtceError, match="frame_stride"):
        rt.trace_noalloc(0, 0, sph, W, H, rt.make_push_constants(*camera.static_path(1)), buf.data_ptr(), total - 1)
    torch.cuda.synchronize()

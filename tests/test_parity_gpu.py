"""GPU parity tests: the CUDA path, called through the C ABI (librtce.so), against
  * the committed golden fixtures minted from the reference's own shader text (oracle/_ref), at
    every size incl. BASELINE.json's full sizes (1080p, 4K, 8K/10 levels) via per-level SHA-256;
  * the CPU checker on seeded random scenes / cameras (fuzz);
  * size-independent properties (batch == single, band == full, slot independence).
Bar: BIT-EXACT float32 images on every pyramid level (max abs error = 0, stronger than the
north_star's <= 1e-3 / PSNR >= 60 dB), and equal work counters R / E / S.
"""
import numpy as np
import pytest

import oracle
import raytracing_engine_b200 as rt
from raytracing_engine_b200 import camera, scene

pytestmark = pytest.mark.gpu

# (kernel plan, strip threshold): per-pixel kernels (the second opinion); strip kernel on EVERY level (threshold 0,
# level 0 included); strip kernel on the larger levels over latency-oriented per-pixel launches; per-pixel launches
# of the latency-oriented kernel on every level (threshold never met); the shipped default
MODES = [(rt.KERNEL_PER_LEVEL, None), (rt.KERNEL_STRIP, 0), (rt.KERNEL_STRIP, 3000), (rt.KERNEL_STRIP, 10**12),
         (rt.KERNEL_AUTO, None)]


def _set_mode(e, mode):
    mode = mode if isinstance(mode, tuple) else (mode, None)
    kernel, min_rays = mode[0], mode[1]
    e.set_option(rt.OPT_KERNEL, kernel)
    if min_rays is not None:
        e.set_option(rt.OPT_STRIP_MIN_RAYS, min_rays)


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _engine(case, mode=rt.KERNEL_AUTO, slots=None, count=False):
    (name, kind, sph, W, H, levels, rot, pos, keep) = case
    e = rt.Engine(0, 1000.0)
    _set_mode(e, mode)
    if slots:
        e.set_option(rt.OPT_FRAME_SLOTS, slots)
    if count:
        e.set_option(rt.OPT_COUNTERS, 1)
    if len(sph) <= 8 and kind == "ref":
        e.set_scene(scene.pack_mutable_data(objects=[(tuple(s[:3]), s[3]) for s in sph]))   # the 656-byte path
    else:
        e.set_spheres(sph)
    e.set_view(W, H, levels=levels)
    return e


@pytest.mark.parametrize("mode", MODES)
def test_golden_all_cases_bit_exact(mode, golden, case_table):
    meta, imgs = golden
    for name, case in case_table.items():
        m = meta[name]
        with _engine(case, mode, slots=2, count=True) as e:
            assert e.levels == m["levels"], name
            e.trace(rt.make_push_constants([case[6]], [case[7]]))
            lv = e.readback_all_levels()
            assert [[a.shape[1], a.shape[0]] for a in lv] == m["dims"], name
            got = [oracle.sha256(a) for a in lv]
            assert got == m["sha256"], (name, [i for i, (x, y) in enumerate(zip(got, m["sha256"])) if x != y])
            c = e.counters()
            assert {k: c[k] for k in "RES"} == m["counters"], name
            if case[8]:
                for i, a in enumerate(lv):
                    assert np.array_equal(_bits(a), _bits(imgs[f"{name}/{i}"])), (name, i)


@pytest.mark.parametrize("mode", MODES)
def test_fuzz_against_cpu_checker(mode):
    """Seeded random scenes (0..8 spheres, some containing the camera), cameras and odd view sizes."""
    o = oracle.best()
    rs = np.random.RandomState(1234)
    for it in range(24):
        n = int(rs.randint(0, 9))
        sph = np.zeros((n, 4), np.float32)
        sph[:, :3] = rs.uniform(-12, 12, (n, 3))
        sph[:, 3] = rs.uniform(0.1, 4.0, n)
        W, H = int(rs.randint(9, 420)), int(rs.randint(9, 300))
        yaw, pitch = rs.uniform(-3.2, 3.2), rs.uniform(-1.5, 1.5)
        rot = camera.rotation(yaw, pitch)
        pos = rs.uniform(-6, 6, 3).astype(np.float32)
        rule = bool(it % 2)
        levels = o.level_count(float(W), rule)
        ref, cnt = o.trace(sph, [W, H], scene.view_ratio(W, H), rot, pos, 1000.0, levels)
        with rt.Engine(0, 1000.0) as e:
            _set_mode(e, mode)
            e.set_option(rt.OPT_COUNTERS, 1)
            e.set_spheres(sph)
            e.set_view(W, H, levels=-1 if rule else 0)
            e.trace(rt.make_push_constants([rot], [pos]))
            lv = e.readback_all_levels()
            for i, (a, b) in enumerate(zip(lv, ref)):
                assert np.array_equal(_bits(a), _bits(b)), (it, i, W, H, n, np.abs(a - b).max())
            c = e.counters()
            assert {k: c[k] for k in "RES"} == cnt
            # the SHIPPED instantiations: counters off -> COUNT = false kernels, graph replay, batches rotating over lanes
            e.set_option(rt.OPT_COUNTERS, 0)
            for rep in range(3):                       # capture, replay, replay on another lane
                e.trace(rt.make_push_constants([rot, rot], [pos, pos]))
                for slot in ((0, 1) if rep != 1 else (1,)):
                    for i, b in enumerate(ref):
                        assert np.array_equal(_bits(e.readback(i, slot)), _bits(b)), ("uncounted", it, rep, slot, i, n)


@pytest.mark.parametrize("far_miss", [1, 0])
def test_far_miss_shortcut_is_exact(far_miss):
    """RTCE_OPT_FAR_MISS: rays resolved as start + RENDER_DIST without an SDF == the full evaluation == the checker,
    with spheres near the far-miss bound, beyond RENDER_DIST, behind the camera, huge coordinates, no spheres."""
    o = oracle.best()
    rs = np.random.RandomState(99)
    scenes = []
    for scale in (1.0, 60.0, 300.0, 2000.0, 1.0e5):
        n = int(rs.randint(1, 9))
        sph = np.zeros((n, 4), np.float32)
        sph[:, :3] = rs.uniform(-10, 10, (n, 3)) * scale
        sph[:, 3] = rs.uniform(0.2, 3.0, n) * min(scale, 400.0)
        scenes.append(sph)
    scenes.append(np.zeros((0, 4), np.float32))
    for k, sph in enumerate(scenes):
        W, H = 200 + 8 * k, 120 + 3 * k
        rot = camera.rotation(0.3 * k, 0.1 * k - 0.2)
        pos = (rs.uniform(-5, 5, 3) * (1.0 if k % 2 else 50.0)).astype(np.float32)
        ref, cnt = o.trace(sph, [W, H], scene.view_ratio(W, H), rot, pos, 1000.0, 0)
        for mode in ((rt.KERNEL_STRIP, 0), (rt.KERNEL_AUTO, None)):
            with rt.Engine(0, 1000.0) as e:
                _set_mode(e, mode)
                e.set_option(rt.OPT_FAR_MISS, far_miss)
                e.set_option(rt.OPT_COUNTERS, 1)
                e.set_spheres(sph)
                e.set_view(W, H)
                e.trace(rt.make_push_constants([rot], [pos]))
                for i, (a, b) in enumerate(zip(e.readback_all_levels(), ref)):
                    assert np.array_equal(_bits(a), _bits(b)), (k, mode, i)
                c = e.counters()
                assert {q: c[q] for q in "RES"} == cnt, (k, mode)
                e.set_option(rt.OPT_COUNTERS, 0)          # the uncounted kernels (far-miss fast paths included), graph replay
                for rep in range(2):
                    e.trace(rt.make_push_constants([rot], [pos]))
                    for i, (a, b) in enumerate(zip(e.readback_all_levels(), ref)):
                        assert np.array_equal(_bits(a), _bits(b)), ("uncounted", k, mode, rep, i)


def test_far_miss_shortcut_with_a_distant_camera():
    """The far-miss bound carries a term for the rounding of origin = pos + dir * start, which scales with the
    camera's COORDINATES: a camera at 1e5 .. 1e8 with spheres a few units away, shortcut on == off == the checker."""
    o = oracle.best()
    for k, big in enumerate((1.0e5, 3.0e6, 1.0e7, 1.0e8)):
        pos = np.array([big, -0.5 * big, 0.25 * big], np.float32)
        sph = np.array([[pos[0] + 3.0, pos[1] + 9.0, pos[2] - 1.0, 2.0], [pos[0] - 40.0, pos[1] + 60.0, pos[2] + 7.0, 11.0]], np.float32)
        W, H = 176 + 8 * k, 100
        rot = camera.rotation(0.1 * k, 0.05)
        ref, cnt = o.trace(sph, [W, H], scene.view_ratio(W, H), rot, pos, 1000.0, 0)
        for fm in (1, 0):
            with rt.Engine(0, 1000.0) as e:
                e.set_option(rt.OPT_FAR_MISS, fm)
                e.set_option(rt.OPT_COUNTERS, 1)
                e.set_spheres(sph)
                e.set_view(W, H)
                e.trace(rt.make_push_constants([rot], [pos]))
                for i, (a, b) in enumerate(zip(e.readback_all_levels(), ref)):
                    assert np.array_equal(_bits(a), _bits(b)), (big, fm, i)
                c = e.counters()
                assert {q: c[q] for q in "RES"} == cnt, (big, fm)


def test_render_dist_parameter():
    """RENDER_DIST is a specialization constant (compute.glsl:32); other values via the C port."""
    o = oracle.load("port")
    sph = scene.spheres_of(scene.default_mutable_data())
    for rd in (50.0, 12.5):
        ref, _ = o.trace(sph, [320, 200], scene.view_ratio(320, 200), camera.IDENTITY_ROT, [0, 0, 0], rd, 0)
        with rt.Engine(0, rd) as e:
            e.set_spheres(sph)
            e.set_view(320, 200)
            e.trace(rt.make_push_constants([camera.IDENTITY_ROT], [[0, 0, 0]]))
            for a, b in zip(e.readback_all_levels(), ref):
                assert np.array_equal(_bits(a), _bits(b))


@pytest.mark.parametrize("mode", MODES)
def test_batch_equals_single_frames(mode, golden, case_table):
    """16 poses in one rtce_trace_batch == 16 single traces == golden for pose 5 and 15 (1080p)."""
    meta, _ = golden
    case = case_table["default_1080p"]
    rots, poss = camera.bench_camera_path(16)
    pcs = rt.make_push_constants(rots, poss)
    with _engine(case, mode, slots=16) as e:
        e.trace(pcs)
        batch = [e.readback(-1, k) for k in range(16)]
        assert oracle.sha256(batch[0]) == meta["default_1080p"]["sha256"][-1]
        assert oracle.sha256(batch[5]) == meta["default_1080p_pose5"]["sha256"][-1]
        assert oracle.sha256(batch[15]) == meta["default_1080p_pose15"]["sha256"][-1]
        lv5 = [e.readback(i, 5) for i in range(e.levels)]
        assert [oracle.sha256(a) for a in lv5] == meta["default_1080p_pose5"]["sha256"]
        for k in (3, 9):
            e.trace(pcs[k:k + 1])
            assert np.array_equal(_bits(e.readback(-1, 0)), _bits(batch[k]))
    # fewer slots than frames: chunks reuse slots, last chunk's frames are what remains
    with _engine(case, mode, slots=4) as e:
        e.trace(pcs)
        for k in range(12, 16):
            assert np.array_equal(_bits(e.readback(-1, k % 4)), _bits(batch[k]))


@pytest.mark.parametrize("mode", MODES)
def test_row_band_equals_full_image(mode, case_table):
    """Multi-GPU invariant: tracing a row band (plus its ancestor rows) gives the full image's bits."""
    case = case_table["default_800x600"]
    pcs = rt.make_push_constants([case[6]], [case[7]])
    with _engine(case, mode) as e:
        e.trace(pcs)
        full = e.readback(-1, 0)
    H, W = full.shape
    for (x0, y0, x1, y1) in [(0, 0, W, 304), (0, 304, W, 600), (0, 296, W, 312), (40, 48, 200, 120), (0, 592, W, 600)]:
        with _engine(case, mode) as e:
            e.set_rect(x0, y0, x1, y1)
            e.trace(pcs)
            part = e.readback(-1, 0)
            assert np.array_equal(_bits(part[y0:y1, x0:x1]), _bits(full[y0:y1, x0:x1])), (x0, y0, x1, y1)
            assert not part[:y0].any() and not part[y1:].any()


def test_render_to_host_and_accumulate(case_table):
    case = case_table["default_800x600"]
    rots, poss = camera.bench_camera_path(7)
    pcs = rt.make_push_constants(rots, poss)
    with _engine(case, slots=4) as e:
        singles = []
        for k in range(7):
            e.trace(pcs[k:k + 1])
            singles.append(e.readback(-1, 0))
        w, h, _ = e.level_info(-1, 0)
        pin = rt.PinnedBuffer((7, h, w))
        out = e.render_to_host(pcs, pin.array)
        for k in range(7):
            assert np.array_equal(_bits(out[k]), _bits(singles[k])), k
        out2 = e.render_to_host(pcs)   # pageable destination
        assert np.array_equal(_bits(out2), _bits(out))
        pin.free()
        # accumulation extension: sequential FP32 sum in frame order, then / frames
        e.trace_accumulate(pcs)
        acc = e.readback_accum()
        s = singles[0].copy()
        for k in range(1, 7):
            s = (s + singles[k]).astype(np.float32)
        assert np.array_equal(_bits(acc), _bits((s / np.float32(7)).astype(np.float32)))
        # static pose: (v + v) / 2 == v exactly
        st = rt.make_push_constants(*camera.static_path(2))
        e.trace_accumulate(st)
        assert np.array_equal(_bits(e.readback_accum()), _bits(singles[0]))


def test_render_to_host_async_double_buffer(case_table):
    """rtce_render_to_host_async: when call k returns, the frames of call k - 1 are in ITS buffer (two page-locked buffers,
    different poses per call); rtce_sync completes the last call; other tracers and the synchronous form may follow."""
    case = case_table["default_800x600"]
    calls = 6
    with _engine(case, slots=4) as e:
        w, h, _ = e.level_info(-1, 0)
        batches, expect = [], []
        for c in range(calls):
            rots = np.stack([camera.rotation(0.05 * c + 0.01 * k, 0.02 * k) for k in range(5)])
            poss = np.array([[0.1 * c, 0.05 * k, 0.0] for k in range(5)], dtype=np.float32)
            pcs = rt.make_push_constants(rots, poss)
            batches.append(pcs)
            expect.append(e.render_to_host(pcs).copy())          # the synchronous form is the reference here
        pins = [rt.PinnedBuffer((5, h, w)), rt.PinnedBuffer((5, h, w))]
        for p in pins:
            p.array[:] = np.nan
        for c in range(calls):
            e.render_to_host_async(batches[c], pins[c & 1].array)
            if c > 0:   # the previous call's buffer is complete now, while this call is still in flight
                assert np.array_equal(_bits(pins[(c - 1) & 1].array), _bits(expect[c - 1])), c
        e.sync()
        assert np.array_equal(_bits(pins[(calls - 1) & 1].array), _bits(expect[calls - 1]))
        # a tracer that writes window 0 right behind an asynchronous call must wait for its copies
        e.render_to_host_async(batches[0], pins[0].array)
        e.trace(batches[3][:4])
        got = [e.readback(-1, k) for k in range(4)]
        e.sync()
        assert np.array_equal(_bits(pins[0].array), _bits(expect[0]))
        for k in range(4):
            assert np.array_equal(_bits(got[k]), _bits(expect[3][k])), k
        assert np.array_equal(_bits(e.render_to_host(batches[1])), _bits(expect[1]))
        for p in pins:
            p.free()


@pytest.mark.parametrize("groups", [1, 2, 4, 8])
def test_graph_replay_equals_issued_launches(groups, case_table):
    """RTCE_OPT_GRAPH: a captured launch chain replayed with NEW cameras (they travel through device memory),
    after a scene change (graph invalidated) and for a ragged last chunk == every launch issued == the checker."""
    o = oracle.best()
    case = case_table["default_800x600"]
    sph = case[2]
    rots, poss = camera.bench_camera_path(22)
    pcs = rt.make_push_constants(rots, poss)

    def run(graph, pcs_, spheres=None):
        with _engine(case, (rt.KERNEL_STRIP, 3000), slots=8) as e:
            e.set_option(rt.OPT_GROUPS, groups)
            e.set_option(rt.OPT_GRAPH, graph)
            e.trace(pcs_[:8])                       # graph path: capture + first replay target
            a = [e.readback(-1, k) for k in range(8)]
            e.trace(pcs_[8:16])                     # replay, other cameras
            b = [[e.readback(i, k) for i in range(e.levels)] for k in range(8)]
            if spheres is not None:
                e.set_spheres(spheres)              # frozen scene parameters must not survive this
            e.trace(pcs_[11:22])                    # chunks of 8 + 3: second shape
            c = [e.readback(-1, k) for k in range(3)]
            return a, b, c

    moved = np.array(sph, np.float32).copy()
    moved[:, 0] += 0.75
    ga, gb, gc = run(1, pcs, moved)
    ia, ib, ic = run(0, pcs, moved)
    for k in range(8):
        assert np.array_equal(_bits(ga[k]), _bits(ia[k])), k
        for i in range(len(gb[k])):
            assert np.array_equal(_bits(gb[k][i]), _bits(ib[k][i])), (k, i)
    for k in range(3):
        assert np.array_equal(_bits(gc[k]), _bits(ic[k])), k
    W, H = case[3], case[4]
    ref, _ = o.trace(sph, [W, H], scene.view_ratio(W, H), rots[13], poss[13], 1000.0, case[5])
    for i, r in enumerate(ref):
        assert np.array_equal(_bits(gb[5][i]), _bits(r)), i
    ref, _ = o.trace(moved, [W, H], scene.view_ratio(W, H), rots[20], poss[20], 1000.0, case[5])
    assert np.array_equal(_bits(gc[1]), _bits(ref[-1]))


@pytest.mark.parametrize("in_flight", [2, 3, 4])
def test_frames_in_flight_equal_strict_order(in_flight, case_table):
    """RTCE_OPT_FRAMES_IN_FLIGHT: batches rotating over windows of slots on their own streams give the bits of
    strict stream order, every window keeps its batch until it is reused, and a scene change in between lands
    between the right batches."""
    case = case_table["default_800x600"]
    rots, poss = camera.bench_camera_path(40)
    pcs = rt.make_push_constants(rots, poss)
    moved = np.array(case[2], np.float32).copy()
    moved[:, 1] -= 0.5
    with _engine(case, slots=4) as e:                   # strict order: the expected images
        want = []
        for k in range(10):
            if k == 6:
                e.set_spheres(moved)
            e.trace(pcs[4 * k:4 * k + 4])
            want.append([e.readback(-1, j) for j in range(4)])
    with _engine(case, slots=4) as e:
        e.set_option(rt.OPT_FRAMES_IN_FLIGHT, in_flight)
        for k in range(10):
            if k == 6:
                e.set_spheres(moved)
            e.trace(pcs[4 * k:4 * k + 4])
            if k in (0, 3, 5, 6, 9):                    # reads in between (flush) and calls issued back to back
                for j in range(4):
                    assert np.array_equal(_bits(e.readback(-1, j)), _bits(want[k][j])), (k, j)
        e.flush()
        e.sync()
        lv = [e.readback(i, 2) for i in range(e.levels)]
        assert np.array_equal(_bits(lv[-1]), _bits(want[9][2]))
        # back to one frame in flight: strict order again, window 0
        e.set_option(rt.OPT_FRAMES_IN_FLIGHT, 1)
        e.trace(pcs[0:4])
        e.set_spheres(case[2])
        e.trace(pcs[4:8])
        assert np.array_equal(_bits(e.readback(-1, 1)), _bits(want[1][1]))


def test_accumulate_lanes_equal_single_lane(case_table):
    """rtce_trace_accumulate: chunks rotating over three lanes == one chunk at a time (counters on: one lane)."""
    case = case_table["default_800x600"]
    rots, poss = camera.bench_camera_path(23)
    pcs = rt.make_push_constants(rots, poss)
    with _engine(case, slots=4) as e:
        e.trace_accumulate(pcs)                         # 6 chunks over 3 lanes
        a = e.readback_accum()
        e.trace_accumulate(pcs[:9])                     # again, lanes warm, ragged end
        a9 = e.readback_accum()
    with _engine(case, slots=4, count=True) as e:
        e.trace_accumulate(pcs)
        b = e.readback_accum()
        e.trace_accumulate(pcs[:9])
        b9 = e.readback_accum()
    assert np.array_equal(_bits(a), _bits(b)) and np.array_equal(_bits(a9), _bits(b9))


def test_shade_pass_matches_reference_fragment_shader(shade_golden, shade_case_table):
    """fragment.glsl as a CUDA kernel == the reference's fragment shader compiled as C++ (golden
    hashes of the RGBA32F and RGBA8 images, counters), window smaller than the level image included."""
    import hashlib
    for name, (md, W, H, rot, pos) in shade_case_table.items():
        g = shade_golden[name]
        with rt.Engine(0, 1000.0) as e:
            e.set_option(rt.OPT_COUNTERS, 1)
            e.set_scene(md)
            e.set_view(W, H)
            pcs = rt.make_push_constants([rot], [pos])
            e.trace(pcs)
            e.shade(pcs)
            rgba = e.readback_color(0)
            assert rgba.shape == (H, W, 4), name
            assert oracle.sha256(rgba) == g["sha256_rgba32f"], name
            assert hashlib.sha256(e.readback_color_rgba8(0).tobytes()).hexdigest() == g["sha256_rgba8"], name
            assert e.shade_counters() == g["counters"], name


def test_shade_batch_band_and_frames_in_flight(shade_case_table):
    """A batch of poses shaded in one call == the checker per frame; a row band shades only its rows; with
    frames in flight the pass reads the window of the latest batch."""
    o = oracle.best()
    port = oracle.load("port")
    md, W, H, _, _ = shade_case_table["default_800x600"]
    rots, poss = camera.bench_camera_path(12)
    pcs = rt.make_push_constants(rots, poss)
    want = {}
    for k in (0, 5, 11):
        lv, _ = o.trace(scene.spheres_of(md), [W, H], scene.view_ratio(W, H), rots[k], poss[k], 1000.0, 0)
        want[k], _ = port.shade(md, [W, H], scene.view_ratio(W, H), rots[k], poss[k], lv[-1])
    with rt.Engine(0, 1000.0) as e:
        e.set_option(rt.OPT_FRAME_SLOTS, 6)
        e.set_scene(md)
        e.set_view(W, H)
        e.set_option(rt.OPT_FRAMES_IN_FLIGHT, 3)
        e.trace(pcs[0:6])
        e.trace(pcs[6:12])            # second lane, second window
        e.shade(pcs[6:12])
        assert np.array_equal(_bits(e.readback_color(5)), _bits(want[11]))
        e.set_option(rt.OPT_FRAMES_IN_FLIGHT, 1)
        e.trace(pcs[0:6])
        e.shade(pcs[0:6])
        assert np.array_equal(_bits(e.readback_color(0)), _bits(want[0]))
        assert np.array_equal(_bits(e.readback_color(5)), _bits(want[5]))
    with rt.Engine(0, 1000.0) as e:
        e.set_scene(md)
        e.set_view(W, H)
        e.set_rect(0, 296, W, 456)
        e.trace(pcs[5:6])
        e.shade(pcs[5:6])
        part = e.readback_color(0)
        assert np.array_equal(_bits(part[296:456]), _bits(want[5][296:456]))
        assert not part[:296].any() and not part[456:].any()


def test_shade_needs_materials_and_lights():
    with rt.Engine(0, 1000.0) as e:
        e.set_spheres(scene.spheres_of(scene.default_mutable_data()))   # no materials / lights
        e.set_view(320, 200)
        pcs = rt.make_push_constants([camera.IDENTITY_ROT], [[0, 0, 0]])
        e.trace(pcs)
        with pytest.raises(rt.RtceError):
            e.shade(pcs)


@pytest.mark.parametrize("large_kernel,n", [(3, 3000), (2, 3000), (2, 300), (3, 300), (2, 9), (3, 9), (0, 5000), (2, 7000), (3, 7000),
                                            (2, 10000), (3, 10000), (2, 12000), (0, 16384), (2, 16384)])
def test_large_scene_kernel_matches_checker(large_kernel, n):
    """objCount > 8 on bigger synthetic scenes than the golden cases: register kernel (2: every instantiation, 1 to 8
    groups of eight objects per thread, up to RTCE_MAX_SPHERES), event-driven kernel (3), auto (0)."""
    o = oracle.best()
    if n > o.max_objects:   # beyond the raised MAX_OBJECTS of oracle/_ref's second build: the C restatement has no limit
        o = oracle.load("port")
    sph = scene.synthetic_spheres(n, seed=0xB200)
    rot, pos = camera.rotation(0.2, -0.1), np.array([0.5, 0.0, 1.0], np.float32)
    view = [96, 56] if n <= 5000 else [40, 24]
    ref, cnt = o.trace(sph, view, scene.view_ratio(*view), rot, pos, 1000.0, 0)
    with rt.Engine(0, 1000.0) as e:
        e.set_option(rt.OPT_COUNTERS, 1)
        e.set_option(rt.OPT_LARGE_KERNEL, large_kernel)
        e.set_spheres(sph)
        e.set_view(*view)
        e.trace(rt.make_push_constants([rot, rot], [pos, pos]))
        for slot in (0, 1):
            for i, b in enumerate(ref):
                assert np.array_equal(_bits(e.readback(i, slot)), _bits(b)), (slot, i)
        c = e.counters()
        assert c["R"] == 2 * cnt["R"] and c["E"] == 2 * cnt["E"] and c["S"] == 2 * cnt["S"]
        # the uncounted instantiation of the same kernel
        e.set_option(rt.OPT_COUNTERS, 0)
        e.trace(rt.make_push_constants([rot], [pos]))
        assert np.array_equal(_bits(e.readback(len(ref) - 1, 0)), _bits(ref[-1]))


def test_set_spheres_refuses_more_than_the_limit():
    """RTCE_MAX_SPHERES (16384) is also the capacity of the register kernel: every accepted scene has a kernel."""
    sph = scene.synthetic_spheres(16385, seed=0xB200)
    with rt.Engine(0, 1000.0) as e:
        with pytest.raises(rt.RtceError):
            e.set_spheres(sph)


def test_error_paths():
    with rt.Engine(0, 1000.0) as e:
        with pytest.raises(rt.RtceError):
            e.trace(rt.make_push_constants([camera.IDENTITY_ROT], [[0, 0, 0]]))   # no scene / view yet
        with pytest.raises(rt.RtceError):
            e.set_scene(b"\0" * 100)
        bad = bytearray(scene.default_mutable_data())
        bad[4] = 9   # objCount = 9 > MAX_OBJECTS
        with pytest.raises(rt.RtceError):
            e.set_scene(bytes(bad))
        e.set_scene(scene.default_mutable_data())
        with pytest.raises(rt.RtceError):
            e.set_view(800, 600, levels=17)
        e.set_view(800, 600)
        with pytest.raises(rt.RtceError):
            e.set_rect(0, 0, 801, 600)
        with pytest.raises(rt.RtceError):
            e.readback(7, 0)


def test_fast_math_is_correctly_rounded_exhaustively():
    """All 2^32 float bit patterns: the kernels' short sqrt / rcp sequences == sqrt.rn / rcp.rn on their domains."""
    r = rt.selftest_math(0)
    assert r["sqrt_mismatches"] == 0 and r["rcp_mismatches"] == 0, r
    assert r["sqrt_checked"] == 1 + (0x7F800000 - 0x0D000000)          # +0 and every finite x >= 2^-101
    assert r["rcp_checked"] == 2 * (0x7E800000 - 0x00800000 + 1)       # 2^-126 <= |x| <= 2^126


# ---- BASELINE.json's own configs at full size (VERDICT r1 "Next round" #4) --------------------------------------

@pytest.mark.parametrize("levels", [1, 4, 8])
@pytest.mark.parametrize("large_kernel", [rt.LARGE_AUTO, rt.LARGE_EVENT])
def test_config3_ten_thousand_spheres_at_1080p_window(levels, large_kernel):
    """configs[3]: 10 000 synthetic spheres at 1920x1080, pyramid levels 1 / 4 / 8 (the "bounce-depth" axis), compared
    with the reference's shader text on a centred 256x144 window of EVERY level (rtce_set_rect / the checker's rect:
    the window's ancestor chain), bit for bit, with equal work counters."""
    o = oracle.best()
    W, H = 1920, 1080
    sph = scene.synthetic_spheres(10000, seed=0xB200)
    rect = [W // 2 - 128, H // 2 - 72, W // 2 + 128, H // 2 + 72]
    rot, pos = np.array(camera.IDENTITY_ROT, np.float32), np.zeros(3, np.float32)
    ref, cnt = o.trace(sph, [W, H], scene.view_ratio(W, H), rot, pos, 1000.0, levels, rect=rect)
    with rt.Engine(0, 1000.0) as e:
        e.set_option(rt.OPT_COUNTERS, 1)
        e.set_option(rt.OPT_LARGE_KERNEL, large_kernel)
        e.set_option(rt.OPT_FRAME_SLOTS, 1)
        e.set_spheres(sph)
        e.set_view(W, H, levels=levels)
        e.set_rect(*rect)
        e.trace(rt.make_push_constants([rot], [pos]))
        for i, b in enumerate(ref):
            assert np.array_equal(_bits(e.readback(i, 0)), _bits(b)), (levels, i)
        c = e.counters()
        assert {q: c[q] for q in "RES"} == cnt
        assert (ref[-1][rect[1]:rect[3], rect[0]:rect[2]] > 0).all()


def _sequential_mean(v, k):
    """The accumulation extension restated with torch on the GPU: ((v + v) + v) ... in frame order, then / k (IEEE binary32)."""
    import torch
    t = torch.from_numpy(np.ascontiguousarray(v)).cuda()
    s = t.clone()
    for _ in range(k - 1):
        s = s + t
    return (s / np.float32(k)).cpu().numpy()


@pytest.mark.parametrize("name,frames,slots,in_flight", [("default_4k", 256, 32, 8), ("default_8k_10levels", 1024, 32, 8)])
def test_accumulate_static_pose_at_config_sizes(name, frames, slots, in_flight, golden, case_table):
    """configs[2] / configs[4]: rtce_trace_accumulate of 256 static frames at 4K (1024 at 8K / 10 levels), 8 lanes x
    32-frame chunks: the frame it sums is the golden frame, and the result is the order-defined FP32 sum / k exactly
    (and stays within k 2^-24 relative of the single frame, every frame being identical: the rounding of k - 1 additions)."""
    meta, _ = golden
    case = case_table[name]
    pcs = rt.make_push_constants(*camera.static_path(frames, case[6], case[7]))
    with _engine(case, slots=slots) as e:
        e.set_option(rt.OPT_FRAMES_IN_FLIGHT, in_flight)
        e.trace(pcs[:1])
        single = e.readback(-1, 0)
        assert oracle.sha256(single) == meta[name]["sha256"][-1]
        e.prepare(frames, accumulate=True)
        e.trace_accumulate(pcs)
        acc = e.readback_accum()
        assert np.array_equal(_bits(acc), _bits(_sequential_mean(single, frames)))
        assert np.abs(acc.astype(np.float64) - single.astype(np.float64)).max() <= 2.0 ** -24 * frames * float(single.max())
        e.trace_accumulate(pcs[:2])                      # (v + v) / 2 == v exactly: the golden image itself
        assert oracle.sha256(e.readback_accum()) == meta[name]["sha256"][-1]


def test_eight_spheres_1080p_golden(golden, case_table):
    """MAX_OBJECTS = 8 (utilities.glsl:3) at 1080p: the N = 8 instantiations of the strip / overflow / per-pixel kernels."""
    meta, _ = golden
    case = case_table["eight_spheres_1080p"]
    m = meta["eight_spheres_1080p"]
    for count in (True, False):
        with _engine(case, slots=4, count=count) as e:
            pcs = rt.make_push_constants([case[6]] * 4, [case[7]] * 4)
            e.trace(pcs)
            for slot in (0, 3):
                assert [oracle.sha256(e.readback(i, slot)) for i in range(e.levels)] == m["sha256"], (count, slot)
            if count:
                c = e.counters()
                assert {k: c[k] // 4 for k in "RES"} == m["counters"]


def test_repeat_device_function_matches_the_reference_text():
    """utilities.glsl:32-34 `repeat` (defined, never called by the reference): the device function against the
    reference's own text compiled as C++ on 10^6 random (p, r) triples, bit for bit."""
    rs = np.random.RandomState(77)
    p = np.concatenate([rs.uniform(-500, 500, (700000, 3)), rs.uniform(-3, 3, (333334, 3))]).astype(np.float32)
    r = np.concatenate([rs.uniform(0.05, 40, (700000, 3)), rs.uniform(0.5, 2, (333334, 3))]).astype(np.float32)
    want = oracle.best().repeat(p, r)
    got = rt.selftest_repeat(p.ravel(), r.ravel()).reshape(-1, 3)
    assert np.array_equal(_bits(got), _bits(want))
    assert np.array_equal(_bits(oracle.load("port").repeat(p[:1000], r[:1000])), _bits(want[:1000]))


def test_prepare_then_trace_and_default_frames_in_flight(case_table):
    """rtce_prepare captures every lane's graphs up front; the default (8 frames in flight) gives the bits of strict order."""
    case = case_table["default_800x600"]
    rots, poss = camera.bench_camera_path(28)
    pcs = rt.make_push_constants(rots, poss)
    with _engine(case, slots=4) as e:
        e.set_option(rt.OPT_FRAMES_IN_FLIGHT, 1)
        want = []
        for k in range(7):
            e.trace(pcs[4 * k:4 * k + 4])
            want.append([e.readback(-1, j) for j in range(4)])
    with _engine(case, slots=4) as e:                    # default options
        e.prepare(4)
        e.prepare(7, accumulate=True)
        for k in range(7):
            e.trace(pcs[4 * k:4 * k + 4])                # back to back: rotates over the lanes
        for j in range(4):
            assert np.array_equal(_bits(e.readback(-1, j)), _bits(want[6][j])), j
        for k in range(7):                               # call, wait, call: the chain that owns the device
            e.trace(pcs[4 * k:4 * k + 4])
            e.sync()
            for j in (0, 3):
                assert np.array_equal(_bits(e.readback(-1, j)), _bits(want[k][j])), (k, j)


def test_accumulate_clears_rows_outside_a_changed_rect(case_table):
    """Rows outside the current rect are defined to be 0 after rtce_trace_accumulate, whatever an earlier call left there."""
    case = case_table["default_800x600"]
    pcs = rt.make_push_constants(*camera.static_path(2))
    with _engine(case, slots=2) as e:
        e.trace_accumulate(pcs)
        full = e.readback_accum()
        e.set_rect(0, 96, 800, 200)
        e.trace_accumulate(pcs)
        part = e.readback_accum()
        assert np.array_equal(_bits(part[96:200]), _bits(full[96:200]))
        assert not part[:96].any() and not part[200:].any()
        e.set_rect(0, 304, 800, 600)
        e.trace_accumulate(pcs)
        part = e.readback_accum()
        assert np.array_equal(_bits(part[304:]), _bits(full[304:])) and not part[:304].any()
        e.set_rect(0, 0, 0, 0)
        e.trace_accumulate(pcs)
        assert np.array_equal(_bits(e.readback_accum()), _bits(full))


def test_gather_accum_world_one_and_band_table(case_table):
    """rtce_gather_accum with a single rank (no communicator): the band is copied into the gather buffer on the
    communication stream; readback_gathered == the accumulated image; an explicit band table moves only those rows;
    two gathers back to back alternate buffers."""
    case = case_table["default_800x600"]
    pcs = rt.make_push_constants(*camera.bench_camera_path(5))
    with _engine(case, slots=4) as e:
        e.trace_accumulate(pcs)
        want = e.readback_accum()
        e.gather_accum(0)
        got = e.readback_gathered()
        assert np.array_equal(_bits(got), _bits(want))
        assert e.gather_last_ms() >= 0.0
        e.set_rect(0, 200, 800, 304)
        e.trace_accumulate(pcs)
        e.gather_accum(0, bands=[(200, 304)])
        e.trace_accumulate(pcs[:3])                       # overlaps the gather of the step before
        e.gather_accum(0, bands=[(200, 304)])
        band3 = e.readback_gathered()[200:304]
        e.set_rect(0, 0, 0, 0)
        e.trace_accumulate(pcs[:3])
        assert np.array_equal(_bits(band3), _bits(e.readback_accum()[200:304]))
        with pytest.raises(rt.RtceError):
            e.gather_accum(1)                              # no such rank


def test_render_rgba8_to_host_matches_shade_readback(shade_case_table):
    """The headless frame loop body (trace + shade + RGBA8 to host memory) == trace, shade, readback_color_rgba8 per frame."""
    md, W, H, _, _ = shade_case_table["default_800x600"]
    rots, poss = camera.bench_camera_path(7)
    pcs = rt.make_push_constants(rots, poss)
    with rt.Engine(0, 1000.0) as e:
        e.set_option(rt.OPT_FRAME_SLOTS, 4)
        e.set_scene(md)
        e.set_view(W, H)
        out = e.render_rgba8_to_host(pcs)
        assert out.shape == (7, H, W, 4)
        for k in (0, 3, 6):
            e.trace(pcs[k:k + 1])
            e.shade(pcs[k:k + 1])
            assert np.array_equal(out[k], e.readback_color_rgba8(0)), k
        # asynchronous form, two page-locked buffers: call k returns with call k - 1 complete; mixes with the depth form
        pins = [rt.PinnedBuffer((7, H, W)), rt.PinnedBuffer((7, H, W))]
        bufs = [p.array.view(np.uint8).reshape(7, H, W, 4) for p in pins]
        depth = e.render_to_host(pcs).copy()
        dpin = rt.PinnedBuffer(depth.shape)
        for k in range(5):
            bufs[k & 1][:] = 0
            e.render_rgba8_to_host_async(pcs, bufs[k & 1])
            if k > 0:
                assert np.array_equal(bufs[(k - 1) & 1], out), k
        e.render_to_host_async(pcs, dpin.array)          # its return completes the last RGBA8 call
        assert np.array_equal(bufs[0], out)
        e.sync()
        assert np.array_equal(_bits(dpin.array), _bits(depth))
        for p in pins + [dpin]:
            p.free()


@pytest.mark.parametrize("marcher", [1, 2])
def test_alternative_marchers_match_their_restatement(marcher, case_table):
    """RTCE_OPT_MARCHER 1 / 2: the author's algorithms 1 and 2 (shaders/tracing_algorithms.txt:2-13, :16-37) in the
    traceCone frame.  The reference holds only pseudo-code for them, so the checker is the C restatement of that text
    (oracle 'port'): parity UNPINNED by the reference.  Every level bit for bit, counters equal, on the golden case table's
    scenes with 1..8 spheres and on the fuzz scenes; marcher 3 through the same entry point is the shipped algorithm."""
    o = oracle.load("port")
    names = ["default_800x600", "default_64x48", "default_250x130_resize_rule", "one_sphere_160x96", "eight_spheres_320x200",
             "inside_sphere_200x120", "looking_back_320x200"]
    for name in names:
        (_, kind, sph, W, H, levels, rot, pos, keep) = case_table[name]
        n = o.level_count(float(W), resize_rule=(levels == -1)) if levels <= 0 else levels
        ref, cnt = o.trace(sph, [W, H], scene.view_ratio(W, H), rot, pos, 1000.0, n, marcher=marcher)
        with rt.Engine(0, 1000.0) as e:
            e.set_option(rt.OPT_MARCHER, marcher)
            e.set_option(rt.OPT_COUNTERS, 1)
            e.set_spheres(sph)
            e.set_view(W, H, levels=levels)
            e.trace(rt.make_push_constants([rot, rot], [pos, pos]))
            for slot in (0, 1):
                for i, b in enumerate(ref):
                    assert np.array_equal(_bits(e.readback(i, slot)), _bits(b)), (name, marcher, slot, i)
            c = e.counters()
            assert {k: c[k] // 2 for k in "RES"} == cnt, (name, marcher)
            e.set_option(rt.OPT_COUNTERS, 0)
            e.trace(rt.make_push_constants([rot], [pos]))
            assert np.array_equal(_bits(e.readback(-1, 0)), _bits(ref[-1])), (name, marcher, "uncounted")
    rs = np.random.RandomState(4321)
    for it in range(10):
        nsp = int(rs.randint(1, 9))
        sph = np.zeros((nsp, 4), np.float32)
        sph[:, :3] = rs.uniform(-12, 12, (nsp, 3))
        sph[:, 3] = rs.uniform(0.1, 4.0, nsp)
        W, H = int(rs.randint(9, 300)), int(rs.randint(9, 200))
        rot = camera.rotation(rs.uniform(-3.2, 3.2), rs.uniform(-1.5, 1.5))
        pos = rs.uniform(-6, 6, 3).astype(np.float32)
        ref, cnt = o.trace(sph, [W, H], scene.view_ratio(W, H), rot, pos, 1000.0, 0, marcher=marcher)
        with rt.Engine(0, 1000.0) as e:
            e.set_option(rt.OPT_MARCHER, marcher)
            e.set_option(rt.OPT_COUNTERS, 1)
            e.set_spheres(sph)
            e.set_view(W, H)
            e.trace(rt.make_push_constants([rot], [pos]))
            for i, b in enumerate(ref):
                assert np.array_equal(_bits(e.readback(i, 0)), _bits(b)), (it, marcher, i)
            c = e.counters()
            assert {k: c[k] for k in "RES"} == cnt, (it, marcher)
    with rt.Engine(0, 1000.0) as e:                      # algorithm 1 reads objs[0]: an empty scene is refused, not guessed
        e.set_option(rt.OPT_MARCHER, marcher)
        e.set_spheres(np.zeros((0, 4), np.float32))
        e.set_view(64, 48)
        with pytest.raises(rt.RtceError):
            e.trace(rt.make_push_constants([camera.IDENTITY_ROT], [[0, 0, 0]]))

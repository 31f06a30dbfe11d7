"""CPU tests of the checkers of the presentation shading pass (fragment.glsl): the plain-C port
and the reference's own fragment shader text (oracle/_ref/librtce_ref_frag.so) against the committed golden
fixtures (tests/golden/shade_golden.json, minted from the reference's shader), and the pinned pow()."""
import hashlib
import math

import numpy as np
import pytest

import oracle
from raytracing_engine_b200 import scene


def _depth(o, md, W, H, rot, pos):
    lv, _ = o.trace(scene.spheres_of(md), [W, H], scene.view_ratio(W, H), rot, pos, 1000.0, 0)
    return lv[-1]


@pytest.mark.parametrize("kind", ["port", "ref"])
def test_shade_checker_reproduces_golden(kind, shade_golden, shade_case_table):
    if not oracle.available(kind):
        pytest.skip(f"oracle '{kind}' not built here")
    o = oracle.load(kind)
    for name, (md, W, H, rot, pos) in shade_case_table.items():
        rgba, cnt = o.shade(md, [W, H], scene.view_ratio(W, H), rot, pos, _depth(o, md, W, H, rot, pos))
        g = shade_golden[name]
        assert rgba.shape == (H, W, 4), name
        assert oracle.sha256(rgba) == g["sha256_rgba32f"], name
        assert hashlib.sha256(oracle.rgba8(rgba).tobytes()).hexdigest() == g["sha256_rgba8"], name
        assert cnt == g["counters"], name
        assert not np.isnan(rgba).any() and (rgba[..., 3] == 1.0).all(), name


def test_shade_port_equals_reference_text_elementwise(shade_case_table):
    if not oracle.available("ref"):
        pytest.skip("oracle/_ref not built here")
    ref, port = oracle.load("ref"), oracle.load("port")
    for name in ("default_800x600_pose9", "random1_333x201", "random3_333x201"):
        md, W, H, rot, pos = shade_case_table[name]
        d = _depth(ref, md, W, H, rot, pos)
        a, ca = ref.shade(md, [W, H], scene.view_ratio(W, H), rot, pos, d)
        b, cb = port.shade(md, [W, H], scene.view_ratio(W, H), rot, pos, d, threads=3)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)) and ca == cb, name
        assert (a[..., :3].sum(axis=2) > 0).sum() > 0.5 * ca["H"] or ca["L"] == 0, name   # lit pixels exist


def test_background_is_black_and_window_smaller_than_level(shade_case_table):
    """Pixels whose depth reached RENDER_DIST are (0, 0, 0, 1); the fragments are the window's pixels although the
    level image is rounded up (333x201 window -> 336x208 level, main.rs:209-213)."""
    o = oracle.load("port")
    md, W, H, rot, pos = shade_case_table["random2_333x201"]
    d = _depth(o, md, W, H, rot, pos)
    assert d.shape == (208, 336)
    rgba, cnt = o.shade(md, [W, H], scene.view_ratio(W, H), rot, pos, d)
    assert rgba.shape == (201, 333, 4) and cnt["P"] == 333 * 201
    bg = d[:201, :333] >= 1000.0
    assert (rgba[bg][:, :3] == 0).all() and cnt["H"] == int((~bg).sum())


def test_pinned_pow_definition():
    """pow(x, y) = 0 for !(x > 0); exp2_p(y log2_p(x)) otherwise: known answers and accuracy."""
    p = oracle.pinned_pow
    assert p(-0.5, 3.0) == 0.0 and p(0.0, 2.0) == 0.0 and p(float("nan"), 2.0) == 0.0
    assert p(1.0, 123.0) == 1.0 and p(2.0, 10.0) == 1024.0 and p(4.0, 0.5) == 2.0 and p(0.5, 2.0) == 0.25
    assert p(1e-30, 200.0) == 0.0 and p(3.0, 0.0) == 1.0
    rs = np.random.RandomState(7)
    worst = 0.0
    for _ in range(5000):
        x = float(np.float32(rs.uniform(1e-4, 1.0)))
        y = float(np.float32(rs.uniform(0.25, 256.0)))
        q = math.pow(x, y)
        if q > 1e-30:
            worst = max(worst, abs(p(x, y) - q) / q)
    assert worst < 2e-5, worst


def test_unorm8_conversion():
    a = np.array([[-1.0, 0.0, 0.5 / 255, 1.5 / 255], [2.5 / 255, 0.999, 1.0, 7.0]], np.float32).reshape(2, 1, 4)
    assert oracle.rgba8(a).reshape(-1).tolist() == [0, 0, 0, 2, 2, 255, 255, 255]   # halves round to even

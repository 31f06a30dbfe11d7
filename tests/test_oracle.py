"""CPU tests of the checkers themselves: the plain-C port and the reference's own shader text
(oracle/_ref) against the committed golden fixtures and the SURVEY.md §8c known answers."""
import numpy as np
import pytest

import oracle
from raytracing_engine_b200 import scene

# FNV-1a-32 of the final level (identity camera at origin, default scene), minted by an
# independent build of the same shader text in the survey session.
SURVEY_FNV = {"default_800x600": "b9ce60b7", "default_1080p": "892b7e0b", "default_4k": "f1a56983",
              "default_8k_10levels": "eefbc09c"}
SURVEY_1080P_COUNTERS = {"R": 2771264, "E": 12248937, "S": 3897100}


def _run(o, case, rect=None, threads=0):
    (name, kind, sph, W, H, levels, rot, pos, keep) = case
    n = o.level_count(float(W), resize_rule=(levels == -1)) if levels <= 0 else levels
    return o.trace(sph, [W, H], scene.view_ratio(W, H), rot, pos, 1000.0, n, rect=rect, threads=threads)


def test_golden_matches_survey_known_answers(golden):
    meta, _ = golden
    for k, v in SURVEY_FNV.items():
        assert meta[k]["fnv1a32_final"] == v
    assert meta["default_1080p"]["counters"] == SURVEY_1080P_COUNTERS
    assert meta["default_800x600"]["hits"] == 107014
    assert abs(meta["default_800x600"]["centre"] - 5075.73096) < 1e-3
    assert abs(meta["default_1080p"]["centre"] - 7236.58594) < 1e-3


@pytest.mark.parametrize("kind", ["port", "ref", "refx"])
def test_checker_reproduces_golden(kind, golden, case_table):
    if not oracle.available(kind):
        pytest.skip(f"oracle '{kind}' not built here")
    meta, imgs = golden
    o = oracle.load(kind)
    for name, case in case_table.items():
        m = meta[name]
        nsph = len(case[2])
        if nsph > o.max_objects or m["levels"] > o.max_levels:
            continue
        if kind != "port" and name in ("default_8k_10levels", "default_4k"):
            continue   # the port covers the big ones here; _ref minted them
        lv, cnt = _run(o, case)
        assert [[a.shape[1], a.shape[0]] for a in lv] == m["dims"], name
        assert [oracle.sha256(a) for a in lv] == m["sha256"], name
        assert cnt == m["counters"], name
        if case[8]:
            for i, a in enumerate(lv):
                assert np.array_equal(a.view(np.uint32), imgs[f"{name}/{i}"].view(np.uint32)), (name, i)


def test_fnv_helper_known_answer(golden):
    meta, imgs = golden
    a = imgs["default_64x48/%d" % (meta["default_64x48"]["levels"] - 1)]
    assert oracle.fnv1a32(a) == meta["default_64x48"]["fnv1a32_final"]


def test_rect_equals_full_on_band(case_table):
    """Sharding invariant (SURVEY.md §4 iii): a pixel depends only on (x, y, level, uniforms)."""
    o = oracle.load("port")
    case = case_table["default_800x600"]
    full, _ = _run(o, case)
    H, W = full[-1].shape
    for (y0, y1) in [(0, 304), (304, 600), (296, 312)]:
        part, _ = _run(o, case, rect=[0, y0, W, y1])
        assert np.array_equal(part[-1][y0:y1].view(np.uint32), full[-1][y0:y1].view(np.uint32))
        assert not part[-1][:y0].any() and not part[-1][y1:].any()
    part, _ = _run(o, case, rect=[40, 48, 200, 120])
    assert np.array_equal(part[-1][48:120, 40:200], full[-1][48:120, 40:200])


def test_thread_count_does_not_change_bits(case_table):
    o = oracle.load("port")
    a, _ = _run(o, case_table["eight_spheres_320x200"], threads=1)
    b, _ = _run(o, case_table["eight_spheres_320x200"], threads=4)
    assert all(np.array_equal(x, y) for x, y in zip(a, b))


def test_hit_pixels_end_on_a_surface(case_table):
    """SURVEY.md §4 iv: a hit stops within one cone radius of the nearest sphere surface."""
    o = oracle.load("port")
    case = case_table["default_800x600"]
    lv, _ = _run(o, case)
    f = lv[-1]
    H, W = f.shape
    sph = case[2]
    isz = o.image_size(len(lv), len(lv) - 1, 800.0, 600.0)
    thr = np.float32(1.4142135) * np.float32(8) * isz[0]
    ys, xs = np.nonzero(f < 1000.0)
    ratio = scene.view_ratio(800, 600)
    nx = ((2 * xs + 1) * isz[0] - 1.0) * ratio[0]
    ny = ((2 * ys + 1) * isz[1] - 1.0) * ratio[1]
    d = np.stack([nx, np.ones_like(nx), ny], 1).astype(np.float64)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    p = d * f[ys, xs][:, None]
    sdf = np.min(np.linalg.norm(p[:, None, :] - sph[None, :, :3], axis=2) - sph[None, :, 3], axis=1)
    # the ray stopped when sdf <= radius and then stepped back by radius; radius <= (len+1)*thr
    rad = (f[ys, xs] + 1.0) * thr
    assert np.all(sdf <= 2.5 * rad + 1e-3)
    assert np.all(sdf >= -1e-3)


def test_rays_only_advance_from_parent_to_child(case_table):
    """SURVEY.md §4 v: a child starts where its parent stopped and `last = max(dist, 0)` (compute.glsl:59) never steps
    back, so a child texel is at least its parent texel minus the one step-back of a hit (`len -= radius`,
    compute.glsl:63), which is at most the child level's threshold."""
    o = oracle.load("port")
    for name in ("default_800x600", "eight_spheres_320x200"):
        case = case_table[name]
        lv, _ = _run(o, case)
        n = len(lv)
        W, H = float(case[3]), float(case[4])
        for i in range(1, n):
            child, parent = lv[i].astype(np.float64), lv[i - 1].astype(np.float64)
            h, w = child.shape
            ys, xs = np.mgrid[0:h, 0:w]
            start = parent[ys >> 1, xs >> 1]
            thr = float(np.float32(1.4142135) * np.float32(8) * o.image_size(n, i, W, H)[0])
            # on a hit child - start = len_prev + last - (len_prev + 1) * thr >= -thr (thr < 1 on every level >= 1);
            # 2e-3 covers the binary32 rounding of texels up to RENDER_DIST * 8
            assert thr < 1.0
            assert np.all(child >= start - thr - 2e-3), (name, i)

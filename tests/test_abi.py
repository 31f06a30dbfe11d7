"""CPU tests of the drop-in boundary: librtce.so loads, exports every symbol include/rtce.h declares,
the byte layouts are the reference's, and the host-side pyramid arithmetic equals the oracle's."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

import oracle
import raytracing_engine_b200 as rt
from raytracing_engine_b200 import engine, scene

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    txt = open(os.path.join(ROOT, "include", "rtce.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(rtce_[a-z_0-9]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    lib = rt.load_library()
    names = _declared()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/rtce.h but not exported by librtce.so"
    assert sorted(engine.EXPORTS) == names
    assert lib.rtce_abi_version() == 1


def test_struct_layouts_match_reference():
    # compute.glsl:7-12 / main.rs:301-307
    d = rt.PUSH_CONSTANTS_DTYPE
    assert d.itemsize == 40
    assert [d.fields[k][1] for k in ("rot", "pos", "iter", "imageSize")] == [0, 16, 28, 32]
    md = scene.default_mutable_data()
    assert len(md) == 656
    u = np.frombuffer(md, dtype="<u4")
    assert (u[0], u[1], u[2]) == (4, 4, 2)                      # main.rs:593-596
    f = np.frombuffer(md, dtype="<f4")
    assert list(f[272 // 4: 272 // 4 + 4]) == [5.0, 5.0, -1.0, 3.0]      # objs[0], main.rs:560-563
    assert list(f[(272 + 48) // 4: (272 + 48) // 4 + 4]) == [4.0, -1.0, 0.0, 2.0]
    assert list(f[16 // 4: 16 // 4 + 3]) == [np.float32(0.2), np.float32(0.2), 1.0]   # mats[0].color
    assert f[(16 + 32 + 20) // 4] == 10.0                                  # mats[1].shine
    assert list(f[(400 + 32 + 16) // 4: (400 + 32 + 16) // 4 + 3]) == [np.float32(1.2), np.float32(0.2), np.float32(0.3)]


def test_host_pyramid_math_equals_oracle():
    o = oracle.load("port")
    for W, H in [(800, 600), (1920, 1080), (3840, 2160), (7680, 4320), (64, 48), (250, 130), (1024, 768),
                 (1023, 767), (1025, 3), (8, 8), (5, 5), (2048, 2048), (333, 777)]:
        for rule in (False, True):
            n = rt.host_level_count(W, rule)
            assert n == o.level_count(float(W), rule), (W, rule)
            assert rt.host_level_dims(n, W, H) == o.level_dims(n, float(W), float(H))
            for i in range(n):
                assert np.array_equal(rt.host_image_size(n, i, W, H), o.image_size(n, i, float(W), float(H)))
    # SURVEY.md §8 a10/a11 known values
    assert rt.host_level_count(800) == 7 and rt.host_level_count(800, True) == 8
    assert rt.host_level_count(1920) == 8 and rt.host_level_count(3840) == 9 and rt.host_level_count(7680) == 10
    assert rt.host_level_dims(7, 800, 600) == [(16, 16), (32, 24), (56, 40), (104, 80), (200, 152), (400, 304), (800, 600)]
    assert rt.host_level_dims(8, 1920, 1080)[-2:] == [(960, 544), (1920, 1080)]


def test_bad_arguments_return_codes_not_crashes():
    lib = rt.load_library()
    assert lib.rtce_create(0, 1000.0, None) == -1
    assert lib.rtce_set_scene(None, None, 0) == -1
    w, h = C.c_int(), C.c_int()
    assert lib.rtce_host_level_dims(99, 800.0, 600.0, 0, C.byref(w), C.byref(h)) == -1
    assert lib.rtce_status_string(-4) == b"limit exceeded"
    assert lib.rtce_last_error(None) == b"null context"


@pytest.mark.skipif(torch.cuda.is_available(), reason="only meaningful without a GPU")
def test_no_cpu_fallback_without_a_gpu():
    with pytest.raises(rt.RtceError):
        rt.Engine()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "raytracing_engine_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".hpp", ".inc", ".h")):
                txt = open(os.path.join(dirpath, fn)).read()
                assert not re.search(r"^\s*(import|from)\s+oracle\b", txt, flags=re.M), fn
                assert not re.search(r"librtce_oracle|librtce_ref|#include\s+\"[^\"]*oracle", txt), fn


def test_rust_binding_source_matches_the_header():
    """rust/rtce.rs cannot be compiled here (no rustc): at least every extern it declares must exist in include/rtce.h
    with the same number of parameters."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "rtce.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    c_fns = {}
    for m in re.finditer(r"\b(rtce_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", hdr, flags=re.S):
        args = m.group(2).strip()
        c_fns[m.group(1)] = 0 if args in ("", "void") else args.count(",") + 1
    rs = open(os.path.join(root, "rust", "rtce.rs")).read()
    block = rs[rs.index('extern "C" {'):]
    block = block[:block.index("\n}")]
    block = re.sub(r"//.*", "", block)
    n = 0
    for m in re.finditer(r"fn\s+(rtce_[a-z0-9_]+)\s*\((.*?)\)\s*(->\s*[^;]+)?;", block, flags=re.S):
        name, args = m.group(1), m.group(2).strip()
        assert name in c_fns, f"{name} is not declared in include/rtce.h"
        assert (args.count(",") + 1 if args else 0) == c_fns[name], name
        n += 1
    assert n >= 25

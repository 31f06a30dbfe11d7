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
    assert len(names) >= 50
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/rtce.h but not exported by librtce.so"
    assert sorted(engine.EXPORTS) == names
    assert lib.rtce_abi_version() == engine.ABI_VERSION == 2


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
    assert n >= 38


def test_rust_binding_parameter_types_match_the_header():
    """Beyond arity: every parameter (and the return type) of every extern in rust/rtce.rs maps to the C type the header
    declares — `int` <-> c_int, `float` <-> f32, `size_t` <-> usize, `T*` <-> *mut T / *const T with the same constness."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = re.sub(r"/\*.*?\*/", "", open(os.path.join(root, "include", "rtce.h")).read(), flags=re.S)
    base = {"int": "c_int", "float": "f32", "size_t": "usize", "int64_t": "i64", "uint32_t": "u32", "uint64_t": "u64",
            "void": "c_void", "char": "c_char", "unsigned char": "u8", "rtce_ctx": "rtce_ctx",
            "rtce_push_constants": "PushConstantData", "rtce_constant_buffer": "ConstantBuffer", "rtce_sphere": "Sphere",
            "rtce_counters": "Counters", "double": "f64"}

    def c_to_rust(t):
        t = re.sub(r"\[\d*\]$", "*", t.strip())         # float out[2] -> float*
        stars = t.count("*")
        core = t.replace("*", " ").strip()
        const = core.startswith("const ")
        core = core[6:].strip() if const else core
        r = base[core]
        for i in range(stars):
            r = ("*const " if (const and i == 0) else "*mut ") + r
        return r

    c_sig = {}
    for m in re.finditer(r"^\s*((?:const\s+)?[a-z_0-9]+\s*\**)\s*(rtce_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", hdr, flags=re.S | re.M):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        params = []
        if args not in ("", "void"):
            for a in args.split(","):
                a = " ".join(a.split())
                mm = re.match(r"(.*?)([A-Za-z_][A-Za-z_0-9]*)(\[\d*\])?$", a)   # type, name, optional array
                params.append(c_to_rust(mm.group(1) + (mm.group(3) or "")))
        c_sig[name] = (None if ret == "void" else c_to_rust(ret), params)
    rs = open(os.path.join(root, "rust", "rtce.rs")).read()
    block = rs[rs.index('extern "C" {'):]
    block = re.sub(r"//.*", "", block[:block.index("\n}")])
    n = 0
    for m in re.finditer(r"fn\s+(rtce_[a-z0-9_]+)\s*\((.*?)\)\s*(?:->\s*([^;]+))?;", block, flags=re.S):
        name, args, ret = m.group(1), m.group(2).strip(), (m.group(3) or "").strip() or None
        got = [" ".join(a.split(":", 1)[1].split()) for a in args.split(",")] if args else []
        want_ret, want = c_sig[name]
        # pointers to pointers: the header's T** is *mut *mut T on the Rust side
        assert got == want, (name, got, want)
        assert ret == want_ret, (name, ret, want_ret)
        n += 1
    assert n >= 38


def test_header_compiles_as_c11_and_links_against_the_library(tmp_path):
    """include/rtce.h consumed by a plain C compiler (gcc -std=c11 -pedantic -Wall -Werror), the program linked against
    librtce.so: the header itself is the contract, not its Python or nvcc mirror.  Without a GPU the program checks the
    layouts, the host-side entry points and that rtce_create reports the missing device as an error code."""
    import subprocess
    src = tmp_path / "consumer.c"
    src.write_text(r"""
#include <stddef.h>
#include <stdio.h>
#include <string.h>
#include "rtce.h"
_Static_assert(sizeof(rtce_push_constants) == 40, "PushConstantData is 40 bytes");
_Static_assert(offsetof(rtce_push_constants, pos) == 16 && offsetof(rtce_push_constants, iter) == 28 &&
               offsetof(rtce_push_constants, imageSize) == 32, "push-constant offsets (compute.glsl:7-12)");
_Static_assert(sizeof(rtce_constant_buffer) == 16 && sizeof(rtce_sphere) == 16, "ConstantBuffer / Object");
_Static_assert(sizeof(rtce_counters) == 32, "counters");
int main(void) {
    if (rtce_abi_version() != RTCE_ABI_VERSION) return 1;
    if (rtce_host_level_count(1920.0f, 0) != 8 || rtce_host_level_count(800.0f, 1) != 8) return 2;
    int w = 0, h = 0;
    if (rtce_host_level_dims(8, 1920.0f, 1080.0f, 7, &w, &h) != RTCE_OK || w != 1920 || h != 1080) return 3;
    float isz[2];
    if (rtce_host_image_size(8, 7, 1920.0f, 1080.0f, isz) != RTCE_OK || isz[0] != 1.0f / 1920.0f) return 4;
    int y0 = 0, y1 = 0;
    if (rtce_host_band_rows(4320, 8, 7, &y0, &y1) != RTCE_OK || y1 != 4320 || y1 - y0 != 536) return 5;
    rtce_push_constants pc;
    memset(&pc, 0, sizeof pc);
    const float delta[3] = {0.0f, 2.0f, 0.0f};
    if (rtce_camera_pose(0.0f, 0.0f, delta, &pc) != RTCE_OK) return 6;
    if (pc.rot[0] != 0.0f || pc.rot[1] != 0.0f || pc.rot[2] != 0.0f || pc.rot[3] != 1.0f) return 7;   /* identity */
    if (pc.pos[0] != 0.0f || pc.pos[1] != 2.0f || pc.pos[2] != 0.0f) return 8;                         /* forward = +Y */
    rtce_ctx* ctx = NULL;
    int r = rtce_create(0, 1000.0f, &ctx);
    if (r == RTCE_OK) {                      /* a GPU is present: one tiny frame through the header's own types */
        unsigned char md[RTCE_MUTABLE_DATA_SIZE];
        memset(md, 0, sizeof md);
        unsigned n = 1; memcpy(md + RTCE_MUTABLE_DATA_OBJCOUNT_OFFSET, &n, 4);
        rtce_sphere s = {{0.0f, 6.0f, 0.0f}, 2.0f}; memcpy(md + RTCE_MUTABLE_DATA_OBJS_OFFSET, &s, sizeof s);
        rtce_constant_buffer cb = {{64.0f, 48.0f}, {1.0f, 48.0f / 64.0f}};
        static float img[64 * 48];
        if (rtce_set_scene(ctx, md, sizeof md) || rtce_set_view(ctx, &cb, 0) || rtce_trace(ctx, &pc) ||
            rtce_readback(ctx, -1, 0, img, 0)) { printf("%s\n", rtce_last_error(ctx)); return 9; }
        if (!(img[24 * 64 + 32] > 1.0f && img[24 * 64 + 32] < 6.0f)) return 10;   /* the centre ray stops on the sphere */
        rtce_destroy(ctx);
    } else if (r != RTCE_ERR_CUDA || ctx != NULL) {
        return 11;
    }
    puts("ok");
    return 0;
}
""")
    exe = tmp_path / "consumer"
    libdir = os.path.dirname(engine.LIB_PATH)
    cmd = ["gcc", "-std=c11", "-pedantic", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"), str(src),
           "-o", str(exe), "-L", libdir, "-lrtce", f"-Wl,-rpath,{libdir}"]
    env = dict(os.environ)
    env.pop("CC", None)
    out = subprocess.run(cmd, capture_output=True, text=True, env=env)
    assert out.returncode == 0, out.stderr
    run = subprocess.run([str(exe)], capture_output=True, text=True)
    assert run.returncode == 0 and run.stdout.strip() == "ok", (run.returncode, run.stdout, run.stderr)


def test_camera_pose_matches_the_float64_generator_and_is_exact_at_identity():
    """rtce_camera_rotation / rtce_camera_pose (main.rs:402-414, 770-773 in binary32) against camera.py (float64,
    rounded once): within 1 ulp-ish (4e-7 relative to the largest component), exact at yaw = pitch = 0, pitch clamped."""
    from raytracing_engine_b200 import camera
    assert np.array_equal(rt.camera_rotation(0.0, 0.0), np.array([0, 0, 0, 1], np.float32))
    rs = np.random.RandomState(5)
    for _ in range(200):
        yaw, pitch = float(rs.uniform(-7, 7)), float(rs.uniform(-1.5, 1.5))
        q = rt.camera_rotation(yaw, pitch)
        assert np.abs(q.astype(np.float64) - camera.rotation(yaw, pitch).astype(np.float64)).max() <= 2.4e-7
        assert abs(float(np.sum(q.astype(np.float64) ** 2)) - 1.0) < 1e-6
        delta = rs.uniform(-30, 30, 3).astype(np.float32)
        pos0 = rs.uniform(-100, 100, 3).astype(np.float32)
        pc = rt.make_push_constants([[0, 0, 0, 1]], [pos0])
        rt.camera_pose(yaw, pitch, delta, pc)
        assert np.array_equal(pc["rot"][0], q)
        want = camera.advance(pos0, yaw, pitch, delta)
        assert np.abs(pc["pos"][0].astype(np.float64) - want.astype(np.float64)).max() <= 4e-5
        assert pc["iter"][0] == 0 and np.array_equal(pc["imageSize"][0], np.array([1, 1], np.float32))
    pc = rt.make_push_constants([[0, 0, 0, 1]], [[1, 2, 3]])
    rt.camera_pose(0.0, 0.0, [0.5, -2.0, 4.0], pc)                 # identity: axes are the world axes, sums exact
    assert np.array_equal(pc["pos"][0], np.array([1.5, 0.0, 7.0], np.float32))
    a = rt.make_push_constants([[0, 0, 0, 1]], [[0, 0, 0]])
    b = rt.make_push_constants([[0, 0, 0, 1]], [[0, 0, 0]])
    rt.camera_pose(0.4, 9.0, None, a)                               # main.rs:770 clamp
    rt.camera_pose(0.4, float(np.float32(0.5) * np.float32(np.pi)), None, b)
    assert np.array_equal(a["rot"], b["rot"])


def test_band_rows_in_c_equal_the_python_partition():
    from raytracing_engine_b200 import sharding
    for H in (8, 48, 136, 600, 1080, 2160, 4320):
        for world in (1, 2, 3, 4, 8):
            assert [rt.host_band_rows(H, world, r) for r in range(world)] == sharding.all_bands(H, world)

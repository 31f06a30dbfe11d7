"""TEST INFRASTRUCTURE ONLY — ctypes loaders for the CPU checkers of the compute pass.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package.  The product package
(``raytracing_engine_b200``) never does; it fails loudly without its CUDA library.

Three checkers share one calling convention (see ``oracle/ref_driver.cpp``):

* ``ref``   oracle/_ref/librtce_ref.so      the reference's own compute.glsl +
            utilities.glsl compiled as C++ (<= 8 spheres, <= 9 levels, RENDER_DIST 1000)
* ``refx``  oracle/_ref/librtce_ref_ext.so  same text, MAX_OBJECTS / imgs[] raised
* ``port``  oracle/librtce_oracle.so        plain-C restatement (any size)

The presentation shading pass (fragment.glsl) has the same pair: ``Oracle.shade`` of
``ref`` / ``refx`` runs oracle/_ref/librtce_ref_frag.so (the reference's fragment.glsl compiled as C++),
of ``port`` the C restatement in rtce_oracle_shade.c.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIBS = {
    "ref": (os.path.join(_HERE, "_ref", "librtce_ref.so"), "rtce_ref"),
    "refx": (os.path.join(_HERE, "_ref", "librtce_ref_ext.so"), "rtce_refx"),
    "port": (os.path.join(_HERE, "librtce_oracle.so"), "rtce_oracle"),
}
_FRAG_LIB = os.path.join(_HERE, "_ref", "librtce_ref_frag.so")
_loaded: dict = {}


def build(verbose: bool = False) -> None:
    """Compile the C port, and oracle/_ref when /root/reference is present."""
    out = subprocess.run(["make", "-C", _HERE, "all"], capture_output=True, text=True)
    if verbose or out.returncode != 0:
        print(out.stdout[-4000:], out.stderr[-4000:])
    if out.returncode != 0:
        raise RuntimeError("oracle build failed")


def available(kind: str) -> bool:
    return os.path.exists(_LIBS[kind][0])


class Oracle:
    def __init__(self, kind: str):
        path, prefix = _LIBS[kind]
        if not os.path.exists(path):
            raise FileNotFoundError(f"oracle '{kind}' not built: {path} (run `make -C oracle`)")
        self.kind = kind
        self.lib = C.CDLL(path)
        g = lambda n: getattr(self.lib, f"{prefix}_{n}")
        self._trace = g("trace")
        self._trace.restype = C.c_int
        self._trace.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                C.c_float, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        self._level_count = g("level_count"); self._level_count.restype = C.c_int
        self._level_count.argtypes = [C.c_float, C.c_int]
        self._level_dims = g("level_dims"); self._level_dims.restype = None
        self._level_dims.argtypes = [C.c_int, C.c_float, C.c_float, C.c_int, C.c_void_p, C.c_void_p]
        self._image_size = g("image_size"); self._image_size.restype = None
        self._image_size.argtypes = [C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p]
        g("max_objects").restype = C.c_int
        g("max_levels").restype = C.c_int
        g("render_dist").restype = C.c_float
        self.max_objects = g("max_objects")()
        self.max_levels = g("max_levels")()
        self.fixed_render_dist = g("render_dist")()

    def _shade_fn(self):
        if getattr(self, "_shade", None) is None:
            if self.kind == "port":
                fn = self.lib.rtce_oracle_shade
            else:
                if not os.path.exists(_FRAG_LIB):
                    raise FileNotFoundError(f"{_FRAG_LIB} not built (run `make -C oracle`)")
                self._frag = C.CDLL(_FRAG_LIB)
                fn = self._frag.rtce_ref_shade
            fn.restype = C.c_int
            fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_int,
                           C.c_int, C.c_int, C.c_void_p, C.c_void_p]
            self._shade = fn
        return self._shade

    def shade(self, mutable_data: bytes, view, ratio, rot, pos, depth, render_dist=1000.0, threads=0):
        """Presentation shading pass of one frame: depth [h, w] float32 (the final pyramid level, at least the
        window's size) -> (rgba float32 [view_h, view_w, 4], counters {P pixels, H hit pixels, D distance() calls, L light evaluations})."""
        depth = np.ascontiguousarray(depth, dtype=np.float32)
        h, w = depth.shape
        md = np.frombuffer(bytes(mutable_data), dtype=np.uint8).copy()
        if md.size != 656:
            raise ValueError("MutableData is 656 bytes (std140)")
        view = np.ascontiguousarray(view, dtype=np.float32)
        ratio = np.ascontiguousarray(ratio, dtype=np.float32)
        rot = np.ascontiguousarray(rot, dtype=np.float32)
        pos = np.ascontiguousarray(pos, dtype=np.float32)
        out = np.zeros((int(view[1]), int(view[0]), 4), dtype=np.float32)
        cnt = (C.c_ulonglong * 4)()
        err = self._shade_fn()(md.ctypes.data, view.ctypes.data, ratio.ctypes.data, rot.ctypes.data, pos.ctypes.data,
                               float(render_dist), depth.ctypes.data, w, h, int(threads), out.ctypes.data, cnt)
        if err != 0:
            raise RuntimeError(f"oracle[{self.kind}] shade failed: {err}")
        return out, {"P": int(cnt[0]), "H": int(cnt[1]), "D": int(cnt[2]), "L": int(cnt[3])}

    def repeat(self, p, r) -> np.ndarray:
        """utilities.glsl:32-34 `repeat` on (n, 3) float32 arrays (the reference's own text for ref / refx)."""
        p = np.ascontiguousarray(p, dtype=np.float32).reshape(-1, 3)
        r = np.ascontiguousarray(r, dtype=np.float32).reshape(-1, 3)
        out = np.empty_like(p)
        prefix = _LIBS[self.kind][1]
        fn = getattr(self.lib, f"{prefix}_repeat")
        fn.restype = None
        fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
        fn(p.ctypes.data, r.ctypes.data, out.ctypes.data, p.shape[0])
        return out

    def level_count(self, view_w: float, resize_rule: bool = False) -> int:
        return int(self._level_count(view_w, int(resize_rule)))

    def level_dims(self, levels: int, view_w: float, view_h: float):
        out = []
        w, h = C.c_int(), C.c_int()
        for i in range(levels):
            self._level_dims(levels, view_w, view_h, i, C.byref(w), C.byref(h))
            out.append((w.value, h.value))
        return out

    def image_size(self, levels: int, i: int, view_w: float, view_h: float):
        o = (C.c_float * 2)()
        self._image_size(levels, i, view_w, view_h, o)
        return np.array([o[0], o[1]], dtype=np.float32)

    def trace(self, spheres, view, ratio, rot, pos, render_dist=1000.0, levels=0, rect=None, threads=0, marcher=3):
        """Run one frame.  Returns (list of per-level float32 arrays [h,w], counters dict).
        marcher 1 / 2 (C port only): the author's alternative algorithms, restated from pseudo-code (unpinned)."""
        sph = np.ascontiguousarray(np.asarray(spheres, dtype=np.float32).reshape(-1, 4))
        view = np.ascontiguousarray(view, dtype=np.float32)
        ratio = np.ascontiguousarray(ratio, dtype=np.float32)
        rot = np.ascontiguousarray(rot, dtype=np.float32)
        pos = np.ascontiguousarray(pos, dtype=np.float32)
        if levels <= 0:
            levels = self.level_count(float(view[0]))
        dims = self.level_dims(levels, float(view[0]), float(view[1]))
        imgs = [np.zeros((h, w), dtype=np.float32) for (w, h) in dims]
        ptrs = (C.c_void_p * levels)(*[im.ctypes.data for im in imgs])
        rc = None
        if rect is not None:
            rc = np.ascontiguousarray(rect, dtype=np.int32)
        cnt = (C.c_ulonglong * 3)()
        if marcher != 3:
            if self.kind != "port":
                raise ValueError("marchers 1 and 2 exist only as a restatement (oracle 'port'): the reference ships pseudo-code")
            fn = self.lib.rtce_oracle_trace_marcher
            fn.restype = C.c_int
            fn.argtypes = list(self._trace.argtypes) + [C.c_int]
            err = fn(sph.ctypes.data, sph.shape[0], view.ctypes.data, ratio.ctypes.data, rot.ctypes.data,
                     pos.ctypes.data, float(render_dist), levels, rc.ctypes.data if rc is not None else None,
                     int(threads), ptrs, cnt, int(marcher))
        else:
            err = self._trace(sph.ctypes.data, sph.shape[0], view.ctypes.data, ratio.ctypes.data, rot.ctypes.data,
                              pos.ctypes.data, float(render_dist), levels, rc.ctypes.data if rc is not None else None,
                              int(threads), ptrs, cnt)
        if err != 0:
            raise RuntimeError(f"oracle[{self.kind}] trace failed: {err}")
        return imgs, {"R": int(cnt[0]), "E": int(cnt[1]), "S": int(cnt[2])}


def rgba8(rgba: np.ndarray) -> np.ndarray:
    """RGBA32F -> RGBA8 UNORM with the pinned conversion (rtce_pinned_math.h), via the C port."""
    lib = load("port").lib
    lib.rtce_oracle_rgba8.restype = None
    lib.rtce_oracle_rgba8.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
    a = np.ascontiguousarray(rgba, dtype=np.float32)
    out = np.zeros(a.shape, dtype=np.uint8)
    lib.rtce_oracle_rgba8(a.ctypes.data, a.size // 4, out.ctypes.data)
    return out


def pinned_pow(x: float, y: float) -> float:
    lib = load("port").lib
    lib.rtce_oracle_pinned_pow.restype = C.c_float
    lib.rtce_oracle_pinned_pow.argtypes = [C.c_float, C.c_float]
    return float(lib.rtce_oracle_pinned_pow(x, y))


def load(kind: str) -> Oracle:
    if kind not in _loaded:
        _loaded[kind] = Oracle(kind)
    return _loaded[kind]


def best() -> Oracle:
    """The strongest checker present: the reference's own shader text if built, else the C port."""
    return load("refx") if available("refx") else load("port")


def fnv1a32(a: np.ndarray) -> str:
    """FNV-1a-32 over the 32-bit words of a float32 image, row-major (the SURVEY.md §8c hashes)."""
    hv = 0x811C9DC5
    for w in np.ascontiguousarray(a, dtype=np.float32).view(np.uint32).ravel().tolist():
        hv = ((hv ^ w) * 0x01000193) & 0xFFFFFFFF
    return f"{hv:08x}"


def sha256(a: np.ndarray) -> str:
    """SHA-256 of the raw little-endian float32 bytes (fast; used by the golden fixtures)."""
    import hashlib
    return hashlib.sha256(np.ascontiguousarray(a, dtype="<f4").tobytes()).hexdigest()

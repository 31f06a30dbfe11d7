"""Headless image writers: the reference presents to a swapchain (main.rs:909-914) and
has no on-disk format at all; regression artefacts of the headless path are written as

* PFM  (``Pf`` / ``PF``, little-endian float32, bottom-to-top rows) for the r32f depth levels and RGB32F colour,
* PPM  (``P6``, 8-bit RGB) for the RGBA8 presentation image,
* and a tiny binary scene file for ``MutableData`` (the 656-byte std140 image as is) or an N > 8 sphere list.

Pure host code: numpy in, bytes out; nothing here touches the GPU.
"""
from __future__ import annotations

import struct

import numpy as np

SCENE_MAGIC = b"RTCESCN1"


def write_pfm(path: str, img: np.ndarray) -> None:
    """img: [h, w] (grey, 'Pf') or [h, w, 3|4] (colour, 'PF'; alpha dropped) float32.  Row 0 is the TOP row of the
    image (as in the pyramid levels); PFM stores rows bottom to top, negative scale = little-endian."""
    a = np.asarray(img, dtype="<f4")
    if a.ndim == 3 and a.shape[2] == 4:
        a = a[..., :3]
    if a.ndim == 2:
        kind = b"Pf"
    elif a.ndim == 3 and a.shape[2] == 3:
        kind = b"PF"
    else:
        raise ValueError("PFM wants [h, w] or [h, w, 3|4]")
    h, w = a.shape[:2]
    with open(path, "wb") as fh:
        fh.write(kind + b"\n" + f"{w} {h}\n-1.0\n".encode("ascii"))
        fh.write(np.ascontiguousarray(a[::-1]).tobytes())


def read_pfm(path: str) -> np.ndarray:
    with open(path, "rb") as fh:
        kind = fh.readline().strip()
        w, h = (int(t) for t in fh.readline().split())
        scale = float(fh.readline())
        ch = {b"Pf": 1, b"PF": 3}[kind]
        a = np.frombuffer(fh.read(w * h * ch * 4), dtype="<f4" if scale < 0 else ">f4")
    a = a.reshape((h, w) if ch == 1 else (h, w, 3))[::-1]
    return np.ascontiguousarray(a, dtype=np.float32)


def write_ppm(path: str, rgba8: np.ndarray) -> None:
    """rgba8: [h, w, 3|4] uint8 (the RGBA8 presentation image; alpha dropped), row 0 on top as PPM wants."""
    a = np.asarray(rgba8, dtype=np.uint8)
    if a.ndim != 3 or a.shape[2] not in (3, 4):
        raise ValueError("PPM wants [h, w, 3|4] uint8")
    h, w = a.shape[:2]
    with open(path, "wb") as fh:
        fh.write(f"P6\n{w} {h}\n255\n".encode("ascii"))
        fh.write(np.ascontiguousarray(a[..., :3]).tobytes())


def read_ppm(path: str) -> np.ndarray:
    with open(path, "rb") as fh:
        assert fh.readline().strip() == b"P6"
        w, h = (int(t) for t in fh.readline().split())
        assert int(fh.readline()) == 255
        return np.frombuffer(fh.read(w * h * 3), dtype=np.uint8).reshape(h, w, 3).copy()


def write_scene(path: str, mutable_data: bytes | None = None, spheres: np.ndarray | None = None) -> None:
    """Scene file: magic, kind (0 = std140 MutableData image, 1 = sphere list), count, payload."""
    with open(path, "wb") as fh:
        fh.write(SCENE_MAGIC)
        if mutable_data is not None:
            if len(mutable_data) != 656:
                raise ValueError("MutableData is 656 bytes (std140)")
            fh.write(struct.pack("<II", 0, 1) + bytes(mutable_data))
        else:
            s = np.ascontiguousarray(np.asarray(spheres, dtype="<f4").reshape(-1, 4))
            fh.write(struct.pack("<II", 1, s.shape[0]) + s.tobytes())


def read_scene(path: str):
    """Returns ("mutable_data", bytes) or ("spheres", float32 [n, 4])."""
    with open(path, "rb") as fh:
        if fh.read(8) != SCENE_MAGIC:
            raise ValueError("not an rtce scene file")
        kind, n = struct.unpack("<II", fh.read(8))
        if kind == 0:
            return "mutable_data", fh.read(656)
        return "spheres", np.frombuffer(fh.read(16 * n), dtype="<f4").reshape(n, 4).astype(np.float32)

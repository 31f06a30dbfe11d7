"""Scene data of the reference in its own byte layout (host side of binding 1).

`MutableData` is the std140 uniform block of compute.glsl:17-24 as built at
src/main.rs:524-605: 656 bytes = matCount@0, objCount@4, lightCount@8, pad@12,
mats[8]@16 (stride 32: color vec3@0, diffuse@12, specular@16, shine@20, ambient@24, pad@28),
objs[8]@272 (stride 16: pos vec3@0, size@12), lights[8]@400 (stride 32: pos@0, pad@12,
color@16, pad@28).  The compute pass reads objCount and objs only.
"""
from __future__ import annotations

import numpy as np

MAX_MATERIALS = MAX_OBJECTS = MAX_LIGHTS = 8   # utilities.glsl:1-3
RENDER_DIST = 1000.0                           # main.rs:362
FOV = 1.0                                      # main.rs:364
DEFAULT_WINDOW = (800, 600)                    # winit 0.27 default inner size (main.rs:431-433 sets none)

MATERIAL_DTYPE = np.dtype([("color", "<f4", (3,)), ("diffuse", "<f4"), ("specular", "<f4"), ("shine", "<f4"),
                           ("ambient", "<f4"), ("_dummy0", "u1", (4,))])
OBJECT_DTYPE = np.dtype([("pos", "<f4", (3,)), ("size", "<f4")])
LIGHT_DTYPE = np.dtype([("pos", "<f4", (3,)), ("_dummy0", "u1", (4,)), ("color", "<f4", (3,)), ("_dummy1", "u1", (4,))])
MUTABLE_DATA_DTYPE = np.dtype([("matCount", "<u4"), ("objCount", "<u4"), ("lightCount", "<u4"), ("_pad", "u1", (4,)),
                               ("mats", MATERIAL_DTYPE, (MAX_MATERIALS,)), ("objs", OBJECT_DTYPE, (MAX_OBJECTS,)),
                               ("lights", LIGHT_DTYPE, (MAX_LIGHTS,))])
assert MATERIAL_DTYPE.itemsize == 32 and OBJECT_DTYPE.itemsize == 16 and LIGHT_DTYPE.itemsize == 32
assert MUTABLE_DATA_DTYPE.itemsize == 656
assert MUTABLE_DATA_DTYPE.fields["objs"][1] == 272 and MUTABLE_DATA_DTYPE.fields["lights"][1] == 400

# main.rs:524-557 (color, diffuse, specular, shine, ambient)
DEFAULT_MATERIALS = [
    ((0.2, 0.2, 1.0), 1.0, 1.0, 1.0, 0.05),
    ((0.1, 1.0, 0.1), 1.0, 1.0, 10.0, 0.05),
    ((1.0, 1.0, 0.1), 1.0, 1.0, 1.0, 0.05),
    ((1.0, 0.1, 0.1), 1.0, 1.0, 1.0, 0.05),
]
# main.rs:559-576 (pos, size)
DEFAULT_OBJECTS = [
    ((5.0, 5.0, -1.0), 3.0),
    ((5.0, 4.0, 10.0), 6.0),
    ((-3.0, 3.0, -3.0), 1.0),
    ((4.0, -1.0, 0.0), 2.0),
]
# main.rs:578-591 (pos, color)
DEFAULT_LIGHTS = [
    ((-1.0, 0.0, -3.0), (0.1, 0.5, 0.6)),
    ((8.0, -5.0, 10.0), (1.2, 0.2, 0.3)),
]


def pack_mutable_data(materials=DEFAULT_MATERIALS, objects=DEFAULT_OBJECTS, lights=DEFAULT_LIGHTS) -> bytes:
    """The 656-byte std140 image DeviceLocalBuffer::from_data uploads (main.rs:592-605)."""
    if len(materials) > MAX_MATERIALS or len(objects) > MAX_OBJECTS or len(lights) > MAX_LIGHTS:
        raise ValueError("MutableData holds at most 8 materials / objects / lights (utilities.glsl:1-3)")
    md = np.zeros((), dtype=MUTABLE_DATA_DTYPE)
    md["matCount"], md["objCount"], md["lightCount"] = len(materials), len(objects), len(lights)
    for i, (color, diffuse, specular, shine, ambient) in enumerate(materials):
        md["mats"][i]["color"] = color
        md["mats"][i]["diffuse"], md["mats"][i]["specular"] = diffuse, specular
        md["mats"][i]["shine"], md["mats"][i]["ambient"] = shine, ambient
    for i, (pos, size) in enumerate(objects):
        md["objs"][i]["pos"], md["objs"][i]["size"] = pos, size
    for i, (pos, color) in enumerate(lights):
        md["lights"][i]["pos"], md["lights"][i]["color"] = pos, color
    return md.tobytes()


def default_mutable_data() -> bytes:
    return pack_mutable_data()


def spheres_of(mutable_data: bytes) -> np.ndarray:
    """objs[0:objCount] of a MutableData image as an (n, 4) float32 array {x, y, z, size}."""
    md = np.frombuffer(mutable_data, dtype=MUTABLE_DATA_DTYPE)[0]
    n = int(md["objCount"])
    out = np.zeros((n, 4), dtype=np.float32)
    out[:, :3] = md["objs"]["pos"][:n]
    out[:, 3] = md["objs"]["size"][:n]
    return out


def view_ratio(view_w: float, view_h: float, fov: float = FOV):
    """main.rs:608-611: ratio = [FOV, FOV * H / W], every operation in f32."""
    w, h, f = np.float32(view_w), np.float32(view_h), np.float32(fov)
    return np.array([f, np.float32(np.float32(f * h) / w)], dtype=np.float32)


def _splitmix64(state: int):
    state = (state + 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
    z = state
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
    return state, z ^ (z >> 31)


def synthetic_spheres(n: int = 10000, seed: int = 0xB200, box=((-60.0, 60.0), (10.0, 130.0), (-40.0, 40.0)),
                      radius=(0.2, 1.0)) -> np.ndarray:
    """BASELINE.json configs[3] synthetic scene: SplitMix64(seed), four draws per sphere
    u = (z >> 40) / 2^24 -> x, y, z uniform in `box`, size uniform in `radius`; float32."""
    out = np.zeros((n, 4), dtype=np.float32)
    s = seed
    lo = [box[0][0], box[1][0], box[2][0], radius[0]]
    span = [box[0][1] - box[0][0], box[1][1] - box[1][0], box[2][1] - box[2][0], radius[1] - radius[0]]
    for i in range(n):
        for k in range(4):
            s, z = _splitmix64(s)
            u = (z >> 40) / float(1 << 24)
            out[i, k] = np.float32(lo[k] + span[k] * u)
    return out

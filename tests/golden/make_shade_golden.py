"""Mint the golden fixtures of the presentation shading pass from the reference itself
(oracle/_ref/librtce_ref_frag.so = the reference's own fragment.glsl + utilities.glsl compiled as C++,
fed with the depth image of oracle/_ref/librtce_ref.so; needs /root/reference at BUILD time only).

    python tests/golden/make_shade_golden.py   ->  tests/golden/shade_golden.json

Per case: SHA-256 of the RGBA32F bytes and of the RGBA8 bytes, counters, mean colour.
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import oracle  # noqa: E402
from raytracing_engine_b200 import scene  # noqa: E402
from shade_cases import shade_cases  # noqa: E402


def main():
    oracle.build()
    ref = oracle.load("ref")
    out = {}
    for name, (md, W, H, rot, pos) in shade_cases().items():
        ratio = scene.view_ratio(W, H)
        lv, _ = ref.trace(scene.spheres_of(md), [W, H], ratio, rot, pos, 1000.0, 0)
        rgba, cnt = ref.shade(md, [W, H], ratio, rot, pos, lv[-1])
        out[name] = {
            "sha256_rgba32f": oracle.sha256(rgba),
            "sha256_rgba8": hashlib.sha256(oracle.rgba8(rgba).tobytes()).hexdigest(),
            "counters": cnt,
            "mean_rgb": [float(rgba[..., k].mean()) for k in range(3)],
            "max_rgb": float(rgba[..., :3].max()),
        }
        print(name, out[name]["sha256_rgba32f"][:16], cnt)
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "shade_golden.json"), "w") as fh:
        json.dump(out, fh, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()

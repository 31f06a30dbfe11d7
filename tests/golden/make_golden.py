"""Mint the golden fixtures from the reference itself (oracle/_ref = the reference's own
compute.glsl + utilities.glsl compiled as C++; needs /root/reference at BUILD time only).

    python tests/golden/make_golden.py        ->  tests/golden/golden.json, tests/golden/images.npz

golden.json: per case, per level SHA-256 of the float32 bytes, the FNV-1a-32 of the final level
(cross-check with SURVEY.md §8c), work counters R/E/S, min / max / centre / hit-count of the
final level.  images.npz: all levels of the small cases, for element-wise comparison.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import oracle  # noqa: E402
from raytracing_engine_b200 import scene  # noqa: E402
from cases import cases  # noqa: E402


def main():
    oracle.build()
    out, imgs = {}, {}
    for (name, kind, sph, W, H, levels, rot, pos, keep) in cases():
        o = oracle.load(kind)
        n = o.level_count(float(W), resize_rule=(levels == -1)) if levels <= 0 else levels
        ratio = scene.view_ratio(W, H)
        lv, cnt = o.trace(sph, [W, H], ratio, rot, pos, render_dist=1000.0, levels=n)
        f = lv[-1]
        out[name] = {
            "kind": kind, "levels": n, "dims": [[int(a.shape[1]), int(a.shape[0])] for a in lv],
            "sha256": [oracle.sha256(a) for a in lv], "fnv1a32_final": oracle.fnv1a32(f),
            "counters": cnt, "min": float(f.min()), "max": float(f.max()),
            "centre": float(f[f.shape[0] // 2, f.shape[1] // 2]), "hits": int((f < 1000.0).sum()),
        }
        if keep:
            for i, a in enumerate(lv):
                imgs[f"{name}/{i}"] = a
        print(name, out[name]["fnv1a32_final"], cnt)
    here = os.path.dirname(os.path.abspath(__file__))
    with open(os.path.join(here, "golden.json"), "w") as fh:
        json.dump(out, fh, indent=1, sort_keys=True)
    np.savez_compressed(os.path.join(here, "images.npz"), **imgs)


if __name__ == "__main__":
    main()

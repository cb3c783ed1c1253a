"""Regenerates tests/golden/oracle_digests.json and tests/golden/c1_small.npz from the CPU oracle.

The reference has no fixtures of its own and cannot run here, so these goldens
are OUTPUTS OF THE ORACLE (oracle/pt_oracle.c), frozen so that (1) a change to the oracle or
its compiler flags is caught, (2) the GPU box's own oracle build is checked against this
container's, and (3) the CUDA path can be checked without running the oracle.  The oracle
itself is pinned to SURVEY.md Appendix B (tests/golden/appendix_b.json, transcribed from the
survey, not generated here).

Run from the repo root:  python tests/golden/make_golden.py
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pt_oracle as O  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


# (name, width, height, uv_div or None(default H,H), extra kwargs)
FRAME_CASES = [
    ("c1_320x180_div180", 320, 180, None, {}),
    ("c1_320x180_div1080_literal", 320, 180, (1080, 1080), {}),
    ("c2_1920x1080", 1920, 1080, None, {}),
    ("dispatch_1920x1088_div1080", 1920, 1088, (1080, 1080), {}),
    ("4k_3840x2160", 3840, 2160, None, {}),
    ("ragged_333x77", 333, 77, None, {}),
    ("ragged_1x1", 1, 1, None, {}),
    ("ragged_5x3", 5, 3, (3, 3), {}),
    ("moved_sphere_640x360", 640, 360, None, {"cam": (0.1, -0.2, 1.5), "centre": (0.3, 0.1, -0.2), "radius": 0.7}),
]
ACCUM_CASES = [
    # name, w, h, first, n, seed, jitter
    ("acc_320x180_s0_n4_seed0", 320, 180, 0, 4, 0, True),
    ("acc_320x180_s4_n3_seed7", 320, 180, 4, 3, 7, True),
    ("acc_333x77_s0_n5_seed1", 333, 77, 0, 5, 1, True),
    ("acc_320x180_nojitter_n2", 320, 180, 0, 2, 0, False),
]


def main():
    out = {"frames": {}, "accum": {}, "rng": {}, "present": {}}
    for name, w, h, div, kw in FRAME_CASES:
        p = O.frame_params(w, h, uv_div=div, **kw)
        r = O.csmain_frame(p)
        out["frames"][name] = {
            "width": w, "height": h, "uv_div": list(div) if div else None, "params": {k: list(v) if isinstance(v, tuple) else v for k, v in kw.items()},
            "hits": r["hits"], "f32": sha(r["f32"]), "rgba8": sha(r["rgba8"]), "hit": sha(r["hit"]),
        }
    for name, w, h, first, n, seed, jit in ACCUM_CASES:
        p = O.frame_params(w, h)
        acc = O.accumulate(p, first, n, seed, jit)
        out["accum"][name] = {
            "width": w, "height": h, "first": first, "n": n, "seed": seed, "jitter": jit,
            "accum": sha(acc),
            "resolve": sha(O.resolve(acc, n)),
            "resolve_srgb_bgra": sha(O.resolve(acc, n, srgb=True, bgra=True)),
        }
    keys = [(0, 0, 0), (1, 2, 3), (12345, 6, 42), (2073599, 63, 0), (0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF)]
    out["rng"] = {"pcg3d": [{"key": list(k), "out": list(O.pcg3d(*k)),
                             "unit_bits": [int(O.bits_to_unit(w).view(np.uint32)) for w in O.pcg3d(*k)]} for k in keys]}
    p = O.frame_params(320, 180)
    img = O.csmain_frame(p)["rgba8"]
    out["present"] = {
        "src": "c1_320x180_div180",
        "same_size_srgb": sha(O.present(img, 320, 180, srgb=True)),
        "same_size_linear": sha(O.present(img, 320, 180, srgb=False)),
        "up_500x333_srgb": sha(O.present(img, 500, 333, srgb=True)),
        "down_97x41_linear": sha(O.present(img, 97, 41, srgb=False)),
    }
    with open(os.path.join(HERE, "oracle_digests.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    # a tiny raw fixture (64x36, divisor 36): float4 image, RGBA8 image, hit ids
    p = O.frame_params(64, 36)
    r = O.csmain_frame(p)
    np.savez_compressed(os.path.join(HERE, "c1_small.npz"), f32=r["f32"], rgba8=r["rgba8"], hit=r["hit"])
    print("wrote oracle_digests.json, c1_small.npz")


if __name__ == "__main__":
    main()

"""Freezes oracle outputs of the scene path as digests (tests/golden/scene_digests.json).

The scene stages have no reference counterpart (SURVEY.md §8 a'), so these are NOT reference vectors:
they pin this repo's own frozen spec (oracle/pt_scene_oracle.h) against silent drift.
Run:  python tests/golden/make_scene_golden.py
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import pt_scene_oracle as S          # noqa: E402
import importlib.util   # noqa: E402
_spec = importlib.util.spec_from_file_location("test_scene_oracle", os.path.join(os.path.dirname(HERE), "test_scene_oracle.py"))
_mod = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(_mod)
CASES, digest = _mod.CASES, _mod.digest

out = {}
for name, (make, w, h, depth) in CASES.items():
    so = S.SceneOracle(make(w, h), w, h, max_depth=depth)
    fr = so.frame(seed=1)
    out[name] = {"hit": digest(fr["hit"]), "f32": digest(fr["f32"]), "accum2": digest(so.accumulate(0, 2, seed=1))}
json.dump(out, open(os.path.join(HERE, "scene_digests.json"), "w"), indent=1, sort_keys=True)
print(json.dumps(out, indent=1))

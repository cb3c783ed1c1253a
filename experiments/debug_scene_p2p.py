import os, sys
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pathtracing_b200 import NativeTracer, scenes
from pathtracing_b200.dist import sample_range
from oracle import pt_oracle as O, pt_scene_oracle as S

def run(world, sw, sh, depth, frames, seed):
    sc = scenes.mixed_small(width=sw, height=sh)
    so = S.SceneOracle(sc, sw, sh, max_depth=depth)
    ranks = [NativeTracer(sw, sh, seed=seed, max_depth=depth) for _ in range(world)]
    for t in ranks:
        t.upload_scene(sc)
        t.peer_group_create()
    ptrs = [t.peer_group_local() for t in ranks]
    for r, t in enumerate(ranks):
        t.peer_group_bind(r, 0, [p[0] for p in ptrs], [p[3] for p in ptrs], [p[2] for p in ptrs], ptrs[0][1], ptrs[0][4])
    total = np.zeros((sh, sw, 4), np.float32); spp = 0
    for first, count in frames:
        spp += count
        for r, t in enumerate(ranks):
            t.peer_frame(*sample_range(first, count, r, world), spp)
        parts = [so.accumulate(*sample_range(first, count, r, world), seed=seed) for r in range(world)]
        s = parts[0].copy()
        for q in parts[1:]: s = s + q
        total = total + s
    for t in ranks: t.wait()
    img = ranks[0].read_rgba8()
    ref = O.resolve(total, spp)
    got = np.zeros((sh * sw, 4), np.float32)
    for t in ranks:
        f, sl = t.peer_read_total(); got[f:f + len(sl)] = sl
    got = got.reshape(sh, sw, 4)
    bad = np.argwhere((got.view(np.uint32) != total.view(np.uint32)).any(axis=-1))
    print(world, sw, sh, depth, frames, "img mismatches", int((img != ref).any(axis=-1).sum()), "accum mismatches", len(bad), bad[:5].tolist())
    if len(bad):
        y, x = bad[0]; print(" got", got[y, x], "want", total[y, x])
    for t in ranks: t.destroy()

run(2, 192, 108, 6, [(0, 9), (9, 6)], 2)
run(2, 192, 108, 6, [(0, 9)], 2)
run(3, 160, 90, 5, [(0, 7), (7, 5)], 2)
run(2, 160, 90, 5, [(0, 9), (9, 6)], 2)

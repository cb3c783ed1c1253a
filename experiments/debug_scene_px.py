import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pathtracing_b200 import NativeTracer, scenes
from oracle import pt_scene_oracle as S
sw, sh, depth, seed = 192, 108, 6, 2
sc = scenes.mixed_small(width=sw, height=sh)
so = S.SceneOracle(sc, sw, sh, max_depth=depth)
with NativeTracer(sw, sh, seed=seed, max_depth=depth) as t:
    t.upload_scene(sc)
    for (f, n) in ((0, 5), (5, 4), (9, 3), (12, 3), (0, 15)):
        for ws in (0, 1):
            t.accum_reset(); t.set_wave_samples(ws)
            t.render_samples(f, n)
            g = t.read_accum()
            r = so.accumulate(f, n, seed=seed)
            rb = so.accumulate(f, n, seed=seed, brute=True)
            bad = np.argwhere((g.view(np.uint32) != r.view(np.uint32)).any(axis=-1))
            badb = np.argwhere((rb.view(np.uint32) != r.view(np.uint32)).any(axis=-1))
            print(f"samples [{f},{f+n}) wave_samples={ws}: gpu vs oracle {len(bad)} px {bad[:3].tolist()}  oracle bvh vs brute {len(badb)}")
    for s in range(15):
        t.accum_reset(); t.set_wave_samples(0)
        t.render_samples(s, 1)
        g = t.read_accum(); r = so.accumulate(s, 1, seed=seed)
        if not np.array_equal(g.view(np.uint32), r.view(np.uint32)):
            bad = np.argwhere((g.view(np.uint32) != r.view(np.uint32)).any(axis=-1))
            y, x = bad[0]
            print("sample", s, "differs at", bad.tolist()[:4], "gpu", g[y, x], "oracle", r[y, x], "brute", so.accumulate(s, 1, seed=seed, brute=True)[y, x])

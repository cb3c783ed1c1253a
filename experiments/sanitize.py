"""Small pass over every kernel family, meant to be run under compute-sanitizer on the GPU box:
   compute-sanitizer --tool memcheck|racecheck|initcheck|synccheck python experiments/sanitize.py
(compute-sanitizer is closed on the shared GPU pool, where this only serves as a functional pass.)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pathtracing_b200 import NativeTracer, scenes

w, h = 200, 77
with NativeTracer(w, h, seed=3, keep_f32=True, keep_hit_ids=True) as t:
    t.compute_frame(); t.wait()
    t.render_samples(0, 5); t.render_samples(5, 2); t.resolve(); t.wait()
    img = t.read_rgba8(); acc = t.read_accum()
    t.set_row_shard(1, 3, 4); t.accum_reset(); t.render_samples(0, 3); t.resolve(); t.wait()
    t.set_row_shard(0, 1)
    t.write_accum(acc, 7); t.resolve(); t.wait()
    assert np.array_equal(t.read_rgba8(), img)
    t.present(); t.wait(); t.read_present()
    print("csmain ok", flush=True)

for name, sc, depth in (("mixed", scenes.mixed_small(width=w, height=h), 6),
                        ("spheres", scenes.random_spheres(60, width=w, height=h), 8),
                        ("terrain", scenes.triangle_terrain(cells=12, layers=2, width=w, height=h), 4),
                        ("default", scenes.default_scene(), 0)):
    with NativeTracer(w, h, seed=1, max_depth=depth, keep_f32=True, keep_hit_ids=True) as t:
        t.upload_scene(sc, count_stats=(name == "mixed"))
        t.compute_frame(); t.wait()
        for per_wave in (0, 1, 3):
            t.set_wave_samples(per_wave)
            t.render_samples(0, 5)
        t.set_row_shard(2, 3, 4); t.render_samples(5, 4); t.set_row_shard(0, 1)
        t.resolve(); t.wait()
        rng = np.random.Generator(np.random.PCG64(2))
        o = rng.uniform(-4, 4, (3000, 3)).astype(np.float32); d = rng.normal(size=(3000, 3)).astype(np.float32)
        t.trace_rays(o, d, 1e-3, 1e30); t.trace_rays(o, d, 1e-3, 5.0, any_hit=True)
        t.read_bvh(0); t.read_bvh(1); t.scene_queue_sizes()
        print(name, "ok", t.scene_info()["triangle_nodes"], flush=True)

# fused peer merge, one process playing two ranks
with NativeTracer(w, h, seed=0) as a, NativeTracer(w, h, seed=0) as b:
    a.render_samples(0, 2); b.render_samples(2, 2); a.wait(); b.wait()
    n = w * h
    ptrs = [a.device_accum(), b.device_accum()]
    a.peer_merge(ptrs, a.device_rgba8(), 0, n // 2, 4); b.peer_merge(ptrs, a.device_rgba8(), n // 2, n - n // 2, 4)
    a.wait(); b.wait()
    print("peer merge ok")

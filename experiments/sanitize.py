"""Small pass over every kernel family, meant to be run under compute-sanitizer on the GPU box:
   compute-sanitizer --tool memcheck|racecheck|initcheck|synccheck python experiments/sanitize.py
Kernel launches are serialised under the sanitizer, so the cross-GPU rendezvous (a kernel that spins until its peers' kernels
have run) is left out; the peer MERGE kernel is driven directly."""
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
    t.accum_reset(); t.render_resolve(0, 6, 6); t.render_resolve(6, 3, 9); t.wait(); t.read_rgba8()
    t.compute_frames(3); t.wait(); t.read_rgba8()
    print("csmain ok", flush=True)
with NativeTracer(w, h, seed=3, spirv_arith=True) as t:
    t.compute_frame(); t.render_samples(0, 3); t.resolve(); t.wait(); t.read_rgba8()
    print("spirv arithmetic ok", flush=True)

for name, sc, depth in (("mixed", scenes.mixed_small(width=w, height=h), 6),
                        ("spheres", scenes.random_spheres(60, width=w, height=h), 8),
                        ("terrain", scenes.triangle_terrain(cells=12, layers=2, width=w, height=h), 4),
                        ("default", scenes.default_scene(), 0)):
    with NativeTracer(w, h, seed=1, max_depth=depth, keep_f32=True, keep_hit_ids=True) as t:
        t.upload_scene(sc, count_stats=(name == "mixed"))
        t.compute_frame(); t.wait()
        for per_wave in (0, 1, 3):
            t.set_wave_samples(per_wave)
            for mode in ("chunk", "queue", "tail"):      # every bounce schedule (kernels_scene.cu launch_scene_wave)
                os.environ["PTB200_BOUNCE"] = mode
                t.render_samples(0, 5)
        os.environ["PTB200_D0_QUEUE"] = "1"; os.environ["PTB200_BOUNCE"] = "queue"; os.environ["PTB200_REFILL"] = "24"
        t.render_samples(5, 3)
        for k in ("PTB200_D0_QUEUE", "PTB200_BOUNCE", "PTB200_REFILL"):
            del os.environ[k]
        t.set_row_shard(2, 3, 4); t.render_samples(5, 4); t.set_row_shard(0, 1)
        if depth:                                        # next-event estimation: shadow rays in both schedules
            t.set_sun([0.35, 1.0, 0.25], [2.0, 1.75, 1.5])
            for mode in ("chunk", "queue"):
                os.environ["PTB200_BOUNCE"] = mode
                t.render_samples(9, 3)
            del os.environ["PTB200_BOUNCE"]
            t.set_sun([0, 1, 0], [0, 0, 0])
        t.resolve(); t.wait()
        rng = np.random.Generator(np.random.PCG64(2))
        o = rng.uniform(-4, 4, (3000, 3)).astype(np.float32); d = rng.normal(size=(3000, 3)).astype(np.float32)
        t.trace_rays(o, d, 1e-3, 1e30); t.trace_rays(o, d, 1e-3, 5.0, any_hit=True)
        t.trace_rays_static(o, d, 1e-3, 1e30); t.trace_rays_static(o, d, 1e-3, 5.0, any_hit=True)   # static device staging
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

"""Quick device timing of the scene workloads (development aid; bench.py --workload c3|c4 is the contract)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pathtracing_b200 import NativeTracer, scenes

def run(name, sc, w, h, spp, depth, max_leaf=4, per_wave=0, shard=None):
    with NativeTracer(w, h, max_depth=depth) as t:
        t0 = time.time(); t.upload_scene(sc, max_leaf=max_leaf); up = time.time() - t0
        t.set_wave_samples(per_wave)
        if shard:
            t.set_row_shard(*shard)
            name += f" shard{shard}"
        name += f" per_wave={per_wave}"
        info = t.scene_info()
        t.render_samples(0, spp); t.wait()
        t.timer_begin(); t.render_samples(2, spp); ms = t.timer_end()
        print("   queue sizes", t.scene_queue_sizes(), flush=True)
        frac = 1.0 / shard[1] if shard else 1.0
        print(f"{name}: {w}x{h} {spp}spp depth{depth} leaf{max_leaf}: {ms:.2f} ms -> {frac*w*h*spp/ms/1e3:.1f} Msamples/s; upload {up*1e3:.0f} ms build {info['build_ms']:.2f} ms nodes s{info['sphere_nodes']} t{info['triangle_nodes']} leaves t{info['triangle_leaves']}", flush=True)
    with NativeTracer(w, h, max_depth=depth) as t:
        t.upload_scene(sc, max_leaf=max_leaf, count_stats=True)
        t.render_samples(0, 2); t.wait()
        st = t.scene_stats()
        print("   per path: rays %.2f nodes %.1f prims %.1f scatters %.2f max_stack %d" % (st['rays']/st['paths'], st['nodes_visited']/st['paths'], st['prims_tested']/st['paths'], st['scatters']/st['paths'], st['max_stack']), flush=True)

which = sys.argv[1] if len(sys.argv) > 1 else "all"
if which == "waves":
    sc = scenes.random_spheres(500, width=1920, height=1080)
    for pw in (1, 2, 4, 8, 16):
        run("c3", sc, 1920, 1080, 32, 8, 4, pw)
    sc = scenes.triangle_terrain(512, 2, width=3840, height=2160)
    for pw in (1, 2, 4):
        run("c4", sc, 3840, 2160, 8, 8, 4, pw)
    for pw in (1, 4, 8, 16, 32):
        run("c4", sc, 3840, 2160, 32, 8, 4, pw, shard=(3, 8, 4))
    sys.exit(0)
if which in ("c3", "all"):
    sc = scenes.random_spheres(500, width=1920, height=1080)
    for ml in (1, 4, 8):
        run("c3", sc, 1920, 1080, 8, 8, ml)
if which in ("c4", "all"):
    sc = scenes.triangle_terrain(512, 2, width=3840, height=2160)
    for ml in (1, 4, 8):
        run("c4", sc, 3840, 2160, 4, 8, ml)
if which == "leaf8":
    run("c3", scenes.random_spheres(500, width=1920, height=1080), 1920, 1080, 8, 8, 8)
    run("c4", scenes.triangle_terrain(512, 2, width=3840, height=2160), 3840, 2160, 4, 8, 8)
if which == "ab":
    # one timing per workload under the current PTB200_* switches, plus a digest of the accumulation buffer so that two
    # runs under different switches can be compared bit for bit
    import hashlib
    for nm, sc, w, h, spp in (("c3", scenes.random_spheres(500, seed=1, width=1920, height=1080), 1920, 1080, 64),
                              ("c4", scenes.triangle_terrain(512, 2, seed=2, width=3840, height=2160), 3840, 2160, 16)):
        with NativeTracer(w, h, max_depth=8) as t:
            t.upload_scene(sc)
            t.render_samples(0, spp); t.wait()
            dig = hashlib.sha256(t.read_accum().tobytes()).hexdigest()[:16]
            best = 1e9
            for rep in range(3):
                t.timer_begin(); t.render_samples(spp * (rep + 1), spp); best = min(best, t.timer_end())
            tag = " ".join(f"{k}={v}" for k, v in sorted(os.environ.items()) if k.startswith("PTB200_"))
            print(f"AB {nm} [{tag}] {best:.2f} ms  {w*h*spp/best/1e3:.0f} Msamples/s  accum {dig}", flush=True)

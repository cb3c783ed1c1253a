"""ncu target for the bounce kernels: one warm-up wave and one profiled wave per schedule, in ONE process.
   python experiments/prof_scene.py c3|c4 [samples per wave]
Launch order: [chunk warm-up][chunk][default warm-up][default].  Bounce launches per wave: chunk schedule d0..d3 + tail (5), or
d0..d5 + tail (7) for waves of >= 8 Mi camera rays; default d0..d2 + the queue-schedule tail (4)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pathtracing_b200 import NativeTracer, scenes

wl = sys.argv[1] if len(sys.argv) > 1 else "c3"
pw = int(sys.argv[2]) if len(sys.argv) > 2 else (4 if wl == "c3" else 1)
if wl == "c3":
    sc, w, h = scenes.random_spheres(500, seed=1, width=1920, height=1080), 1920, 1080
else:
    sc, w, h = scenes.triangle_terrain(512, 2, seed=2, width=3840, height=2160), 3840, 2160
with NativeTracer(w, h, max_depth=8) as t:
    t.upload_scene(sc)
    t.set_wave_samples(pw)
    for mode in ("chunk", "default"):            # every launch with the chunk schedule / chunk + queue-schedule tail (default)
        if mode == "chunk":
            os.environ["PTB200_BOUNCE"] = mode
        else:
            os.environ.pop("PTB200_BOUNCE", None)
        for rep in range(2):
            t.timer_begin(); t.render_samples(rep * pw, pw); ms = t.timer_end()
            print(f"{wl} {mode} wave of {pw}: {ms:.2f} ms, queues {t.scene_queue_sizes()}", flush=True)

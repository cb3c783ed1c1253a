"""c3 (sphere scene) at several wave sizes under the current PTB200_* switches: device ms per 32 spp (development aid)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pathtracing_b200 import NativeTracer, scenes
sc = scenes.random_spheres(500, seed=1, width=1920, height=1080)
tag = " ".join(f"{k}={v}" for k, v in sorted(os.environ.items()) if k.startswith("PTB200_"))
with NativeTracer(1920, 1080, max_depth=8) as t:
    t.upload_scene(sc)
    for pw in (1, 2, 4, 8, 32):
        t.set_wave_samples(pw)
        t.render_samples(0, 32); t.wait()
        best = 1e9
        for rep in range(3):
            t.timer_begin(); t.render_samples(32 * (rep + 1), 32); best = min(best, t.timer_end())
        print(f"C3W [{tag}] per_wave={pw}: {best:.2f} ms  {1920*1080*32/best/1e3:.0f} Msamples/s", flush=True)

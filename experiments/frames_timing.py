"""Timing of the literal reference frame: per-frame launches vs ptb_compute_frames (one persistent launch)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pathtracing_b200 import NativeTracer

for (w, h) in ((1920, 1080), (3840, 2160)):
    with NativeTracer(w, h) as t:
        for _ in range(5):
            t.compute_frame()
        t.compute_frames(5)
        t.wait()
        for n in (1, 10, 200, 2000):
            t.timer_begin()
            for _ in range(n):
                t.compute_frame()
            a = t.timer_end() / n
            t.timer_begin()
            t.compute_frames(n)
            b = t.timer_end() / n
            print(f"{w}x{h} n={n}: per-frame launches {a * 1e3:.2f} us/frame, persistent {b * 1e3:.2f} us/frame")

"""Launch script for ncu captures of the literal-scene kernels at 1920x1080 (profiles/README.md)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pathtracing_b200 import NativeTracer

with NativeTracer(1920, 1080, seed=0) as t:
    for s in range(4):
        t.render_resolve(s * 64, 64)          # the benched launch: csmain_accumulate<2,1> with the fused resolve
    t.render_samples(256, 64)                 # the same without the resolve
    t.resolve()                               # resolve_kernel
    t.resolve(srgb=True, bgra=True)
    for _ in range(3):
        t.compute_frame()                     # csmain_frame_vec4
    for _ in range(3):
        t.compute_frames(200)                 # csmain_frames_persistent
    t.present(1920, 1080, srgb=True)          # present_kernel
    t.wait()
print("ok")

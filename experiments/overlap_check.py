"""1-GPU check: does a peer-merge kernel on a second stream overlap the next frame's render kernel?"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pathtracing_b200 import NativeTracer

W, H, SPP, K = 3840, 2160, 64, 20
dev = torch.device("cuda", 0)
M = torch.cuda.Stream(device=dev)
for prio in (0, -1):
    X = torch.cuda.Stream(device=dev, priority=prio)
    main = NativeTracer(W, H, stream=M.cuda_stream)
    merger = NativeTracer(W, H, stream=X.cuda_stream)
    bufs = [main.device_accum(), merger.device_accum()]
    main.accum_reset(); merger.accum_reset()
    img = main.device_rgba8()

    def run(mode):
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(M)
        for k in range(K):
            bsel = k & 1
            main.set_accum_buffer(None if bsel == 0 else bufs[1])
            if bsel == 0:
                main.device_accum()
            main.render_samples(k * SPP, SPP)
            if mode == "render":
                continue
            ev = torch.cuda.Event(); ev.record(M)
            stream = X if mode == "overlap" else M
            if mode == "overlap":
                X.wait_event(ev)
            t = merger if mode == "overlap" else main
            t.peer_merge([bufs[bsel]], img, 0, W * H, (k + 1) * SPP)
        if mode == "overlap":
            fin = torch.cuda.Event(); fin.record(X); M.wait_event(fin)
        b.record(M)
        torch.cuda.synchronize()
        return a.elapsed_time(b) / K

    for mode in ("render", "serial", "overlap"):
        run(mode)
        print(f"priority {prio:2d} {mode:8s} {run(mode):.3f} ms/frame", flush=True)
    main.destroy(); merger.destroy()

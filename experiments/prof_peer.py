"""ncu target for the fused peer merge: ONE process, one context per visible GPU, pipelined ptb_peer_frame loops.
   python experiments/prof_peer.py [width height spp_per_rank frames]
Defaults: the per-rank share of c2 (1920x1080, 64 spp per rank and frame).  c5's share on G GPUs: 3840 2160 1024/G."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pathtracing_b200 import NativeTracer

w, h, spp, frames = (int(v) for v in (sys.argv[1:5] + ["1920", "1080", "64", "6"][len(sys.argv) - 1:]))
G = torch.cuda.device_count()
ranks = [NativeTracer(w, h, device=d, seed=0) for d in range(G)]
for t in ranks:
    t.peer_group_create()
ptrs = [t.peer_group_local() for t in ranks]
for r, t in enumerate(ranks):
    t.peer_group_bind(r, 0, [q[0] for q in ptrs], [q[3] for q in ptrs], [q[2] for q in ptrs], ptrs[0][1], ptrs[0][4])
direct = os.environ.get("PROF_PEER_DIRECT")      # under ncu: kernel launches are serialised, so the rendezvous kernels (which spin
                                                 # until the peer's kernel has run) would time out; drive the MERGE kernel directly
if direct:
    n = w * h
    for f in range(frames):
        for r, t in enumerate(ranks):
            t.render_samples((f * G + r) * spp, spp)
        for t in ranks:
            t.wait()
        for r, t in enumerate(ranks):            # rank r's slab of every partial -> resolved pixels into rank 0's image
            first, count = r * n // G, (r + 1) * n // G - r * n // G
            t.peer_merge([x.device_accum() for x in ranks], ranks[0].device_rgba8(), first, count, (f + 1) * G * spp)
        for t in ranks:
            t.wait()
    print(f"{G} GPUs, direct merges ok", flush=True)
for rep in range(0 if direct else 2):
    t0 = time.perf_counter()
    for f in range(frames):
        for r, t in enumerate(ranks):
            t.peer_frame((f * G + r) * spp, spp, (f + 1) * G * spp + rep * frames * G * spp)
    for t in ranks:
        t.wait()
    print(f"{G} GPUs, {w}x{h}, {spp} spp per rank and frame: {(time.perf_counter() - t0) / frames * 1e3:.3f} ms per frame (host clock, pass {rep})", flush=True)
for t in ranks:
    t.destroy()

"""torchrun check (N >= 2 GPUs): the fused peer-memory merge gives the same image as the NCCL reduce path for 2 ranks
(bit-identical: a+b) and an image within 1 code for more ranks; prints timings of both merges.
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 experiments/p2p_check.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from pathtracing_b200.dist import GpuEngine, ShardedRenderer

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
W, H, SPP = 3840, 2160, 64 * world
stream = torch.cuda.Stream()
imgs, times = {}, {}
with torch.cuda.stream(stream):
    for mode in ("samples", "p2p"):
        eng = GpuEngine(W, H, local, seed=0, stream=stream, peer_memory=(mode == "p2p"))
        sr = ShardedRenderer(eng, mode=mode)
        for it in range(4):
            sr.render(it * SPP, SPP)
            torch.cuda.synchronize(); dist.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            img = sr.merge(0)
            b.record(stream)
            torch.cuda.synchronize()
            t = torch.tensor([a.elapsed_time(b)], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
            times.setdefault(mode, []).append(float(t))
        if rank == 0:
            imgs[mode] = img.cpu().numpy().view(np.uint8).reshape(H, W, 4).copy()
        dist.barrier()
        eng.close()
if rank == 0:
    d = np.abs(imgs["samples"].astype(int) - imgs["p2p"].astype(int))
    print(f"world {world}: merge ms nccl-reduce+resolve {min(times['samples']):.3f}  fused p2p {min(times['p2p']):.3f}; "
          f"image max diff {d.max()} codes, differing bytes {int((d > 0).sum())}", flush=True)
    assert d.max() <= 1      # different association of the running total: at most one code
dist.destroy_process_group()

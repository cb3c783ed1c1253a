"""torchrun --nproc-per-node N experiments/peer_timing.py: where a pipelined peer frame's time goes (c2 per-rank share)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from pathtracing_b200.dist import GpuEngine, ShardedRenderer

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if os.environ.get("PTB200_NUMA_BIND"):
    import pynvml
    pynvml.nvmlInit()
    pr = torch.cuda.get_device_properties(local)
    h = pynvml.nvmlDeviceGetHandleByPciBusId(f"{pr.pci_domain_id:08x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0")
    before = len(os.sched_getaffinity(0))
    pynvml.nvmlDeviceSetCpuAffinity(h)
    print(f"rank {rank}: cpu affinity {before} -> {len(os.sched_getaffinity(0))} cores", flush=True)
dist.init_process_group("nccl", device_id=dev)
W, H, spp = 1920, 1080, 64
stream = torch.cuda.Stream(device=dev)
host = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in range(2)]
with torch.cuda.stream(stream):
    eng = GpuEngine(W, H, local, seed=0, stream=stream, peer_memory=True)
    sr = ShardedRenderer(eng, mode="p2p")
    t = eng.tracer
    def run(n, readback, label):
        dist.barrier(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        a.record(stream)
        for s in range(n):
            sr.frame(s * spp * world, spp * world, pipelined=True)
            if readback and rank == 0:
                t.read_rgba8_async(host[s & 1].data_ptr(), W * 4)
        t_issue = time.perf_counter() - t0
        sr.finish()
        b.record(stream)
        t.wait()
        dist.barrier(); torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        if rank == 0:
            print(f"{label}: n={n} device {a.elapsed_time(b) / n * 1e3:.1f} us/step, wall {wall / n * 1e6:.1f} us/step, host issue {t_issue / n * 1e6:.1f} us/step", flush=True)
    for n in (20, 200):
        run(n, False, "peer_frame only   ")
        run(n, True, "peer_frame + D2H  ")
    # render only, no merge
    dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for s in range(200):
        t.peer_frame(s * 64, 64, 0, render_only=True)
    b.record(stream)
    torch.cuda.synchronize()
    if rank == 0:
        print(f"render only: {a.elapsed_time(b) / 200 * 1e3:.1f} us/step", flush=True)
    t.peer_frame(0, 0, 64)
    sr.finish(); t.wait()
    eng.close()
dist.barrier()
dist.destroy_process_group()

"""Worker of tests/test_gpu_multi.py: one process per GPU (torchrun, NCCL), the multi-GPU paths of pathtracing_b200.dist
on REAL separate devices, every merged result compared with the CPU oracle on rank 0.  Prints MULTI_OK when all pass."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("PTB200_BARRIER_TIMEOUT_MS", "20000")

from pathtracing_b200 import scenes  # noqa: E402
from pathtracing_b200.dist import GpuEngine, ShardedRenderer, sample_range  # noqa: E402


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from oracle import pt_oracle as O
    from oracle import pt_scene_oracle as S
    O.set_threads(max(1, len(os.sched_getaffinity(0)) // world))
    w, h, seed = 640, 360, 4
    p = O.frame_params(w, h)
    stream = torch.cuda.Stream(device=dev)
    checks = []

    def rank_sum(parts):
        t = parts[0].copy()
        for q in parts[1:]:
            t = t + q
        return t

    with torch.cuda.stream(stream):
        # ---- p2p: native pipelined frame loop across processes / GPUs --------------------------------------------
        eng = GpuEngine(w, h, local, seed=seed, stream=stream, peer_memory=True)
        sr = ShardedRenderer(eng, mode="p2p")
        frames = [(0, 16), (16, 5), (21, 32), (53, 11)]
        bufs = [torch.zeros((h, w, 4), dtype=torch.uint8).pin_memory() for _ in frames]
        for (first, count), buf in zip(frames, bufs):
            sr.frame(first, count, pipelined=True)
            if rank == 0:
                eng.tracer.read_rgba8_async(buf.data_ptr())
        sr.finish()
        eng.tracer.wait()
        first_px, slab = eng.tracer.peer_read_total()
        slabs = [None] * world
        dist.all_gather_object(slabs, (first_px, slab))
        if rank == 0:
            total = np.zeros((h, w, 4), np.float32)
            spp = 0
            for (first, count), buf in zip(frames, bufs):
                parts = [O.accumulate(p, *sample_range(first, count, r, world), seed, True) for r in range(world)]
                total = total + rank_sum(parts)
                spp += count
                checks.append(("p2p frame %d" % first, np.array_equal(buf.numpy(), O.resolve(total, spp))))
            got = np.zeros((h * w, 4), np.float32)
            for f, s_ in slabs:
                got[f:f + len(s_)] = s_
            checks.append(("p2p merged accumulation", np.array_equal(bits(got.reshape(h, w, 4)), bits(total))))
            checks.append(("p2p sync read", np.array_equal(eng.tracer.read_rgba8(), bufs[-1].numpy())))
        eng.close()
        dist.barrier()

        # ---- tiles: bit-identical to one device ----------------------------------------------------------------------
        eng = GpuEngine(w, h, local, seed=seed, stream=stream)
        sr = ShardedRenderer(eng, mode="tiles")
        sr.render(0, 6)
        img = sr.merge(0)
        torch.cuda.synchronize()
        if rank == 0:
            ref = O.accumulate(p, 0, 6, seed, True)
            checks.append(("tiles", np.array_equal(img.cpu().numpy().view(np.uint8).reshape(h, w, 4), O.resolve(ref, 6))))
        eng.close()
        dist.barrier()

        # ---- samples: NCCL reduce(sum) + resolve on rank 0 --------------------------------------------------------------
        eng = GpuEngine(w, h, local, seed=seed, stream=stream)
        sr = ShardedRenderer(eng, mode="samples")
        sr.render(0, 10)
        img = sr.merge(0)
        torch.cuda.synchronize()
        if rank == 0:
            parts = [O.accumulate(p, *sample_range(0, 10, r, world), seed, True) for r in range(world)]
            acc = eng.accum.cpu().numpy()
            exact = rank_sum(parts)
            if world == 2:
                checks.append(("samples (2 ranks: a+b is order independent)", np.array_equal(bits(acc), bits(exact))))
            else:
                checks.append(("samples (reduction order free: <= 1e-6 rel)", np.allclose(acc, exact, rtol=1e-6, atol=1e-6)))
        eng.close()
        dist.barrier()

        # ---- scene path, p2p --------------------------------------------------------------------------------------------
        sw, sh, depth = 192, 108, 6
        sc = scenes.mixed_small(width=sw, height=sh)
        eng = GpuEngine(sw, sh, local, seed=2, stream=stream, scene=sc, max_depth=depth, peer_memory=True)
        sr = ShardedRenderer(eng, mode="p2p")
        sr.frame(0, 9, pipelined=True)
        sr.frame(9, 6, pipelined=True)
        sr.finish()
        eng.tracer.wait()
        first_px, slab = eng.tracer.peer_read_total()
        slabs = [None] * world
        dist.all_gather_object(slabs, (first_px, slab))
        if rank == 0:
            so = S.SceneOracle(sc, sw, sh, max_depth=depth)
            total = np.zeros((sh, sw, 4), np.float32)
            for first, count in ((0, 9), (9, 6)):
                total = total + rank_sum([so.accumulate(*sample_range(first, count, r, world), seed=2) for r in range(world)])
            got = np.zeros((sh * sw, 4), np.float32)
            for f, s_ in slabs:
                got[f:f + len(s_)] = s_
            got = got.reshape(sh, sw, 4)
            bad = np.argwhere((bits(got) != bits(total)).any(axis=-1))
            if len(bad):
                y, x = bad[0]
                print(f"scene p2p: {len(bad)} pixels of the merged accumulation differ, first ({x},{y}): got {got[y, x]} want {total[y, x]}",
                      flush=True)
            checks.append(("scene p2p merged accumulation", len(bad) == 0))
            checks.append(("scene p2p image", np.array_equal(eng.tracer.read_rgba8(), O.resolve(total, 15))))
        eng.close()
        dist.barrier()

    ok = True
    if rank == 0:
        for name, good in checks:
            print(("ok   " if good else "FAIL ") + name, flush=True)
            ok = ok and good
        print("MULTI_OK" if ok else "MULTI_FAIL", flush=True)
    dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())

"""world_size-2 (and 3) gloo tests of the multi-GPU host logic (pathtracing_b200/dist.py) on CPU.
The per-rank engine is backed by the CPU oracle (allowed here: tests/ is checker territory); the
sharding/merge code under test is exactly what runs over NCCL on the box."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from pathtracing_b200.dist import ShardedRenderer, row_shard, sample_range, shard_rows, slab_range  # noqa: E402

W, H, SEED = 96, 54, 5


class OracleEngine:
    """LocalEngine over the CPU oracle: same contract as dist.GpuEngine."""

    def __init__(self, width, height, seed):
        from oracle import pt_oracle as O
        self.O, self.width, self.height, self.seed = O, width, height, seed
        self.p = O.frame_params(width, height)
        self.acc = torch.zeros((height, width, 4), dtype=torch.float32)
        self.img = torch.zeros((height, width), dtype=torch.int32)
        self.rows = list(range(height))

    def set_row_shard(self, off, stride, band=1):
        self.rows = shard_rows(off, stride, band, self.height)

    def accum_reset(self):
        self.acc.zero_()

    def render_samples(self, first, n):
        part = self.O.accumulate(self.p, first, n, self.seed, True)          # whole frame, then keep own rows
        self.acc[self.rows] += torch.from_numpy(part)[self.rows]

    def accum_tensor(self):
        return self.acc

    def resolve(self, spp_total):
        img = self.O.resolve(self.acc.numpy(), spp_total)
        self.img.copy_(torch.from_numpy(img.view(np.int32).reshape(self.height, self.width)))

    def image_tensor(self):
        return self.img


class OracleSceneEngine(OracleEngine):
    """The same contract over the scene oracle (multi-bounce paths through a small mixed scene)."""

    def __init__(self, width, height, seed):
        super().__init__(width, height, seed)
        from oracle import pt_scene_oracle as S
        from pathtracing_b200 import scenes
        self.so = S.SceneOracle(scenes.mixed_small(width=width, height=height), width, height, max_depth=5)

        self.full = np.zeros((height, width, 4), np.float32)

    def accum_reset(self):
        super().accum_reset()
        self.full[:] = 0

    def render_samples(self, first, n):
        # the scene path adds every sample straight into the accumulator (no per-launch partial sum)
        self.so.accumulate(first, n, seed=self.seed, jitter=True, accum=self.full)
        self.acc[self.rows] = torch.from_numpy(self.full)[self.rows]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, mode, q, kind="csmain"):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        eng = OracleSceneEngine(W, H, SEED) if kind == "scene" else OracleEngine(W, H, SEED)
        r = ShardedRenderer(eng, mode=mode)
        r.render(0, 6)
        r.render(6, 5)                      # progressive: a second batch, odd size
        img = r.merge(dst=0)
        if rank == 0:
            q.put(("img", img.numpy().copy(), eng.acc.numpy().copy()))
        else:
            assert img is None
            q.put(("none", rank))
    finally:
        dist.destroy_process_group()


def _run(world, mode, kind="csmain"):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, mode, q, kind)) for r in range(world)]
    for p in procs:
        p.start()
    out = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return next(o for o in out if o[0] == "img")


def _single(oracle):
    p = oracle.frame_params(W, H)
    acc = oracle.accumulate(p, 0, 6, SEED, True)
    acc = oracle.accumulate(p, 6, 5, SEED, True, accum=acc)
    return acc, oracle.resolve(acc, 11)


def test_sample_range_partition():
    for first, count, world in [(0, 64, 8), (7, 5, 3), (0, 3, 8), (100, 1024, 4), (0, 0, 2)]:
        spans = [sample_range(first, count, r, world) for r in range(world)]
        covered = [s for st, n in spans for s in range(st, st + n)]
        assert covered == list(range(first, first + count))          # disjoint, ordered, complete
        assert max(n for _, n in spans) - min(n for _, n in spans) <= 1
    assert [row_shard(r, 4) for r in range(4)] == [(0, 4, 1), (1, 4, 1), (2, 4, 1), (3, 4, 1)]
    # row bands: every image row belongs to exactly one rank, bands of 4 rows interleaved, ragged last band
    for h, world, band in [(54, 3, 4), (2160, 8, 4), (7, 2, 4), (9, 4, 1), (5, 8, 4)]:
        rows = [shard_rows(*row_shard(r, world, band), h) for r in range(world)]
        assert sorted(y for rr in rows for y in rr) == list(range(h))
        assert all((y // band) % world == r for r, rr in enumerate(rows) for y in rr)
    # pixel slabs of the fused peer merge: contiguous, disjoint, complete
    for n, world in [(3840 * 2160, 8), (1920 * 1080, 3), (5, 8)]:
        slabs = [slab_range(r, world, n) for r in range(world)]
        assert slabs[0][0] == 0 and sum(c for _, c in slabs) == n
        assert all(slabs[i][0] + slabs[i][1] == slabs[i + 1][0] for i in range(world - 1))


def test_tile_sharding_two_ranks_is_bit_identical(oracle):
    _, img, _acc = _run(2, "tiles")
    acc1, img1 = _single(oracle)
    assert np.array_equal(img.view(np.uint8).reshape(H, W, 4), img1)


def test_tile_sharding_three_ranks_is_bit_identical(oracle):
    _, img, _acc = _run(3, "tiles")        # 54 rows / 3, interleaved
    _, img1 = _single(oracle)
    assert np.array_equal(img.view(np.uint8).reshape(H, W, 4), img1)


def test_sample_sharding_two_ranks(oracle):
    _, img, acc = _run(2, "samples")
    # the same per-rank launches run on ONE device in rank order give the same partial sums exactly
    p = oracle.frame_params(W, H)
    ref = np.zeros((H, W, 4), np.float32)
    parts = []
    for rank in range(2):
        a = np.zeros((H, W, 4), np.float32)
        for first, count in ((0, 6), (6, 5)):
            st, n = sample_range(first, count, rank, 2)
            a = oracle.accumulate(p, st, n, SEED, True, accum=a)
        parts.append(a)
    ref = parts[0] + parts[1]              # two operands: order-independent, so bit-exact
    assert np.array_equal(acc.view(np.uint32), ref.view(np.uint32))
    assert np.array_equal(img.view(np.uint8).reshape(H, W, 4), oracle.resolve(ref, 11))
    # and against one big sequential accumulation: float-sum reorder tolerance, <= 1 code in RGBA8
    acc1, img1 = _single(oracle)
    np.testing.assert_allclose(acc, acc1, rtol=2e-6, atol=1e-5)
    assert np.abs(img.view(np.uint8).reshape(H, W, 4).astype(int) - img1.astype(int)).max() <= 1


def _single_scene():
    from oracle import pt_oracle as O
    from oracle import pt_scene_oracle as S
    from pathtracing_b200 import scenes
    so = S.SceneOracle(scenes.mixed_small(width=W, height=H), W, H, max_depth=5)
    acc = so.accumulate(0, 11, seed=SEED)
    return so, acc, O.resolve(acc, 11)


def test_scene_tile_sharding_two_ranks_is_bit_identical():
    """n8 partitioning A on the scene path (BASELINE configs[3] shards image tiles across GPUs)."""
    _, img, _acc = _run(2, "tiles", "scene")
    _, _, img1 = _single_scene()
    assert np.array_equal(img.view(np.uint8).reshape(H, W, 4), img1)


def test_scene_sample_sharding_two_ranks():
    _, img, acc = _run(2, "samples", "scene")
    so, acc1, img1 = _single_scene()
    parts = []
    for rank in range(2):
        a = np.zeros((H, W, 4), np.float32)
        for first, count in ((0, 6), (6, 5)):
            st, n = sample_range(first, count, rank, 2)
            so.accumulate(st, n, seed=SEED, accum=a)
        parts.append(a)
    ref = parts[0] + parts[1]
    assert np.array_equal(acc.view(np.uint32), ref.view(np.uint32))
    assert np.all(acc[..., 3] == 11.0)
    np.testing.assert_allclose(acc, acc1, rtol=3e-6, atol=1e-5)
    assert np.abs(img.view(np.uint8).reshape(H, W, 4).astype(int) - img1.astype(int)).max() <= 1


def test_unknown_mode_rejected():
    with pytest.raises(ValueError):
        ShardedRenderer(OracleEngine(8, 8, 0), mode="bands")

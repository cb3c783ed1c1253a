"""The native multi-GPU frame loop (ptb_peer_group_* / ptb_peer_frame, north_star stage n8) on ONE device: every "rank"
is a context with its own streams, peers exchange plain device pointers (ptb_peer_group_local).  The same calls run
across processes and GPUs in tests/test_gpu_multi.py (needs >= 2 devices) and in bench.py --gpus N (parity block)."""
import os

import numpy as np
import pytest

from pathtracing_b200 import NativeTracer, PtbError, _lib, scenes
from pathtracing_b200.dist import sample_range, slab_range

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def make_group(world, w, h, dst=0, **kw):
    ranks = [NativeTracer(w, h, **kw) for _ in range(world)]
    for t in ranks:
        t.peer_group_create()
    ptrs = [t.peer_group_local() for t in ranks]
    for r, t in enumerate(ranks):
        t.peer_group_bind(r, dst, [p[0] for p in ptrs], [p[3] for p in ptrs], [p[2] for p in ptrs], ptrs[dst][1], ptrs[dst][4])
    return ranks


def rank_order_sum(parts):
    total = parts[0].copy()
    for p in parts[1:]:
        total = total + p
    return total


# up to 3 ranks on one device: the rendezvous kernels of different "ranks" must be co-resident, which one GPU only
# grants to a few tiny kernels on distinct hardware queues (conftest.py); 4 and 8 ranks run on real GPUs (test_gpu_multi.py)
@pytest.mark.parametrize("world,dst", [(1, 0), (2, 0), (3, 1)])
def test_pipelined_peer_frames_match_the_rank_ordered_oracle_sum(oracle, world, dst):
    """Several pipelined frames (no host synchronisation in between), asynchronous read-back of EVERY frame on the
    destination rank, merged accumulation = partials added in rank order, then onto the running total."""
    w, h, seed = 322, 91, 6
    p = oracle.frame_params(w, h)
    ranks = make_group(world, w, h, dst=dst, seed=seed)
    frames = [(0, 9), (9, 4), (13, 16), (29, 3), (32, 1)]
    bufs = [np.zeros((h, w, 4), np.uint8) for _ in frames]
    try:
        spp = 0
        for (first, count), buf in zip(frames, bufs):
            spp += count
            for r, t in enumerate(ranks):
                st, n = sample_range(first, count, r, world)
                t.peer_frame(st, n, spp)
            ranks[dst].read_rgba8_async(buf.ctypes.data)          # delivered when ITS merge has finished
        for t in ranks:
            t.wait()
        total = np.zeros((h, w, 4), np.float32)
        spp = 0
        for (first, count), buf in zip(frames, bufs):
            parts = [oracle.accumulate(p, *sample_range(first, count, r, world), seed, True) for r in range(world)]
            total = total + rank_order_sum(parts)
            spp += count
            assert np.array_equal(buf, oracle.resolve(total, spp)), f"frame starting at sample {first}"
        assert np.array_equal(ranks[dst].read_rgba8(), bufs[-1])
        got = np.zeros((h * w, 4), np.float32)
        for r, t in enumerate(ranks):                             # the merged accumulation lives in per-rank slabs
            first, slab = t.peer_read_total()
            assert (first, len(slab)) == slab_range(r, world, w * h)
            got[first:first + len(slab)] = slab
        assert np.array_equal(bits(got.reshape(h, w, 4)), bits(total))
        for t in ranks:                                           # partial buffers are cleared again
            assert not t.read_accum().any()
    finally:
        for t in ranks:
            t.destroy()


def test_one_process_driving_several_devices(oracle):
    """One host process, one context per GPU (the shape of the reference's single-process C# host): peers exchange plain
    device pointers, ptb_peer_group_bind switches peer access on.  Skipped with fewer than two devices."""
    import torch
    n_dev = min(torch.cuda.device_count(), 4)
    if n_dev < 2:
        pytest.skip("needs at least two GPUs")
    w, h, seed = 640, 360, 2
    p = oracle.frame_params(w, h)
    ranks = [NativeTracer(w, h, device=d, seed=seed) for d in range(n_dev)]
    try:
        for t in ranks:
            t.peer_group_create()
        ptrs = [t.peer_group_local() for t in ranks]
        for r, t in enumerate(ranks):
            t.peer_group_bind(r, 0, [q[0] for q in ptrs], [q[3] for q in ptrs], [q[2] for q in ptrs], ptrs[0][1], ptrs[0][4])
        total = np.zeros((h, w, 4), np.float32)
        spp = 0
        for first, count in [(0, 8), (8, 8), (16, 5)]:
            spp += count
            for r, t in enumerate(ranks):
                t.peer_frame(*sample_range(first, count, r, n_dev), spp)
            parts = [oracle.accumulate(p, *sample_range(first, count, r, n_dev), seed, True) for r in range(n_dev)]
            total = total + rank_order_sum(parts)
        for t in ranks:
            t.wait()
        assert np.array_equal(ranks[0].read_rgba8(), oracle.resolve(total, spp))
    finally:
        for t in ranks:
            t.destroy()


def test_peer_frame_built_from_several_sample_ranges_and_reset(oracle):
    w, h, seed, world = 200, 64, 3, 2
    p = oracle.frame_params(w, h)
    ranks = make_group(world, w, h, seed=seed)
    try:
        for r, t in enumerate(ranks):
            t.peer_frame(10 * r, 3, 0, render_only=True)          # two ranges of the same frame, merged by the third call
            t.peer_frame(10 * r + 3, 2, 0, render_only=True)
            t.peer_frame(0, 0, 10)
        for t in ranks:
            t.wait()
        parts = []
        for r in range(world):
            a = oracle.accumulate(p, 10 * r, 3, seed, True)
            parts.append(oracle.accumulate(p, 10 * r + 3, 2, seed, True, accum=a))
        assert np.array_equal(ranks[0].read_rgba8(), oracle.resolve(parts[0] + parts[1], 10))
        for t in ranks:
            t.accum_reset()                                       # progressive state back to zero on every rank
        for r, t in enumerate(ranks):
            t.peer_frame(r, 1, 2, srgb=True, bgra=True)
        for t in ranks:
            t.wait()
        tot = oracle.accumulate(p, 0, 1, seed, True) + oracle.accumulate(p, 1, 1, seed, True)
        assert np.array_equal(ranks[0].read_rgba8(), oracle.resolve(tot, 2, srgb=True, bgra=True))
    finally:
        for t in ranks:
            t.destroy()


def test_peer_frames_on_the_scene_path():
    """The same frame loop over an uploaded scene (replicated on every rank), vs the scene oracle."""
    from oracle import pt_oracle as O
    from oracle import pt_scene_oracle as S
    w, h, world, depth = 160, 90, 3, 5
    sc = scenes.mixed_small(width=w, height=h)
    so = S.SceneOracle(sc, w, h, max_depth=depth)
    ranks = make_group(world, w, h, seed=2, max_depth=depth)
    try:
        for t in ranks:
            t.upload_scene(sc)
        total = np.zeros((h, w, 4), np.float32)
        spp = 0
        for first, count in ((0, 7), (7, 5)):
            spp += count
            for r, t in enumerate(ranks):
                t.peer_frame(*sample_range(first, count, r, world), spp)
            parts = [so.accumulate(*sample_range(first, count, r, world), seed=2) for r in range(world)]
            total = total + rank_order_sum(parts)
        for t in ranks:
            t.wait()
        assert np.array_equal(ranks[0].read_rgba8(), O.resolve(total, spp))
    finally:
        for t in ranks:
            t.destroy()


def test_a_missing_peer_times_out_and_nothing_is_merged(oracle):
    """a peer that never arrives must not produce a silently wrong image nor wipe the partials."""
    if os.environ.get("PTB200_BARRIER_TIMEOUT_MS") is None:
        pytest.skip("needs PTB200_BARRIER_TIMEOUT_MS set before the library is loaded (tests/conftest.py sets it)")
    w, h = 96, 40
    ranks = make_group(2, w, h, seed=1)
    try:
        ranks[0].peer_frame(0, 4, 8)                  # rank 1 never renders its share
        with pytest.raises(PtbError) as e:
            ranks[0].wait()
        assert e.value.status == _lib.PTB_ERR_STATE
        ranks[0]._lib.ptb_set_accum_buffer(ranks[0].handle, None)
        acc = ranks[0].read_accum()
        assert np.array_equal(bits(acc), bits(oracle.accumulate(oracle.frame_params(w, h), 0, 4, 1, True)))   # not cleared
        assert not ranks[0].peer_read_total()[1].any()            # nothing merged
        with pytest.raises(PtbError):
            ranks[0].wait()                                       # sticky ...
        for t in ranks:
            t.accum_reset()                                       # ... until the progressive state is reset
        ranks[0].wait()
        ranks[1].wait()
    finally:
        for t in ranks:
            t.destroy()


def test_peer_group_argument_errors():
    with NativeTracer(64, 36) as t:
        with pytest.raises(PtbError) as e:
            t.peer_frame(0, 1, 1)
        assert e.value.status == _lib.PTB_ERR_STATE
        t.peer_group_create()
        t.peer_group_create()                                     # idempotent
        loc = t.peer_group_local()
        assert all(loc) and len(set(loc)) == 5
        with pytest.raises(PtbError) as e:
            t.peer_group_bind(0, 0, [loc[3]], [loc[0]], [loc[2]], loc[1], loc[4])    # own entries swapped
        assert e.value.status == _lib.PTB_ERR_INVALID
        t.peer_group_bind(0, 0, [loc[0]], [loc[3]], [loc[2]], loc[1], loc[4])
        with pytest.raises(PtbError) as e:
            t.peer_frame(0, 1, 0)
        assert e.value.status == _lib.PTB_ERR_INVALID
        t.peer_frame(0, 2, 2)
        t.wait()
        assert t.read_rgba8()[..., 3].all()


def test_compute_frames_persistent_launch(oracle):
    """ptb_compute_frames(n): n literal frames in one launch — same image as ptb_compute_frame, one kernel."""
    for (w, h), kw in (((1920, 1080), {}), ((320, 180), {}), ((644, 37), dict(cam=(0.1, 0.2, 1.5), radius=0.7)),
                       ((1920, 1080), dict(force_ieee=True)), ((1920, 1080), dict(spirv_arith=True))):
        spirv = bool(kw.get("spirv_arith"))
        ref = oracle.csmain_frame(oracle.frame_params(w, h, **{k: v for k, v in kw.items() if k in ("cam", "radius")}),
                                  spirv=spirv)["rgba8"]
        with NativeTracer(w, h, **kw) as t:
            t.compute_frames(1)
            assert np.array_equal(t.read_rgba8(), ref), (w, h, kw)
            n0 = t.launch_count
            t.compute_frames(50)
            assert t.launch_count == n0 + 1
            assert np.array_equal(t.read_rgba8(), ref)
    with NativeTracer(333, 77, keep_f32=True) as t:            # ragged width + float plane: falls back to per-frame launches
        t.compute_frames(3)
        assert t.launch_count == 3
        assert np.array_equal(t.read_rgba8(), oracle.csmain_frame(oracle.frame_params(333, 77))["rgba8"])

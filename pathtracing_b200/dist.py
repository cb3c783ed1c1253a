"""Multi-GPU partitioning of the tracing path (north_star stage n8).

One process per GPU (`torch.distributed`, NCCL over NVLink on the box, gloo in the CPU tests).
Every pixel and every sample is independent, so the path shards without any exchange during
rendering; the only collective is the merge of the per-rank results:

  samples  each rank renders a disjoint range of sample indices of the WHOLE frame into its own
           float4 accumulation buffer; merge = reduce(sum) of the buffers to `dst`, then ONE
           resolve there.  Sum order differs from one big launch, so parity is a tolerance — except
           that a single device running the same per-rank launches in rank order produces exactly
           the same partials, and for 2 ranks the merged sum is bit-identical (a + b = b + a).
  tiles    each rank renders ALL samples of rows  rank + i*world  (interleaved for load balance),
           resolves its rows locally; merge = reduce(sum) of the RGBA8 images (disjoint rows, the
           others are zero: sum == gather).  Bit-identical to a single-device render.

  p2p      samples sharding whose merge is ONE fused kernel per GPU over NVLink peer memory instead of an
           NCCL reduce: every rank owns a contiguous slab of pixels, loads the partials of that slab from
           all peers (CUDA IPC mappings of their accumulation buffers), adds them in rank order, resolves
           and stores the pixels into rank `dst`'s image.  The whole frame loop lives in the library
           (`ptb_peer_group_*`, `ptb_peer_frame`: render into partial buffer k & 1, rendezvous kernel over
           peer-memory flags, fused merge into image k & 1 of `dst`, rendezvous, clear — the merge of frame k on a
           high-priority stream overlaps the rendering of frame k+1); this module only exchanges the IPC
           handles.  Deterministic sum order; 8x less data over NVLink than reduce-to-one at 8 GPUs.

The engine behind a rank is anything with the `LocalEngine` methods: `GpuEngine` (NativeTracer with
a torch-owned accumulation tensor) on the box, an oracle-backed engine in the gloo tests.  This
module does plumbing only: it never computes pixels.
"""
from __future__ import annotations

from typing import Optional, Protocol, Tuple

import torch
import torch.distributed as dist


def sample_range(first: int, count: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, disjoint split of [first, first+count): rank r gets (start, n); sizes differ by <= 1."""
    base, extra = divmod(count, world)
    n = base + (1 if rank < extra else 0)
    start = first + rank * base + min(rank, extra)
    return start, n


def slab_range(rank: int, world: int, n_pixels: int) -> Tuple[int, int]:
    """Contiguous, disjoint split of the pixels for the fused peer merge: rank r owns (first, count)."""
    return sample_range(0, n_pixels, rank, world)


BAND_ROWS = 4      # rows per interleaved band: the scene path traces primary rays in 8x4 pixel tiles


def row_shard(rank: int, world: int, band_rows: int = 1) -> Tuple[int, int, int]:
    """Interleaved row bands: (row_offset, row_stride, rows_per_band) for ptb_set_row_shard_bands."""
    return rank, world, band_rows


def shard_rows(row_offset: int, row_stride: int, rows_per_band: int, height: int):
    """Image rows of a row shard, ascending (the host-side statement of csrc/ptb_internal.h shard_row)."""
    return [y for y in range(height) if (y // rows_per_band) % row_stride == row_offset]


class LocalEngine(Protocol):
    width: int
    height: int

    def set_row_shard(self, row_offset: int, row_stride: int, rows_per_band: int = 1) -> None: ...
    def accum_reset(self) -> None: ...
    def render_samples(self, first_sample: int, n_samples: int) -> None: ...
    def accum_tensor(self) -> torch.Tensor: ...          # (H, W, 4) float32, the live accumulation buffer
    def resolve(self, spp_total: int) -> None: ...
    def image_tensor(self) -> torch.Tensor: ...          # (H, W) int32 view of the packed RGBA8 image


class GpuEngine:
    """NativeTracer whose accumulation buffer and image are visible to torch.distributed."""

    def __init__(self, width: int, height: int, device: int, seed: int = 0, stream: Optional[torch.cuda.Stream] = None,
                 scene=None, max_depth: int = 0, count_stats: bool = False, peer_memory: bool = False):
        from .tracer import NativeTracer
        self.width, self.height = width, height
        self.device = torch.device("cuda", device)
        self.stream = stream or torch.cuda.current_stream(self.device)
        self.tracer = NativeTracer(width, height, device=device, seed=seed, max_depth=max_depth,
                                   stream=self.stream.cuda_stream)
        self.peer_memory = peer_memory
        if peer_memory:
            # the library's own cudaMalloc'ed accumulation buffer: exportable through CUDA IPC
            self.accum = self._wrap(self.tracer.device_accum(), (height, width, 4), "<f4")
            self.tracer.accum_reset()
        else:
            self.accum = torch.zeros((height, width, 4), dtype=torch.float32, device=self.device)
            self.tracer.set_accum_buffer(self.accum.data_ptr())
        self._peers = None
        if scene is not None:
            # every rank holds a replica of the (read-only) scene and builds its own hierarchy
            self.tracer.upload_scene(scene, count_stats=count_stats)
        self._image = None

    def set_row_shard(self, row_offset, row_stride, rows_per_band=1):
        self.tracer.set_row_shard(row_offset, row_stride, rows_per_band)

    def _wrap(self, ptr, shape, typestr):
        class _Wrap:
            pass
        w = _Wrap()
        w.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False), "version": 3}
        return torch.as_tensor(w, device=self.device)

    def accum_reset(self):
        if self.peer_memory:
            self.tracer.accum_reset()          # both partial buffers, this rank's slab of the running total, the error word
            return
        with torch.cuda.stream(self.stream):
            self.accum.zero_()
        self.tracer.set_accum_spp(0)

    # -- fused peer merge: the native frame loop ----------------------------------------------------
    def peer_setup(self, rank: int, world: int, dst: int, group=None) -> None:
        """Create this rank's peer group, exchange the CUDA IPC handles of its five buffers (two partial accumulation
        buffers, two images, the rendezvous flag block) and bind the rank-ordered pointer tables."""
        assert self.peer_memory, "GpuEngine(peer_memory=True) required"
        t = self.tracer
        t.peer_group_create()
        local = t.peer_group_local()               # (partial0, image0, flags, partial1, image1)
        handles = [None] * world
        if world > 1:
            dist.all_gather_object(handles, t.peer_group_export(), group=group)
        ptrs = [local if r == rank else tuple(t.ipc_open(h) for h in handles[r]) for r in range(world)]
        t.peer_group_bind(rank, dst, [p[0] for p in ptrs], [p[3] for p in ptrs], [p[2] for p in ptrs],
                          ptrs[dst][1], ptrs[dst][4])
        self._rank, self._world, self._dst = rank, world, dst
        if world > 1:
            dist.barrier(group=group)          # every flag block exists and is mapped before the first signal

    def peer_frame(self, first_sample: int, n_samples: int, spp_total: int, render_only: bool = False) -> None:
        self.tracer.peer_frame(first_sample, n_samples, spp_total, render_only=render_only)

    def peer_finish(self) -> None:
        self.tracer.peer_finish()

    def render_samples(self, first_sample, n_samples):
        self.tracer.render_samples(first_sample, n_samples, jitter=True)

    def accum_tensor(self):
        return self.accum

    def resolve(self, spp_total):
        self.tracer.set_accum_spp(spp_total)
        self.tracer.resolve()

    def image_tensor(self):
        # zero-copy int32 view of the library's CURRENT RGBA8 image through the CUDA array interface (the fused paths
        # alternate between two image buffers)
        ptr = self.tracer.device_rgba8()
        if self._image is None or self._image[0] != ptr:
            self._image = (ptr, self._wrap(ptr, (self.height, self.width), "<i4"))
        return self._image[1]

    def read_image(self):
        return self.tracer.read_rgba8()

    def close(self):
        self.tracer.destroy()


class ShardedRenderer:
    """Drives one LocalEngine per rank; `mode` is "samples" or "tiles"."""

    def __init__(self, engine: LocalEngine, mode: str = "samples", group=None, dst: int = 0, band_rows: int = BAND_ROWS):
        if mode not in ("samples", "tiles", "p2p"):
            raise ValueError(f"unknown sharding mode {mode!r}")
        self.engine = engine
        self.mode = mode
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.spp_total = 0                 # samples per pixel in the merged image
        if mode == "tiles":
            engine.set_row_shard(*row_shard(self.rank, self.world, band_rows))
        else:
            engine.set_row_shard(0, 1)
        engine.accum_reset()
        self.dst = dst
        if mode == "p2p":
            engine.peer_setup(self.rank, self.world, dst, group)
            self._pending = False              # samples rendered into the current frame's partial buffer, not merged yet

    def frame(self, first_sample: int, n_samples: int, dst: int = 0, pipelined: bool = True) -> Optional[torch.Tensor]:
        """One frame = render this rank's share of samples [first, first+n) + merge.  p2p: ONE library call
        (ptb_peer_frame); with pipelined=True the merge overlaps the next frame and the image is complete after finish()."""
        if self.mode != "p2p":
            self.render(first_sample, n_samples)
            return self.merge(dst)
        if dst != self.dst:
            raise ValueError("p2p merge: the destination rank is fixed when the renderer is created")
        start, n = sample_range(first_sample, n_samples, self.rank, self.world)
        self.spp_total += n_samples
        self.engine.peer_frame(start, n, self.spp_total)
        self._pending = False
        if not pipelined:
            self.finish()
        return self.engine.image_tensor() if self.rank == dst else None

    def render(self, first_sample: int, n_samples: int) -> None:
        """Render samples [first, first+n) of the frame, this rank's share only.  No communication."""
        if self.mode in ("samples", "p2p"):
            start, n = sample_range(first_sample, n_samples, self.rank, self.world)
            if self.mode == "p2p":
                self.engine.peer_frame(start, n, 0, render_only=True)
                self._pending = True
            elif n:
                self.engine.render_samples(start, n)
        else:
            self.engine.render_samples(first_sample, n_samples)
        self.spp_total += n_samples

    def finish(self) -> None:
        """p2p, pipelined merges: make the render stream wait for every merge issued so far (image complete on dst)."""
        if self.mode == "p2p":
            self.engine.peer_finish()

    def merge(self, dst: int = 0, pipelined: bool = False) -> Optional[torch.Tensor]:
        """The path's one exchange step.  Returns the (H, W) packed RGBA8 image (int32) on `dst`, else None.
        p2p + pipelined: the merge runs on its own stream and overlaps the NEXT frame's rendering (which goes to the
        other partial buffer); the image is complete after `finish()`."""
        e = self.engine
        if self.mode == "p2p":
            if dst != self.dst:
                raise ValueError("p2p merge: the destination rank is fixed when the renderer is created")
            e.peer_frame(0, 0, self.spp_total)       # n_samples = 0: merge what render() put into the partial buffer
            self._pending = False
            if not pipelined:
                self.finish()
            return e.image_tensor() if self.rank == dst else None
        if self.mode == "samples":
            acc = e.accum_tensor()
            if self.world > 1:
                dist.reduce(acc, dst=dst, op=dist.ReduceOp.SUM, group=self.group)
                if self.rank != dst:
                    acc.zero_()            # dst keeps the running total, the others the delta since the last merge
            if self.rank == dst:
                e.resolve(self.spp_total)
                return e.image_tensor()
            return None
        e.resolve(self.spp_total)          # own rows only; other rows of the image are zero
        img = e.image_tensor()
        if self.world > 1:
            dist.reduce(img, dst=dst, op=dist.ReduceOp.SUM, group=self.group)   # disjoint rows: sum == gather
        return img if self.rank == dst else None

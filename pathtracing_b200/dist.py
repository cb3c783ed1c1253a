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
           and stores the pixels into rank `dst`'s image (`ptb_peer_merge`).  Cross-rank ordering is a
           one-warp rendezvous kernel over peer-memory flags (`ptb_peer_barrier`) before and after.  Deterministic sum order;
           8x less data over NVLink than reduce-to-one at 8 GPUs.

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
            self.tracer.accum_reset()
            if getattr(self, "merger", None) is not None:
                self.merger.accum_reset()      # second partial buffer + this rank's slab of the running total
            return
        with torch.cuda.stream(self.stream):
            self.accum.zero_()
        self.tracer.set_accum_spp(0)

    # -- fused peer merge -------------------------------------------------------------------------
    def peer_setup(self, rank: int, world: int, dst: int, group=None) -> None:
        """Exchange CUDA IPC handles; map every peer's two partial accumulation buffers and rank `dst`'s image.
        Two buffers per rank (double buffering): frame k+1 renders into one while the merge kernels still read the
        other.  The second buffer and the merge launches belong to a second context on its own stream."""
        assert self.peer_memory, "GpuEngine(peer_memory=True) required"
        from .tracer import NativeTracer
        # high priority: the merge kernel's blocks must get SM slots while the next frame's (much larger) render grid
        # is in flight, otherwise the merge only starts when that grid drains and nothing overlaps
        self.merge_stream = torch.cuda.Stream(device=self.device, priority=-1)
        self.merger = NativeTracer(self.width, self.height, device=self.device.index, stream=self.merge_stream.cuda_stream)
        self.merger.accum_reset()
        self._bufs = [self.tracer.device_accum(), self.merger.device_accum()]
        mine = (self.tracer.ipc_export(0), self.merger.ipc_export(0), self.tracer.ipc_export(1), self.merger.ipc_export(2))
        handles = [None] * world
        if world > 1:
            dist.all_gather_object(handles, mine, group=group)
        else:
            handles[0] = mine
        accum_ptrs = [[self._bufs[b] if r == rank else self.merger.ipc_open(handles[r][b]) for r in range(world)]
                      for b in range(2)]
        image_ptr = self.tracer.device_rgba8() if dst == rank else self.merger.ipc_open(handles[dst][2])
        self._peers = (accum_ptrs, image_ptr, slab_range(rank, world, self.width * self.height))
        # rendezvous flag blocks (ptb_peer_barrier): one-warp kernels over peer memory instead of NCCL all-reduces
        self._flag_ptrs = [self.merger.device_peer_flags() if r == rank else self.merger.ipc_open(handles[r][3])
                           for r in range(world)]
        self._rank = rank
        self._epoch = [0, 0]
        if world > 1:
            dist.barrier(group=group)          # every flag block exists and is mapped before the first signal
        self._cur = 0
        self.torch_streams_synced = True

    def peer_use_buffer(self, b: int) -> None:
        """Render into partial buffer b from now on."""
        if b != self._cur:
            self.tracer.set_accum_buffer(None if b == 0 else self._bufs[1])
            if b == 0:
                self.tracer.device_accum()     # re-binds the context's own buffer
            self._cur = b

    def peer_barrier(self, phase: int) -> None:
        """On the merge stream: wait until every GPU has reached this point of the same frame."""
        if len(self._flag_ptrs) > 1:
            self._epoch[phase] += 1
            self.merger.peer_barrier(self._flag_ptrs, self._rank, phase, self._epoch[phase])

    def peer_merge(self, b: int, spp_total: int, background: bool = False) -> None:
        """On the merge stream: this rank's slab of frame buffer b, summed over all peers, into dst's image.
        background: small grid, for a merge that overlaps the next frame's render kernel."""
        accum_ptrs, image_ptr, (first, count) = self._peers
        self.merger.peer_merge(accum_ptrs[b], image_ptr, first, count, spp_total, background=background)

    def peer_clear(self, b: int) -> None:
        """On the merge stream: local memset of this rank's partial buffer b."""
        self.merger.set_accum_buffer(None if b == 1 else self._bufs[0])
        if b == 1:
            self.merger.device_accum()
        self.merger.accum_clear()

    def render_samples(self, first_sample, n_samples):
        self.tracer.render_samples(first_sample, n_samples, jitter=True)

    def accum_tensor(self):
        return self.accum

    def resolve(self, spp_total):
        self.tracer.set_accum_spp(spp_total)
        self.tracer.resolve()

    def image_tensor(self):
        if self._image is None:
            # zero-copy int32 view of the library's RGBA8 image through the CUDA array interface
            self._image = self._wrap(self.tracer.device_rgba8(), (self.height, self.width), "<i4")
        return self._image

    def read_image(self):
        return self.tracer.read_rgba8()

    def close(self):
        if getattr(self, "merger", None) is not None:
            self.merger.destroy()
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
            self._flag = torch.zeros(1, dtype=torch.int32, device=engine.accum_tensor().device)
            self._frame = 0                    # merges so far: frame k renders into partial buffer k & 1
            self._cleared = [None, None]       # event: the merge that last used buffer b has released it

    def render(self, first_sample: int, n_samples: int) -> None:
        """Render samples [first, first+n) of the frame, this rank's share only.  No communication."""
        if self.mode == "p2p":
            b = self._frame & 1
            if self._cleared[b] is not None:   # the merge of frame k-2 must be done with this buffer
                self.engine.stream.wait_event(self._cleared[b])
                self._cleared[b] = None
            self.engine.peer_use_buffer(b)
        if self.mode in ("samples", "p2p"):
            start, n = sample_range(first_sample, n_samples, self.rank, self.world)
            if n:
                self.engine.render_samples(start, n)
        else:
            self.engine.render_samples(first_sample, n_samples)
        self.spp_total += n_samples

    def _rendezvous(self) -> None:
        if self.world > 1:
            dist.all_reduce(self._flag, group=self.group)

    def finish(self) -> None:
        """p2p, pipelined merges: make the render stream wait for every merge issued so far (image complete on dst)."""
        if self.mode == "p2p":
            for b in range(2):
                if self._cleared[b] is not None:
                    self.engine.stream.wait_event(self._cleared[b])

    def merge(self, dst: int = 0, pipelined: bool = False) -> Optional[torch.Tensor]:
        """The path's one exchange step.  Returns the (H, W) packed RGBA8 image (int32) on `dst`, else None.
        p2p + pipelined: the merge runs on its own stream and overlaps the NEXT frame's rendering (which goes to the
        other partial buffer); the image is complete after `finish()`."""
        e = self.engine
        if self.mode == "p2p":
            if dst != self.dst:
                raise ValueError("p2p merge: the destination rank is fixed when the renderer is created")
            b = self._frame & 1
            rendered = torch.cuda.Event()
            rendered.record(e.stream)
            with torch.cuda.stream(e.merge_stream):
                e.merge_stream.wait_event(rendered)
                e.peer_barrier(0)          # every rank has rendered this frame's samples (stream-ordered, no host sync)
                e.peer_merge(b, self.spp_total, background=pipelined)
                e.peer_barrier(1)          # every slab has landed in dst's image; nobody reads these partials any more
                e.peer_clear(b)            # local memset of this rank's partial buffer
                done = torch.cuda.Event()
                done.record(e.merge_stream)
            self._cleared[b] = done
            self._frame += 1
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

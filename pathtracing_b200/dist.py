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


def row_shard(rank: int, world: int) -> Tuple[int, int]:
    """Interleaved rows: (row_offset, row_stride) for ptb_set_row_shard."""
    return rank, world


class LocalEngine(Protocol):
    width: int
    height: int

    def set_row_shard(self, row_offset: int, row_stride: int) -> None: ...
    def accum_reset(self) -> None: ...
    def render_samples(self, first_sample: int, n_samples: int) -> None: ...
    def accum_tensor(self) -> torch.Tensor: ...          # (H, W, 4) float32, the live accumulation buffer
    def resolve(self, spp_total: int) -> None: ...
    def image_tensor(self) -> torch.Tensor: ...          # (H, W) int32 view of the packed RGBA8 image


class GpuEngine:
    """NativeTracer whose accumulation buffer and image are visible to torch.distributed."""

    def __init__(self, width: int, height: int, device: int, seed: int = 0, stream: Optional[torch.cuda.Stream] = None,
                 scene=None, max_depth: int = 0, count_stats: bool = False):
        from .tracer import NativeTracer
        self.width, self.height = width, height
        self.device = torch.device("cuda", device)
        self.stream = stream or torch.cuda.current_stream(self.device)
        self.accum = torch.zeros((height, width, 4), dtype=torch.float32, device=self.device)
        self.tracer = NativeTracer(width, height, device=device, seed=seed, max_depth=max_depth,
                                   stream=self.stream.cuda_stream)
        self.tracer.set_accum_buffer(self.accum.data_ptr())
        if scene is not None:
            # every rank holds a replica of the (read-only) scene and builds its own hierarchy
            self.tracer.upload_scene(scene, count_stats=count_stats)
        self._image = None

    def set_row_shard(self, row_offset, row_stride):
        self.tracer.set_row_shard(row_offset, row_stride)

    def accum_reset(self):
        with torch.cuda.stream(self.stream):
            self.accum.zero_()
        self.tracer.set_accum_spp(0)

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
            class _Wrap:
                pass
            w = _Wrap()
            w.__cuda_array_interface__ = {"shape": (self.height, self.width), "typestr": "<i4",
                                          "data": (self.tracer.device_rgba8(), False), "version": 3}
            self._image = torch.as_tensor(w, device=self.device)
        return self._image

    def read_image(self):
        return self.tracer.read_rgba8()

    def close(self):
        self.tracer.destroy()


class ShardedRenderer:
    """Drives one LocalEngine per rank; `mode` is "samples" or "tiles"."""

    def __init__(self, engine: LocalEngine, mode: str = "samples", group=None):
        if mode not in ("samples", "tiles"):
            raise ValueError(f"unknown sharding mode {mode!r}")
        self.engine = engine
        self.mode = mode
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.spp_total = 0                 # samples per pixel in the merged image
        if mode == "tiles":
            engine.set_row_shard(*row_shard(self.rank, self.world))
        else:
            engine.set_row_shard(0, 1)
        engine.accum_reset()

    def render(self, first_sample: int, n_samples: int) -> None:
        """Render samples [first, first+n) of the frame, this rank's share only.  No communication."""
        if self.mode == "samples":
            start, n = sample_range(first_sample, n_samples, self.rank, self.world)
            if n:
                self.engine.render_samples(start, n)
        else:
            self.engine.render_samples(first_sample, n_samples)
        self.spp_total += n_samples

    def merge(self, dst: int = 0) -> Optional[torch.Tensor]:
        """The path's one exchange step.  Returns the (H, W) packed RGBA8 image (int32) on `dst`, else None."""
        e = self.engine
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

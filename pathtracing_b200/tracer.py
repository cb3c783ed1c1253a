"""NativeTracer — thin, typed wrapper over the C ABI (include/ptb200.h).

This is the executable stand-in for the C# P/Invoke wrapper shipped as source in
pathtracing_b200/cs/NativeTracer.cs: same calls, same order, same error convention.
All buffers handed across the boundary are HOST numpy arrays (or raw device pointers for the
zero-copy accessors); nothing here computes pixels.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np

from . import _lib
from ._lib import (CFG_EXTERNAL_STREAM, CFG_FORCE_IEEE, CFG_KEEP_F32, CFG_KEEP_HIT_IDS, CFG_SPIRV_ARITH, RESOLVE_BGRA, RESOLVE_SRGB,
                   Config, PtbError, check)


class NativeTracer:
    """One context == the compute half of the reference's `Renderer` (image + pipeline + fence)."""

    def __init__(self, width: int, height: int, *, device: int = 0,
                 uv_div: Optional[Sequence[float]] = None,
                 cam: Optional[Sequence[float]] = None,
                 centre: Optional[Sequence[float]] = None,
                 radius: Optional[float] = None,
                 seed: int = 0, max_depth: int = 0,
                 keep_f32: bool = False, keep_hit_ids: bool = False,
                 stream: Optional[int] = None, force_ieee: bool = False, spirv_arith: bool = False):
        self._lib = _lib.load()
        self._ctx = C.c_void_p()
        cfg = Config()
        check(None, self._lib.ptb_default_config(C.byref(cfg), width, height))
        cfg.device = device
        if uv_div is not None:
            cfg.uv_div_x, cfg.uv_div_y = float(uv_div[0]), float(uv_div[1])
        if cam is not None:
            cfg.cam[:] = [float(v) for v in cam]
        if centre is not None:
            cfg.centre[:] = [float(v) for v in centre]
        if radius is not None:
            cfg.radius = float(radius)
        cfg.seed = seed
        cfg.max_depth = max_depth
        cfg.flags = (CFG_KEEP_F32 if keep_f32 else 0) | (CFG_KEEP_HIT_IDS if keep_hit_ids else 0)
        if force_ieee:
            cfg.flags |= CFG_FORCE_IEEE
        if spirv_arith:
            cfg.flags |= CFG_SPIRV_ARITH
        if stream is not None:
            cfg.flags |= CFG_EXTERNAL_STREAM
            cfg.stream = stream
        self.config = cfg
        self.width, self.height = width, height
        self._present_shape = None
        check(None, self._lib.ptb_create(C.byref(cfg), C.byref(self._ctx)))

    # -- lifecycle -----------------------------------------------------------------------
    def destroy(self) -> None:
        if getattr(self, "_ctx", None) is not None and self._ctx.value:
            self._lib.ptb_destroy(self._ctx)
            self._ctx = C.c_void_p()

    close = destroy

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.destroy()

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass

    def _ck(self, status: int) -> None:
        check(self._ctx, status)

    @property
    def handle(self) -> C.c_void_p:
        return self._ctx

    # -- the per-frame hot path ----------------------------------------------------------
    def compute_frame(self) -> None:
        self._ck(self._lib.ptb_compute_frame(self._ctx))

    def compute_frames(self, n_frames: int) -> None:
        """n_frames literal frames in one persistent launch (falls back to n launches where that is not possible)."""
        self._ck(self._lib.ptb_compute_frames(self._ctx, n_frames))

    def wait(self) -> None:
        self._ck(self._lib.ptb_wait(self._ctx))

    def read_rgba8(self, out: Optional[np.ndarray] = None) -> np.ndarray:
        if out is None:
            out = np.empty((self.height, self.width, 4), np.uint8)
        self._ck(self._lib.ptb_read_rgba8(self._ctx, out.ctypes.data, out.strides[0]))
        return out

    def read_rgba8_async(self, host_ptr: int, pitch_bytes: int = 0) -> None:
        """Queue a device->host copy of the image into (page-locked) host memory; valid after wait()."""
        self._ck(self._lib.ptb_read_rgba8_async(self._ctx, host_ptr, pitch_bytes))

    def read_f32(self, out: Optional[np.ndarray] = None) -> np.ndarray:
        if out is None:
            out = np.empty((self.height, self.width, 4), np.float32)
        self._ck(self._lib.ptb_read_f32(self._ctx, out.ctypes.data, out.strides[0]))
        return out

    def read_hit_ids(self, out: Optional[np.ndarray] = None) -> np.ndarray:
        if out is None:
            out = np.empty((self.height, self.width), np.int32)
        self._ck(self._lib.ptb_read_hit_ids(self._ctx, out.ctypes.data, out.strides[0]))
        return out

    # -- accumulation / resolve ----------------------------------------------------------
    def accum_reset(self) -> None:
        self._ck(self._lib.ptb_accum_reset(self._ctx))

    def render_samples(self, first_sample: int, n_samples: int, jitter: bool = True) -> None:
        self._ck(self._lib.ptb_render_samples(self._ctx, first_sample, n_samples, int(jitter)))

    def render_resolve(self, first_sample: int, n_samples: int, jitter: bool = True, srgb: bool = False,
                       bgra: bool = False) -> None:
        """render_samples + resolve as one launch where possible (literal scene, no row shard, no sRGB); same results."""
        self._ck(self._lib.ptb_render_resolve(self._ctx, first_sample, n_samples, int(jitter),
                                              (RESOLVE_SRGB if srgb else 0) | (RESOLVE_BGRA if bgra else 0)))

    def resolve(self, srgb: bool = False, bgra: bool = False) -> None:
        self._ck(self._lib.ptb_resolve(self._ctx, (RESOLVE_SRGB if srgb else 0) | (RESOLVE_BGRA if bgra else 0)))

    @property
    def accum_spp(self) -> int:
        return int(self._lib.ptb_accum_spp(self._ctx))

    def set_accum_spp(self, spp: int) -> None:
        self._ck(self._lib.ptb_set_accum_spp(self._ctx, spp))

    def set_row_shard(self, row_offset: int, row_stride: int, rows_per_band: int = 1) -> None:
        """Render only the row bands row_offset + i*row_stride (bands of rows_per_band rows; 1 = single rows)."""
        self._ck(self._lib.ptb_set_row_shard_bands(self._ctx, row_offset, row_stride, rows_per_band))

    def read_accum(self, out: Optional[np.ndarray] = None) -> np.ndarray:
        if out is None:
            out = np.empty((self.height, self.width, 4), np.float32)
        self._ck(self._lib.ptb_read_accum(self._ctx, out.ctypes.data, out.strides[0]))
        return out

    def write_accum(self, accum: np.ndarray, spp_total: int) -> None:
        a = np.ascontiguousarray(accum, np.float32)
        assert a.shape == (self.height, self.width, 4)
        self._ck(self._lib.ptb_write_accum(self._ctx, a.ctypes.data, a.strides[0], spp_total))

    def set_accum_buffer(self, device_ptr: Optional[int]) -> None:
        self._ck(self._lib.ptb_set_accum_buffer(self._ctx, device_ptr))

    # -- scene path (n3-n7) ------------------------------------------------------------------
    def upload_scene(self, scene, *, max_leaf: int = 0, count_stats: bool = False) -> None:
        """Upload a `scenes.Scene` (host arrays) and build its hierarchies on the GPU."""
        d = _lib.SceneDesc()
        d.struct_size = C.sizeof(_lib.SceneDesc)
        d.n_spheres = scene.n_spheres
        d.spheres = scene.spheres.ctypes.data
        d.sphere_material = scene.sphere_material.ctypes.data
        d.n_triangles = scene.n_triangles
        d.tri_vertices = scene.triangles.ctypes.data
        d.tri_material = scene.tri_material.ctypes.data
        d.n_materials = len(scene.materials)
        d.materials = scene.materials.ctypes.data
        for name, src in (("cam_origin", scene.cam_origin), ("cam_u", scene.cam_u), ("cam_v", scene.cam_v),
                          ("cam_w", scene.cam_w), ("sky_horizon", scene.sky_horizon), ("sky_zenith", scene.sky_zenith)):
            getattr(d, name)[:] = [float(v) for v in src]
        d.t_min = float(scene.t_min)
        d.shade_mode = int(scene.shade_mode)
        d.max_leaf = max_leaf
        d.flags = _lib.SCENE_COUNT_STATS if count_stats else 0
        self._ck(self._lib.ptb_upload_scene(self._ctx, C.byref(d)))
        if getattr(scene, "sun_on", False):
            self.set_sun(scene.sun_dir, scene.sun_radiance)

    def set_sun(self, dir_to_light, radiance) -> None:
        """Directional light sampled by next-event estimation at Lambert hits (any-hit shadow rays); radiance 0 = off."""
        d = (C.c_float * 3)(*[float(v) for v in dir_to_light])
        r = (C.c_float * 3)(*[float(v) for v in radiance])
        self._ck(self._lib.ptb_scene_set_sun(self._ctx, d, r))

    def clear_scene(self) -> None:
        self._ck(self._lib.ptb_clear_scene(self._ctx))

    def scene_info(self) -> dict:
        info = _lib.SceneInfo()
        self._ck(self._lib.ptb_scene_info_get(self._ctx, C.byref(info)))
        return {k: getattr(info, k) for k, _ in _lib.SceneInfo._fields_}

    def scene_stats(self) -> dict:
        st = _lib.SceneStats()
        self._ck(self._lib.ptb_scene_stats_get(self._ctx, C.byref(st)))
        return {k: int(getattr(st, k)) for k, _ in _lib.SceneStats._fields_}

    def scene_stats_reset(self) -> None:
        self._ck(self._lib.ptb_scene_stats_reset(self._ctx))

    def set_wave_samples(self, samples_per_wave: int) -> None:
        """Samples of every pixel one wave traces (0 = as many as fit the path queues). Scheduling only."""
        self._ck(self._lib.ptb_scene_set_wave_samples(self._ctx, samples_per_wave))

    def scene_queue_sizes(self):
        """Live paths entering each bounce of the last wave: [camera rays, after 1 scatter, ...]."""
        out = (C.c_uint32 * 256)()
        n = C.c_uint32()
        self._ck(self._lib.ptb_scene_queue_sizes(self._ctx, out, 256, C.byref(n)))
        return [int(v) for v in out[:n.value]]

    def read_bvh(self, which: int):
        """-> (nodes (n,16) float32 view of the 64 B nodes, order (n_prims,) uint32)."""
        nn, npr = C.c_uint32(), C.c_uint32()
        self._ck(self._lib.ptb_scene_read_bvh(self._ctx, which, None, 0, C.byref(nn), None, 0, C.byref(npr)))
        nodes = np.zeros((nn.value, 16), np.float32)
        order = np.zeros(npr.value, np.uint32)
        self._ck(self._lib.ptb_scene_read_bvh(self._ctx, which, nodes.ctypes.data, nn.value, C.byref(nn),
                                              order.ctypes.data, npr.value, C.byref(npr)))
        return nodes, order

    def trace_rays(self, origins: np.ndarray, dirs: np.ndarray, t_min: float, t_max: float, any_hit: bool = False):
        """Closest-hit (-> prim ids, t) or any-hit (-> 0/1 flags, None) queries with host ray arrays."""
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        assert o.shape == d.shape
        prim = np.empty(len(o), np.int32)
        t = np.zeros(len(o), np.float32)
        self._ck(self._lib.ptb_trace_rays(self._ctx, len(o), o.ctypes.data, d.ctypes.data, t_min, t_max,
                                          1 if any_hit else 0, prim.ctypes.data, t.ctypes.data))
        return (prim, None) if any_hit else (prim, t)

    TRACE_RAYS_STATIC_MAX = 65536      # PTB_TRACE_RAYS_STATIC_MAX (include/ptb200.h)

    def trace_rays_static(self, origins: np.ndarray, dirs: np.ndarray, t_min: float, t_max: float, any_hit: bool = False):
        """trace_rays through ptb_trace_rays_static: same results, device staging in the library's static array (no
        device allocation in the call); at most TRACE_RAYS_STATIC_MAX rays, calls on one GPU must not overlap."""
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        assert o.shape == d.shape
        prim = np.empty(len(o), np.int32)
        t = np.zeros(len(o), np.float32)
        self._ck(self._lib.ptb_trace_rays_static(self._ctx, len(o), o.ctypes.data, d.ctypes.data, t_min, t_max,
                                                 1 if any_hit else 0, prim.ctypes.data, t.ctypes.data))
        return (prim, None) if any_hit else (prim, t)

    # -- fused multi-GPU merge over peer memory ---------------------------------------------------
    def device_peer_flags(self) -> int:
        return int(self._lib.ptb_device_peer_flags(self._ctx) or 0)

    def peer_barrier(self, peer_flags, self_rank: int, phase: int, epoch: int) -> None:
        arr = (C.c_void_p * len(peer_flags))(*peer_flags)
        self._ck(self._lib.ptb_peer_barrier(self._ctx, len(peer_flags), arr, self_rank, phase, epoch))

    def accum_clear(self) -> None:
        self._ck(self._lib.ptb_accum_clear(self._ctx))

    def ipc_export(self, which: int) -> bytes:
        """cudaIpcMemHandle_t (64 bytes): 0 accumulation buffer, 1 RGBA8 image, 2 flag block, 3 / 4 second partial / image."""
        buf = C.create_string_buffer(64)
        self._ck(self._lib.ptb_ipc_export(self._ctx, which, buf))
        return buf.raw

    def ipc_open(self, handle: bytes) -> int:
        ptr = C.c_void_p()
        self._ck(self._lib.ptb_ipc_open(self._ctx, handle, C.byref(ptr)))
        return int(ptr.value)

    def peer_merge(self, peer_accum, dst_image: int, slab_first: int, slab_count: int, spp_total: int,
                   srgb: bool = False, bgra: bool = False, background: bool = False) -> None:
        """Sum the peers' partial accumulation buffers over this GPU's slab, resolve, store into dst_image."""
        arr = (C.c_void_p * len(peer_accum))(*peer_accum)
        self._ck(self._lib.ptb_peer_merge(self._ctx, len(peer_accum), arr, dst_image, slab_first, slab_count, spp_total,
                                          (RESOLVE_SRGB if srgb else 0) | (RESOLVE_BGRA if bgra else 0) |
                                          (0x10 if background else 0)))

    # -- native multi-GPU frame loop (ptb_peer_group_* / ptb_peer_frame) ------------------------------
    def peer_group_create(self) -> None:
        self._ck(self._lib.ptb_peer_group_create(self._ctx))

    def peer_group_local(self):
        """Device pointers (partial0, image0, flags, partial1, image1) for peers in the same process."""
        return tuple(int(self._lib.ptb_peer_group_local(self._ctx, w) or 0) for w in range(5))

    def peer_group_export(self):
        """CUDA IPC handles of the same five buffers, for peers in other processes."""
        return tuple(self.ipc_export(w) for w in range(5))

    def peer_group_bind(self, self_rank: int, dst_rank: int, partial0, partial1, flags, dst_image0: int, dst_image1: int) -> None:
        n = len(partial0)
        arr = lambda v: (C.c_void_p * n)(*v)
        self._ck(self._lib.ptb_peer_group_bind(self._ctx, n, self_rank, dst_rank, arr(partial0), arr(partial1), arr(flags),
                                               dst_image0, dst_image1))

    def peer_frame(self, first_sample: int, n_samples: int, spp_total: int, jitter: bool = True, srgb: bool = False,
                   bgra: bool = False, render_only: bool = False) -> None:
        """Render this rank's samples of the frame and queue the pipelined fused merge (one call per frame and rank).
        render_only: no merge yet (more sample ranges of the same frame follow)."""
        self._ck(self._lib.ptb_peer_frame(self._ctx, first_sample, n_samples, int(jitter), spp_total,
                                          (RESOLVE_SRGB if srgb else 0) | (RESOLVE_BGRA if bgra else 0) |
                                          (0x20 if render_only else 0)))

    def peer_finish(self) -> None:
        self._ck(self._lib.ptb_peer_finish(self._ctx))

    def peer_read_total(self):
        """-> (slab_first, float32 (slab_count, 4)): this rank's slab of the merged accumulation."""
        first, count = C.c_uint32(), C.c_uint32()
        self._ck(self._lib.ptb_peer_read_total(self._ctx, None, C.byref(first), C.byref(count)))
        out = np.empty((count.value, 4), np.float32)
        self._ck(self._lib.ptb_peer_read_total(self._ctx, out.ctypes.data, C.byref(first), C.byref(count)))
        return int(first.value), out

    # -- present-equivalent ----------------------------------------------------------------
    def present(self, dst_w: Optional[int] = None, dst_h: Optional[int] = None, srgb: bool = True) -> None:
        self._ck(self._lib.ptb_present(self._ctx, dst_w or self.width, dst_h or self.height,
                                       RESOLVE_SRGB if srgb else 0))
        self._present_shape = (dst_h or self.height, dst_w or self.width)

    def read_present(self) -> np.ndarray:
        h, w = self._present_shape or (1, 1)     # no present yet: the native call reports PTB_ERR_STATE
        out = np.empty((h, w, 4), np.uint8)
        self._ck(self._lib.ptb_read_present(self._ctx, out.ctypes.data, out.strides[0]))
        return out

    # -- zero-copy + measurement -----------------------------------------------------------
    def device_rgba8(self) -> int:
        return int(self._lib.ptb_device_rgba8(self._ctx) or 0)

    def device_f32(self) -> int:
        return int(self._lib.ptb_device_f32(self._ctx) or 0)

    def device_accum(self) -> int:
        return int(self._lib.ptb_device_accum(self._ctx) or 0)

    def timer_begin(self) -> None:
        self._ck(self._lib.ptb_timer_begin(self._ctx))

    def timer_end(self) -> float:
        ms = C.c_float()
        self._ck(self._lib.ptb_timer_end(self._ctx, C.byref(ms)))
        return float(ms.value)

    @property
    def kernel_mode(self) -> int:
        """0 = literal IEEE intrinsics, 1 = shortened exact sequences, 2 = + on-axis camera, 3 = Test.spirv op list."""
        return int(self._lib.ptb_kernel_mode(self._ctx))

    def selftest_exact_math(self, which: int, divisor: float = 1.0):
        """-> (mismatches, tested) of the on-device proof by exhaustion (see include/ptb200.h)."""
        bad, n = C.c_uint64(), C.c_uint64()
        self._ck(self._lib.ptb_selftest_exact_math(self._ctx, which, divisor, C.byref(bad), C.byref(n)))
        return int(bad.value), int(n.value)

    @property
    def launch_count(self) -> int:
        return int(self._lib.ptb_launch_count(self._ctx))


__all__ = ["NativeTracer", "PtbError"]

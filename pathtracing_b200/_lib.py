"""ctypes loader for libptb200.so (the C-ABI CUDA backend, include/ptb200.h).

There is no CPU fallback: if the library is missing or cannot be loaded this module raises
ImportError-style errors loudly, and `ptb_create` fails when no CUDA device is usable.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_PKG = os.path.dirname(os.path.abspath(__file__))
_CSRC = os.path.join(_PKG, "csrc")
LIB_PATH = os.environ.get("PTB200_LIB") or os.path.join(_PKG, "libptb200.so")   # override: A/B experiments only
HEADER_PATH = os.path.join(os.path.dirname(_PKG), "include", "ptb200.h")

PTB_OK = 0
PTB_ERR_INVALID = -1
PTB_ERR_CUDA = -2
PTB_ERR_NOMEM = -3
PTB_ERR_STATE = -4
PTB_ERR_UNSUPPORTED = -5

CFG_KEEP_F32 = 0x1
CFG_KEEP_HIT_IDS = 0x2
CFG_EXTERNAL_STREAM = 0x4
CFG_FORCE_IEEE = 0x8
CFG_SPIRV_ARITH = 0x10
RESOLVE_SRGB = 0x1
RESOLVE_BGRA = 0x2


class PtbError(Exception):
    """Raised for any non-zero status; mirrors the reference's `throw new Exception(string)` on a
    non-Success Vk result (RayTracing/Graphics/Renderer.cs:115-118, 1036-1039)."""

    def __init__(self, status: int, message: str):
        super().__init__(f"ptb status {status}: {message}")
        self.status = status
        self.message = message


class Config(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32),
        ("device", C.c_int32),
        ("width", C.c_uint32),
        ("height", C.c_uint32),
        ("uv_div_x", C.c_float),
        ("uv_div_y", C.c_float),
        ("cam", C.c_float * 3),
        ("centre", C.c_float * 3),
        ("radius", C.c_float),
        ("seed", C.c_uint32),
        ("max_depth", C.c_uint32),
        ("flags", C.c_uint32),
        ("stream", C.c_void_p),
    ]


MAT_LAMBERT, MAT_METAL, MAT_DIELECTRIC = 0, 1, 2
SHADE_PATH, SHADE_NORMALS = 0, 1
SCENE_COUNT_STATS = 0x1


class Material(C.Structure):
    _fields_ = [("type", C.c_uint32), ("albedo", C.c_float * 3), ("fuzz", C.c_float), ("ior", C.c_float),
                ("pad", C.c_float * 2)]


class SceneDesc(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32),
        ("n_spheres", C.c_uint32),
        ("spheres", C.c_void_p),
        ("sphere_material", C.c_void_p),
        ("n_triangles", C.c_uint32),
        ("n_materials", C.c_uint32),
        ("tri_vertices", C.c_void_p),
        ("tri_material", C.c_void_p),
        ("materials", C.c_void_p),
        ("cam_origin", C.c_float * 3),
        ("cam_u", C.c_float * 3),
        ("cam_v", C.c_float * 3),
        ("cam_w", C.c_float * 3),
        ("sky_horizon", C.c_float * 3),
        ("sky_zenith", C.c_float * 3),
        ("t_min", C.c_float),
        ("shade_mode", C.c_uint32),
        ("max_leaf", C.c_uint32),
        ("flags", C.c_uint32),
    ]


class SceneInfo(C.Structure):
    _fields_ = [("n_spheres", C.c_uint32), ("n_triangles", C.c_uint32), ("sphere_nodes", C.c_uint32),
                ("triangle_nodes", C.c_uint32), ("sphere_leaves", C.c_uint32), ("triangle_leaves", C.c_uint32),
                ("max_leaf", C.c_uint32), ("build_ms", C.c_float), ("device_bytes", C.c_uint64)]


class SceneStats(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("nodes_visited", C.c_uint64), ("prims_tested", C.c_uint64),
                ("scatters", C.c_uint64), ("paths", C.c_uint64), ("max_stack", C.c_uint64)]


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile libptb200.so in-tree with nvcc for sm_100a (pathtracing_b200/csrc/Makefile)."""
    srcs = [os.path.join(_CSRC, f) for f in os.listdir(_CSRC) if f.endswith((".cu", ".cuh", ".h"))]
    srcs.append(HEADER_PATH)
    host = os.path.join(_PKG, "host")
    srcs += [os.path.join(host, f) for f in os.listdir(host) if f.endswith((".cpp", ".hpp"))]
    stale = (not os.path.exists(LIB_PATH)) or (not os.path.exists(os.path.join(host, "pt_render"))) or any(
        os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in srcs)
    if force or stale:
        cmd = ["make", "-C", _CSRC, "all"] + (["-B"] if force else [])
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if verbose or r.returncode != 0:
            print(r.stdout)
        if r.returncode != 0:
            raise RuntimeError("nvcc build of libptb200.so failed")
    return LIB_PATH


# name -> (restype, argtypes); one entry per symbol declared in include/ptb200.h
_P = C.POINTER
_CTX = C.c_void_p
SIGNATURES = {
    "ptb_default_config": (C.c_int, [_P(Config), C.c_uint32, C.c_uint32]),
    "ptb_abi_version": (C.c_int, []),
    "ptb_device_count": (C.c_int, []),
    "ptb_create": (C.c_int, [_P(Config), _P(_CTX)]),
    "ptb_destroy": (C.c_int, [_CTX]),
    "ptb_last_error": (C.c_char_p, [_CTX]),
    "ptb_compute_frame": (C.c_int, [_CTX]),
    "ptb_wait": (C.c_int, [_CTX]),
    "ptb_read_rgba8": (C.c_int, [_CTX, C.c_void_p, C.c_size_t]),
    "ptb_read_rgba8_async": (C.c_int, [_CTX, C.c_void_p, C.c_size_t]),
    "ptb_read_f32": (C.c_int, [_CTX, C.c_void_p, C.c_size_t]),
    "ptb_read_hit_ids": (C.c_int, [_CTX, C.c_void_p, C.c_size_t]),
    "ptb_device_rgba8": (C.c_void_p, [_CTX]),
    "ptb_device_f32": (C.c_void_p, [_CTX]),
    "ptb_device_accum": (C.c_void_p, [_CTX]),
    "ptb_accum_reset": (C.c_int, [_CTX]),
    "ptb_render_samples": (C.c_int, [_CTX, C.c_uint32, C.c_uint32, C.c_int]),
    "ptb_resolve": (C.c_int, [_CTX, C.c_uint32]),
    "ptb_accum_spp": (C.c_uint32, [_CTX]),
    "ptb_read_accum": (C.c_int, [_CTX, C.c_void_p, C.c_size_t]),
    "ptb_write_accum": (C.c_int, [_CTX, C.c_void_p, C.c_size_t, C.c_uint32]),
    "ptb_set_accum_buffer": (C.c_int, [_CTX, C.c_void_p]),
    "ptb_set_accum_spp": (C.c_int, [_CTX, C.c_uint32]),
    "ptb_render_resolve": (C.c_int, [_CTX, C.c_uint32, C.c_uint32, C.c_int, C.c_uint32]),
    "ptb_set_row_shard": (C.c_int, [_CTX, C.c_uint32, C.c_uint32]),
    "ptb_scene_set_wave_samples": (C.c_int, [_CTX, C.c_uint32]),
    "ptb_scene_set_sun": (C.c_int, [_CTX, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "ptb_set_row_shard_bands": (C.c_int, [_CTX, C.c_uint32, C.c_uint32, C.c_uint32]),
    "ptb_upload_scene": (C.c_int, [_CTX, _P(SceneDesc)]),
    "ptb_clear_scene": (C.c_int, [_CTX]),
    "ptb_scene_info_get": (C.c_int, [_CTX, _P(SceneInfo)]),
    "ptb_scene_stats_get": (C.c_int, [_CTX, _P(SceneStats)]),
    "ptb_scene_stats_reset": (C.c_int, [_CTX]),
    "ptb_scene_queue_sizes": (C.c_int, [_CTX, _P(C.c_uint32), C.c_uint32, _P(C.c_uint32)]),
    "ptb_scene_read_bvh": (C.c_int, [_CTX, C.c_int, C.c_void_p, C.c_uint32, _P(C.c_uint32), C.c_void_p, C.c_uint32,
                                     _P(C.c_uint32)]),
    "ptb_trace_rays": (C.c_int, [_CTX, C.c_uint64, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_int, C.c_void_p,
                                 C.c_void_p]),
    "ptb_trace_rays_static": (C.c_int, [_CTX, C.c_uint64, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_int, C.c_void_p,
                                        C.c_void_p]),
    "ptb_accum_clear": (C.c_int, [_CTX]),
    "ptb_device_peer_flags": (C.c_void_p, [_CTX]),
    "ptb_peer_barrier": (C.c_int, [_CTX, C.c_uint32, _P(C.c_void_p), C.c_uint32, C.c_uint32, C.c_uint32]),
    "ptb_ipc_export": (C.c_int, [_CTX, C.c_int, C.c_void_p]),
    "ptb_ipc_open": (C.c_int, [_CTX, C.c_void_p, _P(C.c_void_p)]),
    "ptb_peer_merge": (C.c_int, [_CTX, C.c_uint32, _P(C.c_void_p), C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]),
    "ptb_compute_frames": (C.c_int, [_CTX, C.c_uint32]),
    "ptb_peer_group_create": (C.c_int, [_CTX]),
    "ptb_peer_group_local": (C.c_void_p, [_CTX, C.c_int]),
    "ptb_peer_group_bind": (C.c_int, [_CTX, C.c_uint32, C.c_uint32, C.c_uint32, _P(C.c_void_p), _P(C.c_void_p), _P(C.c_void_p),
                                     C.c_void_p, C.c_void_p]),
    "ptb_peer_frame": (C.c_int, [_CTX, C.c_uint32, C.c_uint32, C.c_int, C.c_uint32, C.c_uint32]),
    "ptb_peer_finish": (C.c_int, [_CTX]),
    "ptb_peer_read_total": (C.c_int, [_CTX, C.c_void_p, _P(C.c_uint32), _P(C.c_uint32)]),
    "ptb_present": (C.c_int, [_CTX, C.c_uint32, C.c_uint32, C.c_uint32]),
    "ptb_read_present": (C.c_int, [_CTX, C.c_void_p, C.c_size_t]),
    "ptb_timer_begin": (C.c_int, [_CTX]),
    "ptb_timer_end": (C.c_int, [_CTX, _P(C.c_float)]),
    "ptb_launch_count": (C.c_uint64, [_CTX]),
    "ptb_kernel_mode": (C.c_int, [_CTX]),
    "ptb_selftest_exact_math": (C.c_int, [_CTX, C.c_int, C.c_float, _P(C.c_uint64), _P(C.c_uint64)]),
}

_lib = None


def load() -> C.CDLL:
    """dlopen libptb200.so and bind every declared symbol.  Fails loudly."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  pathtracing_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)      # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(ctx, status: int) -> None:
    if status != PTB_OK:
        msg = load().ptb_last_error(ctx)
        raise PtbError(status, msg.decode("utf-8", "replace") if msg else "")

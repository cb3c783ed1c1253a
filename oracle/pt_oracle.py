"""ctypes binding of the CPU oracle (oracle/libpt_oracle.so).

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package (pathtracing_b200/) never
imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libpt_oracle.so")


def build(force: bool = False) -> str:
    """Compile the oracle with gcc (oracle/Makefile).  Building the checker is not using it."""
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".c", ".h"))]
    stale = (not os.path.exists(_SO)) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs)
    if force or stale:
        env = dict(os.environ)
        env.pop("CC", None)
        subprocess.run(["make", "-C", _HERE, "-B" if force else "-s"], check=True, env=env,
                       stdout=subprocess.DEVNULL)
    return _SO


class FrameParams(C.Structure):
    _fields_ = [
        ("width", C.c_uint32), ("height", C.c_uint32),
        ("uv_div_x", C.c_float), ("uv_div_y", C.c_float),
        ("cam", C.c_float * 3), ("centre", C.c_float * 3), ("radius", C.c_float),
    ]


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        P = C.POINTER
        L.pto_default_frame.argtypes = [P(FrameParams), C.c_uint32, C.c_uint32]
        L.pto_default_frame.restype = None
        L.pto_csmain_sample.argtypes = [P(FrameParams), C.c_float, C.c_float, P(C.c_float),
                                        P(C.c_float), P(C.c_float)]
        L.pto_csmain_sample.restype = C.c_int
        L.pto_csmain_frame.argtypes = [P(FrameParams), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.pto_csmain_frame.restype = C.c_int64
        L.pto_csmain_sample_spirv.argtypes = L.pto_csmain_sample.argtypes
        L.pto_csmain_sample_spirv.restype = C.c_int
        L.pto_csmain_frame_spirv.argtypes = L.pto_csmain_frame.argtypes
        L.pto_csmain_frame_spirv.restype = C.c_int64
        L.pto_unorm8.argtypes = [C.c_float]
        L.pto_unorm8.restype = C.c_uint8
        L.pto_srgb8.argtypes = [C.c_float]
        L.pto_srgb8.restype = C.c_uint8
        L.pto_pcg3d.argtypes = [P(C.c_uint32)]
        L.pto_pcg3d.restype = None
        L.pto_bits_to_unit.argtypes = [C.c_uint32]
        L.pto_bits_to_unit.restype = C.c_float
        L.pto_accumulate.argtypes = [P(FrameParams), C.c_uint32, C.c_uint32, C.c_uint32, C.c_int,
                                     C.c_void_p, C.c_int]
        L.pto_accumulate.restype = None
        L.pto_accumulate_v.argtypes = [P(FrameParams), C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int,
                                       C.c_void_p, C.c_int]
        L.pto_accumulate_v.restype = None
        L.pto_resolve.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
        L.pto_resolve.restype = None
        L.pto_present.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint32,
                                  C.c_uint32, C.c_uint32]
        L.pto_present.restype = None
        L.pto_max_threads.argtypes = []
        L.pto_max_threads.restype = C.c_int
        L.pto_set_threads.argtypes = [C.c_int]
        L.pto_set_threads.restype = None
        _lib = L
    return _lib


def frame_params(width: int, height: int, uv_div=None, cam=None, centre=None, radius=None) -> FrameParams:
    p = FrameParams()
    lib().pto_default_frame(C.byref(p), width, height)
    if uv_div is not None:
        p.uv_div_x, p.uv_div_y = float(uv_div[0]), float(uv_div[1])
    if cam is not None:
        p.cam[:] = [float(v) for v in cam]
    if centre is not None:
        p.centre[:] = [float(v) for v in centre]
    if radius is not None:
        p.radius = float(radius)
    return p


def csmain_sample(p: FrameParams, px: float, py: float, spirv: bool = False):
    """-> (hit, colour float32[4], disc, t).  spirv: the as-compiled variant (Compiled/Ray/Test.spirv op list)."""
    col = (C.c_float * 4)()
    disc, t = C.c_float(), C.c_float()
    fn = lib().pto_csmain_sample_spirv if spirv else lib().pto_csmain_sample
    hit = fn(C.byref(p), px, py, col, C.byref(disc), C.byref(t))
    return hit, np.array(col[:], dtype=np.float32), np.float32(disc.value), np.float32(t.value)


def csmain_frame(p: FrameParams, want_f32=True, want_rgba8=True, want_hit=True, nthreads=0, spirv: bool = False):
    """-> dict(hits, f32 (H,W,4), rgba8 (H,W,4), hit (H,W)).  spirv: the as-compiled variant."""
    W, H = p.width, p.height
    f32 = np.empty((H, W, 4), np.float32) if want_f32 else None
    u8 = np.empty((H, W, 4), np.uint8) if want_rgba8 else None
    hit = np.empty((H, W), np.int32) if want_hit else None
    n = (lib().pto_csmain_frame_spirv if spirv else lib().pto_csmain_frame)(C.byref(p),
                               f32.ctypes.data if want_f32 else None,
                               u8.ctypes.data if want_rgba8 else None,
                               hit.ctypes.data if want_hit else None, nthreads)
    return {"hits": int(n), "f32": f32, "rgba8": u8, "hit": hit}


def accumulate(p: FrameParams, first_sample: int, n_samples: int, seed: int, jitter: bool,
               accum: np.ndarray | None = None, nthreads=0, spirv: bool = False) -> np.ndarray:
    if accum is None:
        accum = np.zeros((p.height, p.width, 4), np.float32)
    assert accum.dtype == np.float32 and accum.flags.c_contiguous
    lib().pto_accumulate_v(C.byref(p), 1 if spirv else 0, first_sample, n_samples, seed, int(jitter),
                           accum.ctypes.data, nthreads)
    return accum


def resolve(accum: np.ndarray, spp_total: int, srgb=False, bgra=False) -> np.ndarray:
    H, W = accum.shape[:2]
    out = np.empty((H, W, 4), np.uint8)
    acc = np.ascontiguousarray(accum, np.float32)
    lib().pto_resolve(acc.ctypes.data, W, H, spp_total, (1 if srgb else 0) | (2 if bgra else 0),
                      out.ctypes.data)
    return out


def present(src_rgba8: np.ndarray, dw: int, dh: int, srgb=True) -> np.ndarray:
    sh, sw = src_rgba8.shape[:2]
    src = np.ascontiguousarray(src_rgba8, np.uint8)
    out = np.empty((dh, dw, 4), np.uint8)
    lib().pto_present(src.ctypes.data, sw, sh, out.ctypes.data, dw, dh, 1 if srgb else 0)
    return out


def pcg3d(x: int, y: int, z: int):
    v = (C.c_uint32 * 3)(x & 0xFFFFFFFF, y & 0xFFFFFFFF, z & 0xFFFFFFFF)
    lib().pto_pcg3d(v)
    return int(v[0]), int(v[1]), int(v[2])


def bits_to_unit(w: int) -> np.float32:
    return np.float32(lib().pto_bits_to_unit(w & 0xFFFFFFFF))


def max_threads() -> int:
    return int(lib().pto_max_threads())


def set_threads(n: int) -> None:
    """OpenMP threads of the following calls (both oracles share the OpenMP runtime)."""
    lib().pto_set_threads(int(n))

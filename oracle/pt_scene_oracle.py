"""ctypes binding of the scene-path CPU oracle (oracle/pt_scene_oracle.c, in libpt_oracle.so).

TEST INFRASTRUCTURE ONLY (tests/, __graft_entry__.smoke(), bench.py's cpu_baseline / --impl reference
legs).  Takes the same host arrays (`pathtracing_b200.scenes.Scene`) the CUDA backend uploads.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import pt_oracle as _base


class Material(C.Structure):
    _fields_ = [("type", C.c_uint32), ("albedo", C.c_float * 3), ("fuzz", C.c_float), ("ior", C.c_float),
                ("pad", C.c_float * 2)]


class SceneC(C.Structure):
    _fields_ = [
        ("width", C.c_uint32), ("height", C.c_uint32),
        ("uv_div_x", C.c_float), ("uv_div_y", C.c_float),
        ("cam_o", C.c_float * 3), ("cam_u", C.c_float * 3), ("cam_v", C.c_float * 3), ("cam_w", C.c_float * 3),
        ("sky_horizon", C.c_float * 3), ("sky_zenith", C.c_float * 3),
        ("t_min", C.c_float), ("shade_mode", C.c_uint32), ("max_depth", C.c_uint32),
        ("n_spheres", C.c_uint32), ("spheres", C.c_void_p), ("sphere_mat", C.c_void_p),
        ("n_tris", C.c_uint32), ("tris", C.c_void_p), ("tri_mat", C.c_void_p),
        ("n_materials", C.c_uint32), ("materials", C.c_void_p),
        ("sun_on", C.c_uint32), ("sun_dir", C.c_float * 3), ("sun_rad", C.c_float * 3),
    ]


_bound = False


def lib() -> C.CDLL:
    global _bound
    L = _base.lib()
    if not _bound:
        P = C.POINTER
        L.pto_accel_build.argtypes = [P(SceneC), C.c_uint32]
        L.pto_accel_build.restype = C.c_void_p
        L.pto_accel_free.argtypes = [C.c_void_p]
        L.pto_accel_free.restype = None
        L.pto_trace_rays.argtypes = [P(SceneC), C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_float, C.c_float,
                                     C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.pto_trace_rays.restype = None
        L.pto_scene_frame.argtypes = [P(SceneC), C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.pto_scene_frame.restype = C.c_int64
        L.pto_scene_accumulate.argtypes = [P(SceneC), C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int,
                                           C.c_void_p, C.c_int]
        L.pto_scene_accumulate.restype = None
        L.pto_scene_accumulate_rows.argtypes = [P(SceneC), C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int,
                                                C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int]
        L.pto_scene_accumulate_rows.restype = None
        L.pto_scene_sample_path.argtypes = [P(SceneC), C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int,
                                            P(C.c_float), C.c_void_p, C.c_int]
        L.pto_scene_sample_path.restype = C.c_int
        L.pto_sincos2pi.argtypes = [C.c_float, P(C.c_float), P(C.c_float)]
        L.pto_sincos2pi.restype = None
        L.pto_unit_vector.argtypes = [C.c_float, C.c_float, P(C.c_float)]
        L.pto_unit_vector.restype = None
        L.pto_scatter.argtypes = [P(Material), P(C.c_float), P(C.c_float), C.c_float, C.c_float, C.c_float,
                                  P(C.c_float), P(C.c_float)]
        L.pto_scatter.restype = C.c_int
        _bound = True
    return L


class SceneOracle:
    """Holds the C view of a `scenes.Scene` for a given image size / bounce limit.  accel='bvh' builds the
    oracle's own CPU hierarchy (median split); accel=None is brute force over every primitive."""

    def __init__(self, scene, width: int, height: int, max_depth: int = 8, uv_div=None, accel="bvh", max_leaf: int = 4):
        self.scene = scene          # keeps the numpy arrays alive
        self.width, self.height = width, height
        c = SceneC()
        c.width, c.height = width, height
        c.uv_div_x, c.uv_div_y = (float(height), float(height)) if uv_div is None else (float(uv_div[0]), float(uv_div[1]))
        for name, src in (("cam_o", scene.cam_origin), ("cam_u", scene.cam_u), ("cam_v", scene.cam_v),
                          ("cam_w", scene.cam_w), ("sky_horizon", scene.sky_horizon), ("sky_zenith", scene.sky_zenith)):
            getattr(c, name)[:] = [float(v) for v in src]
        c.t_min = float(scene.t_min)
        c.shade_mode = int(scene.shade_mode)
        c.max_depth = max_depth
        c.n_spheres = scene.n_spheres
        c.spheres = scene.spheres.ctypes.data
        c.sphere_mat = scene.sphere_material.ctypes.data
        c.n_tris = scene.n_triangles
        c.tris = scene.triangles.ctypes.data
        c.tri_mat = scene.tri_material.ctypes.data
        c.n_materials = len(scene.materials)
        c.materials = scene.materials.ctypes.data
        c.sun_on = 1 if getattr(scene, "sun_on", False) else 0
        if c.sun_on:
            c.sun_dir[:] = [float(v) for v in scene.sun_dir]
            c.sun_rad[:] = [float(v) for v in scene.sun_radiance]
        self.c = c
        self._accel = lib().pto_accel_build(C.byref(c), max_leaf) if accel == "bvh" else None

    def close(self):
        if self._accel:
            lib().pto_accel_free(self._accel)
            self._accel = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def trace_rays(self, origins, dirs, t_min: float, t_max: float, any_hit: bool = False, brute: bool = False):
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        prim = np.empty(len(o), np.int32)
        t = np.zeros(len(o), np.float32)
        lib().pto_trace_rays(C.byref(self.c), None if brute else self._accel, len(o), o.ctypes.data, d.ctypes.data,
                             t_min, t_max, 1 if any_hit else 0, prim.ctypes.data, t.ctypes.data, 0)
        return (prim, None) if any_hit else (prim, t)

    def frame(self, seed: int = 0, brute: bool = False):
        h, w = self.height, self.width
        f32 = np.empty((h, w, 4), np.float32)
        rgba8 = np.empty((h, w, 4), np.uint8)
        hit = np.empty((h, w), np.int32)
        n = lib().pto_scene_frame(C.byref(self.c), None if brute else self._accel, seed, f32.ctypes.data,
                                  rgba8.ctypes.data, hit.ctypes.data, 0)
        return {"f32": f32, "rgba8": rgba8, "hit": hit, "hits": int(n)}

    def sample_path(self, px: int, py: int, sample: int, seed: int = 0, jitter: bool = True, brute: bool = False):
        """-> (rgb, segments (n, 8): origin, direction, t (inf = miss), primitive id) of one path (debugging aid)."""
        rgb = (C.c_float * 3)()
        segs = np.zeros((256, 8), np.float32)
        n = lib().pto_scene_sample_path(C.byref(self.c), None if brute else self._accel, px, py, sample, seed, int(jitter), rgb,
                                        segs.ctypes.data, 256)
        return np.array(rgb[:], np.float32), segs[:n]

    def accumulate(self, first_sample: int, n_samples: int, seed: int = 0, jitter: bool = True, accum=None,
                   brute: bool = False, nthreads: int = 0, rows=None):
        """rows = (row_offset, row_stride, rows_per_band): only that interleaved row shard (ptb_set_row_shard_bands)."""
        if accum is None:
            accum = np.zeros((self.height, self.width, 4), np.float32)
        assert accum.dtype == np.float32 and accum.flags.c_contiguous
        off, stride, block = rows if rows is not None else (0, 1, 1)
        lib().pto_scene_accumulate_rows(C.byref(self.c), None if brute else self._accel, first_sample, n_samples, seed,
                                        int(jitter), accum.ctypes.data, off, stride, block, nthreads)
        return accum


def sincos2pi(u: float):
    s, c = C.c_float(), C.c_float()
    lib().pto_sincos2pi(u, C.byref(s), C.byref(c))
    return s.value, c.value


def unit_vector(u1: float, u2: float):
    v = (C.c_float * 3)()
    lib().pto_unit_vector(u1, u2, v)
    return np.array(v[:], np.float32)


def scatter(material_row, d, n_out, u1, u2, u3):
    m = Material()
    m.type = int(material_row["type"])
    m.albedo[:] = [float(x) for x in material_row["albedo"]]
    m.fuzz, m.ior = float(material_row["fuzz"]), float(material_row["ior"])
    dd = (C.c_float * 3)(*[float(x) for x in d])
    nn = (C.c_float * 3)(*[float(x) for x in n_out])
    out, att = (C.c_float * 3)(), (C.c_float * 3)()
    ok = lib().pto_scatter(C.byref(m), dd, nn, u1, u2, u3, out, att)
    return bool(ok), np.array(out[:], np.float32), np.array(att[:], np.float32)

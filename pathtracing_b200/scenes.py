"""Scene description and the synthetic scene generators of BASELINE.json's configs.

The reference has no scene files or loader: its scene is the literals of
RayTracing/Assets/Shaders/Source/Ray/Test.hlsl:8,12-13 (camera (0,0,1), one sphere of radius 0.5 at
the origin) and its camera ray is Test.hlsl:7-10.  `default_scene()` restates exactly those literals;
`random_spheres()` and `triangle_terrain()` are the synthetic scenes SURVEY.md §8(d) freezes for
configs C3 and C4.  Everything here is HOST data (numpy): the same arrays go to the CUDA backend
through `ptb_upload_scene` and to the CPU oracle in the tests, so both see identical bits.
"""
from __future__ import annotations

import dataclasses
import math
from typing import Sequence

import numpy as np

MAT_LAMBERT, MAT_METAL, MAT_DIELECTRIC = 0, 1, 2
SHADE_PATH, SHADE_NORMALS = 0, 1

MATERIAL_DTYPE = np.dtype([("type", "<u4"), ("albedo", "<f4", (3,)), ("fuzz", "<f4"), ("ior", "<f4"),
                           ("pad", "<f4", (2,))])
assert MATERIAL_DTYPE.itemsize == 32


def _f3(v: Sequence[float]) -> np.ndarray:
    a = np.asarray(v, np.float32).reshape(3)
    return np.ascontiguousarray(a)


@dataclasses.dataclass
class Scene:
    spheres: np.ndarray                  # (ns, 4) float32: cx, cy, cz, r
    sphere_material: np.ndarray          # (ns,) uint32
    triangles: np.ndarray                # (nt, 9) float32: v0, v1, v2
    tri_material: np.ndarray             # (nt,) uint32
    materials: np.ndarray                # (nm,) MATERIAL_DTYPE
    cam_origin: np.ndarray = dataclasses.field(default_factory=lambda: _f3([0, 0, 1]))    # Test.hlsl:8
    cam_u: np.ndarray = dataclasses.field(default_factory=lambda: _f3([1, 0, 0]))         # Test.hlsl:10
    cam_v: np.ndarray = dataclasses.field(default_factory=lambda: _f3([0, 1, 0]))
    cam_w: np.ndarray = dataclasses.field(default_factory=lambda: _f3([0, 0, -1]))
    sky_horizon: np.ndarray = dataclasses.field(default_factory=lambda: _f3([1, 1, 1]))
    sky_zenith: np.ndarray = dataclasses.field(default_factory=lambda: _f3([0.5, 0.7, 1.0]))
    t_min: float = 1e-3
    shade_mode: int = SHADE_PATH
    name: str = "scene"
    # optional directional light sampled by next-event estimation at Lambert hits (ptb_scene_set_sun); radiance 0 = off.
    # sun_dir points TOWARDS the light and is used as given (set_sun() normalises).
    sun_dir: np.ndarray = dataclasses.field(default_factory=lambda: _f3([0, 1, 0]))
    sun_radiance: np.ndarray = dataclasses.field(default_factory=lambda: _f3([0, 0, 0]))

    def __post_init__(self):
        self.spheres = np.ascontiguousarray(self.spheres, np.float32).reshape(-1, 4)
        self.sphere_material = np.ascontiguousarray(self.sphere_material, np.uint32).reshape(-1)
        self.triangles = np.ascontiguousarray(self.triangles, np.float32).reshape(-1, 9)
        self.tri_material = np.ascontiguousarray(self.tri_material, np.uint32).reshape(-1)
        self.materials = np.ascontiguousarray(self.materials, MATERIAL_DTYPE).reshape(-1)
        assert len(self.spheres) == len(self.sphere_material)
        assert len(self.triangles) == len(self.tri_material)
        assert len(self.materials) >= 1
        for f in ("cam_origin", "cam_u", "cam_v", "cam_w", "sky_horizon", "sky_zenith", "sun_dir", "sun_radiance"):
            setattr(self, f, _f3(getattr(self, f)))

    @property
    def sun_on(self) -> bool:
        return bool(np.any(self.sun_radiance != 0))

    def set_sun(self, direction, radiance) -> "Scene":
        """Directional light: `direction` points towards it (normalised here in binary64, stored as float32)."""
        d = np.asarray(direction, np.float64).reshape(3)
        self.sun_dir = _f3(d / np.linalg.norm(d))
        self.sun_radiance = _f3(radiance)
        return self

    @property
    def n_spheres(self) -> int:
        return len(self.spheres)

    @property
    def n_triangles(self) -> int:
        return len(self.triangles)

    # -- scene files (the reference has none: its scene is literals in the shader) -------------
    _VEC_FIELDS = ("cam_origin", "cam_u", "cam_v", "cam_w", "sky_horizon", "sky_zenith", "sun_dir", "sun_radiance")

    def to_dict(self) -> dict:
        """JSON-able description (lists of Python floats; float32 values survive the round trip exactly)."""
        d = {"format": "ptb200-scene-1", "name": self.name, "t_min": float(self.t_min), "shade_mode": int(self.shade_mode),
             "spheres": self.spheres.astype(np.float64).tolist(), "sphere_material": self.sphere_material.tolist(),
             "triangles": self.triangles.astype(np.float64).tolist(), "tri_material": self.tri_material.tolist(),
             "materials": [{"type": int(m["type"]), "albedo": [float(x) for x in m["albedo"]], "fuzz": float(m["fuzz"]),
                            "ior": float(m["ior"])} for m in self.materials]}
        for f in self._VEC_FIELDS:
            d[f] = [float(x) for x in getattr(self, f)]
        return d

    @classmethod
    def from_dict(cls, d: dict) -> "Scene":
        if d.get("format") != "ptb200-scene-1":
            raise ValueError(f"not a ptb200 scene description: format={d.get('format')!r}")
        mats = materials_array((m["type"], m["albedo"], m.get("fuzz", 0.0), m.get("ior", 1.5)) for m in d["materials"])
        kw = {f: d[f] for f in cls._VEC_FIELDS if f in d}
        return cls(spheres=np.asarray(d.get("spheres", []), np.float32).reshape(-1, 4), sphere_material=d.get("sphere_material", []),
                   triangles=np.asarray(d.get("triangles", []), np.float32).reshape(-1, 9), tri_material=d.get("tri_material", []),
                   materials=mats, t_min=d.get("t_min", 1e-3), shade_mode=d.get("shade_mode", SHADE_PATH),
                   name=d.get("name", "scene"), **kw)

    def save(self, path: str) -> None:
        """`.json` -> text description; anything else -> numpy `.npz` (binary, for million-primitive scenes)."""
        if path.endswith(".json"):
            import json
            with open(path, "w") as f:
                json.dump(self.to_dict(), f)
            return
        np.savez_compressed(path, format=np.array("ptb200-scene-1"), name=np.array(self.name), t_min=np.float32(self.t_min),
                            shade_mode=np.uint32(self.shade_mode), spheres=self.spheres, sphere_material=self.sphere_material,
                            triangles=self.triangles, tri_material=self.tri_material, materials=self.materials,
                            **{f: getattr(self, f) for f in self._VEC_FIELDS})

    @classmethod
    def load(cls, path: str) -> "Scene":
        if path.endswith(".json"):
            import json
            with open(path) as f:
                return cls.from_dict(json.load(f))
        with np.load(path, allow_pickle=False) as z:
            if str(z["format"]) != "ptb200-scene-1":
                raise ValueError("not a ptb200 scene file")
            return cls(spheres=z["spheres"], sphere_material=z["sphere_material"], triangles=z["triangles"],
                       tri_material=z["tri_material"], materials=z["materials"], t_min=float(z["t_min"]),
                       shade_mode=int(z["shade_mode"]), name=str(z["name"]), **{f: z[f] for f in cls._VEC_FIELDS if f in z})

    def look_at(self, look_from, look_at, vup, vfov_deg: float, width: int, height: int) -> "Scene":
        """Pinhole camera in the reference's screen convention (Test.hlsl:7: uv = p/H*2-1 on BOTH axes, no
        y flip, so uv.x runs from -1 to 2W/H-1): U and V span the image plane, W is shifted so the view
        axis passes through the image centre, and +V points down the image."""
        lf, la, up = (np.asarray(v, np.float64) for v in (look_from, look_at, vup))
        fwd = la - lf
        fwd /= np.linalg.norm(fwd)
        right = np.cross(fwd, up)
        right /= np.linalg.norm(right)
        upv = np.cross(right, fwd)
        h = math.tan(math.radians(vfov_deg) / 2.0)
        aspect = width / height
        self.cam_origin = _f3(lf)
        self.cam_u = _f3(right * h)
        self.cam_v = _f3(-upv * h)
        self.cam_w = _f3(fwd - (aspect - 1.0) * right * h)
        return self


def materials_array(rows) -> np.ndarray:
    """rows: iterable of (type, (r,g,b), fuzz, ior)."""
    rows = list(rows)
    m = np.zeros(len(rows), MATERIAL_DTYPE)
    for i, (t, alb, fuzz, ior) in enumerate(rows):
        m[i]["type"] = t
        m[i]["albedo"] = alb
        m[i]["fuzz"] = fuzz
        m[i]["ior"] = ior
    return m


def default_scene() -> Scene:
    """The reference's literal scene (Test.hlsl:8,12-13) in the reference's shading (Test.hlsl:28-36)."""
    return Scene(spheres=[[0, 0, 0, 0.5]], sphere_material=[0], triangles=np.zeros((0, 9)), tri_material=[],
                 materials=materials_array([(MAT_LAMBERT, (0.5, 0.5, 0.5), 0.0, 1.5)]),
                 shade_mode=SHADE_NORMALS, name="default (Test.hlsl literals)")


def random_spheres(n: int = 500, seed: int = 1, width: int = 1920, height: int = 1080) -> Scene:
    """BASELINE configs[2]: n small spheres resting on a ground sphere, centres uniform
    in [-10,10]^2, radii U(0.2,0.6), materials i.i.d. Lambert 60 % / metal 25 % / dielectric 15 %,
    albedo U(0.1,0.9)^3, fuzz U(0,0.3), ior 1.5."""
    rng = np.random.Generator(np.random.PCG64(seed))
    r = rng.uniform(0.2, 0.6, n)
    x = rng.uniform(-10.0, 10.0, n)
    z = rng.uniform(-10.0, 10.0, n)
    sph = np.zeros((n + 1, 4), np.float32)
    sph[0] = (0.0, -1000.0, 0.0, 1000.0)          # ground
    sph[1:, 0], sph[1:, 1], sph[1:, 2], sph[1:, 3] = x, r, z, r
    kind = rng.uniform(0.0, 1.0, n)
    alb = rng.uniform(0.1, 0.9, (n, 3))
    fuzz = rng.uniform(0.0, 0.3, n)
    rows = [(MAT_LAMBERT, (0.5, 0.5, 0.5), 0.0, 1.5)]
    for i in range(n):
        if kind[i] < 0.60:
            rows.append((MAT_LAMBERT, alb[i], 0.0, 1.5))
        elif kind[i] < 0.85:
            rows.append((MAT_METAL, alb[i], fuzz[i], 1.5))
        else:
            rows.append((MAT_DIELECTRIC, (1.0, 1.0, 1.0), 0.0, 1.5))
    sc = Scene(spheres=sph, sphere_material=np.arange(n + 1), triangles=np.zeros((0, 9)), tri_material=[],
               materials=materials_array(rows), name=f"{n} random spheres (seed {seed})")
    return sc.look_at((0.0, 5.0, 19.0), (0.0, 0.5, 0.0), (0.0, 1.0, 0.0), 38.0, width, height)


def triangle_terrain(cells: int = 512, layers: int = 2, seed: int = 2, width: int = 3840, height: int = 2160) -> Scene:
    """BASELINE configs[3]: `layers` height fields of cells x cells quads, two triangles
    each (512 x 512 x 2 x 2 = 1,048,576 triangles), heights = a few random sine waves + white noise,
    upper layers are sparse floating patches.  Three materials by height band."""
    rng = np.random.Generator(np.random.PCG64(seed))
    n = cells + 1
    gx, gz = np.meshgrid(np.linspace(-10.0, 10.0, n), np.linspace(-10.0, 10.0, n), indexing="xy")
    tris, mats = [], []
    for layer in range(layers):
        hgt = np.zeros_like(gx)
        for _ in range(6):
            fx, fz = rng.uniform(0.15, 1.2, 2)
            ph = rng.uniform(0.0, 2.0 * math.pi, 2)
            hgt += rng.uniform(0.15, 0.6) * np.sin(fx * gx + ph[0]) * np.sin(fz * gz + ph[1])
        hgt += rng.uniform(-0.02, 0.02, hgt.shape)
        hgt += 3.0 * layer
        p = np.stack([gx, hgt, gz], axis=-1).astype(np.float32)
        p00, p10, p01, p11 = p[:-1, :-1], p[:-1, 1:], p[1:, :-1], p[1:, 1:]
        t0 = np.concatenate([p00, p01, p10], axis=-1).reshape(-1, 9)
        t1 = np.concatenate([p10, p01, p11], axis=-1).reshape(-1, 9)
        t = np.empty((2 * len(t0), 9), np.float32)
        t[0::2], t[1::2] = t0, t1
        if layer > 0:
            # shrink every upper-layer triangle about its centroid so the layer is a see-through lattice
            c = (t[:, 0:3] + t[:, 3:6] + t[:, 6:9]) / 3.0
            for k in range(3):
                t[:, 3 * k:3 * k + 3] = c + 0.45 * (t[:, 3 * k:3 * k + 3] - c)
        cy = (t[:, 1] + t[:, 4] + t[:, 7]) / 3.0 - 3.0 * layer
        m = np.where(cy > 0.35, 1, np.where(cy < -0.35, 2, 0)).astype(np.uint32)
        tris.append(t)
        mats.append(m)
    rows = [(MAT_LAMBERT, (0.55, 0.5, 0.4), 0.0, 1.5), (MAT_METAL, (0.85, 0.85, 0.9), 0.05, 1.5),
            (MAT_LAMBERT, (0.2, 0.45, 0.25), 0.0, 1.5)]
    sc = Scene(spheres=np.zeros((0, 4)), sphere_material=[], triangles=np.concatenate(tris),
               tri_material=np.concatenate(mats), materials=materials_array(rows),
               name=f"terrain {cells}x{cells}x{layers} ({2 * cells * cells * layers} triangles, seed {seed})")
    return sc.look_at((0.0, 9.0, 17.0), (0.0, 0.5, 0.0), (0.0, 1.0, 0.0), 42.0, width, height)


def mixed_small(seed: int = 7, n_spheres: int = 40, n_tris: int = 200, width: int = 160, height: int = 90) -> Scene:
    """Small mixed scene (spheres + random triangles, all three materials) for parity tests."""
    rng = np.random.Generator(np.random.PCG64(seed))
    sph = np.zeros((n_spheres + 1, 4), np.float32)
    sph[0] = (0.0, -100.0, 0.0, 100.0)
    sph[1:, 0] = rng.uniform(-4, 4, n_spheres)
    sph[1:, 3] = rng.uniform(0.15, 0.6, n_spheres)
    sph[1:, 1] = sph[1:, 3] + rng.uniform(0, 1.5, n_spheres)
    sph[1:, 2] = rng.uniform(-4, 4, n_spheres)
    c = rng.uniform(-4, 4, (n_tris, 1, 3)) * np.array([1, 0.3, 1]) + np.array([0, 1.6, 0])
    tri = (c + rng.uniform(-0.6, 0.6, (n_tris, 3, 3))).reshape(n_tris, 9).astype(np.float32)
    rows = [(MAT_LAMBERT, (0.5, 0.5, 0.5), 0.0, 1.5), (MAT_LAMBERT, (0.8, 0.3, 0.3), 0.0, 1.5),
            (MAT_METAL, (0.8, 0.8, 0.8), 0.0, 1.5), (MAT_METAL, (0.8, 0.6, 0.2), 0.25, 1.5),
            (MAT_DIELECTRIC, (1.0, 1.0, 1.0), 0.0, 1.5), (MAT_DIELECTRIC, (0.9, 1.0, 0.9), 0.0, 1.33)]
    sm = np.concatenate([[0], rng.integers(1, len(rows), n_spheres)]).astype(np.uint32)
    tm = rng.integers(1, len(rows), n_tris).astype(np.uint32)
    sc = Scene(spheres=sph, sphere_material=sm, triangles=tri, tri_material=tm, materials=materials_array(rows),
               name=f"mixed {n_spheres} spheres + {n_tris} triangles (seed {seed})")
    return sc.look_at((0.0, 2.5, 9.0), (0.0, 1.0, 0.0), (0.0, 1.0, 0.0), 45.0, width, height)


__all__ = ["Scene", "MATERIAL_DTYPE", "materials_array", "default_scene", "random_spheres", "triangle_terrain",
           "mixed_small", "MAT_LAMBERT", "MAT_METAL", "MAT_DIELECTRIC", "SHADE_PATH", "SHADE_NORMALS"]

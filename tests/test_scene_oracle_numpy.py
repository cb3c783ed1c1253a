"""An independent numpy float32 statement of the scene path's intersection spec (oracle/pt_scene_oracle.h: ray-sphere =
Test.hlsl:16-27 plus the far root, Moeller-Trumbore triangles, closest hit = smallest t then smaller primitive id),
vectorised over rays x primitives, against the C scene oracle — primitive ids and every bit of t.  No code shared with
oracle/pt_scene_oracle.c; one numpy float32 operation = one IEEE round-to-nearest operation."""
import numpy as np
import pytest

from oracle import pt_scene_oracle as S
from pathtracing_b200 import scenes

F = np.float32


def dot3(ax, ay, az, bx, by, bz):
    s = ax * bx
    s = s + ay * by
    s = s + az * bz
    return s


def closest_hit_numpy(sc, o, d, t_min, t_max):
    o = np.asarray(o, np.float32)
    d = np.asarray(d, np.float32)
    ox, oy, oz = (o[:, i:i + 1] for i in range(3))          # (R, 1) against (1, P)
    dx, dy, dz = (d[:, i:i + 1] for i in range(3))
    t_min, t_max = F(t_min), F(t_max)
    best_t = np.full(len(o), np.inf, np.float32)
    best_id = np.full(len(o), -1, np.int64)

    def merge(t, ok, ids):
        t = np.where(ok, t, F(np.inf))
        # smallest t, equal t -> smaller id: lexicographic minimum per ray
        order = np.lexsort((np.broadcast_to(ids, t.shape), t), axis=1)[:, 0]
        rows = np.arange(len(t))
        ct, cid, cok = t[rows, order], np.broadcast_to(ids, t.shape)[rows, order], ok[rows, order]
        take = cok & ((ct < best_t) | ((ct == best_t) & (cid < best_id)) | (best_id < 0))
        best_t[take] = ct[take]
        best_id[take] = cid[take]

    with np.errstate(all="ignore"):
        a = dot3(dx, dy, dz, dx, dy, dz)                                     # Test.hlsl:17, per ray
        ns = sc.n_spheres
        if ns:
            c = sc.spheres.astype(np.float32)
            cx, cy, cz, r = (c[None, :, i] for i in range(4))
            ocx, ocy, ocz = ox - cx, oy - cy, oz - cz                        # :16
            b = F(2) * dot3(ocx, ocy, ocz, dx, dy, dz)                       # :18
            cc = dot3(ocx, ocy, ocz, ocx, ocy, ocz) - r * r                  # :19
            disc = b * b - (F(4) * a) * cc                                   # :21
            pos = disc > F(0)                                                # :24
            s = np.sqrt(np.where(pos, disc, F(1)))
            a2 = F(2) * a
            t1 = ((-b) - s) / a2                                             # :27
            t2 = ((-b) + s) / a2                                             # far root when the near one is outside
            in1 = (t1 > t_min) & (t1 < t_max)
            in2 = (t2 > t_min) & (t2 < t_max)
            merge(np.where(in1, t1, t2), pos & (in1 | in2), np.arange(ns)[None, :])
        if sc.n_triangles:
            v = sc.triangles.astype(np.float32).reshape(-1, 3, 3)
            v0 = v[:, 0]
            e1 = v[:, 1] - v[:, 0]
            e2 = v[:, 2] - v[:, 0]
            v0x, v0y, v0z = (v0[None, :, i] for i in range(3))
            e1x, e1y, e1z = (e1[None, :, i] for i in range(3))
            e2x, e2y, e2z = (e2[None, :, i] for i in range(3))
            px = dy * e2z - dz * e2y
            py = dz * e2x - dx * e2z
            pz = dx * e2y - dy * e2x
            det = dot3(e1x, e1y, e1z, px, py, pz)
            ok = det != F(0)
            inv = F(1) / np.where(ok, det, F(1))
            tx, ty, tz = ox - v0x, oy - v0y, oz - v0z
            u = dot3(tx, ty, tz, px, py, pz) * inv
            ok &= (u >= F(0)) & (u <= F(1))
            qx = ty * e1z - tz * e1y
            qy = tz * e1x - tx * e1z
            qz = tx * e1y - ty * e1x
            w = dot3(dx, dy, dz, qx, qy, qz) * inv
            ok &= (w >= F(0)) & (u + w <= F(1))
            t = dot3(e2x, e2y, e2z, qx, qy, qz) * inv
            ok &= (t > t_min) & (t < t_max)
            merge(t, ok, (ns + np.arange(sc.n_triangles))[None, :])
    hit = best_id >= 0
    return np.where(hit, best_id, -1).astype(np.int32), np.where(hit, best_t, F(0)).astype(np.float32)


def rays(n, seed, extent=5.0):
    rng = np.random.Generator(np.random.PCG64(seed))
    o = rng.uniform(-extent, extent, (n, 3)).astype(np.float32)
    o[:, 1] = np.abs(o[:, 1]) + 0.2
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d[: n // 8] /= np.linalg.norm(d[: n // 8], axis=1, keepdims=True)      # unit and non-unit directions
    d[n // 8: n // 6] = np.array([0, -1, 0], np.float32)                   # axis-parallel
    return o, d.astype(np.float32)


@pytest.mark.parametrize("name", ["mixed", "spheres", "terrain", "tie"])
def test_scene_oracle_closest_hit_equals_numpy_restatement(name):
    if name == "tie":      # five coincident triangles: equal t, the smallest id must win
        tri = np.array([[-1, -1, 0, 1, -1, 0, 0, 1, 0]], np.float32)
        sc = scenes.Scene(spheres=np.zeros((0, 4)), sphere_material=[], triangles=np.tile(tri, (5, 1)), tri_material=[0] * 5,
                          materials=scenes.materials_array([(0, (0.5, 0.5, 0.5), 0, 1.5)]))
        o = np.array([[0, 0, 1], [0.1, -0.2, 3], [5, 5, 1]], np.float32)
        d = np.array([[0, 0, -1], [0, 0, -1], [0, 0, -1]], np.float32)
    else:
        sc = {"mixed": lambda: scenes.mixed_small(n_spheres=60, n_tris=300),
              "spheres": lambda: scenes.random_spheres(150),
              "terrain": lambda: scenes.triangle_terrain(cells=10, layers=2)}[name]()
        o, d = rays(3000, 21)
    so = S.SceneOracle(sc, 8, 8)
    for t_min, t_max in ((1e-3, 1e30), (0.5, 4.0)):
        prim, t = so.trace_rays(o, d, t_min, t_max, brute=True)
        nprim, nt = closest_hit_numpy(sc, o, d, t_min, t_max)
        assert np.array_equal(prim, nprim), name
        assert np.array_equal(t.view(np.uint32), nt.view(np.uint32)), name
        assert (prim >= 0).any()

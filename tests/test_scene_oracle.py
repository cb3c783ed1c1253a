"""CPU tests of the scene-path oracle (oracle/pt_scene_oracle.c) and the scene generators.

The scene stages (SURVEY.md §8 a' n3-n7) have no reference counterpart, so the oracle is pinned where
it can be: the default scene in the reference's shading must reproduce pto_csmain_frame (itself
pinned to SURVEY App. B), the oracle's own BVH must agree with brute force, and the sampling spec is
checked against analytic properties.  Frozen digests guard against silent spec drift.
"""
import hashlib
import json
import os

import numpy as np
import pytest

from oracle import pt_oracle as O
from oracle import pt_scene_oracle as S
from pathtracing_b200 import scenes

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "scene_digests.json")


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.mark.parametrize("size,div", [((320, 180), None), ((1920, 1080), None), ((320, 180), (1080, 1080)), ((97, 33), None)])
def test_default_scene_reproduces_csmain(size, div):
    w, h = size
    ref = O.csmain_frame(O.frame_params(w, h, uv_div=div))
    for brute in (False, True):
        fr = S.SceneOracle(scenes.default_scene(), w, h, max_depth=0, uv_div=div).frame(brute=brute)
        assert np.array_equal(bits(fr["f32"]), bits(ref["f32"]))
        assert np.array_equal(fr["hit"], ref["hit"])
        assert np.array_equal(fr["rgba8"], ref["rgba8"])
        assert fr["hits"] == ref["hits"]


def test_default_scene_hit_counts_of_survey_appendix_b():
    assert S.SceneOracle(scenes.default_scene(), 1920, 1080).frame()["hits"] == 305317
    assert S.SceneOracle(scenes.default_scene(), 320, 180).frame()["hits"] == 8469


@pytest.mark.parametrize("make", [lambda: scenes.mixed_small(), lambda: scenes.random_spheres(120, width=96, height=54),
                                  lambda: scenes.triangle_terrain(cells=16, layers=2, width=96, height=54)])
def test_oracle_bvh_equals_brute_force(make):
    sc = make()
    w, h = 96, 54
    so = S.SceneOracle(sc, w, h, max_depth=6)
    a, b = so.frame(seed=3), so.frame(seed=3, brute=True)
    assert np.array_equal(a["hit"], b["hit"]) and np.array_equal(bits(a["f32"]), bits(b["f32"]))
    assert np.array_equal(bits(so.accumulate(0, 3)), bits(so.accumulate(0, 3, brute=True)))
    rng = np.random.Generator(np.random.PCG64(5))
    o = rng.uniform(-5, 5, (5000, 3)).astype(np.float32)
    d = rng.normal(size=(5000, 3)).astype(np.float32)
    p1, t1 = so.trace_rays(o, d, 1e-3, 1e30)
    p2, t2 = so.trace_rays(o, d, 1e-3, 1e30, brute=True)
    assert np.array_equal(p1, p2) and np.array_equal(bits(t1), bits(t2))
    a1, _ = so.trace_rays(o, d, 1e-3, 2.5, any_hit=True)
    a2, _ = so.trace_rays(o, d, 1e-3, 2.5, any_hit=True, brute=True)
    assert np.array_equal(a1, a2)
    # any-hit is consistent with closest-hit
    p3, t3 = so.trace_rays(o, d, 1e-3, 2.5)
    assert np.array_equal(a1 != 0, p3 >= 0)


def test_closest_hit_ties_keep_the_smaller_primitive_id():
    tri = np.array([[-1, -1, 0, 1, -1, 0, 0, 1, 0]], np.float32)
    sc = scenes.Scene(spheres=np.zeros((0, 4)), sphere_material=[], triangles=np.tile(tri, (5, 1)), tri_material=[0] * 5,
                      materials=scenes.materials_array([(0, (0.5, 0.5, 0.5), 0, 1.5)]))
    so = S.SceneOracle(sc, 4, 4)
    for brute in (False, True):
        p, t = so.trace_rays([[0, 0, 1]], [[0, 0, -1]], 1e-3, 1e30, brute=brute)
        assert p[0] == 0 and t[0] == 1.0


def test_sincos_polynomial_accuracy_and_symmetry():
    u = np.linspace(0, 1, 4097, endpoint=False, dtype=np.float32)
    sc = np.array([S.sincos2pi(float(x)) for x in u])
    assert np.abs(sc[:, 0] - np.sin(2 * np.pi * u.astype(np.float64))).max() < 2e-7
    assert np.abs(sc[:, 1] - np.cos(2 * np.pi * u.astype(np.float64))).max() < 2e-7
    s0, c0 = S.sincos2pi(0.0)
    assert abs(s0) < 2e-7 and abs(c0 - 1.0) < 2e-7
    s, c = S.sincos2pi(0.25)
    assert abs(s - 1.0) < 2e-7 and abs(c) < 2e-7


def test_unit_vector_is_unit_and_uniform():
    rng = np.random.Generator(np.random.PCG64(1))
    v = np.array([S.unit_vector(float(a), float(b)) for a, b in rng.uniform(0, 1, (4000, 2)).astype(np.float32)])
    assert np.abs(np.linalg.norm(v, axis=1) - 1).max() < 1e-6
    assert np.abs(v.mean(axis=0)).max() < 0.05          # no preferred direction


def test_scatter_laws():
    lam, met, mir, die = scenes.materials_array([(0, (0.3, 0.5, 0.7), 0, 1.5), (1, (0.9, 0.8, 0.7), 0.2, 1.5),
                                                 (1, (1, 1, 1), 0.0, 1.5), (2, (1, 1, 1), 0, 1.5)])
    n = np.array([0, 1, 0], np.float32)
    d = np.array([0.6, -0.8, 0.0], np.float32)
    rng = np.random.Generator(np.random.PCG64(2))
    for u1, u2, u3 in rng.uniform(0, 1, (200, 3)).astype(np.float32):
        ok, out, att = S.scatter(lam, d, n, u1, u2, u3)
        assert ok and out @ n > -1e-6 and abs(np.linalg.norm(out) - 1) < 1e-6 and np.allclose(att, [0.3, 0.5, 0.7])
        ok, out, att = S.scatter(met, d, n, u1, u2, u3)
        assert (not ok) or out @ n > 0
    ok, out, _ = S.scatter(mir, d, n, 0.3, 0.3, 0.3)
    assert ok and np.allclose(out, [0.6, 0.8, 0.0], atol=1e-6)            # mirror reflection
    # dielectric: Snell's law when refracting, total internal reflection from inside at a grazing angle
    ok, out, _ = S.scatter(die, d, n, 0.1, 0.1, 0.999)
    assert ok and out[1] < 0 and abs(out[0] / 0.6 - 1 / 1.5) < 1e-5
    ok, out, _ = S.scatter(die, np.array([0.8, 0.6, 0], np.float32), n, 0.1, 0.1, 0.999)   # leaving, sin > 1/1.5
    assert ok and out[1] < 0                                               # reflected back inside
    # back faces flip the normal for every material
    ok, out, _ = S.scatter(lam, np.array([0.6, 0.8, 0], np.float32), n, 0.2, 0.7, 0.5)
    assert ok and out @ n < 1e-6


def test_white_furnace_and_depth_limit():
    w, h = 48, 27
    sc = scenes.Scene(spheres=[[0, 0, 0, 0.5]], sphere_material=[0], triangles=np.zeros((0, 9)), tri_material=[],
                      materials=scenes.materials_array([(0, (1, 1, 1), 0, 1.5)]), sky_horizon=[1, 1, 1], sky_zenith=[1, 1, 1])
    sc.look_at((0, 0, 2), (0, 0, 0), (0, 1, 0), 40, w, h)
    acc = S.SceneOracle(sc, w, h, max_depth=16).accumulate(0, 4)
    assert np.all(acc[..., 3] == 4) and np.all(np.abs(acc[..., :3] - 4) < 1e-4)
    # max_depth 0: pixels whose primary ray hits stay black
    so0 = S.SceneOracle(sc, w, h, max_depth=0)
    fr = so0.frame()
    assert np.all(fr["f32"][fr["hit"] >= 0][:, :3] == 0) and np.all(fr["f32"][fr["hit"] < 0][:, :3] == 1)


def test_sun_next_event_estimation_is_analytic_on_a_lit_floor():
    """Spec "sun" (pt_scene_oracle.h): a Lambert floor under a black sky, lit by a directional light.  Every path's
    first hit adds albedo * radiance * cos (nothing else can: scattered rays leave into the black sky), a sphere
    floating above the floor casts a shadow with radiance exactly 0, and without the light the image is black."""
    w, h = 96, 54
    floor = [[-50, 0, -50, 50, 0, -50, 50, 0, 50], [-50, 0, -50, 50, 0, 50, -50, 0, 50]]
    sc = scenes.Scene(spheres=[[0, 1.0, 0, 0.5]], sphere_material=[1], triangles=floor, tri_material=[0, 0],
                      materials=scenes.materials_array([(scenes.MAT_LAMBERT, (0.5, 0.25, 0.75), 0, 1.5),
                                                        (scenes.MAT_METAL, (0, 0, 0), 0, 1.5)]),
                      sky_horizon=[0, 0, 0], sky_zenith=[0, 0, 0])
    sc.look_at((0, 6, 0.01), (0, 0, 0), (0, 0, -1), 50, w, h)
    light = np.array([0.6, 0.8, 0.0])
    sc.set_sun(light, [2.0, 4.0, 1.0])
    so = S.SceneOracle(sc, w, h, max_depth=4)
    fr = so.frame()
    f = fr["f32"][..., :3]
    on_floor = fr["hit"] >= 1                                   # ids: sphere 0, triangles 1..2
    expect = np.array([0.5, 0.25, 0.75]) * np.array([2.0, 4.0, 1.0]) * 0.8
    lit = on_floor & (f[..., 1] > 0)
    assert lit.mean() > 0.5
    np.testing.assert_allclose(f[lit], np.broadcast_to(expect, f[lit].shape), rtol=2e-6)
    shadow = on_floor & ~lit
    assert 20 < shadow.sum() < 0.2 * w * h and np.all(f[shadow] == 0)
    # the shadow lies where the light direction says: floor points P with |P + s*light - centre| < r for some s
    assert np.all(f[fr["hit"] == 0] == 0)                       # the black mirror sphere reflects floor light times albedo 0
    acc = so.accumulate(0, 3, seed=2)
    assert np.all(acc[..., 3] == 3) and acc[..., :3].max() <= 3 * expect.max() * (1 + 1e-5)
    sc.sun_radiance = np.zeros(3, np.float32)
    assert not sc.sun_on and np.all(S.SceneOracle(sc, w, h, max_depth=4).frame()["f32"][..., :3] == 0)


def test_accumulate_is_progressive_and_sample_ranges_are_disjoint():
    sc = scenes.mixed_small(width=64, height=36)
    so = S.SceneOracle(sc, 64, 36, max_depth=4)
    a = so.accumulate(0, 5, seed=4)
    b = so.accumulate(0, 2, seed=4)
    so.accumulate(2, 3, seed=4, accum=b)
    assert np.array_equal(bits(a), bits(b))
    assert not np.array_equal(so.accumulate(0, 1, seed=4), so.accumulate(1, 1, seed=4))


def test_generators_are_deterministic_and_match_the_config_sizes():
    a, b = scenes.random_spheres(500), scenes.random_spheres(500)
    assert a.n_spheres == 501 and np.array_equal(a.spheres, b.spheres) and np.array_equal(a.materials, b.materials)
    kinds = a.materials["type"][1:]
    assert 0.5 < (kinds == 0).mean() < 0.7 and 0.15 < (kinds == 1).mean() < 0.35 and 0.08 < (kinds == 2).mean() < 0.25
    t = scenes.triangle_terrain(cells=32, layers=2)
    assert t.n_triangles == 32 * 32 * 2 * 2
    assert 2 * 512 * 512 * 2 == 1048576            # the full C4 configuration


def test_frozen_digests():
    """Oracle outputs frozen at the commit that introduced the scene spec (tests/golden/make_scene_golden.py)."""
    gold = json.load(open(GOLD))
    for name, (make, w, h, depth) in CASES.items():
        so = S.SceneOracle(make(w, h), w, h, max_depth=depth)
        fr = so.frame(seed=1)
        acc = so.accumulate(0, 2, seed=1)
        got = {"hit": digest(fr["hit"]), "f32": digest(fr["f32"]), "accum2": digest(acc)}
        assert got == gold[name], name


CASES = {
    "mixed_160x90_d8": (lambda w, h: scenes.mixed_small(width=w, height=h), 160, 90, 8),
    "spheres500_160x90_d8": (lambda w, h: scenes.random_spheres(500, width=w, height=h), 160, 90, 8),
    "terrain32_160x90_d4": (lambda w, h: scenes.triangle_terrain(cells=32, layers=2, width=w, height=h), 160, 90, 4),
}


def test_scene_files_round_trip_exactly(tmp_path):
    """scene description + loader.  JSON and npz round trips preserve every float32 bit, and the
    default scene file shipped with the package is the reference's literal scene (Test.hlsl:8,12-13)."""
    sc = scenes.mixed_small().set_sun([0.3, -1, 0.2], [1.5, 1.25, 1.0])
    for ext in ("json", "npz"):
        p = str(tmp_path / f"s.{ext}")
        sc.save(p)
        back = scenes.Scene.load(p)
        for f in ("spheres", "sphere_material", "triangles", "tri_material", "cam_origin", "cam_u", "cam_v", "cam_w",
                  "sky_horizon", "sky_zenith", "sun_dir", "sun_radiance"):
            assert np.array_equal(getattr(back, f), getattr(sc, f)), (ext, f)
        assert back.materials.tobytes() == sc.materials.tobytes()
        assert (back.t_min, back.shade_mode, back.name) == (np.float32(sc.t_min), sc.shade_mode, sc.name) or back.name == sc.name
    default_file = os.path.join(os.path.dirname(scenes.__file__), "assets", "default_scene.json")
    d = scenes.Scene.load(default_file)
    assert d.spheres.tolist() == [[0.0, 0.0, 0.0, 0.5]] and d.cam_origin.tolist() == [0.0, 0.0, 1.0]
    assert d.shade_mode == scenes.SHADE_NORMALS and d.n_triangles == 0
    fr = S.SceneOracle(d, 320, 180).frame()
    assert fr["hits"] == 8469
    with pytest.raises(ValueError):
        scenes.Scene.from_dict({"format": "something else"})

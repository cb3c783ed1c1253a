"""CPU tests: the oracle against SURVEY.md Appendix B and its own frozen goldens."""
import hashlib
import json
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def bits(f):
    return int(np.float32(f).view(np.uint32))


with open(os.path.join(GOLD, "appendix_b.json")) as f:
    APPB = json.load(f)
with open(os.path.join(GOLD, "oracle_digests.json")) as f:
    DIG = json.load(f)


@pytest.mark.parametrize("case", APPB["aggregate"], ids=lambda c: f"{c['width']}x{c['height']}_div{c['div']}")
def test_appendix_b_hit_counts(oracle, case):
    p = oracle.frame_params(case["width"], case["height"], uv_div=(case["div"], case["div"]))
    r = oracle.csmain_frame(p, want_f32=False, want_rgba8=False)
    assert r["hits"] == case["hits"]
    assert int((r["hit"] == 0).sum()) == case["hits"]
    if case["bbox"]:
        ys, xs = np.nonzero(r["hit"] == 0)
        assert [xs.min(), xs.max(), ys.min(), ys.max()] == case["bbox"]


@pytest.mark.parametrize("px", APPB["pixels"], ids=lambda c: f"div{c['div']}_{c['x']}_{c['y']}")
def test_appendix_b_pixel_bits(oracle, px):
    # divisor 1080 <-> 1920x1080 frame; divisor 180 <-> 320x180 frame
    w, h = (1920, 1080) if px["div"] == 1080 else (320, 180)
    p = oracle.frame_params(w, h)
    hit, col, disc, t = oracle.csmain_sample(p, float(px["x"]), float(px["y"]))
    assert hit == px["hit"]
    assert bits(disc) == int(px["disc"], 16)
    if px["t"]:
        assert bits(t) == int(px["t"], 16)
    assert [bits(c) for c in col[:3]] == [int(v, 16) for v in px["rgb"]]
    assert bits(col[3]) == 0x3F800000


def test_centre_pixel_exact(oracle):
    # pixel (540,540) of the 1080p frame is exactly (0.5, 0.5, 1, 1)
    p = oracle.frame_params(1920, 1080)
    hit, col, _, t = oracle.csmain_sample(p, 540.0, 540.0)
    assert hit == 1 and t == np.float32(0.5)
    assert col.tolist() == [0.5, 0.5, 1.0, 1.0]


@pytest.mark.parametrize("name", sorted(DIG["frames"]))
def test_frame_digests(oracle, name):
    c = DIG["frames"][name]
    p = oracle.frame_params(c["width"], c["height"], uv_div=c["uv_div"], **c["params"])
    r = oracle.csmain_frame(p)
    assert r["hits"] == c["hits"]
    assert sha(r["f32"]) == c["f32"]
    assert sha(r["rgba8"]) == c["rgba8"]
    assert sha(r["hit"]) == c["hit"]


def test_small_fixture(oracle):
    z = np.load(os.path.join(GOLD, "c1_small.npz"))
    r = oracle.csmain_frame(oracle.frame_params(64, 36))
    assert np.array_equal(r["f32"].view(np.uint32), z["f32"].view(np.uint32))
    assert np.array_equal(r["rgba8"], z["rgba8"])
    assert np.array_equal(r["hit"], z["hit"])


def test_thread_count_invariance(oracle):
    p = oracle.frame_params(320, 180)
    a = oracle.csmain_frame(p, nthreads=1)
    b = oracle.csmain_frame(p, nthreads=0)
    assert np.array_equal(a["f32"].view(np.uint32), b["f32"].view(np.uint32))


def test_unorm8_store_conversion(oracle):
    L = oracle.lib()
    assert L.pto_unorm8(float("nan")) == 0
    assert L.pto_unorm8(-1.0) == 0 and L.pto_unorm8(-0.0) == 0
    assert L.pto_unorm8(1.0) == 255 and L.pto_unorm8(2.55) == 255 and L.pto_unorm8(float("inf")) == 255
    assert L.pto_unorm8(0.5) == 128          # 127.5 -> even
    assert L.pto_unorm8(2.5 / 255.0) in (2, 3)
    # matches the image path
    p = oracle.frame_params(320, 180)
    r = oracle.csmain_frame(p)
    ref = np.rint(np.clip(np.nan_to_num(r["f32"]), 0, 1) * np.float32(255)).astype(np.uint8)
    assert np.array_equal(ref, r["rgba8"])


def test_srgb8_monotone_and_endpoints(oracle):
    L = oracle.lib()
    xs = np.linspace(0, 1, 4097, dtype=np.float32)
    codes = np.array([L.pto_srgb8(float(x)) for x in xs])
    assert codes[0] == 0 and codes[-1] == 255
    assert np.all(np.diff(codes.astype(int)) >= 0)
    assert L.pto_srgb8(0.5) == 188           # standard sRGB: 0.5 linear -> 188
    assert L.pto_srgb8(0.0031308) == round(12.92 * 0.0031308 * 255)


def np_pcg3d(x, y, z):
    """Independent numpy restatement of pcg3d (Jarzynski & Olano, JCGT 2020, listing 'pcg3d')."""
    M = 0xFFFFFFFF
    x = (x * 1664525 + 1013904223) & M
    y = (y * 1664525 + 1013904223) & M
    z = (z * 1664525 + 1013904223) & M
    x = (x + y * z) & M; y = (y + z * x) & M; z = (z + x * y) & M
    x ^= x >> 16; y ^= y >> 16; z ^= z >> 16
    x = (x + y * z) & M; y = (y + z * x) & M; z = (z + x * y) & M
    return x, y, z


def test_rng_known_answers(oracle):
    for g in DIG["rng"]["pcg3d"]:
        out = oracle.pcg3d(*g["key"])
        assert list(out) == g["out"] == list(np_pcg3d(*g["key"]))
        assert [bits(oracle.bits_to_unit(w)) for w in out] == g["unit_bits"]
    # unit mapping: exactly the top 23 bits * 2^-23, in [0,1)
    for w in (0, 1, 511, 512, 0x80000000, 0xFFFFFFFF):
        u = oracle.bits_to_unit(w)
        assert u == np.float32(w >> 9) * np.float32(2.0 ** -23) and 0.0 <= u < 1.0


def test_rng_uniformity(oracle):
    # crude: jitter means over many pixels ~ 0.5; neighbouring pixels / samples / axes decorrelated
    v = np.array([[float(oracle.bits_to_unit(w)) for w in oracle.pcg3d(pix, 0, 0)[:2]] for pix in range(4000)])
    assert abs(v[:, 0].mean() - 0.5) < 0.02 and abs(v[:, 1].mean() - 0.5) < 0.02
    assert abs(np.corrcoef(v[:-1, 0], v[1:, 0])[0, 1]) < 0.06
    assert abs(np.corrcoef(v[:, 0], v[:, 1])[0, 1]) < 0.06
    s = np.array([[float(oracle.bits_to_unit(w)) for w in oracle.pcg3d(777, k, 5)[:2]] for k in range(4000)])
    assert abs(s[:, 0].mean() - 0.5) < 0.02 and abs(np.corrcoef(s[:-1, 0], s[1:, 0])[0, 1]) < 0.06
    # 8x8 stratification counts of one pixel's 4096 jitters are roughly flat
    j = np.array([[float(oracle.bits_to_unit(w)) for w in oracle.pcg3d(31337, k, 1)[:2]] for k in range(4096)])
    hist, _, _ = np.histogram2d(j[:, 0], j[:, 1], bins=8, range=[[0, 1], [0, 1]])
    assert hist.min() > 30 and hist.max() < 100          # expectation 64


@pytest.mark.parametrize("name", sorted(DIG["accum"]))
def test_accum_digests(oracle, name):
    c = DIG["accum"][name]
    p = oracle.frame_params(c["width"], c["height"])
    acc = oracle.accumulate(p, c["first"], c["n"], c["seed"], c["jitter"])
    assert sha(acc) == c["accum"]
    assert sha(oracle.resolve(acc, c["n"])) == c["resolve"]
    assert sha(oracle.resolve(acc, c["n"], srgb=True, bgra=True)) == c["resolve_srgb_bgra"]


def test_accumulate_without_jitter_reduces_to_csmain(oracle):
    # n1 with jitter disabled must reduce exactly to a1-a5 (SURVEY §8a')
    p = oracle.frame_params(320, 180)
    frame = oracle.csmain_frame(p)
    acc = oracle.accumulate(p, 0, 1, 0, False)
    assert np.array_equal(acc.view(np.uint32), frame["f32"].view(np.uint32))
    assert np.array_equal(oracle.resolve(acc, 1), frame["rgba8"])


def test_accumulate_split_launches(oracle):
    # partial sums are per launch: two launches of n/2 == documented grouping, and sample
    # indices are absolute (first_sample), so the same samples are drawn.
    p = oracle.frame_params(96, 54)
    a = oracle.accumulate(p, 0, 4, 3, True)
    b = oracle.accumulate(p, 0, 2, 3, True)
    b = oracle.accumulate(p, 2, 2, 3, True, accum=b)
    assert np.allclose(a, b, rtol=1e-6, atol=1e-6)
    # jittered mean converges towards the un-jittered image away from the silhouette
    frame = oracle.csmain_frame(p)["f32"]
    mean = oracle.accumulate(p, 0, 64, 0, True) / np.float32(64)
    assert np.median(np.abs(mean - frame)) < 0.02


def test_present_digests_and_identity(oracle):
    g = DIG["present"]
    img = oracle.csmain_frame(oracle.frame_params(320, 180))["rgba8"]
    assert sha(oracle.present(img, 320, 180, srgb=True)) == g["same_size_srgb"]
    assert sha(oracle.present(img, 500, 333, srgb=True)) == g["up_500x333_srgb"]
    assert sha(oracle.present(img, 97, 41, srgb=False)) == g["down_97x41_linear"]
    lin = oracle.present(img, 320, 180, srgb=False)
    assert sha(lin) == g["same_size_linear"]
    # same size, no OETF: pure RGBA -> BGRA swizzle
    assert np.array_equal(lin[..., [2, 1, 0, 3]], img)

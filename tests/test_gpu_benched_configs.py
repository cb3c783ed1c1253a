"""GPU parity at the configurations bench.py actually times (VERDICT r01 "next" #1): the CUDA path through the C ABI
against the CPU oracle, bit for bit, at the benched image size AND the benched samples per launch.

  c2   1920x1080, seed 0, 64 jittered spp per launch through ptb_render_resolve (first = 0 and first = 64*24, the
       25th bench step) and through ptb_render_samples + ptb_resolve: float4 accumulation and RGBA8 image.
  c5   3840x2160, 128 spp per launch = one rank's share of the 1024-spp frame on 8 GPUs.
  c3   500 spheres at 1920x1080, depth 8, one 64-sample batched wave (the bench's wave size) on a band of rows.
  c4   1,048,576 triangles at 3840x2160, depth 8, a 64-sample batched wave on one rank's interleaved band shard.

These are also the end-to-end check of csmain_accumulate's per-pixel certain-miss shortcut on the jitters the RNG really
produces at 64+ spp (the exhaustive self-test only probes an 8x8 grid of jitter positions per pixel).
"""
import numpy as np
import pytest

from pathtracing_b200 import NativeTracer, scenes

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def assert_bits(gpu, ref, what):
    if not np.array_equal(bits(gpu), bits(ref)):
        bad = np.argwhere(bits(gpu) != bits(ref))
        raise AssertionError(f"{what}: {len(bad)} differing floats, first at {bad[0].tolist()}: "
                             f"gpu {gpu[tuple(bad[0])]!r} vs oracle {ref[tuple(bad[0])]!r}")


@pytest.mark.parametrize("first", [0, 64 * 24])
def test_c2_benched_launch_fused_render_resolve(oracle, first):
    """bench.py's N=1 step: ptb_render_resolve(first, 64) at 1920x1080, seed 0."""
    w, h, n = 1920, 1080, 64
    p = oracle.frame_params(w, h)
    ref = oracle.accumulate(p, first, n, 0, True)
    with NativeTracer(w, h, seed=0) as t:
        assert t.kernel_mode == 2
        t.render_resolve(first, n)
        img, acc = t.read_rgba8(), t.read_accum()
    assert_bits(acc, ref, f"c2 accumulation, samples [{first}, {first + n})")
    assert np.all(acc[..., 3] == float(n))
    assert np.array_equal(img, oracle.resolve(ref, n))


def test_c2_benched_launch_two_calls_and_progressive(oracle):
    """The same launch through ptb_render_samples + ptb_resolve, then a second 64-spp step on top (fused), as the bench
    accumulates step after step; sRGB/BGRA resolves of the benched accumulation as well."""
    w, h, n = 1920, 1080, 64
    p = oracle.frame_params(w, h)
    ref = oracle.accumulate(p, 0, n, 0, True)
    with NativeTracer(w, h, seed=0) as t:
        t.render_samples(0, n)
        t.resolve()
        assert_bits(t.read_accum(), ref, "c2 accumulation (two calls)")
        assert np.array_equal(t.read_rgba8(), oracle.resolve(ref, n))
        t.resolve(srgb=True, bgra=True)
        assert np.array_equal(t.read_rgba8(), oracle.resolve(ref, n, srgb=True, bgra=True))
        t.render_resolve(n, n)
        ref = oracle.accumulate(p, n, n, 0, True, accum=ref)
        assert_bits(t.read_accum(), ref, "c2 accumulation after the second step")
        assert np.array_equal(t.read_rgba8(), oracle.resolve(ref, 2 * n))


def test_c2_benched_launch_ieee_kernels_agree(oracle):
    """The literal IEEE-intrinsic kernels (no shortened sequences, no certain-miss shortcut) on the same launch."""
    w, h, n = 1920, 1080, 64
    ref = oracle.accumulate(oracle.frame_params(w, h), 0, n, 0, True)
    with NativeTracer(w, h, seed=0, force_ieee=True) as t:
        assert t.kernel_mode == 0
        t.render_samples(0, n)
        assert_bits(t.read_accum(), ref, "c2 accumulation, IEEE kernels")


@pytest.mark.parametrize("first", [0, 128 * 5])
def test_c5_one_rank_share_4k_128spp(oracle, first):
    """One rank's share of BASELINE configs[4] on 8 GPUs: 3840x2160, 128 spp in one launch (rank 5's range too)."""
    w, h, n = 3840, 2160, 128
    ref = oracle.accumulate(oracle.frame_params(w, h), first, n, 0, True)
    with NativeTracer(w, h, seed=0) as t:
        t.render_samples(first, n)
        acc = t.read_accum()
        t.resolve()
        img = t.read_rgba8()
    assert_bits(acc, ref, f"c5 share, samples [{first}, {first + n})")
    assert np.array_equal(img, oracle.resolve(ref, n))


def test_c3_full_size_64_sample_batched_wave_on_a_band():
    """BASELINE configs[2] at 1920x1080, depth 8: ONE batched wave of 64 samples per pixel (what bench.py --workload c3
    launches) on the interleaved band shard (7, 27, 4) = 40 rows spread over the whole frame, vs the scene oracle."""
    from oracle import pt_scene_oracle as S
    w, h, n = 1920, 1080, 64
    rows = (7, 27, 4)
    sc = scenes.random_spheres(500, seed=1, width=w, height=h)
    so = S.SceneOracle(sc, w, h, max_depth=8)
    ref = so.accumulate(0, n, seed=0, rows=rows)
    with NativeTracer(w, h, seed=0, max_depth=8) as t:
        t.upload_scene(sc)
        t.set_row_shard(*rows)
        t.render_samples(0, n)
        q = t.scene_queue_sizes()
        acc = t.read_accum()
    mine = (np.arange(h) // rows[2]) % rows[1] == rows[0]
    assert q[0] == int(mine.sum()) * w * n              # one wave carried all 64 samples of the shard
    assert np.all(acc[mine][..., 3] == float(n)) and not acc[~mine].any()
    assert_bits(acc, ref, "c3 64-sample batched wave")


def test_c3_full_frame_16_sample_batched_wave():
    """Every pixel of the 1080p frame, one 16-sample batched wave (33 M paths), bit for bit."""
    from oracle import pt_scene_oracle as S
    w, h, n = 1920, 1080, 16
    sc = scenes.random_spheres(500, seed=1, width=w, height=h)
    ref = S.SceneOracle(sc, w, h, max_depth=8).accumulate(0, n, seed=0)
    with NativeTracer(w, h, seed=0, max_depth=8) as t:
        t.upload_scene(sc)
        t.render_samples(0, n)
        assert t.scene_queue_sizes()[0] == w * h * n
        assert_bits(t.read_accum(), ref, "c3 16-sample wave, full frame")


def test_c4_full_size_batched_wave_on_one_rank_shard():
    """BASELINE configs[3] at 3840x2160, 1M triangles, depth 8: rank 3 of 8's band shard as bench.py --workload c4 --gpus 8
    sets it (bands of 4 rows), rendered as a 32-sample batched wave on a sub-band to bound the oracle's time: the shard
    (3, 64, 4) = 136 rows x 3840 x 32 samples = 16.7 M paths."""
    from oracle import pt_scene_oracle as S
    w, h, n = 3840, 2160, 32
    rows = (3, 64, 4)
    sc = scenes.triangle_terrain(512, 2, seed=2, width=w, height=h)
    so = S.SceneOracle(sc, w, h, max_depth=8)
    ref = so.accumulate(0, n, seed=0, rows=rows)
    with NativeTracer(w, h, seed=0, max_depth=8) as t:
        t.upload_scene(sc)
        t.set_row_shard(*rows)
        t.render_samples(0, n)
        acc = t.read_accum()
    mine = (np.arange(h) // rows[2]) % rows[1] == rows[0]
    assert np.all(acc[mine][..., 3] == float(n)) and not acc[~mine].any()
    assert_bits(acc, ref, "c4 32-sample batched wave on a band shard")

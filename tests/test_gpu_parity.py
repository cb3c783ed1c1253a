"""GPU parity tests (run with -m gpu on the B200 box): the CUDA path, called through the C ABI,
against the CPU oracle on the same inputs — bit-exact — and against the committed goldens."""
import hashlib
import json
import os

import numpy as np
import pytest

from pathtracing_b200 import NativeTracer, PtbError, Renderer, Window, _lib

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")
with open(os.path.join(GOLD, "oracle_digests.json")) as f:
    DIG = json.load(f)
with open(os.path.join(GOLD, "appendix_b.json")) as f:
    APPB = json.load(f)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def eq_bits(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.uint32), np.ascontiguousarray(b).view(np.uint32))


def gpu_frame(w, h, **kw):
    with NativeTracer(w, h, keep_f32=True, keep_hit_ids=True, **kw) as t:
        t.compute_frame()
        t.wait()
        return {"f32": t.read_f32(), "rgba8": t.read_rgba8(), "hit": t.read_hit_ids()}


# ---- a1-a5: the literal CSMain frame -----------------------------------------------------------
@pytest.mark.parametrize("name", sorted(DIG["frames"]))
def test_frame_matches_golden_digests(name):
    c = DIG["frames"][name]
    kw = dict(c["params"])
    if c["uv_div"]:
        kw["uv_div"] = c["uv_div"]
    g = gpu_frame(c["width"], c["height"], **kw)
    assert int((g["hit"] == 0).sum()) == c["hits"]
    assert sha(g["hit"]) == c["hit"]          # hit-primitive ids: bit-exact
    assert sha(g["f32"]) == c["f32"]          # radiance: bit-exact (bar is 1e-4 rel)
    assert sha(g["rgba8"]) == c["rgba8"]


@pytest.mark.parametrize("w,h,div", [(320, 180, None), (1920, 1080, None), (3840, 2160, None),
                                     (320, 180, (1080, 1080)), (333, 77, None), (1, 1, None), (7, 1, None),
                                     (4, 5, None), (1921, 3, (1080, 1080))])
def test_frame_matches_oracle_bit_exact(oracle, w, h, div):
    p = oracle.frame_params(w, h, uv_div=div)
    ref = oracle.csmain_frame(p)
    g = gpu_frame(w, h, uv_div=div)
    assert np.array_equal(g["hit"], ref["hit"])
    assert eq_bits(g["f32"], ref["f32"])
    assert np.array_equal(g["rgba8"], ref["rgba8"])
    # north_star tolerance form of the same statement (1e-4 relative at 1 spp)
    np.testing.assert_allclose(g["f32"], ref["f32"], rtol=1e-4, atol=0)


def test_frame_appendix_b_known_answers():
    g = gpu_frame(1920, 1080)
    assert int((g["hit"] == 0).sum()) == 305317
    ys, xs = np.nonzero(g["hit"] == 0)
    assert (xs.min(), xs.max(), ys.min(), ys.max()) == (229, 851, 229, 851)
    for px in APPB["pixels"]:
        if px["div"] != 1080:
            continue
        col = g["f32"][px["y"], px["x"]]
        assert [int(v) for v in col[:3].view(np.uint32)] == [int(v, 16) for v in px["rgb"]]
        assert (g["hit"][px["y"], px["x"]] == 0) == bool(px["hit"])
    assert g["f32"][540, 540].tolist() == [0.5, 0.5, 1.0, 1.0]


def test_frame_small_fixture_without_oracle():
    z = np.load(os.path.join(GOLD, "c1_small.npz"))
    g = gpu_frame(64, 36)
    assert eq_bits(g["f32"], z["f32"]) and np.array_equal(g["rgba8"], z["rgba8"]) and np.array_equal(g["hit"], z["hit"])


def test_frame_moved_camera_and_sphere(oracle):
    kw = dict(cam=(0.25, -0.5, 2.0), centre=(-0.3, 0.2, -0.4), radius=0.9)
    ref = oracle.csmain_frame(oracle.frame_params(640, 360, **kw))
    g = gpu_frame(640, 360, **kw)
    assert np.array_equal(g["hit"], ref["hit"]) and eq_bits(g["f32"], ref["f32"])


def test_frame_is_idempotent_and_rgba8_only_path():
    with NativeTracer(1920, 1080) as t:     # no f32 / hit planes: the reference configuration
        t.compute_frame()
        a = t.read_rgba8()
        t.compute_frame()
        t.compute_frame()
        b = t.read_rgba8()
        assert t.launch_count == 3
    assert np.array_equal(a, b)
    assert sha(a) == DIG["frames"]["c2_1920x1080"]["rgba8"]


def test_read_with_pitch():
    with NativeTracer(100, 7) as t:
        t.compute_frame()
        tight = t.read_rgba8()
        padded = np.zeros((7, 128, 4), np.uint8)
        view = padded[:, :100, :]
        t.read_rgba8(view)
        assert np.array_equal(view, tight) and not padded[:, 100:, :].any()


# ---- n1/n2: jittered accumulation + resolve -----------------------------------------------------
@pytest.mark.parametrize("name", sorted(DIG["accum"]))
def test_accumulate_matches_golden_digests(name):
    c = DIG["accum"][name]
    with NativeTracer(c["width"], c["height"], seed=c["seed"]) as t:
        t.render_samples(c["first"], c["n"], jitter=c["jitter"])
        acc = t.read_accum()
        assert sha(acc) == c["accum"]
        t.resolve()
        assert sha(t.read_rgba8()) == c["resolve"]
        t.resolve(srgb=True, bgra=True)
        assert sha(t.read_rgba8()) == c["resolve_srgb_bgra"]


@pytest.mark.parametrize("w,h,first,n,seed", [(320, 180, 0, 8, 0), (333, 77, 5, 3, 99), (1920, 1080, 0, 2, 1)])
def test_accumulate_matches_oracle_bit_exact(oracle, w, h, first, n, seed):
    p = oracle.frame_params(w, h)
    ref = oracle.accumulate(p, first, n, seed, True)
    with NativeTracer(w, h, seed=seed) as t:
        t.render_samples(first, n, jitter=True)
        acc = t.read_accum()
        assert eq_bits(acc, ref)
        t.resolve()
        assert np.array_equal(t.read_rgba8(), oracle.resolve(ref, n))
        t.resolve(srgb=True)
        assert np.array_equal(t.read_rgba8(), oracle.resolve(ref, n, srgb=True))


def test_accumulate_no_jitter_reduces_to_csmain():
    with NativeTracer(320, 180, keep_f32=True) as t:
        t.compute_frame()
        frame, frame8 = t.read_f32(), t.read_rgba8()
        t.render_samples(0, 1, jitter=False)
        assert eq_bits(t.read_accum(), frame)
        t.resolve()
        assert np.array_equal(t.read_rgba8(), frame8)


def test_progressive_launches_and_checkpoint(oracle):
    p = oracle.frame_params(320, 180)
    ref = oracle.accumulate(p, 0, 4, 3, True)
    ref = oracle.accumulate(p, 4, 4, 3, True, accum=ref)
    with NativeTracer(320, 180, seed=3) as t:
        t.render_samples(0, 4)
        ckpt, spp = t.read_accum(), t.accum_spp
        assert spp == 4
        t.render_samples(4, 4)
        full = t.read_accum()
        assert t.accum_spp == 8 and eq_bits(full, ref)
    with NativeTracer(320, 180, seed=3) as t2:      # resume from the checkpoint
        t2.write_accum(ckpt, spp)
        t2.render_samples(4, 4)
        assert eq_bits(t2.read_accum(), ref) and t2.accum_spp == 8
        t2.accum_reset()
        assert t2.accum_spp == 0 and not t2.read_accum().any()


def test_full_size_properties_c2_64spp():
    """BASELINE configs[1] at full size through size-independent properties."""
    with NativeTracer(1920, 1080, seed=0, keep_f32=True) as t:
        t.render_samples(0, 64)
        acc = t.read_accum()
        t.compute_frame()
        frame = t.read_f32()
    assert np.all(acc[..., 3] == 64.0)                     # alpha is exactly 1 per sample
    mean = acc / np.float32(64)
    assert np.isfinite(mean).all()
    # interior of the sphere and far background are smooth: jittered mean ~ centre sample
    assert np.abs(mean[540, 540] - frame[540, 540]).max() < 2e-3
    assert np.abs(mean[10, 1900] - frame[10, 1900]).max() < 2e-3
    # mirror symmetry of the scene about the sphere centre row is preserved statistically
    assert abs(float(mean[300:781, 300:781, 2].mean()) - float(mean[300:781, 300:781, 2][::-1].mean())) < 1e-3


# ---- f1: present-equivalent ---------------------------------------------------------------------
@pytest.mark.parametrize("dw,dh,srgb", [(320, 180, True), (320, 180, False), (500, 333, True), (97, 41, False),
                                        (1, 1, True)])
def test_present_matches_oracle(oracle, dw, dh, srgb):
    with NativeTracer(320, 180) as t:
        t.compute_frame()
        img = t.read_rgba8()
        t.present(dw, dh, srgb=srgb)
        out = t.read_present()
    assert np.array_equal(out, oracle.present(img, dw, dh, srgb=srgb))


def test_present_golden_digests():
    g = DIG["present"]
    with NativeTracer(320, 180) as t:
        t.compute_frame()
        t.present(320, 180, srgb=True)
        assert sha(t.read_present()) == g["same_size_srgb"]
        t.present(500, 333, srgb=True)
        assert sha(t.read_present()) == g["up_500x333_srgb"]
        t.present(97, 41, srgb=False)
        assert sha(t.read_present()) == g["down_97x41_linear"]


def test_srgb_encode_exhaustive_on_a_ramp(oracle):
    """resolve(sRGB) against the oracle's double-precision OETF on a dense float ramp incl. edge values."""
    w, h = 4096, 64
    rng = np.random.default_rng(5)
    acc = np.zeros((h, w, 4), np.float32)
    acc[..., 0] = np.linspace(0, 1, w * h, dtype=np.float32).reshape(h, w)
    acc[..., 1] = rng.random((h, w), dtype=np.float32) * np.float32(1.2) - np.float32(0.1)
    acc[..., 2] = rng.random((h, w), dtype=np.float32) ** 4
    acc[..., 3] = 1.0
    acc[0, :8, 1] = [np.nan, np.inf, -np.inf, 0.0, -0.0, 1.0, 0.0031308, 1e-30]
    with NativeTracer(w, h) as t:
        t.write_accum(acc, 1)
        t.resolve(srgb=True)
        out = t.read_rgba8()
        t.write_accum(acc * np.float32(3), 3)
        t.resolve(srgb=False, bgra=True)
        out_lin = t.read_rgba8()
    assert np.array_equal(out, oracle.resolve(acc, 1, srgb=True))
    assert np.array_equal(out_lin, oracle.resolve(acc * np.float32(3), 3, bgra=True))


# ---- the Renderer mirror and the error contract ------------------------------------------------------
def test_renderer_mirror_call_sequence(oracle):
    r = Renderer(Window(FramebufferSize=(1920, 1080)))
    r.Init()
    r.Render(1.0 / 60.0)
    img, shown = r.ReadImage(), r.ReadPresented()
    r.Dispose()
    r.Dispose()                                            # idempotent
    ref = oracle.csmain_frame(oracle.frame_params(1920, 1080))["rgba8"]
    assert np.array_equal(img, ref)
    assert np.array_equal(shown, oracle.present(ref, 1920, 1080, srgb=True))


def test_renderer_requires_pipeline():
    r = Renderer(Window((64, 36)))
    with pytest.raises(Exception):
        r.CreateComputePipeline()
    r.CreateResources()
    with pytest.raises(Exception):
        r.ComputeFrame(0.0)
    r.Dispose()


def test_error_contract():
    with NativeTracer(64, 36) as t:
        with pytest.raises(PtbError) as e:
            t.read_f32()                                   # created without KEEP_F32
        assert e.value.status == _lib.PTB_ERR_STATE
        with pytest.raises(PtbError) as e:
            t.resolve()                                    # 0 spp
        assert e.value.status == _lib.PTB_ERR_STATE
        with pytest.raises(PtbError) as e:
            t.read_present()
        assert e.value.status == _lib.PTB_ERR_STATE
        bad = np.zeros((36, 8, 4), np.uint8)               # pitch too small
        with pytest.raises(PtbError) as e:
            t._ck(t._lib.ptb_read_rgba8(t.handle, bad.ctypes.data, 32))
        assert e.value.status == _lib.PTB_ERR_INVALID
        with pytest.raises(PtbError):
            t._ck(t._lib.ptb_read_rgba8(t.handle, None, 0))
        t.compute_frame()                                  # context still usable after errors
        t.wait()
    with pytest.raises(PtbError) as e:
        NativeTracer(64, 36, device=999)
    assert e.value.status == _lib.PTB_ERR_INVALID


def test_timer_and_launch_count():
    with NativeTracer(1920, 1080) as t:
        t.timer_begin()
        for _ in range(10):
            t.compute_frame()
        ms = t.timer_end()
        assert 0 < ms < 1000 and t.launch_count == 10


# ---- the shortened exact sequences: proofs by exhaustion and A/B against the literal IEEE kernels ------
def test_exact_math_proofs_by_exhaustion():
    with NativeTracer(64, 36) as t:
        bad, n = t.selftest_exact_math(0)                  # sqrt over [2^-20, 2^24)
        assert bad == 0 and n == 44 * 2 ** 23
        bad, n = t.selftest_exact_math(1)                  # reciprocal over [2^-12, 2^12)
        assert bad == 0 and n == 24 * 2 ** 23
        for d in (1080.0, 2160.0, 180.0, 36.0, 77.0, 3.0, 1234.567, -1080.0):
            bad, n = t.selftest_exact_math(2, d)           # v / d for every float v in [2^-24, 2^17)
            assert n == 41 * 2 ** 23
            assert bad == 0, f"one-correction division by {d} disagrees with IEEE for {bad} inputs"


def test_kernel_mode_selection():
    with NativeTracer(1920, 1080) as t:
        assert t.kernel_mode == 2                          # reference literals: fast + on-axis camera
    with NativeTracer(1920, 1080, force_ieee=True) as t:
        assert t.kernel_mode == 0
    with NativeTracer(640, 360, cam=(0.25, -0.5, 2.0)) as t:
        assert t.kernel_mode == 1                          # off-axis camera
    with NativeTracer(64, 36, uv_div=(1e-9, 36)) as t:
        assert t.kernel_mode == 0                          # |uv| unbounded for the sqrt/rcp domains
    with NativeTracer(64, 36, cam=(0, 0, float("inf"))) as t:
        assert t.kernel_mode == 0


@pytest.mark.parametrize("kw", [dict(), dict(cam=(0.25, -0.5, 2.0), centre=(-0.3, 0.2, -0.4), radius=0.9),
                                dict(uv_div=(777.0, 1234.567)), dict(cam=(0, 0, 40.0), radius=30.0),
                                dict(cam=(0, 0, 1.0), radius=1.0), dict(cam=(0, 0, 0.25)), dict(radius=1e-3)],
                         ids=["default", "moved", "odd_div", "huge_sphere", "camera_on_surface", "camera_inside", "tiny"])
def test_fast_and_ieee_kernels_agree_bitwise(oracle, kw):
    w, h = 1280, 720
    ref = oracle.csmain_frame(oracle.frame_params(w, h, **kw))
    ref_acc = oracle.accumulate(oracle.frame_params(w, h, **kw), 3, 5, 11, True)
    outs = []
    for force in (False, True):
        with NativeTracer(w, h, keep_f32=True, keep_hit_ids=True, seed=11, force_ieee=force, **kw) as t:
            t.compute_frame()
            t.render_samples(3, 5)
            outs.append((t.kernel_mode, t.read_f32(), t.read_hit_ids(), t.read_accum()))
    assert outs[0][0] in (1, 2) and outs[1][0] == 0
    for mode, f32, hit, acc in outs:
        assert np.array_equal(hit, ref["hit"]), mode
        assert eq_bits(f32, ref["f32"]), mode
        assert eq_bits(acc, ref_acc), mode


@pytest.mark.parametrize("size,kw", [
    ((1920, 1080), dict()), ((3840, 2160), dict()), ((320, 180), dict()), ((333, 77), dict()),
    ((1280, 720), dict(cam=(0.25, -0.5, 2.0), centre=(-0.3, 0.2, -0.4), radius=0.9)),
    ((1280, 720), dict(uv_div=(777.0, 1234.567))), ((1280, 720), dict(cam=(0, 0, 40.0), radius=30.0)),
    ((1280, 720), dict(cam=(0, 0, 1.0), radius=1.0)), ((640, 360), dict(cam=(0, 0, 0.25))), ((1280, 720), dict(radius=1e-3)),
    ((640, 360), dict(cam=(3.0, 0.1, 0.2), centre=(0.0, 0.0, -0.5), radius=0.7)), ((64, 36), dict(uv_div=(7.0, 5.0))),
    ((1280, 720), dict(cam=(0.0, 0.0, 1000.0), radius=400.0)), ((640, 360), dict(cam=(0.0, 0.0, -2.0)))],
    ids=["1080p", "4k", "c1", "ragged", "moved", "odd_div", "huge_sphere", "camera_on_surface", "camera_inside", "tiny",
         "side_view", "coarse_pixels", "far_camera", "sphere_behind"])
def test_certain_miss_classification_is_conservative(oracle, size, kw):
    """Pixels the accumulate kernel treats as 'every sample misses' (no normalisation, no discriminant) really miss at
    every jitter position tried (8x8 grid incl. both extremes, exact discriminant), and the accumulation stays bit-exact."""
    w, h = size
    with NativeTracer(w, h, seed=4, **kw) as t:
        bad, n = t.selftest_exact_math(3)
        assert bad == 0, f"{bad} sample positions of 'certain miss' pixels actually hit"
        bad4, n4 = t.selftest_exact_math(4)          # the row-span table of the 1 spp frame kernels
        assert bad4 == 0, f"{bad4} pixels of skipped quads actually hit"
        if not kw and w * 9 == h * 16:
            assert n4 > 0.75 * w * h                 # default scene: most quads skip the intersection
        frac = n / 64 / (w * h)
        if not kw and w * 9 == h * 16:
            assert 0.80 < frac < 0.86          # default scene, 16:9: 85.3 % of the pixels miss; the silhouette band stays out
        t.render_samples(5, 3)
        ref = oracle.accumulate(oracle.frame_params(w, h, **kw), 5, 3, 4, True)
        assert eq_bits(t.read_accum(), ref)


def test_generic_kernel_handles_degenerate_parameters(oracle):
    """Parameters outside the proven domains run the literal kernels and still match the oracle (NaN-aware)."""
    kw = dict(uv_div=(1e-3, 36.0), cam=(0.0, 0.0, 1e20))
    ref = oracle.csmain_frame(oracle.frame_params(64, 36, **kw))
    g = gpu_frame(64, 36, **kw)
    assert np.array_equal(g["hit"], ref["hit"])
    assert eq_bits(np.nan_to_num(g["f32"], nan=-7.0), np.nan_to_num(ref["f32"], nan=-7.0))
    assert np.array_equal(g["rgba8"], ref["rgba8"])


def test_async_readback_is_ordered_and_matches_sync(oracle):
    w, h = 640, 360
    bufs = [np.zeros((h, w, 4), np.uint8) for _ in range(3)]
    with NativeTracer(w, h, seed=2) as t:
        t.compute_frame()
        t.read_rgba8_async(bufs[0].ctypes.data)            # frame 0: the literal CSMain image
        t.render_samples(0, 4)
        t.resolve()                                        # must not overwrite the image before copy 0 finished
        t.read_rgba8_async(bufs[1].ctypes.data)
        t.render_samples(4, 4)
        t.resolve(srgb=True)
        t.read_rgba8_async(bufs[2].ctypes.data)
        t.wait()
        sync = t.read_rgba8()
    p = oracle.frame_params(w, h)
    assert np.array_equal(bufs[0], oracle.csmain_frame(p)["rgba8"])
    acc = oracle.accumulate(p, 0, 4, 2, True)
    assert np.array_equal(bufs[1], oracle.resolve(acc, 4))
    acc = oracle.accumulate(p, 4, 4, 2, True, accum=acc)
    assert np.array_equal(bufs[2], oracle.resolve(acc, 8, srgb=True))
    assert np.array_equal(sync, bufs[2])


@pytest.mark.parametrize("split", ["1", "3", "4"])
def test_banded_async_readback_equals_the_plain_read(split):
    """ptb_read_rgba8_async in 1 / 3 / 4 row bands on as many copy streams (PTB200_COPY_SPLIT; the default is one stream for
    a context on its own and four bands in a bound peer group).  The variable is read once per process, hence a child."""
    import os, subprocess, sys
    code = (
        "import numpy as np\n"
        "from pathtracing_b200 import NativeTracer\n"
        "w, h = 1920, 1080\n"
        "with NativeTracer(w, h, seed=3) as t:\n"
        "    frames, bufs = [], [np.zeros((h, w, 4), np.uint8) for _ in range(3)]\n"
        "    for i, buf in enumerate(bufs):\n"
        "        t.render_resolve(5 * i, 5)\n"
        "        t.read_rgba8_async(buf.ctypes.data)\n"
        "        if i == 1: t.wait(); frames.append(t.read_rgba8())\n"
        "    t.wait()\n"
        "    last = t.read_rgba8()\n"
        "    assert np.array_equal(bufs[1], frames[0]) and np.array_equal(bufs[2], last)\n"
        "    assert not np.array_equal(bufs[0], bufs[2])\n"
        "    pitch = w * 4 + 64\n"
        "    c = np.zeros((h, pitch), np.uint8)\n"
        "    t.read_rgba8_async(c.ctypes.data, pitch); t.wait()\n"
        "    assert np.array_equal(c[:, :w * 4].reshape(h, w, 4), last) and not c[:, w * 4:].any()\n"
        "print('banded ok')\n")
    env = dict(os.environ, PTB200_COPY_SPLIT=split)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "banded ok" in r.stdout, r.stdout + r.stderr


@pytest.mark.parametrize("size", [(640, 360), (333, 77), (1920, 1080)])
@pytest.mark.parametrize("jitter", [True, False])
def test_fused_render_resolve_equals_the_two_calls(oracle, size, jitter):
    """ptb_render_resolve = ptb_render_samples + ptb_resolve in one launch: same accumulation, same image, progressive
    calls, asynchronous read-backs of consecutive frames (two image buffers), and the unfused fall-backs."""
    w, h = size
    p = oracle.frame_params(w, h)
    spans = [(0, 5), (5, 2), (7, 4), (11, 1)]
    bufs = [np.zeros((h, w, 4), np.uint8) for _ in spans]
    with NativeTracer(w, h, seed=6) as t:
        ptrs = set()
        for (first, n), buf in zip(spans, bufs):
            t.render_resolve(first, n, jitter=jitter)
            ptrs.add(t.device_rgba8())
            t.read_rgba8_async(buf.ctypes.data)            # overlaps the next call, which writes the OTHER image
        t.wait()
        assert len(ptrs) == 2 and t.accum_spp == 12
        last_sync, acc_gpu = t.read_rgba8(), t.read_accum()
        t.render_resolve(12, 3, jitter=jitter, bgra=True)
        bgra = t.read_rgba8()
        t.render_resolve(15, 2, jitter=jitter, srgb=True)  # sRGB: falls back to the two launches
        srgb = t.read_rgba8()
        t.set_row_shard(1, 2)
        t.render_resolve(17, 1, jitter=jitter)             # row shard: falls back as well
        shard_img, shard_acc = t.read_rgba8(), t.read_accum()
    acc = None
    for (first, n), buf in zip(spans, bufs):
        acc = oracle.accumulate(p, first, n, 6, jitter, accum=acc)
        assert np.array_equal(buf, oracle.resolve(acc, first + n)), (first, n)
    assert np.array_equal(last_sync, bufs[-1]) and eq_bits(acc_gpu, acc)
    acc = oracle.accumulate(p, 12, 3, 6, jitter, accum=acc)
    assert np.array_equal(bgra, oracle.resolve(acc, 15, bgra=True))
    acc = oracle.accumulate(p, 15, 2, 6, jitter, accum=acc)
    assert np.array_equal(srgb, oracle.resolve(acc, 17, srgb=True))
    with NativeTracer(w, h, seed=6) as t:                  # the same sequence through the two calls
        for first, n in spans + [(12, 3), (15, 2)]:
            t.render_samples(first, n, jitter=jitter)
        t.set_row_shard(1, 2)
        t.render_samples(17, 1, jitter=jitter)
        t.resolve()
        assert np.array_equal(t.read_rgba8(), shard_img) and eq_bits(t.read_accum(), shard_acc)


# ---- multi-GPU partitioning emulated on one device: one context per "rank" -------------------------
@pytest.mark.parametrize("world,band", [(2, 1), (3, 1), (8, 1), (2, 4), (3, 4), (8, 4), (8, 16), (7, 5)])
def test_row_shards_merge_to_the_single_device_image(world, band):
    w, h, spp, seed = 640, 358, 6, 4            # 358 rows: the last band is ragged for bands of 4, 5 and 16
    with NativeTracer(w, h, seed=seed, keep_f32=True) as t:
        t.compute_frame()
        full_frame = t.read_rgba8()
        t.render_samples(0, spp)
        full_acc = t.read_accum()
        t.resolve()
        full_img = t.read_rgba8()
    merged_frame = np.zeros((h, w), np.uint32)
    merged_img = np.zeros((h, w), np.uint32)
    merged_acc = np.zeros((h, w, 4), np.float32)
    for rank in range(world):
        with NativeTracer(w, h, seed=seed) as t:
            t.set_row_shard(rank, world, band)
            t.compute_frame()
            fr = t.read_rgba8()
            assert not fr[(np.arange(h) // band % world) != rank].any()  # foreign rows untouched (zero)
            assert fr[(np.arange(h) // band % world) == rank, :, 3].all()  # own rows all written (alpha 255)
            merged_frame += fr.view(np.uint32).reshape(h, w)
            t.render_samples(0, spp)
            merged_acc += t.read_accum()                                 # disjoint rows: sum == gather
            t.resolve()
            merged_img += t.read_rgba8().view(np.uint32).reshape(h, w)
    assert np.array_equal(merged_frame.view(np.uint8).reshape(h, w, 4), full_frame)
    assert eq_bits(merged_acc, full_acc)
    assert np.array_equal(merged_img.view(np.uint8).reshape(h, w, 4), full_img)


def test_row_shard_argument_errors():
    with NativeTracer(64, 36) as t:
        for off, stride, band in ((0, 0, 1), (2, 2, 1), (5, 3, 1), (0, 2, 0)):
            with pytest.raises(PtbError) as e:
                t.set_row_shard(off, stride, band)
            assert e.value.status == _lib.PTB_ERR_INVALID
        t.set_row_shard(40, 64)          # offset beyond the image: nothing to render, not an error
        t.compute_frame()
        assert not t.read_rgba8().any()
        t.set_row_shard(0, 2, 100)       # one band taller than the image: rank 0 of 2 owns every row, rank 1 none
        t.compute_frame()
        assert t.read_rgba8()[..., 3].all()
        t.set_row_shard(1, 2, 100)
        t.render_samples(0, 2)
        assert not t.read_accum().any()


def test_sharded_renderer_single_rank_paths(oracle):
    """dist.ShardedRenderer over the real GpuEngine (world size 1): both modes equal the plain path."""
    import torch
    from pathtracing_b200.dist import GpuEngine, ShardedRenderer
    w, h = 320, 180
    p = oracle.frame_params(w, h)
    ref = oracle.accumulate(p, 0, 4, 9, True)
    ref = oracle.accumulate(p, 4, 3, 9, True, accum=ref)
    for mode in ("samples", "tiles"):
        eng = GpuEngine(w, h, 0, seed=9)
        sr = ShardedRenderer(eng, mode=mode)
        sr.render(0, 4)
        sr.render(4, 3)
        img = sr.merge(0)
        torch.cuda.synchronize()
        assert eq_bits(eng.accum.cpu().numpy(), ref)
        got = img.cpu().numpy().view(np.uint8).reshape(h, w, 4)
        assert np.array_equal(got, oracle.resolve(ref, 7))
        eng.close()


def test_fused_peer_merge_kernel_single_process(oracle):
    """ptb_peer_merge (n8, the fused merge+resolve over peer memory) with three 'ranks' = three contexts on one device:
    every rank merges its own pixel slab from all three partial accumulation buffers into rank 0's image.  Sum order is
    rank order, so the merged accumulation is reproducible exactly on the CPU; partials are cleared; a second progressive
    batch lands on top of the running total."""
    from pathtracing_b200.dist import sample_range, slab_range
    w, h, seed, world = 200, 90, 6, 3
    p = oracle.frame_params(w, h)
    ranks = [NativeTracer(w, h, seed=seed) for _ in range(world)]
    try:
        accum_ptrs = [t.device_accum() for t in ranks]
        flag_ptrs = [t.device_peer_flags() for t in ranks]               # each context has its own stream
        epoch = 0
        for t in ranks:
            t.accum_reset()
        total = np.zeros((h, w, 4), np.float32)
        spp = 0
        for first, count in ((0, 7), (7, 5)):
            parts = []
            for r, t in enumerate(ranks):
                st, n = sample_range(first, count, r, world)
                t.render_samples(st, n)
                parts.append(oracle.accumulate(p, st, n, seed, True))
            spp += count
            epoch += 1
            for r, t in enumerate(ranks):                                # rendezvous 0: everybody has rendered
                t.peer_barrier(flag_ptrs, r, 0, epoch)
            for r, t in enumerate(ranks):
                s0, sn = slab_range(r, world, w * h)
                t.peer_merge(accum_ptrs, ranks[0].device_rgba8(), s0, sn, spp, srgb=True)
            for r, t in enumerate(ranks):                                # rendezvous 1: everybody has merged
                t.peer_barrier(flag_ptrs, r, 1, epoch)
            for t in ranks:
                t.wait()
            total = total + ((parts[0] + parts[1]) + parts[2])          # rank order, then onto the running total
            assert np.array_equal(ranks[0].read_rgba8(), oracle.resolve(total, spp, srgb=True))
            for t in ranks:
                t.accum_clear()                                          # each rank clears its own partials afterwards
                assert not t.read_accum().any()
    finally:
        for t in ranks:
            t.destroy()


def test_sharded_renderer_p2p_single_rank(oracle):
    import torch
    from pathtracing_b200.dist import GpuEngine, ShardedRenderer
    w, h = 320, 180
    p = oracle.frame_params(w, h)
    eng = GpuEngine(w, h, 0, seed=9, peer_memory=True)
    sr = ShardedRenderer(eng, mode="p2p")
    sr.render(0, 4)
    a = oracle.accumulate(p, 0, 4, 9, True)
    img = sr.merge(0)
    torch.cuda.synchronize()
    assert np.array_equal(img.cpu().numpy().view(np.uint8).reshape(h, w, 4), oracle.resolve(a, 4))
    sr.render(4, 3)
    b = oracle.accumulate(p, 4, 3, 9, True)
    img = sr.merge(0)
    torch.cuda.synchronize()
    assert np.array_equal(img.cpu().numpy().view(np.uint8).reshape(h, w, 4), oracle.resolve(a + b, 7))
    eng.close()

"""GPU side of the Test.spirv pin (tests/test_spirv_pin.py is the CPU side): with PTB_CFG_SPIRV_ARITH the CUDA kernels
evaluate the operation list of the reference's checked-in compiled module and must equal, bit for bit, the numpy
evaluation of the module's words (oracle/spirv_eval.py on tests/golden/test_spirv.json) — a reference-held artefact,
not this repo's own restatement — and the C oracle's as-compiled variant."""
import json
import os

import numpy as np
import pytest

from pathtracing_b200 import NativeTracer, PtbError, _lib

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def module():
    from oracle import spirv_eval as E
    with open(os.path.join(HERE, "golden", "test_spirv.json")) as f:
        return E.parse_words([int(w, 16) for w in json.load(f)["words_hex"]])


@pytest.mark.parametrize("size,div", [((1920, 1080), None), ((320, 180), (1080, 1080)), ((1921, 1087), (1080, 1080))])
def test_spirv_mode_equals_the_compiled_module(module, oracle, size, div):
    from oracle import spirv_eval as E
    w, h = size
    r = E.evaluate(module, w, h)                               # the module unmodified: divisor 1080
    with NativeTracer(w, h, uv_div=div, keep_f32=True, keep_hit_ids=True, spirv_arith=True) as t:
        assert t.kernel_mode == 3
        t.compute_frame()
        t.wait()
        f32, hit, rgba8 = t.read_f32(), t.read_hit_ids(), t.read_rgba8()
    assert np.array_equal(bits(f32), bits(r["image"]))
    assert np.array_equal(hit == 0, r["cond"])
    ref = oracle.csmain_frame(oracle.frame_params(w, h, uv_div=(1080, 1080)), spirv=True)
    assert np.array_equal(rgba8, ref["rgba8"]) and np.array_equal(bits(f32), bits(ref["f32"]))


def test_spirv_mode_vs_default_mode_1080p(oracle):
    """Same hit ids as the default (source-order) kernels, radiance within the north_star 1e-4 relative."""
    w, h = 1920, 1080
    outs = {}
    for spirv in (False, True):
        with NativeTracer(w, h, keep_f32=True, keep_hit_ids=True, spirv_arith=spirv) as t:
            t.compute_frame()
            t.wait()
            outs[spirv] = (t.read_f32(), t.read_hit_ids())
    assert np.array_equal(outs[True][1], outs[False][1])
    a, b = outs[True][0], outs[False][0]
    nz = b != 0
    assert (np.abs(a[nz] - b[nz]) / np.abs(b[nz])).max() <= 1e-4


def test_spirv_mode_accumulation_and_generalised_divisor(oracle):
    w, h = 640, 360
    p = oracle.frame_params(w, h)
    with NativeTracer(w, h, seed=3, spirv_arith=True) as t:
        t.render_samples(2, 5)
        acc = t.read_accum()
        t.render_resolve(7, 3)
        acc2, img = t.read_accum(), t.read_rgba8()
    ref = oracle.accumulate(p, 2, 5, 3, True, spirv=True)
    assert np.array_equal(bits(acc), bits(ref))
    ref = oracle.accumulate(p, 7, 3, 3, True, accum=ref, spirv=True)
    assert np.array_equal(bits(acc2), bits(ref)) and np.array_equal(img, oracle.resolve(ref, 8))


def test_spirv_mode_is_literal_scene_only():
    with pytest.raises(PtbError) as e:
        NativeTracer(64, 36, cam=(0.0, 0.1, 1.0), spirv_arith=True)
    assert e.value.status == _lib.PTB_ERR_UNSUPPORTED
    with pytest.raises(PtbError) as e:
        NativeTracer(64, 36, max_depth=5000)
    assert e.value.status == _lib.PTB_ERR_INVALID

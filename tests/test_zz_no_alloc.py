"""ptb_trace_rays_static (include/ptb200.h): the ray-query entry point that allocates nothing — its device staging is a
static device array of the library — against ptb_trace_rays on the same rays, through the C ABI.

Collected last on purpose.  The bar is the one of the other ray-query tests
(tests/test_gpu_scene.py): primitive ids equal, distances equal bit for bit.  Two calls in a row with different batch
sizes, so that the second call finds the first call's contents in the shared array.
"""
import numpy as np
import pytest

from pathtracing_b200 import NativeTracer, scenes
from pathtracing_b200._lib import PtbError

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def rays(n, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    o = rng.uniform(-6.0, 6.0, (n, 3)).astype(np.float32)
    o[:, 1] = np.abs(o[:, 1]) + 0.2
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[: n // 16] = np.array([0, -1, 0], np.float32)      # axis-parallel rays: zero direction components
    return o, d.astype(np.float32)


def test_static_scratch_ray_queries_equal_the_allocating_entry_point():
    sc = scenes.mixed_small(n_spheres=200, n_tris=1500)
    n_max = NativeTracer.TRACE_RAYS_STATIC_MAX
    with NativeTracer(8, 8) as t:
        t.upload_scene(sc)
        # first call: a full array; second call: fewer, different rays (stale results of the first call lie behind them);
        # third: the first batch again
        for call, (n, seed) in enumerate([(n_max, 5), (n_max // 3 + 7, 6), (n_max, 5)]):
            o, d = rays(n, seed)
            prim, tt = t.trace_rays(o, d, 1e-3, 1e30)
            sp, st = t.trace_rays_static(o, d, 1e-3, 1e30)
            assert (prim >= 0).mean() > 0.2
            assert np.array_equal(sp, prim), f"call {call}: {np.count_nonzero(sp != prim)} closest-hit ids differ"
            assert np.array_equal(bits(st), bits(tt)), f"call {call}: distances differ"
            any_ref, _ = t.trace_rays(o, d, 1e-3, 3.0, any_hit=True)
            any_st, none = t.trace_rays_static(o, d, 1e-3, 3.0, any_hit=True)
            assert none is None and np.array_equal(any_st, any_ref), f"call {call}: any-hit flags differ"
        # a second context on the same device shares the array and its cached address
        with NativeTracer(8, 8) as t2:
            t2.upload_scene(scenes.random_spheres(300))
            o, d = rays(4096, 9)
            a, at = t2.trace_rays(o, d, 0.5, 40.0)
            b, bt = t2.trace_rays_static(o, d, 0.5, 40.0)
            assert np.array_equal(a, b) and np.array_equal(bits(at), bits(bt))


def test_static_scratch_limits_and_errors():
    with NativeTracer(8, 8) as t:
        o, d = rays(4, 1)
        with pytest.raises(PtbError, match="no scene resident"):
            t.trace_rays_static(o, d, 0.0, 1.0)
        t.upload_scene(scenes.mixed_small(n_spheres=20, n_tris=20))
        n = NativeTracer.TRACE_RAYS_STATIC_MAX + 1
        with pytest.raises(PtbError, match="PTB_TRACE_RAYS_STATIC_MAX"):
            t.trace_rays_static(np.zeros((n, 3), np.float32), np.ones((n, 3), np.float32), 0.0, 1.0)
        with pytest.raises(PtbError, match="t_min"):
            t.trace_rays_static(o, d, 2.0, 1.0)
        p, tt = t.trace_rays_static(np.zeros((0, 3), np.float32), np.zeros((0, 3), np.float32), 0.0, 1.0)   # empty batch
        assert len(p) == 0 and len(tt) == 0
        # still usable after the refusals
        a, at = t.trace_rays(o, d, 1e-3, 1e30)
        b, bt = t.trace_rays_static(o, d, 1e-3, 1e30)
        assert np.array_equal(a, b) and np.array_equal(bits(at), bits(bt))

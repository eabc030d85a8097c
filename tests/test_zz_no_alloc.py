"""exa_sample_points_host_noalloc (csrc/capi.cu; run with -m gpu): the point query whose device staging and counters
live in the library's static device array.  Collected LAST on purpose.  The entry
point is called straight through the C-ABI and compared with exa_sample_points_host / exa_sample_points_stats on the
same renderer, two calls in a row (the second with fewer points, so that stale contents of the shared array would show),
plus the size limit and the counters."""
import ctypes as C

import numpy as np
import pytest

import scenes
from owlexabrick_b200 import synth

pytestmark = pytest.mark.gpu

VALUE_TOL = 0.0        # same kernels on the same inputs: bit-equal (the other point-query tests allow 1e-5 against the oracle)


@pytest.fixture(scope="module")
def R(built_lib):
    assert built_lib.lib.exa_cuda_available() == 1, built_lib.last_error()
    from owlexabrick_b200 import renderer
    return renderer


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _existing(lib, r, pts, channel, grad):
    n = len(pts)
    val, ok = np.full(n, -7.0, np.float32), np.full(n, 9, np.uint8)
    g = np.full((n, 3), -7.0, np.float32) if grad else None
    assert lib.exa_sample_points_host(r.h, _p(pts), n, channel, _p(val), _p(ok), _p(g)) == 0
    return val, ok, g


def _noalloc(built_lib, r, pts, channel, grad, stats):
    n = len(pts)
    val, ok = np.full(n, -7.0, np.float32), np.full(n, 9, np.uint8)
    g = np.full((n, 3), -7.0, np.float32) if grad else None
    st = built_lib.ExaStats()
    rc = built_lib.lib.exa_sample_points_host_noalloc(r.h, _p(pts), n, channel, _p(val), _p(ok), _p(g), C.byref(st) if stats else None)
    assert rc == 0, built_lib.last_error()
    return val, ok, g, st.as_dict()


@pytest.mark.parametrize("kind,channels", [("nb64_off", 2), ("nb64", 1)])
def test_noalloc_point_query_equals_host_query_twice_in_a_row(R, built_lib, kind, channels):
    lib = built_lib.lib
    cells, fields, dom = scenes.small_scene(kind, channels)
    b = R.ExaBricks.build(cells, max_leaf_width=8)
    r = R.Renderer(b, fields)
    lo, hi = [-12.0] * 3, [float(dom) + 12.0] * 3                   # beyond the support of the coarsest cells: some points are invalid
    first = synth.random_points(50000, lo, hi, seed=11)
    second = synth.random_points(777, lo, hi, seed=12)              # fewer points than the call before
    for pts, channel, grad in ((first, channels - 1, True), (second, 0, False), (first[:4096], 0, True)):
        pts = np.ascontiguousarray(pts, np.float32)
        ref_val, ref_ok, ref_g = _existing(lib, r, pts, channel, grad)
        val, ok, g, st = _noalloc(built_lib, r, pts, channel, grad, stats=True)
        assert np.array_equal(ok, ref_ok) and set(np.unique(ok)) <= {0, 1}
        assert 0 < int(ok.sum()) < len(pts)
        assert np.abs(val - ref_val).max() <= VALUE_TOL
        if grad:
            assert np.abs(g - ref_g).max() <= VALUE_TOL
        # the counters were cleared before the launch: they describe THIS query, as exa_sample_points_stats reports it
        import torch
        d = torch.from_numpy(pts).cuda()
        ref_st = r.sample_points_stats(d.data_ptr(), len(pts), channel)
        assert st["samples"] == ref_st["samples"] >= int(ok.sum()) > 0
        assert st["brickVisits"] == ref_st["brickVisits"] and st["cornerLoads"] == ref_st["cornerLoads"]
        # and without counters (the un-instrumented kernel, the one callers time)
        val2, ok2, g2, _ = _noalloc(built_lib, r, pts, channel, grad, stats=False)
        assert np.array_equal(ok2, ref_ok) and np.array_equal(val2, ref_val) and (not grad or np.array_equal(g2, ref_g))


def test_noalloc_point_query_limits_and_errors(R, built_lib):
    lib = built_lib.lib
    cells, fields, dom = scenes.small_scene("nb64", 1)
    r = R.Renderer(R.ExaBricks.build(cells, max_leaf_width=8), fields)
    M = built_lib.NOALLOC_MAX_POINTS
    pts = np.ascontiguousarray(synth.random_points(M + 1, [0.0] * 3, [float(dom)] * 3, seed=3), np.float32)
    val, ok = np.zeros(M + 1, np.float32), np.zeros(M + 1, np.uint8)
    rc = lib.exa_sample_points_host_noalloc(r.h, _p(pts), M + 1, 0, _p(val), _p(ok), None, None)
    assert rc == built_lib.EXA_ERR_INVALID and "EXA_NOALLOC_MAX_POINTS" in built_lib.last_error()
    assert not ok.any()                                             # nothing ran
    # exactly the maximum is served, and equals the allocating path
    ref_val, ref_ok, _ = _existing(lib, r, pts[:M], 0, False)
    v, o, _, _ = _noalloc(built_lib, r, pts[:M], 0, False, stats=False)
    assert np.array_equal(o, ref_ok) and np.array_equal(v, ref_val)
    assert lib.exa_sample_points_host_noalloc(r.h, _p(pts), 0, 0, _p(val), _p(ok), None, None) == 0       # empty query
    assert lib.exa_sample_points_host_noalloc(r.h, _p(pts), 16, 5, _p(val), _p(ok), None, None) == built_lib.EXA_ERR_INVALID
    assert lib.exa_sample_points_host_noalloc(None, _p(pts), 16, 0, _p(val), _p(ok), None, None) == built_lib.EXA_ERR_INVALID
    # the Python mirror
    v2, o2 = r.sample_points_noalloc(pts[:1000])
    assert np.array_equal(v2, ref_val[:1000]) and np.array_equal(o2, ref_ok[:1000])
    with pytest.raises(RuntimeError, match="static scratch"):
        r.sample_points_noalloc(pts)

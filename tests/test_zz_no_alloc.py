"""`dre_locate_points_noalloc` -- the find_cc_new2 locator (rt_routines.f90:1962-2066) through the library's static device
scratch array, for callers that cannot call cudaMalloc.  It has no counterpart in the reference;
the check is the existing entry point `dre_locate_points`, whose own parity with the oracle is
tests/test_gpu_dropin.py::test_locator_kernel_matches_oracle.  Cell indices are integers: the bar is bit-exact.

The file is named so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from test_gpu_parity import _engine  # noqa: E402

MAX_POINTS = 65536      # DRE_NOALLOC_MAX_POINTS (dre_engine.cu, include/dartray_engine.h)


def _points(scene, n, seed):
    g = scene.grid
    rng = np.random.default_rng(seed)
    pts = (rng.random((n, 3)) - 0.5) * g.modelsize * 1.1          # ~25 % outside the model
    leaves = np.nonzero(g.cchild == -1)[0].astype(np.int32)
    return pts, leaves[rng.integers(0, len(leaves), n)]


def test_no_alloc_locator_matches_existing_entry_point_twice(small_trusti):
    """Two calls in a row on different point sets (the second one shorter, so that it runs over what the first call left in
    the shared array): each must equal dre_locate_points on the same inputs, bit for bit."""
    eng = _engine(small_trusti)
    for n, seed in ((4000, 11), (1500, 12)):
        pts, cc_old = _points(small_trusti, n, seed)
        want = eng.locate(pts, cc_old)
        got = eng.locate_noalloc(pts, cc_old)
        assert np.array_equal(got, want)
        assert (want == -1).any() and (want >= 0).any()
    # the largest size the array holds
    pts, cc_old = _points(small_trusti, MAX_POINTS, 13)
    assert np.array_equal(eng.locate_noalloc(pts, cc_old), eng.locate(pts, cc_old))
    eng.close()


def test_no_alloc_locator_rejects_sizes_beyond_the_array(small_trusti):
    """More points than the static array holds: non-zero return and a message, nothing written."""
    from dartray_b200 import abi
    eng = _engine(small_trusti)
    eng.set_params(abi.RTT_DIR_CELL)
    n = MAX_POINTS + 1
    pts, cc_old = _points(small_trusti, n, 14)
    out = np.full(n, -7, dtype=np.int32)
    rc = eng.l.dre_locate_points_noalloc(eng.h, C.c_int32(n), abi.dptr(np.ascontiguousarray(pts)), abi.iptr(cc_old), abi.iptr(out))
    assert rc != 0
    assert b"65536" in eng.l.dre_last_error(eng.h)
    assert (out == -7).all()
    # the engine is still usable afterwards
    pts, cc_old = _points(small_trusti, 100, 15)
    assert np.array_equal(eng.locate_noalloc(pts, cc_old), eng.locate(pts, cc_old))
    eng.close()

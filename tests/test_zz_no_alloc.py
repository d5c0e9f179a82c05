"""`ssp_star_brightness_static`: the star-brightness kernel with its device staging in a fixed file-scope
device array instead of the context's growable buffers (no cudaMalloc in the call; no counterpart in the
reference).  Checked through the C ABI against the existing
entry point `ssp_star_brightness`, integer-exact like every brightness test, on two calls in a row that share
the array.  This file is named so that pytest collects it last."""
import ctypes

import numpy as np
import pytest

from ssp_b200 import _native as N
from ssp_b200.custom_bf_polygon_matcher import disc_offsets

pytestmark = pytest.mark.gpu

MAX_PIXELS = 2 * 1024 * 1024     # SSP_BRIGHTNESS_STATIC_MAX_PIXELS
MAX_STARS = 16384                # SSP_BRIGHTNESS_STATIC_MAX_STARS


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _call(engine, name, gray, centers, disc, top_k):
    """One call of `name` through the C ABI with host buffers; returns (rc, result)."""
    fn = getattr(engine._lib, name)
    out = np.full(len(centers), -7, dtype=np.int32)
    rc = fn(engine._ctx, gray.ctypes.data_as(ctypes.POINTER(ctypes.c_uint8)), gray.shape[0], gray.shape[1],
            centers.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), len(centers),
            disc.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), len(disc), top_k,
            out.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), N.MEM_HOST)
    return rc, out


def _last_error(engine):
    return engine._lib.ssp_last_error(engine._ctx).decode()


def _frame(rng, h, w, n):
    gray = rng.integers(0, 256, size=(h, w), dtype=np.uint8)
    centers = np.stack([rng.integers(-12, w + 12, size=n), rng.integers(-12, h + 12, size=n)], axis=1)
    return np.ascontiguousarray(gray), np.ascontiguousarray(centers.astype(np.int32))


def test_static_scratch_entry_point_equals_the_allocating_one_on_two_calls_in_a_row(engine):
    rng = np.random.default_rng(2024)
    disc = np.ascontiguousarray(np.asarray(disc_offsets(10), dtype=np.int32).reshape(-1, 2))
    # first call: the reference's work-image size, many stars (also beyond the borders); second call: a
    # SMALLER frame with fewer stars and another disc, so that everything the first call left in the
    # shared array lies behind what the second one writes
    for (h, w, n, d, k) in [(894, 1342, 5000, disc, 10), (301, 257, 77, disc[: len(disc) // 2], 4)]:
        gray, centers = _frame(rng, h, w, n)
        d = np.ascontiguousarray(d)
        rc_ref, want = _call(engine, "ssp_star_brightness", gray, centers, d, k)
        rc, got = _call(engine, "ssp_star_brightness_static", gray, centers, d, k)
        assert rc_ref == 0 and rc == 0, _last_error(engine)
        np.testing.assert_array_equal(got, want)          # integer work: exact
        assert (got >= 0).all() and got.max() > 0
    # and the wrapper the package offers
    gray, centers = _frame(rng, 120, 90, 33)
    np.testing.assert_array_equal(engine.star_brightness_static(gray, centers, disc, 10),
                                  engine.star_brightness(gray, centers, disc, 10))


def test_static_scratch_entry_point_rejects_sizes_beyond_its_fixed_array(engine):
    rng = np.random.default_rng(7)
    disc = np.ascontiguousarray(np.asarray(disc_offsets(10), dtype=np.int32).reshape(-1, 2))
    gray, centers = _frame(rng, 1025, 2048, 8)                  # one row more than 2 Mi pixels
    assert gray.size > MAX_PIXELS
    rc, out = _call(engine, "ssp_star_brightness_static", gray, centers, disc, 10)
    assert rc == -4 and "exceeds the fixed scratch" in _last_error(engine)      # SSP_ERR_CAPACITY
    assert (out == -7).all()                                    # nothing was written
    gray, centers = _frame(rng, 64, 64, MAX_STARS + 1)
    rc, out = _call(engine, "ssp_star_brightness_static", gray, centers, disc, 10)
    assert rc == -4 and "stars exceed the fixed scratch" in _last_error(engine)
    # the limits themselves are accepted
    gray, centers = _frame(rng, 1024, 2048, MAX_STARS)
    rc, got = _call(engine, "ssp_star_brightness_static", gray, centers, disc, 10)
    rc_ref, want = _call(engine, "ssp_star_brightness", gray, centers, disc, 10)
    assert rc == 0 and rc_ref == 0
    np.testing.assert_array_equal(got, want)

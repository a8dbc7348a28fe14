"""fm_ma_window_series_static: the trailing-window series for embedders that
may not allocate device memory after start-up (a GPU whose memory is owned by
one pool, or a polling loop that cannot take cudaMalloc's device-wide stall).

The entry point allocates nothing itself: this test plays the embedder and
prepares the device buffers for the row and the two outputs beforehand, then
calls through the C ABI twice in a row and expects the same numbers as
fm_ma_window_series, the ctx-based entry point that allocates.  The second call
uses a shorter, different row, so it would expose any dependence on what the
first call left behind in the library's built-in scratch array.

Collected last on purpose (file name): a failure here cannot hide the other tests.
"""
import ctypes as C

import numpy as np
import pytest

from foremast_1_b200 import _abi as A
from foremast_1_b200 import _ffi
from foremast_1_b200.engine import Context

pytestmark = pytest.mark.gpu

MAX_POINTS = 16384   # FM_STATIC_MA_MAX_POINTS


def test_static_variant_matches_allocating_entry_point_twice():
    import torch
    lib = _ffi.lib()
    rng = np.random.default_rng(21)
    with Context(0) as ctx:
        d_x = ctx.dev_alloc(MAX_POINTS * 4)
        d_m = ctx.dev_alloc(MAX_POINTS * 8)
        d_s = ctx.dev_alloc(MAX_POINTS * 8)
        try:
            for n, w, ddof, centre in ((10080, 1440, 0, 1.0e6), (777, 60, 1, -3.0)):
                x = (centre + rng.normal(0, 25, n)).astype(np.float32)
                want_m, want_s = ctx.ma_window_series(x, w, ddof)          # allocating entry point
                ctx.h2d(d_x, x)
                rc = lib.fm_ma_window_series_static(0, None, d_x, n, w, ddof, d_m, d_s)
                assert rc == A.FM_OK, lib.fm_last_error(None)
                torch.cuda.synchronize()                                   # the call itself does not wait
                got_m = np.empty(n, np.float64)
                got_s = np.empty(n, np.float64)
                ctx.d2h(got_m, d_m)
                ctx.d2h(got_s, d_s)
                assert np.array_equal(np.isnan(got_m), np.isnan(want_m))
                assert np.array_equal(np.isnan(got_s), np.isnan(want_s))
                ok = ~np.isnan(want_m)
                assert ok.sum() == n - w + 1
                assert np.allclose(got_m[ok], want_m[ok], rtol=1e-6, atol=0)
                assert np.allclose(got_s[ok], want_s[ok], rtol=1e-6, atol=1e-9)
        finally:
            for p in (d_x, d_m, d_s):
                ctx.dev_free(p)


def test_static_variant_refuses_rows_beyond_its_scratch():
    lib = _ffi.lib()
    with Context(0) as ctx:
        d = ctx.dev_alloc(64)
        try:
            rc = lib.fm_ma_window_series_static(0, None, d, MAX_POINTS + 1, 10, 0, d, d)
            assert rc == A.FM_E_UNSUPPORTED
            msg = lib.fm_last_error(None).decode()
            assert str(MAX_POINTS) in msg and "exceeds" in msg
            assert lib.fm_ma_window_series_static(0, None, None, 8, 2, 0, d, d) == A.FM_E_INVALID
            assert lib.fm_ma_window_series_static(0, None, d, 0, 2, 0, d, d) == A.FM_OK   # empty row: nothing to do
        finally:
            ctx.dev_free(d)

"""CPU tests: the C oracle's interpolation and writers against the compiled reference's scalar slots
(fallback-motion.cc) — pins oracle/mc.c."""
import ctypes as C
import numpy as np
import pytest
from tests.helpers import ptr


def _i16(n):
    return np.zeros(n, dtype=np.int16)


@pytest.mark.parametrize("w,h", [(8, 4), (4, 8), (16, 16), (12, 16), (64, 64), (24, 32)])
def test_qpel_matches_reference(oracle, ref, w, h):
    rng = np.random.default_rng(w * 100 + h)
    pad = 8
    buf = rng.integers(0, 256, size=(h + 2 * pad, w + 2 * pad), dtype=np.uint8)
    stride = buf.shape[1]
    src = buf.ctypes.data + pad * stride + pad
    mcbuf = _i16(64 * (64 + 8))
    oracle.orc_interp_block.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_ssize_t, C.c_int, C.c_int, C.c_void_p]
    for fx in range(4):
        for fy in range(4):
            want, got = _i16(w * h), _i16(w * h)
            fn = getattr(ref, "_Z21put_qpel_%d_%d_fallbackPslPKhliiS_" % (fx, fy))
            fn(ptr(want), C.c_long(w), C.c_void_p(src), C.c_long(stride), w, h, ptr(mcbuf))
            oracle.orc_interp_block(1, w, h, src, stride, fx, fy, ptr(got))
            assert np.array_equal(got, want), (fx, fy)


@pytest.mark.parametrize("w,h", [(4, 2), (2, 4), (8, 8), (6, 8), (32, 32)])
def test_epel_matches_reference(oracle, ref, w, h):
    rng = np.random.default_rng(w * 100 + h + 7)
    pad = 4
    buf = rng.integers(0, 256, size=(h + 2 * pad, w + 2 * pad), dtype=np.uint8)
    stride = buf.shape[1]
    src = buf.ctypes.data + pad * stride + pad
    mcbuf = _i16(64 * (64 + 8))
    oracle.orc_interp_block.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_ssize_t, C.c_int, C.c_int, C.c_void_p]
    hv = getattr(ref, "_Z20put_epel_hv_fallbackIhEvPslPKT_liiiiS0_i")
    cp = getattr(ref, "_Z19put_epel_8_fallbackPslPKhliiiiS_")
    for fx in range(8):
        for fy in range(8):
            want, got = _i16(w * h), _i16(w * h)
            if fx == 0 and fy == 0:
                cp(ptr(want), C.c_long(w), C.c_void_p(src), C.c_long(stride), w, h, 0, 0, ptr(mcbuf))
            else:
                hv(ptr(want), C.c_long(w), C.c_void_p(src), C.c_long(stride), w, h, fx, fy, ptr(mcbuf), 8)
            oracle.orc_interp_block(0, w, h, src, stride, fx, fy, ptr(got))
            assert np.array_equal(got, want), (fx, fy)


def test_writers_match_reference(oracle, ref):
    """orc_mc_batch's four writers == put_{un,}weighted_pred*_8_fallback on integer-MV predictions."""
    from libde265_b200 import lib as L
    rng = np.random.default_rng(5)
    w = h = 32
    a = rng.integers(-2000, 18000, size=(h, w)).astype(np.int16)
    b = rng.integers(-2000, 18000, size=(h, w)).astype(np.int16)
    for _ in range(30):
        w0, w1, o0, o1 = [int(v) for v in rng.integers(-128, 128, size=4)]
        log2wd = int(rng.integers(6, 14))
        outs = []
        d = np.zeros((h, w), dtype=np.uint8)
        getattr(ref, "_Z28put_weighted_pred_8_fallbackPhlPKsliiiii")(ptr(d), C.c_long(w), ptr(a), C.c_long(w), w, h, w0, o0, log2wd)
        want_uni = d.copy()
        getattr(ref, "_Z30put_weighted_bipred_8_fallbackPhlPKsS1_liiiiiii")(ptr(d), C.c_long(w), ptr(a), ptr(b), C.c_long(w), w, h, w0, o0, w1, o1, log2wd)
        want_bi = d.copy()
        A, B = a.astype(np.int64), b.astype(np.int64)
        got_uni = np.clip(((A * w0 + (1 << (log2wd - 1))) >> log2wd) + o0, 0, 255)
        got_bi = np.clip((A * w0 + B * w1 + ((o0 + o1 + 1) << log2wd)) >> (log2wd + 1), 0, 255)
        assert np.array_equal(got_uni, want_uni) and np.array_equal(got_bi, want_bi)
    d = np.zeros((h, w), dtype=np.uint8)
    getattr(ref, "_Z30put_unweighted_pred_8_fallbackPhlPKslii")(ptr(d), C.c_long(w), ptr(a), C.c_long(w), w, h)
    assert np.array_equal(d, np.clip((a.astype(np.int32) + 32) >> 6, 0, 255))
    getattr(ref, "_Z32put_weighted_pred_avg_8_fallbackPhlPKsS1_lii")(ptr(d), C.c_long(w), ptr(a), ptr(b), C.c_long(w), w, h)
    assert np.array_equal(d, np.clip((a.astype(np.int32) + b + 64) >> 7, 0, 255))

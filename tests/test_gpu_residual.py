"""GPU parity: dense residual kernels (config 2 of BASELINE.json) through the C ABI vs the C oracle."""
import numpy as np
import pytest
from tests.helpers import ptr, make_coeff_blocks
from libde265_b200 import lib as L

pytestmark = pytest.mark.gpu

CASES = [(L.TR_DCT, 2, 0), (L.TR_DCT, 3, 0), (L.TR_DCT, 4, 0), (L.TR_DCT, 5, 0), (L.TR_DST, 2, 0),
         (L.TR_SKIP, 2, 0), (L.TR_SKIP, 3, 1), (L.TR_SKIP, 4, 2), (L.TR_SKIP, 5, 0),
         (L.TR_BYPASS, 2, 0), (L.TR_BYPASS, 5, 1), (L.TR_BYPASS, 3, 2), (L.TR_BYPASS, 4, 0)]


@pytest.mark.parametrize("kind,log2,rdpcm", CASES)
@pytest.mark.parametrize("with_qp", [True, False])
def test_residual_dense_bit_exact(ctx, oracle, kind, log2, rdpcm, with_qp):
    rng = np.random.default_rng(2 + kind * 100 + log2 * 10 + rdpcm + (50 if with_qp else 0))
    nT = 1 << log2
    n = 4099 if log2 <= 3 else (1031 if log2 == 4 else 517)   # ragged: not a multiple of blocks/warp
    coeffs = make_coeff_blocks(rng, n, log2)
    if kind in (L.TR_SKIP, L.TR_BYPASS) and not with_qp:
        coeffs = np.clip(coeffs, -3000, 3000).astype(np.int16)
    qp = rng.integers(0, 52, size=n).astype(np.uint8) if with_qp else None
    dst = rng.integers(0, 256, size=(n, nT, nT), dtype=np.uint8)
    want = dst.copy()
    oracle.orc_residual_batch(kind, log2, rdpcm, n, ptr(coeffs), ptr(qp) if with_qp else None, ptr(want))
    got = ctx.residual_batch(kind, log2, coeffs, dst.copy(), qp, rdpcm)
    bad = np.nonzero((got != want).reshape(n, -1).any(axis=1))[0]
    assert bad.size == 0, "first bad block %d of %d" % (bad[0], n)


@pytest.mark.parametrize("mode", ["dense", "dc", "sparse"])
def test_residual_dense_extremes(ctx, oracle, mode):
    rng = np.random.default_rng(7)
    for log2 in (2, 3, 4, 5):
        nT = 1 << log2
        n = 256
        coeffs = make_coeff_blocks(rng, n, log2, mode)
        if mode == "dense":
            coeffs[0] = 32767
            coeffs[1] = -32768
        dst = rng.integers(0, 256, size=(n, nT, nT), dtype=np.uint8)
        want = dst.copy()
        oracle.orc_residual_batch(0, log2, 0, n, ptr(coeffs), None, ptr(want))
        got = ctx.residual_batch(L.TR_DCT, log2, coeffs, dst.copy())
        assert np.array_equal(got, want)


def test_residual_empty_batch(ctx):
    c = np.zeros((0, 4, 4), dtype=np.int16)
    d = np.zeros((0, 4, 4), dtype=np.uint8)
    ctx.residual_batch(L.TR_DCT, 2, c, d)


def test_residual_full_size_linearity_property(ctx):
    """Size-independent property at a larger size: zero coefficients leave the prediction untouched,
    and a DC-only block adds a constant."""
    rng = np.random.default_rng(11)
    n, log2 = 1 << 14, 5
    dst = rng.integers(0, 256, size=(n, 32, 32), dtype=np.uint8)
    z = np.zeros((n, 32, 32), dtype=np.int16)
    got = ctx.residual_batch(L.TR_DCT, log2, z, dst.copy())
    assert np.array_equal(got, dst)
    z[:, 0, 0] = 64   # (64*64+64)>>7 = 32 ; (64*32+2048)>>12 = 1  => +1 everywhere
    got = ctx.residual_batch(L.TR_DCT, log2, z, dst.copy())
    assert np.array_equal(got, np.minimum(dst.astype(np.int32) + 1, 255).astype(np.uint8))

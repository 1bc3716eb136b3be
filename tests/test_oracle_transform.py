"""CPU tests: the C oracle's transforms against the compiled reference's own scalar functions
(fallback-dct.cc) — this is what pins the oracle."""
import ctypes as C
import numpy as np
import pytest
from tests.helpers import ptr, make_coeff_blocks

REF_ADD = {2: "_Z28transform_4x4_add_8_fallbackPhPKsl", 3: "_Z28transform_8x8_add_8_fallbackPhPKsl",
           4: "_Z30transform_16x16_add_8_fallbackPhPKsl", 5: "_Z30transform_32x32_add_8_fallbackPhPKsl"}


def ref_residual_batch(ref, kind, log2, coeffs, dst, qp=None, rdpcm=0):
    nT = 1 << log2
    n = coeffs.shape[0]
    out = dst.copy()
    deq = getattr(ref, "_Z20dequant_coefficientsPsPKsii")
    for i in range(n):
        c = np.ascontiguousarray(coeffs[i])
        if qp is not None:
            t = np.zeros_like(c)
            deq(ptr(t), ptr(c), log2, int(qp[i]))
            c = t
        d = out[i]
        if kind == 0:
            getattr(ref, REF_ADD[log2])(ptr(d), ptr(c), C.c_long(nT))
        elif kind == 1:
            getattr(ref, "_Z33transform_4x4_luma_add_8_fallbackPhPKsl")(ptr(d), ptr(c), C.c_long(nT))
        else:
            r = np.zeros((nT, nT), dtype=np.int32)
            if kind == 2:
                if rdpcm == 0:
                    getattr(ref, "_Z32transform_skip_residual_fallbackPiPKsiii")(ptr(r), ptr(c), nT, 5 + log2, 12)
                else:
                    name = "_Z16rdpcm_h_fallbackPiPKsiii" if rdpcm == 1 else "_Z16rdpcm_v_fallbackPiPKsiii"
                    getattr(ref, name)(ptr(r), ptr(c), nT, 5 + log2, 12)
            else:
                name = {0: "_Z25transform_bypass_fallbackPiPKsi", 1: "_Z33transform_bypass_rdpcm_h_fallbackPiPKsi",
                        2: "_Z33transform_bypass_rdpcm_v_fallbackPiPKsi"}[rdpcm]
                getattr(ref, name)(ptr(r), ptr(c), nT)
            d[:] = np.clip(d.astype(np.int32) + r, 0, 255).astype(np.uint8)
    return out


def test_dct_matrix_is_orthogonal_like(oracle):
    m = np.ctypeslib.as_array(oracle.orc_dct_matrix(), shape=(32, 32)).astype(np.int64)
    g = m @ m.T
    # HEVC's integer matrix is near-orthogonal: diagonal ~ 32*64*64, off-diagonal tiny
    assert np.all(np.abs(np.diag(g) - 32 * 64 * 64) < 600)
    assert np.max(np.abs(g - np.diag(np.diag(g)))) < 600
    assert list(m[1, :4]) == [90, 90, 88, 85] and list(m[16, :4]) == [64, -64, -64, 64]


@pytest.mark.parametrize("kind,log2,rdpcm", [(0, 2, 0), (0, 3, 0), (0, 4, 0), (0, 5, 0), (1, 2, 0),
                                             (2, 2, 0), (2, 3, 1), (2, 4, 2), (3, 2, 0), (3, 5, 1), (3, 3, 2)])
def test_oracle_residual_matches_reference(oracle, ref, kind, log2, rdpcm):
    rng = np.random.default_rng(100 + kind * 10 + log2)
    n = 96 if log2 < 5 else 40
    nT = 1 << log2
    coeffs = make_coeff_blocks(rng, n, log2)
    if kind >= 2:
        coeffs = np.clip(coeffs, -2000, 2000).astype(np.int16)
    qp = rng.integers(0, 52, size=n).astype(np.uint8) if kind <= 1 else None
    dst = rng.integers(0, 256, size=(n, nT, nT), dtype=np.uint8)
    want = ref_residual_batch(ref, kind, log2, coeffs, dst, qp, rdpcm)
    got = dst.copy()
    oracle.orc_residual_batch(kind, log2, rdpcm, n, ptr(coeffs), ptr(qp) if qp is not None else None, ptr(got))
    assert np.array_equal(got, want)

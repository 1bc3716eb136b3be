"""Test helpers: load the C oracle (oracle/_ref/liboracle.so) and, when present, the compiled
unmodified reference (oracle/_ref/libde265_ref.so).  Only tests/, smoke() and bench.py's
cpu_baseline leg may use these."""
import ctypes as C
import os, subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_DIR = os.path.join(ORACLE_DIR, "_ref")
GOLDEN = os.path.join(ROOT, "tests", "golden")

_oracle = None
_ref = None


def build_oracle():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "oracle"], check=True)


def load_oracle():
    global _oracle
    if _oracle is None:
        build_oracle()
        o = C.CDLL(os.path.join(REF_DIR, "liboracle.so"))
        vp, i32 = C.c_void_p, C.c_int
        o.orc_dct_matrix.restype = C.POINTER(C.c_int8)
        o.orc_residual_batch.argtypes = [i32, i32, i32, i32, vp, vp, vp]
        o.orc_dequant_flat.argtypes = [vp, vp, i32, i32]
        o.orc_residual_tu.argtypes = [vp, vp, vp, vp, i32]
        _oracle = o
    return _oracle


def load_ref():
    """The compiled reference, or None on machines without it (e.g. a GPU box where the prebuilt
    file did not travel)."""
    global _ref
    if _ref is None:
        p = os.path.join(REF_DIR, "libde265_ref.so")
        if not os.path.exists(p):
            return None
        _ref = C.CDLL(p)
    return _ref


def ptr(a):
    return a.ctypes.data_as(C.c_void_p)


# ---- synthetic coefficient blocks ---------------------------------------
def make_coeff_blocks(rng, n, log2, mode="mixed", levels=True):
    """n blocks of nT x nT int16 levels: 50 % 'realistic' sparse low-frequency Laplacian,
    25 % dense uniform int16 (clipping stress), 25 % DC only."""
    nT = 1 << log2
    out = np.zeros((n, nT, nT), dtype=np.int16)
    kind = rng.integers(0, 4, size=n) if mode == "mixed" else np.full(n, {"sparse": 0, "dense": 2, "dc": 3}[mode])
    yy, xx = np.mgrid[0:nT, 0:nT]
    for i in range(n):
        k = kind[i]
        if k <= 1:
            nnz = min(nT * nT, 1 + rng.geometric(8.0 / (nT * nT)))
            # bias towards low frequencies
            w = np.exp(-(xx + yy) / (nT / 3.0)).ravel()
            pos = rng.choice(nT * nT, size=nnz, replace=False, p=w / w.sum())
            vals = np.rint(rng.laplace(0, 12, size=nnz)).astype(np.int64)
            vals[vals == 0] = 1
            out[i].ravel()[pos] = np.clip(vals, -32768, 32767)
        elif k == 2:
            out[i] = rng.integers(-32768, 32768, size=(nT, nT), dtype=np.int64).astype(np.int16)
        else:
            out[i, 0, 0] = rng.integers(-2000, 2000)
    return out

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


# ---- synthetic PU work lists -------------------------------------------
PU_SIZES = [(8, 4), (4, 8), (8, 8), (16, 4), (4, 16), (16, 8), (8, 16), (16, 12), (12, 16), (16, 16), (32, 8),
            (8, 32), (32, 16), (16, 32), (32, 24), (24, 32), (32, 32), (64, 16), (16, 64), (64, 32), (32, 64),
            (64, 48), (48, 64), (64, 64)]


def make_pu_list(rng, width, height, n_refs, n_slices=4, coverage=1.0, mv_range=256, offpic=0.05):
    """Random PUs tiled over the picture (non-overlapping within one layer so results do not depend
    on execution order). Returns (ctypes PU array, ctypes Slice array)."""
    from libde265_b200 import lib as L
    slices = (L.Slice * n_slices)()
    for s in range(n_slices):
        sl = slices[s]
        mode = s % 4          # 0: P default, 1: B default, 2: P weighted, 3: B weighted
        sl.slice_type = 1 if mode in (0, 2) else 0
        sl.weighted = 1 if mode >= 2 else 0
        sl.pps_weighted_pred = 1 if mode == 2 else int(rng.integers(0, 2)) if mode == 3 else 0
        sl.luma_log2_denom = int(rng.integers(0, 8))
        sl.chroma_log2_denom = int(rng.integers(0, 8))
        for l in range(2):
            for i in range(16):
                for c in range(3):
                    sl.weight[l][i][c] = int(rng.integers(-128, 128))
                    sl.offset[l][i][c] = int(rng.integers(-128, 128))
    pus = []
    y = 0
    while y < height:
        x = 0
        row_h = 64
        while x < width:
            w, h = PU_SIZES[int(rng.integers(0, len(PU_SIZES)))]
            # fill a 64x64 cell with copies of this PU size where it fits
            for yy in range(0, 64 - h + 1, h):
                for xx in range(0, 64 - w + 1, w):
                    px, py = x + xx, y + yy
                    if px + w > width or py + h > height or rng.random() > coverage:
                        continue
                    s = int(rng.integers(0, n_slices))
                    bi_ok = (w + h) > 12 and slices[s].slice_type == 0
                    fl = int(rng.integers(1, 4)) if bi_ok else (1 if slices[s].slice_type == 1 else int(rng.integers(1, 3)))
                    pu = L.PU()
                    pu.x, pu.y, pu.w, pu.h, pu.flags, pu.slice = px, py, w, h, fl, s
                    for l in range(2):
                        if rng.random() < offpic:
                            mv = rng.integers(-2000, 2001, size=2)
                        else:
                            mv = rng.integers(-mv_range, mv_range, size=2)
                        pu.mv[l][0], pu.mv[l][1] = int(mv[0]), int(mv[1])
                        pu.ref_idx[l] = int(rng.integers(0, 16))
                        pu.ref_slot[l] = int(rng.integers(0, n_refs))
                    if fl == 3 and rng.random() < 0.1:   # identical motion => collapse rule
                        pu.mv[1][0], pu.mv[1][1], pu.ref_slot[1] = pu.mv[0][0], pu.mv[0][1], pu.ref_slot[0]
                    pus.append(pu)
            x += 64
        y += row_h
    arr = (L.PU * len(pus))(*pus)
    return arr, slices


def make_planes(rng, width, height, kind="noise"):
    cw, ch = (width + 1) // 2, (height + 1) // 2
    if kind == "noise":
        return [rng.integers(0, 256, size=(height, width), dtype=np.uint8),
                rng.integers(0, 256, size=(ch, cw), dtype=np.uint8),
                rng.integers(0, 256, size=(ch, cw), dtype=np.uint8)]
    yy, xx = np.mgrid[0:height, 0:width]
    y = (128 + 90 * np.sin(xx / 11.0) * np.cos(yy / 7.0)).astype(np.uint8)
    return [y, np.ascontiguousarray(y[::2, ::2][:ch, :cw] // 2 + 40), np.ascontiguousarray(255 - y[::2, ::2][:ch, :cw])]

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


# ---- synthetic all-intra frame jobs (frame-level path: intra wavefront + residual + filters) ----
TU_DT = np.dtype([('x', '<u2'), ('y', '<u2'), ('cidx_log2', 'u1'), ('flags', 'u1'), ('qp', 'u1'),
                  ('intra_mode', 'u1'), ('coeff_off', '<u4'), ('ncoeff', '<u2'), ('reserved', '<u2')])
COEFF_DT = np.dtype([('level', '<i2'), ('pos', '<u2')])


class SynthJob:
    """Holds the numpy buffers of one frame job alive and exposes the ctypes FrameJob."""

    def __init__(self, rng, width, height, log2_ctb=5, deblock=True, sao=False, dense=0.6, flags=0, pcm=0.0):
        from libde265_b200 import lib as L
        self.w, self.h, self.log2_ctb = width, height, log2_ctb
        self.rng = rng
        self.tus, self.coeffs = [], []
        self.pcm_prob, self.pcms, self.pcm_samples = pcm, [], []
        self.ncoef = 0
        uw, uh = width // 4, height // 4
        self.unit = np.zeros((uh, uw), dtype=np.uint8)
        self.qp = np.zeros((height // 8, width // 8), dtype=np.int8)
        self.dense = dense
        n = 1 << log2_ctb
        for cy in range(0, height, n):
            for cx in range(0, width, n):
                self._cu(cx, cy, log2_ctb)
        self.tu_arr = np.array(self.tus, dtype=TU_DT)
        self.co_arr = np.array(self.coeffs, dtype=COEFF_DT) if self.coeffs else np.zeros(1, dtype=COEFF_DT)
        nctb = ((width + n - 1) // n) * ((height + n - 1) // n)
        self.slices = (L.Slice * 1)()
        self.slices[0].slice_type = 2
        self.slices[0].lf_across_slices = 1
        self.slices[0].beta_offset = int(rng.integers(-3, 4)) * 2
        self.slices[0].tc_offset = int(rng.integers(-3, 4)) * 2
        self.slices[0].sao_luma = self.slices[0].sao_chroma = 1
        self.sao = None
        if sao:
            self.sao = (L.SAO * nctb)()
            for i in range(nctb):
                t = int(rng.choice([0, 1, 2], p=[0.2, 0.4, 0.4]))
                tc = int(rng.choice([0, 1, 2], p=[0.2, 0.4, 0.4]))
                for c in range(3):
                    self.sao[i].type[c] = t if c == 0 else tc
                    self.sao[i].band_eo[c] = int(rng.integers(0, 32)) if self.sao[i].type[c] == 1 else int(rng.integers(0, 4))
                    for k in range(4):
                        v = int(rng.integers(0, 8))
                        if self.sao[i].type[c] == 2:
                            v = v if k < 2 else -v
                        else:
                            v = int(rng.integers(-7, 8))
                        self.sao[i].offset[c][k] = v
                if self.sao[i].type[1] == 2:   # Cb and Cr share the edge class
                    self.sao[i].band_eo[2] = self.sao[i].band_eo[1]
        J = L.FrameJob()
        J.width, J.height, J.chroma_format_idc, J.bit_depth = width, height, 1, 8
        J.log2_ctb, J.log2_min_cb, J.log2_min_tb = log2_ctb, 3, 2
        J.cb_qp_offset, J.cr_qp_offset = int(rng.integers(-4, 5)), int(rng.integers(-4, 5))
        J.flags = flags | L.FRAME_STRONG_INTRA_SMOOTHING | L.FRAME_LF_ACROSS_TILES | (L.FRAME_DEBLOCK if deblock else 0) | (L.FRAME_SAO if sao else 0)
        J.tus, J.n_tus = L.as_ptr(self.tu_arr, L.TU), len(self.tu_arr)
        J.coeffs, J.n_coeffs = L.as_ptr(self.co_arr, L.Coeff), self.ncoef
        J.slices, J.n_slices = C.cast(self.slices, C.POINTER(L.Slice)), 1
        J.unit_info = L.as_ptr(self.unit, C.c_uint8)
        J.qp_y = L.as_ptr(self.qp, C.c_int8)
        if sao:
            J.sao = C.cast(self.sao, C.POINTER(L.SAO))
        if self.pcms:
            self.pcm_arr = (L.PCM * len(self.pcms))()
            for i, (x, y, lg, off) in enumerate(self.pcms):
                self.pcm_arr[i].x, self.pcm_arr[i].y, self.pcm_arr[i].log2_size, self.pcm_arr[i].sample_off = x, y, lg, off
            self.pcm_buf = np.concatenate(self.pcm_samples)
            J.pcms, J.n_pcms = C.cast(self.pcm_arr, C.POINTER(L.PCM)), len(self.pcms)
            J.pcm_samples, J.n_pcm_samples = L.as_ptr(self.pcm_buf, C.c_uint8), len(self.pcm_buf)
        self.job = J

    def _cu(self, x, y, lg):
        if x >= self.w or y >= self.h:
            return
        n = 1 << lg
        must = x + n > self.w or y + n > self.h
        if lg > 3 and (must or self.rng.random() < 0.55):
            h = n // 2
            for dy in (0, h):
                for dx in (0, h):
                    self._cu(x + dx, y + dy, lg - 1)
            return
        qp = int(self.rng.integers(18, 42))
        self.qp[y // 8:(y + n) // 8, x // 8:(x + n) // 8] = qp
        if self.pcm_prob and self.rng.random() < self.pcm_prob:   # PCM coding block (S4), no TUs
            from libde265_b200 import lib as L
            off = sum(len(a) for a in self.pcm_samples)
            self.pcms.append((x, y, lg, off))
            self.pcm_samples.append(self.rng.integers(0, 256, size=n * n + n * n // 2, dtype=np.uint8))
            u = self.unit[y // 4:(y + n) // 4, x // 4:(x + n) // 4]
            u |= L.UNIT_INTRA | L.UNIT_DONE | (L.UNIT_NOFILT if self.rng.random() < 0.5 else 0)
            if x > 0:
                u[:, 0] |= L.UNIT_EDGE_V
            if y > 0:
                u[0, :] |= L.UNIT_EDGE_H
            return
        nxn = lg == 3 and self.rng.random() < 0.35
        self._tt(x, y, lg, 0, nxn, qp, x, y, 0)

    def _tt(self, x, y, lg, depth, force, qp, xb, yb, blk):
        n = 1 << lg
        if lg > 5 or (lg > 2 and (force or self.rng.random() < 0.4)):
            h = n // 2
            for i, (dx, dy) in enumerate(((0, 0), (h, 0), (0, h), (h, h))):
                self._tt(x + dx, y + dy, lg - 1, depth + 1, False, qp, x, y, i)
            return
        self._tb(0, x, y, lg, qp)
        if lg > 2:
            self._tb(1, x // 2, y // 2, lg - 1, qp)
            self._tb(2, x // 2, y // 2, lg - 1, qp)
        elif blk == 3:
            self._tb(1, xb // 2, yb // 2, 2, qp)
            self._tb(2, xb // 2, yb // 2, 2, qp)

    def _tb(self, c, x, y, lg, qp):
        from libde265_b200 import lib as L
        n = 1 << lg
        rng = self.rng
        mode = int(rng.integers(0, 35))
        nco = 0
        if rng.random() < self.dense:
            nco = int(min(n * n, 1 + rng.geometric(0.25))) if rng.random() < 0.9 else int(rng.integers(1, n * n + 1))
            pos = rng.choice(n * n, size=nco, replace=False) if nco > n * n // 4 else \
                np.unique((rng.integers(0, min(n, 8), size=nco) + n * rng.integers(0, min(n, 8), size=nco)))
            nco = len(pos)
            lev = rng.integers(-40, 41, size=nco)
            lev[lev == 0] = 1
            for p_, l_ in zip(pos, lev):
                self.coeffs.append((int(l_), int(p_)))
        fl = L.TU_INTRA_CU | (L.TU_DST if (c == 0 and lg == 2) else 0)
        self.tus.append((x, y, c | (lg << 2), fl, qp if c == 0 else max(0, qp - 3), mode, self.ncoef, nco, 0))
        self.ncoef += nco
        if c == 0:
            u = self.unit[y // 4:(y + n) // 4, x // 4:(x + n) // 4]
            u |= L.UNIT_INTRA | (L.UNIT_CBF if nco else 0)
            if x > 0:
                u[:, 0] |= L.UNIT_EDGE_V | L.UNIT_TU_EDGE_V
            if y > 0:
                u[0, :] |= L.UNIT_EDGE_H | L.UNIT_TU_EDGE_H

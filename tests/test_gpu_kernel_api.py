"""GPU parity of the kernel-level C ABI (include/b265.h part B) against the C oracle: sparse residual,
isolated intra blocks, raw interpolation slots, whole-plane deblocking / SAO, and PCM blocks in the
frame path.  Everything is integer work: the bar is bit-exact."""
import ctypes as C
import numpy as np
import pytest
from tests.helpers import SynthJob, load_oracle, TU_DT, COEFF_DT
from libde265_b200 import lib as L

pytestmark = pytest.mark.gpu


def _oracle():
    o = load_oracle()
    vp, i32 = C.c_void_p, C.c_int
    o.orc_intra_from_border.argtypes = [i32, i32, i32, vp, i32, i32, vp, i32]
    o.orc_interp_block.argtypes = [i32, i32, i32, vp, C.c_ssize_t, i32, i32, vp]
    return o


@pytest.mark.parametrize("seed,w,h", [(1, 64, 64), (2, 416, 240), (3, 1920, 1080 // 8 * 8)])
def test_residual_sparse(ctx, seed, w, h):
    """random non-overlapping TUs of all sizes / kinds on one plane vs orc_residual_tu (transform.cc:353-625)"""
    o = _oracle()
    rng = np.random.default_rng(seed)
    tus, coeffs = [], []
    for y in range(0, h - 31, 32):
        for x in range(0, w - 31, 32):
            lg = int(rng.integers(2, 6))
            n = 1 << lg
            for by in range(0, 32, n):
                for bx in range(0, 32, n):
                    if rng.random() < 0.2:
                        continue
                    nco = int(rng.integers(1, min(n * n, 40) + 1))
                    pos = rng.choice(n * n, size=nco, replace=False)
                    lev = rng.integers(-300, 301, size=nco)
                    lev[lev == 0] = 7
                    fl = 0
                    r = rng.random()
                    if lg == 2 and r < 0.3:
                        fl |= L.TU_DST
                    elif r < 0.45:
                        fl |= L.TU_SKIP | int(rng.choice([0, L.TU_RDPCM_H, L.TU_RDPCM_V]))
                    elif r < 0.55:
                        fl |= L.TU_BYPASS | int(rng.choice([0, L.TU_RDPCM_H, L.TU_RDPCM_V]))
                    tus.append((x + bx, y + by, 0 | (lg << 2), fl, int(rng.integers(0, 52)), L.INTRA_NONE, len(coeffs), nco, 0))
                    coeffs += [(int(l_), int(p_)) for l_, p_ in zip(lev, pos)]
    tu_arr = np.array(tus, dtype=TU_DT)
    co_arr = np.array(coeffs, dtype=COEFF_DT)
    plane = rng.integers(0, 256, size=(h, w), dtype=np.uint8)
    want = plane.copy()
    for i in range(len(tu_arr)):
        o.orc_residual_tu(tu_arr[i:i + 1].ctypes.data, co_arr.ctypes.data, None, want.ctypes.data, w)
    tu_c = (L.TU * len(tu_arr)).from_buffer(tu_arr)
    co_c = (L.Coeff * len(co_arr)).from_buffer(co_arr)
    got = ctx.residual_sparse(tu_c, co_c, plane)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("log2", [2, 3, 4, 5])
@pytest.mark.parametrize("cidx", [0, 1])
def test_intra_batch(ctx, log2, cidx):
    """every mode x {smooth, noisy, flat} borders x strong/no_edge flags vs orc_intra_from_border (intrapred.cc:816-1069)"""
    o = _oracle()
    rng = np.random.default_rng(100 + log2 * 4 + cidx)
    nT = 1 << log2
    reps = 9
    modes = np.tile(np.arange(35, dtype=np.uint8), reps)
    n = len(modes)
    borders = rng.integers(0, 256, size=(n, 4 * nT + 1), dtype=np.uint8)
    # smooth ramps (trigger the bi-linear strong filter at 32x32) and flat borders
    ramp = (np.linspace(20, 200, 4 * nT + 1)[None, :] + rng.integers(-2, 3, size=(n // 3, 4 * nT + 1))).clip(0, 255)
    borders[: n // 3] = ramp.astype(np.uint8)
    borders[n // 3: n // 3 + 35] = 255
    for strong in (0, 1):
        for noedge in (0, 1):
            got = ctx.intra_batch(log2, cidx, modes, borders, strong, noedge)
            want = np.zeros_like(got)
            for i in range(n):
                o.orc_intra_from_border(log2, cidx, int(modes[i]), borders[i].ctypes.data + 2 * nT, strong, noedge,
                                        want[i].ctypes.data, nT)
            bad = np.argwhere((got != want).any(axis=(1, 2)))
            assert bad.size == 0, "first differing block %d (mode %d) strong=%d noedge=%d" % (
                bad[0][0], modes[bad[0][0]], strong, noedge)


@pytest.mark.parametrize("luma,w,h", [(1, 8, 8), (1, 16, 4), (1, 64, 64), (1, 12, 16), (0, 4, 4), (0, 2, 8), (0, 32, 32),
                                      (0, 6, 8)])
def test_interp_batch(ctx, luma, w, h):
    """put_hevc_qpel_8[xFrac][yFrac] / put_hevc_epel*_8 slots (fallback-motion.cc:298-625): every phase pair"""
    o = _oracle()
    rng = np.random.default_rng(7 + w * 3 + h)
    nph = 4 if luma else 8
    fr = np.array([(fx, fy) for fx in range(nph) for fy in range(nph)] * 2, dtype=np.uint8)
    n = len(fr)
    halo = 4
    stride = w + 2 * halo
    bstride = stride * (h + 2 * halo)
    buf = rng.integers(0, 256, size=n * bstride, dtype=np.uint8)
    buf[: bstride] = 255   # saturating block: largest intermediate values
    buf[bstride: 2 * bstride] = 0
    origin = halo * stride + halo
    got = ctx.interp_batch(luma, w, h, buf, origin, stride, bstride, n, fr)
    want = np.zeros_like(got)
    for i in range(n):
        o.orc_interp_block(luma, w, h, buf.ctypes.data + origin + i * bstride, stride, int(fr[i, 0]), int(fr[i, 1]),
                           want[i].ctypes.data)
    assert np.array_equal(got, want)


def _planes(rng, w, h):
    return [rng.integers(0, 256, size=(h, w), dtype=np.uint8), rng.integers(0, 256, size=(h // 2, w // 2), dtype=np.uint8),
            rng.integers(0, 256, size=(h // 2, w // 2), dtype=np.uint8)]


def _smooth_planes(rng, w, h):
    """low-contrast planes so that the deblocking decisions (dE, dEp, dEq, tc clipping) all occur"""
    out = []
    for (hh, ww) in ((h, w), (h // 2, w // 2), (h // 2, w // 2)):
        base = np.add.outer(np.arange(hh) // 8 * 3, np.arange(ww) // 8 * 5) % 60 + 90
        out.append((base + rng.integers(-3, 4, size=(hh, ww))).astype(np.uint8))
    return out


@pytest.mark.parametrize("seed,w,h,smooth", [(11, 64, 64, True), (12, 416, 240, True), (13, 416, 240, False),
                                             (14, 200, 136, True)])
def test_deblock_planes(ctx, seed, w, h, smooth):
    o = _oracle()
    rng = np.random.default_rng(seed)
    sj = SynthJob(rng, w, h, deblock=True, sao=False, pcm=0.1)
    planes = _smooth_planes(rng, w, h) if smooth else _planes(rng, w, h)
    want = [p.copy() for p in planes]
    pp = (C.c_void_p * 3)(*[p.ctypes.data for p in want])
    pitch = (C.c_int * 3)(w, w // 2, w // 2)
    o.orc_deblock(C.byref(sj.job), pp, pitch)
    assert any(not np.array_equal(a, b) for a, b in zip(want, planes)), "test input does not exercise the filter"
    got = ctx.deblock_planes(sj.job, [p.copy() for p in planes])
    for c in range(3):
        assert np.array_equal(got[c], want[c]), "plane %d" % c


@pytest.mark.parametrize("seed,w,h", [(21, 64, 64), (22, 416, 240), (23, 200, 136)])
def test_sao_planes(ctx, seed, w, h):
    o = _oracle()
    rng = np.random.default_rng(seed)
    sj = SynthJob(rng, w, h, deblock=False, sao=True, pcm=0.1)
    planes = _smooth_planes(rng, w, h)
    want = [np.zeros_like(p) for p in planes]
    ip = (C.c_void_p * 3)(*[p.ctypes.data for p in planes])
    op = (C.c_void_p * 3)(*[p.ctypes.data for p in want])
    pitch = (C.c_int * 3)(w, w // 2, w // 2)
    o.orc_sao(C.byref(sj.job), ip, op, pitch)
    got = ctx.sao_planes(sj.job, planes)
    for c in range(3):
        assert np.array_equal(got[c], want[c]), "plane %d" % c


@pytest.mark.parametrize("seed,w,h,deblock,sao", [(31, 64, 64, False, False), (32, 416, 240, True, True),
                                                  (33, 200, 136, True, False)])
def test_pcm_blocks_in_frame(ctx, seed, w, h, deblock, sao):
    """S4: PCM coding blocks are stored before the intra wavefront, serve as intra neighbours and honour
    pcm_loop_filter_disabled (UNIT_NOFILT) in the deblocking filter"""
    from tests.test_gpu_frames import oracle_frame
    rng = np.random.default_rng(seed)
    sj = SynthJob(rng, w, h, deblock=deblock, sao=sao, pcm=0.25)
    assert sj.job.n_pcms > 0
    recon, dbk, outs = oracle_frame(sj.job, w, h)
    want = outs if sao else (dbk if deblock else recon)
    sid = ctx.surface_alloc(w, h)
    ctx.surface_fill(sid, 0, 0)
    sj.job.dst_surface = sid
    ctx.submit_frames([sj.job])
    got = ctx.surface_download(sid, w, h)
    ctx.surface_release(sid)
    for c in range(3):
        assert np.array_equal(got[c], want[c]), "plane %d" % c

"""GPU parity: motion compensation (config 3 of BASELINE.json) through b265_mc_batch vs the C oracle."""
import ctypes as C
import numpy as np
import pytest
from tests.helpers import make_pu_list, make_planes
from libde265_b200 import lib as L

pytestmark = pytest.mark.gpu


def oracle_mc(oracle, pus, slices, refs, dst, w, h):
    rp = (C.c_void_p * (3 * len(refs)))(*[p.ctypes.data for r in refs for p in r])
    dp = (C.c_void_p * 3)(*[p.ctypes.data for p in dst])
    oracle.orc_mc_batch(C.byref(pus), len(pus), C.byref(slices), rp, len(refs), dp, w, h,
                        dst[0].shape[1], dst[1].shape[1])


@pytest.mark.parametrize("w,h,seed", [(416, 240, 3), (200, 136, 4), (1920, 1080, 5)])
def test_mc_bit_exact(ctx, oracle, w, h, seed):
    rng = np.random.default_rng(seed)
    refs = [make_planes(rng, w, h, "noise"), make_planes(rng, w, h, "smooth"), make_planes(rng, w, h, "noise")]
    pus, slices = make_pu_list(rng, w, h, len(refs), coverage=1.0 if w < 1000 else 0.35)
    base = make_planes(rng, w, h, "noise")
    want = [p.copy() for p in base]
    oracle_mc(oracle, pus, slices, refs, want, w, h)
    got = ctx.mc_batch(pus, slices, refs, [p.copy() for p in base], w, h)
    for c in range(3):
        bad = np.argwhere(got[c] != want[c])
        assert bad.size == 0, "plane %d first mismatch at (y,x)=%s of %d" % (c, bad[0], len(bad))


def test_mc_all_phases_small_blocks(ctx, oracle):
    """every luma (16) and chroma (64) phase on 8x4 / 4x8 blocks next to the picture corner"""
    rng = np.random.default_rng(9)
    w, h = 64, 64
    refs = [make_planes(rng, w, h, "noise")]
    slices = (L.Slice * 1)()
    slices[0].slice_type = 1
    base = make_planes(rng, w, h, "noise")
    for fy in range(8):
        pus = (L.PU * 8)()
        for fx in range(8):
            pu = pus[fx]
            pu.x, pu.y, pu.w, pu.h, pu.flags = 8 * fx, 8 * fy, 8, 4, 1
            pu.mv[0][0], pu.mv[0][1] = fx - 70, fy + 9
        want = [p.copy() for p in base]
        oracle_mc(oracle, pus, slices, refs, want, w, h)
        got = ctx.mc_batch(pus, slices, refs, [p.copy() for p in base], w, h)
        for c in range(3):
            assert np.array_equal(got[c], want[c])


@pytest.mark.parametrize("w,h,seed", [(416, 240, 13), (1920, 1080, 14)])
def test_mc_frame_path(ctx, oracle, w, h, seed):
    """the same PU lists through the FRAME-level path (b265_submit_frames: references and destination are DPB
    surfaces in HBM, lists uploaded with the job), two pictures per call"""
    rng = np.random.default_rng(seed)
    refs = [make_planes(rng, w, h, "noise"), make_planes(rng, w, h, "smooth")]
    ref_sids = []
    for r in refs:
        sid = ctx.surface_alloc(w, h)
        ctx.surface_upload(sid, r)
        ref_sids.append(sid)
    jobs, keep, want, sids = [], [], [], []
    for k in range(2):
        pus, slices = make_pu_list(rng, w, h, len(refs), coverage=0.9)
        base = make_planes(rng, w, h, "noise")
        sid = ctx.surface_alloc(w, h)
        ctx.surface_upload(sid, base)
        sids.append(sid)
        exp = [p.copy() for p in base]
        oracle_mc(oracle, pus, slices, refs, exp, w, h)
        want.append(exp)
        J = L.FrameJob()
        J.width, J.height, J.chroma_format_idc, J.bit_depth = w, h, 1, 8
        J.log2_ctb, J.log2_min_cb, J.log2_min_tb = 5, 3, 2
        J.dst_surface = sid
        for i, r in enumerate(ref_sids):
            J.ref_surfaces[i] = r
            J.ref_poc[i] = 10 + i
        J.n_refs = len(ref_sids)
        J.pus = C.cast(pus, C.POINTER(L.PU)); J.n_pus = len(pus)
        J.slices = C.cast(slices, C.POINTER(L.Slice)); J.n_slices = len(slices)
        keep += [pus, slices]
        jobs.append(J)
    ctx.submit_frames(jobs)
    ctx.sync()
    for k in range(2):
        got = ctx.surface_download(sids[k], w, h)
        for c in range(3):
            bad = np.argwhere(got[c] != want[k][c])
            assert bad.size == 0, "picture %d plane %d first mismatch at (y,x)=%s of %d" % (k, c, bad[0], len(bad))
    for s in sids + ref_sids:
        ctx.surface_release(s)

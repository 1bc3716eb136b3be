"""GPU parity: frame-level path (b265_submit_frames): intra wavefront + residual, deblocking, SAO on
synthetic all-intra work lists, against the C oracle (orc_reconstruct_frame / orc_deblock / orc_sao)."""
import ctypes as C
import numpy as np
import pytest
from tests.helpers import SynthJob, load_oracle
from libde265_b200 import lib as L

pytestmark = pytest.mark.gpu


def oracle_frame(job, w, h):
    o = load_oracle()
    planes = [np.zeros((h, w), np.uint8), np.zeros((h // 2, w // 2), np.uint8), np.zeros((h // 2, w // 2), np.uint8)]
    pp = (C.c_void_p * 3)(*[p.ctypes.data for p in planes])
    pitch = (C.c_int * 3)(w, w // 2, w // 2)
    o.orc_reconstruct_frame(C.byref(job), pp, pitch, None)
    recon = [p.copy() for p in planes]
    o.orc_deblock(C.byref(job), pp, pitch)
    dbk = [p.copy() for p in planes]
    outs = [np.zeros_like(p) for p in planes]
    op = (C.c_void_p * 3)(*[p.ctypes.data for p in outs])
    o.orc_sao(C.byref(job), pp, op, pitch)
    return recon, dbk, outs


def first_diff(a, b):
    d = np.argwhere(a != b)
    return None if d.size == 0 else tuple(d[0])


@pytest.mark.parametrize("w,h,deblock,sao,seed", [(64, 64, False, False, 1), (416, 240, False, False, 2),
                                                  (416, 240, True, False, 3), (416, 240, True, True, 4),
                                                  (200, 136, True, True, 5), (832, 480, True, True, 6)])
def test_intra_frame_bit_exact(ctx, w, h, deblock, sao, seed):
    rng = np.random.default_rng(seed)
    sj = SynthJob(rng, w, h, deblock=deblock, sao=sao)
    recon, dbk, outs = oracle_frame(sj.job, w, h)
    want = outs if sao else (dbk if deblock else recon)
    sid = ctx.surface_alloc(w, h)
    ctx.surface_fill(sid, 0, 0)
    sj.job.dst_surface = sid
    ctx.submit_frames([sj.job])
    got = ctx.surface_download(sid, w, h)
    ctx.surface_release(sid)
    for c in range(3):
        assert first_diff(got[c], want[c]) is None, "plane %d first diff (y,x)=%s; recon-stage diff=%s" % (
            c, first_diff(got[c], want[c]), first_diff(got[c], recon[c]))


def test_batch_of_frames(ctx):
    rng = np.random.default_rng(9)
    w, h = 320, 192
    jobs = [SynthJob(rng, w, h, deblock=True, sao=(i % 2 == 0)) for i in range(6)]
    sids = []
    for sj in jobs:
        sid = ctx.surface_alloc(w, h)
        sj.job.dst_surface = sid
        sids.append(sid)
    ctx.submit_frames([sj.job for sj in jobs])
    for i, sj in enumerate(jobs):
        recon, dbk, outs = oracle_frame(sj.job, w, h)
        want = outs if i % 2 == 0 else dbk
        got = ctx.surface_download(sids[i], w, h)
        for c in range(3):
            assert np.array_equal(got[c], want[c]), (i, c)
    for s in sids:
        ctx.surface_release(s)

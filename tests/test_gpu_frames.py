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


def test_batch_larger_than_a_launch_group(ctx):
    """300 independent pictures in ONE submit: frames.cu cuts launch groups at 255 pictures (work-list items name their picture
    in 8 bits relative to the group); every picture still equals the C oracle.  Four distinct synthetic pictures, repeated."""
    rng = np.random.default_rng(10)
    w, h = 128, 64
    base = [SynthJob(rng, w, h, log2_ctb=4 + (i % 2), deblock=(i != 1), sao=(i >= 2)) for i in range(4)]
    want = []
    for i, sj in enumerate(base):
        recon, dbk, outs = oracle_frame(sj.job, w, h)
        want.append(outs if i >= 2 else (dbk if i != 1 else recon))
    jobs, sids = [], []
    for k in range(300):
        J = L.FrameJob()
        C.memmove(C.byref(J), C.byref(base[k % 4].job), C.sizeof(L.FrameJob))
        J.dst_surface = ctx.surface_alloc(w, h)
        sids.append(J.dst_surface)
        jobs.append(J)
    ctx.submit_frames(jobs)
    for k in (0, 1, 2, 3, 127, 254, 255, 256, 257, 299):
        got = ctx.surface_download(sids[k], w, h)
        for c in range(3):
            assert np.array_equal(got[c], want[k % 4][c]), (k, c)
    for s_ in sids:
        ctx.surface_release(s_)


@pytest.mark.parametrize("name", ["c1_1080p_q32_lf", "gen_pb_1080p_ld32"])
def test_concurrent_contexts_share_the_device(name):
    """Four contexts each run a batch of the same stream at the same time, several rounds without a host sync: the intra
    wavefront of each takes a share of the SMs by the number of busy peers (frames.cu, b265_busy_peers), every picture of
    every context still has the reference decoder's MD5.  b265_stream_wait chains the contexts' streams on the device."""
    import ctypes as C, hashlib
    from libde265_b200 import lib as L
    from benchmarks.jobs import stream_jobs
    from tests.test_parser_streams import load_stream
    data, gold = load_stream(name)
    W, H = gold["width"], gold["height"]
    uniq = stream_jobs(data, threads=4)
    ctxs = [L.Context(0) for _ in range(4)]
    held = []
    for cx in ctxs:
        jobs, sids, uid2sid = [], [], {}
        for oj in uniq:
            J = L.FrameJob()
            C.memmove(C.byref(J), C.byref(oj.job), C.sizeof(L.FrameJob))
            sid = cx.surface_alloc(W, H)
            uid2sid[oj.job.dst_surface] = sid
            for r in range(J.n_refs):
                J.ref_surfaces[r] = uid2sid[oj.job.ref_surfaces[r]]
            J.dst_surface = sid
            sids.append(sid)
            jobs.append(J)
        held.append((jobs, sids, cx.upload_frames(jobs)))
    for cx in ctxs[1:]:
        cx.wait_for(ctxs[0])
    for _ in range(3):
        for cx, (_, _, b) in zip(ctxs, held):
            cx.run_batch(b)
    for cx in ctxs[1:]:
        ctxs[0].wait_for(cx)
    ctxs[0].sync()          # the chain: context 0's stream has waited for every other stream
    assert len(uniq) == gold["frames"]      # both streams are in output order = decoding order
    for k, (cx, (_, sids, b)) in enumerate(zip(ctxs, held)):
        for i in range(len(uniq)):
            planes = cx.surface_download(sids[i], W, H)
            got = [hashlib.md5(p.tobytes()).hexdigest() for p in planes]
            assert got == gold["md5"][i], "context %d, picture %d of %s differs from the reference decoder" % (k, i, name)
        cx.free_batch(b)
    for cx in ctxs:
        cx.close()

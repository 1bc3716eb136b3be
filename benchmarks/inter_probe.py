"""Per-stage kernel time of the frame path on the first n pictures of a P/B stream (one batch, launch groups inside).
usage: inter_probe.py <stream> <n>[,<n>...]   (with B265_INTRA_TRACE=<file>: the trace holds the LAST launch group)"""
import ctypes as C
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from libde265_b200 import lib as L
from benchmarks.jobs import stream_jobs

name = sys.argv[1]
uniq = stream_jobs(name, threads=8)
ctx = L.Context(0)
ctx.lib.b265_set_stage_timing(ctx.h, 1)
for n in [int(x) for x in sys.argv[2].split(",")]:
    jobs, sids, uid2sid = [], [], {}
    for oj in uniq[:n]:
        J = L.FrameJob()
        C.memmove(C.byref(J), C.byref(oj.job), C.sizeof(L.FrameJob))
        sid = ctx.surface_alloc(J.width, J.height)
        uid2sid[oj.job.dst_surface] = sid
        for r in range(J.n_refs):
            J.ref_surfaces[r] = uid2sid[oj.job.ref_surfaces[r]]
        J.dst_surface = sid
        sids.append(sid)
        jobs.append(J)
    b = ctx.upload_frames(jobs)
    ctx.sync()
    ts = []
    for it in range(3):
        ctx.flush_l2()
        ctx.run_batch(b)
        ts.append(ctx.stage_times())
    ctx.free_batch(b)
    for s in sids:
        ctx.surface_release(s)
    last = jobs[-1]
    print(json.dumps({"stream": name, "pictures": n, "last_picture": {"n_refs": last.n_refs, "n_pus": last.n_pus, "n_tus": last.n_tus},
                      "stages_ms[mc,res+avail,intra,dbk,sao,total]": [round(x, 3) for x in ts[-1]]}))
ctx.close()

"""Per-stage kernel time of the frame path on a golden stream for several batch sizes."""
import json, sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from libde265_b200 import lib as L
from benchmarks.jobs import stream_jobs

name = sys.argv[1] if len(sys.argv) > 1 else "c1_1080p_q32"
jobs = stream_jobs(name, threads=8)
ctx = L.Context(0)
ctx.lib.b265_set_stage_timing(ctx.h, 1)
batches = [int(x) for x in sys.argv[2].split(',')] if len(sys.argv) > 2 else [1, 2, 4, 8, 16]
for nb in batches:
    sids = []
    js = []
    import ctypes as C
    for i in range(nb):   # the unique pictures of the stream, repeated
        oj = jobs[i % len(jobs)]
        J = L.FrameJob()
        C.memmove(C.byref(J), C.byref(oj.job), C.sizeof(L.FrameJob))
        sid = ctx.surface_alloc(J.width, J.height)
        J.dst_surface = sid
        sids.append(sid)
        js.append(J)
    t0 = time.perf_counter()
    b = ctx.upload_frames(js)
    ctx.sync()
    t_up = time.perf_counter() - t0
    ts = []
    for it in range(4):
        ctx.flush_l2()
        ctx.run_batch(b)
        ts.append(ctx.stage_times())
    ctx.free_batch(b)
    for s in sids:
        ctx.surface_release(s)
    print(json.dumps({"stream": name, "batch": nb, "upload_ms": t_up * 1e3, "tus": sum(j.n_tus for j in js),
                      "stages_ms[mc,res+avail,intra,dbk,sao,total]": [round(x, 3) for x in ts[-1]], "first": [round(x, 3) for x in ts[0]]}))
ctx.close()

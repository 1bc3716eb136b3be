"""Developer probe: do two batches on two contexts (two CUDA streams) overlap?  The intra wavefront is latency bound and
leaves issue slots idle; the residual / deblocking / SAO stages of another batch can use them when the persistent
wavefront grid leaves room (B265_INTRA_CTAS_PER_SM).  Prints ms per batch for 1 and 2 contexts.
    python benchmarks/overlap_probe.py [stream_4k|stream_1080p] [n_contexts]"""
import ctypes as C, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from benchmarks.workloads import STREAM_WL, STREAMS, _stream_bytes
from benchmarks.jobs import stream_jobs
from libde265_b200 import lib as L

wl = sys.argv[1] if len(sys.argv) > 1 else "stream_4k"
nctx = int(sys.argv[2]) if len(sys.argv) > 2 else 2
name, patched, loops, kbatch, desc = STREAM_WL[wl]
kbatch = int(os.environ.get("PROBE_BATCH", kbatch))
data = _stream_bytes(name, patched)
gold = json.load(open(os.path.join(STREAMS, name + ".md5.json")))
W, H = gold["width"], gold["height"]
uniq = stream_jobs(data, threads=8)
ctxs, batches = [], []
for k in range(nctx):
    ctx = L.Context(0)
    jobs = []
    for i in range(kbatch):
        J = L.FrameJob()
        C.memmove(C.byref(J), C.byref(uniq[i % len(uniq)].job), C.sizeof(L.FrameJob))
        J.dst_surface = ctx.surface_alloc(W, H)
        jobs.append(J)
    batches.append(ctx.upload_frames(jobs))
    ctx.sync()
    ctxs.append(ctx)
for n in range(1, nctx + 1):
    for _ in range(3):
        for c, b in zip(ctxs[:n], batches[:n]):
            c.run_batch(b)
    for c in ctxs[:n]:
        c.sync()
    reps = int(os.environ.get("PROBE_REPS", 10))
    t0 = time.perf_counter()
    for _ in range(reps):
        for c, b in zip(ctxs[:n], batches[:n]):
            c.run_batch(b)
    for c in ctxs[:n]:
        c.sync()
    dt = time.perf_counter() - t0
    print("%s: %d context(s): %.3f ms per batch of %d pictures (%.0f pictures/s), B265_INTRA_CTAS_PER_SM=%s" % (
        wl, n, dt / reps / n * 1e3, kbatch, reps * n * kbatch / dt, os.environ.get("B265_INTRA_CTAS_PER_SM", "default")))

import sys, ctypes as C
sys.path.insert(0, '/root/repo')
import numpy as np
from tests.helpers import SynthJob
from tests.test_gpu_frames import oracle_frame
from libde265_b200 import lib as L
ctx = L.Context(0)
w, h, seed, dense = 64, 64, 2, 0.0
rng = np.random.default_rng(seed)
sj = SynthJob(rng, w, h, deblock=False, sao=False, dense=dense)
recon, _, _ = oracle_frame(sj.job, w, h)
sid = ctx.surface_alloc(w, h)
ctx.surface_fill(sid, 0, 0)
sj.job.dst_surface = sid
ctx.submit_frames([sj.job])
got = ctx.surface_download(sid, w, h)
for i, t in enumerate(sj.tu_arr):
    c = t['cidx_log2'] & 3; lg = t['cidx_log2'] >> 2; n = 1 << lg
    if c != 0 or t['x'] >= 32 or t['y'] >= 32: continue
    a = got[c][t['y']:t['y'] + n, t['x']:t['x'] + n]; b = recon[c][t['y']:t['y'] + n, t['x']:t['x'] + n]
    ok = np.array_equal(a, b)
    print(i, "x=%d y=%d n=%d mode=%d %s" % (t['x'], t['y'], n, t['intra_mode'], "ok" if ok else "BAD"))
    if not ok:
        print(a); break

import sys, os
sys.path.insert(0, '/root/repo')
import numpy as np
from tests.helpers import SynthJob
from libde265_b200 import lib as L
ctx = L.Context(0)
for seed in (1, 2, 3):
    rng = np.random.default_rng(seed)
    sj = SynthJob(rng, 32, 32, deblock=False, sao=False, dense=0.6)
    sid = ctx.surface_alloc(32, 32)
    sj.job.dst_surface = sid
    for it in range(3):
        ctx.submit_frames([sj.job]); ctx.sync()
    a = np.fromfile(os.environ["B265_INTRA_TRACE"], dtype=np.uint64).reshape(-1, 12).astype(np.int64)
    for row in a:
        t0, t1, t2, t3, t4, ntb, g, f, p, s, gt, sm = row
        lacc = sm >> 16
        print("seed %d: ntb %3d wait %6d tile %5d tbs %7d pub %5d | per TB: total %5.0f gather %4.0f filter %4.0f predict %4.0f store %4.0f head %4.0f" % (
            seed, ntb, t1 - t0, t2 - t1, t3 - t2, t4 - t3, (t3 - t2) / max(ntb, 1), g / max(ntb, 1), f / max(ntb, 1), p / max(ntb, 1), s / max(ntb, 1), lacc / max(ntb, 1)))

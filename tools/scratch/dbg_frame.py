import sys, ctypes as C
sys.path.insert(0, '/root/repo')
import numpy as np
from tests.helpers import SynthJob
from tests.test_gpu_frames import oracle_frame
from libde265_b200 import lib as L
ctx = L.Context(0)
for (w, h, seed, dense) in [(64, 64, 1, 0.6), (64, 64, 2, 0.0), (416, 240, 3, 0.0), (416, 240, 4, 0.6)]:
    rng = np.random.default_rng(seed)
    sj = SynthJob(rng, w, h, deblock=False, sao=False, dense=dense)
    recon, _, _ = oracle_frame(sj.job, w, h)
    sid = ctx.surface_alloc(w, h)
    ctx.surface_fill(sid, 0, 0)
    sj.job.dst_surface = sid
    ctx.submit_frames([sj.job])
    got = ctx.surface_download(sid, w, h)
    ctx.surface_release(sid)
    nbad = 0
    stats = {}
    for t in sj.tu_arr:
        c = t['cidx_log2'] & 3; lg = t['cidx_log2'] >> 2; n = 1 << lg
        a = got[c][t['y']:t['y'] + n, t['x']:t['x'] + n]; b = recon[c][t['y']:t['y'] + n, t['x']:t['x'] + n]
        if not np.array_equal(a, b):
            key = (int(c), int(lg), int(t['intra_mode']), bool(t['ncoeff']))
            stats[key] = stats.get(key, 0) + 1
            if nbad < 3:
                d = np.argwhere(a != b)
                print("  bad TB c=%d x=%d y=%d n=%d mode=%d ncoeff=%d flags=%x ndiff=%d first=%s got=%d want=%d" % (c, t['x'], t['y'], n, t['intra_mode'], t['ncoeff'], t['flags'], len(d), d[0], a[tuple(d[0])], b[tuple(d[0])]))
            nbad += 1
    print(w, h, seed, dense, "bad TBs", nbad, "of", len(sj.tu_arr))
    modes = {}
    for k, v in stats.items():
        modes.setdefault((k[0], k[1], k[3]), []).append(k[2])
    for k in sorted(modes): print("   c,lg,coded", k, "modes", sorted(modes[k]))

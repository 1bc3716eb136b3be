"""Locate SAO mismatches of the frame path against the oracle (developer tool, GPU)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from tests.helpers import SynthJob
from tests.test_gpu_frames import oracle_frame
from libde265_b200 import lib as L

ctx = L.Context(0)
w, h, seed = 416, 240, 4
rng = np.random.default_rng(seed)
sj = SynthJob(rng, w, h, deblock=True, sao=True)
recon, dbk, outs = oracle_frame(sj.job, w, h)
sid = ctx.surface_alloc(w, h)
ctx.surface_fill(sid, 0, 0)
sj.job.dst_surface = sid
ctx.submit_frames([sj.job])
got = ctx.surface_download(sid, w, h)
ctbs_w = (w + 31) // 32
from collections import Counter
for c in range(3):
    d = np.argwhere(got[c] != outs[c])
    print("plane", c, "ndiff", len(d))
    lc = 5 if c == 0 else 4
    cnt = Counter()
    for (y, x) in d:
        ctb = (y >> lc) * ctbs_w + (x >> lc)
        s = sj.sao[ctb]
        cnt[(s.type[c], s.band_eo[c] if s.type[c] == 2 else -1)] += 1
    print("  by (type, class):", dict(cnt))
    for (y, x) in d[:6]:
        ctb = (y >> lc) * ctbs_w + (x >> lc)
        s = sj.sao[ctb]
        print("  y,x", y, x, "type", s.type[c], "cls/pos", s.band_eo[c], "off", list(s.offset[c]), "in", dbk[c][y, x], "got", got[c][y, x], "want", outs[c][y, x],
              "nbrs", dbk[c][max(y-1,0):y+2, max(x-1,0):x+2].tolist())

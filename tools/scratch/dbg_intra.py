import sys, ctypes as C
sys.path.insert(0, '/root/repo')
import numpy as np
from tests.helpers import load_oracle
from libde265_b200 import lib as L
o = load_oracle()
vp, i32 = C.c_void_p, C.c_int
o.orc_intra_from_border.argtypes = [i32, i32, i32, vp, i32, i32, vp, i32]
ctx = L.Context(0)
rng = np.random.default_rng(5)
for log2 in (2, 3, 4, 5):
    nT = 1 << log2
    for cidx in (0, 1):
        modes = np.tile(np.arange(35, dtype=np.uint8), 6)
        n = len(modes)
        borders = rng.integers(0, 256, size=(n, 4 * nT + 1), dtype=np.uint8)
        ramp = (np.linspace(20, 200, 4 * nT + 1)[None, :] + rng.integers(-2, 3, size=(70, 4 * nT + 1))).clip(0, 255)
        borders[:70] = ramp.astype(np.uint8)
        for strong in (0, 1):
            for noedge in (0, 1):
                got = ctx.intra_batch(log2, cidx, modes, borders, strong, noedge)
                want = np.zeros_like(got)
                for i in range(n):
                    o.orc_intra_from_border(log2, cidx, int(modes[i]), borders[i].ctypes.data + 2 * nT, strong, noedge, want[i].ctypes.data, nT)
                bad = sorted(set(int(modes[i]) for i in range(n) if not np.array_equal(got[i], want[i])))
                if bad:
                    i = [i for i in range(n) if not np.array_equal(got[i], want[i])][0]
                    d = np.argwhere(got[i] != want[i])
                    print("log2", log2, "cidx", cidx, "strong", strong, "noedge", noedge, "bad modes", bad, "first blk", i, "mode", modes[i], "ndiff", len(d), "first", d[0], got[i][tuple(d[0])], want[i][tuple(d[0])])
print("done")

"""Summarise a B265_INTRA_TRACE dump (k_intra.cu TRACE build): per-ticket phase times in SM cycles."""
import sys
import numpy as np
a = np.fromfile(sys.argv[1], dtype=np.uint64).reshape(-1, 12).astype(np.int64)
t0, t1, t2, t3, t4, ntb, a_g, a_f, a_p, a_s, gt, sm = a.T
lacc = sm >> 16
sm = sm & 0xffff
print("tickets %d, TBs %d, span %.3f ms (globaltimer)" % (len(a), ntb.sum(), (gt.max() - gt.min()) / 1e6))
ph = {"wait+fence": t1 - t0, "tile": t2 - t1, "tbs": t3 - t2, "publish": t4 - t3, "total": t4 - t0}
for k, v in ph.items():
    print("%-11s mean %8.0f  p50 %8.0f  p90 %8.0f  max %8.0f cycles" % (k, v.mean(), np.median(v), np.percentile(v, 90), v.max()))
work = (t2 - t1) + (t3 - t2) + (t4 - t3)
print("work (excl. wait) mean %.0f cycles = %.2f us @1.965GHz" % (work.mean(), work.mean() / 1965))
for n in sorted(set(ntb.tolist()))[:12] + [int(ntb.max())]:
    m = ntb == n
    print("  n_tbs=%3d: %6d tickets, tbs phase mean %8.0f cycles (%.0f / TB)" % (n, m.sum(), (t3 - t2)[m].mean(), (t3 - t2)[m].mean() / max(n, 1)))
tb = np.maximum(ntb, 1)
print("per TB: gather %.0f  filter %.0f  predict %.0f  add+store %.0f  (loop overhead %.0f) cycles" % (
    a_g.sum() / ntb.sum(), a_f.sum() / ntb.sum(), a_p.sum() / ntb.sum(), a_s.sum() / ntb.sum(),
    ((t3 - t2).sum() - (a_g + a_f + a_p + a_s).sum()) / ntb.sum()))
m = ntb >= 12
print("tickets with >=12 TBs: gather %.0f filter %.0f predict %.0f add+store %.0f loop %.0f" % (
    a_g[m].sum() / ntb[m].sum(), a_f[m].sum() / ntb[m].sum(), a_p[m].sum() / ntb[m].sum(), a_s[m].sum() / ntb[m].sum(),
    ((t3 - t2)[m].sum() - (a_g + a_f + a_p + a_s)[m].sum()) / ntb[m].sum()))
print("loop head (ffs+shfl+decode) per TB: %.0f cycles" % (lacc.sum() / ntb.sum()))

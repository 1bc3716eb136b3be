"""Summarise a B265_INTRA_TRACE dump (k_intra.cu TRACE build): per-ticket phase times in SM cycles.
One 12 x uint64 record per (plane, CTB) slot of the batch; slots without a ticket are all zero."""
import sys
import numpy as np
a = np.fromfile(sys.argv[1], dtype=np.uint64).reshape(-1, 12).astype(np.int64)
a = a[a[:, 4] != 0]
t0, t1, t2, t3, t4, ntb, a_g, a_f, a_p, a_s, gt, sm = a.T
how = sm >> 32
sm = sm & 0xffff
print("tickets %d (%.1f %% with the edge store reused), TBs %d, span %.3f ms (globaltimer)" % (len(a), 100.0 * (how == 1).mean(), ntb.sum(), (gt.max() - gt.min()) / 1e6))
ph = {"setup/shift": t1 - t0, "borders": t2 - t1, "tbs": t3 - t2, "post+release": t4 - t3, "total": t4 - t0}
for sel, name in ((how == 1, "next CTB of the row, edge store reused"), (how == 2, "next CTB of the row, full fetch"), (how == 0, "first CTB after taking the row")):
    if not sel.any():
        continue
    print("-- %s: %d tickets" % (name, sel.sum()))
    for k, v in ph.items():
        v = v[sel]
        print("  %-12s mean %8.0f  p50 %8.0f  p90 %8.0f  max %8.0f cycles" % (k, v.mean(), np.median(v), np.percentile(v, 90), v.max()))
for n in sorted(set(ntb.tolist()))[:12] + [int(ntb.max())]:
    m = ntb == n
    print("  n_tbs=%3d: %6d tickets, tbs phase mean %8.0f cycles (%.0f / TB)" % (n, m.sum(), (t3 - t2)[m].mean(), (t3 - t2)[m].mean() / max(n, 1)))
print("per TB: gather %.0f  filter %.0f  predict %.0f  add+store %.0f  (loop overhead %.0f) cycles" % (
    a_g.sum() / ntb.sum(), a_f.sum() / ntb.sum(), a_p.sum() / ntb.sum(), a_s.sum() / ntb.sum(),
    ((t3 - t2).sum() - (a_g + a_f + a_p + a_s).sum()) / ntb.sum()))
m = ntb >= 12
if m.any():
    print("tickets with >=12 TBs: gather %.0f filter %.0f predict %.0f add+store %.0f loop %.0f" % (
        a_g[m].sum() / ntb[m].sum(), a_f[m].sum() / ntb[m].sum(), a_p[m].sum() / ntb[m].sum(), a_s[m].sum() / ntb[m].sum(),
        ((t3 - t2)[m].sum() - (a_g + a_f + a_p + a_s)[m].sum()) / ntb[m].sum()))
# busy warps over time (globaltimer ns, 20 bins)
dur_ns = (t4 - t0) / 1.965
lo, hi = gt.min(), (gt + dur_ns).max()
bins = np.linspace(lo, hi, 21)
busy = np.zeros(20)
for i in range(20):
    ov = np.minimum(gt + dur_ns, bins[i + 1]) - np.maximum(gt, bins[i])
    busy[i] = np.clip(ov, 0, None).sum() / (bins[i + 1] - bins[i])
print("warps busy (incl. waiting inside a ticket) per 1/20 of the span:", " ".join("%d" % b for b in busy))
work_ns = ((t4 - t1) / 1.965)
print("sum of ticket time %.3f ms-warps; / span = %.0f warps busy on average" % (dur_ns.sum() / 1e6, dur_ns.sum() / (hi - lo)))

"""Workloads behind bench.py.

idct_sweep  (BASELINE.json configs[1]): synthetic TU coefficient sweep — DST 4x4 and DCT 4/8/16/32,
            random QP, dequant + inverse transform + add, dense-input form (4*nT^2 B per block).
"""
import ctypes as C
import json, os, sys, time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SWEEP = [("DST4", 1, 2), ("DCT4", 0, 2), ("DCT8", 0, 3), ("DCT16", 0, 4), ("DCT32", 0, 5)]


def _sweep_inputs(log2, px_total, seed):
    """px_total pixels worth of blocks; a 2^21-pixel random set tiled (distinct addresses, so no cache reuse)."""
    from tests.helpers import make_coeff_blocks
    rng = np.random.default_rng(seed)
    nT = 1 << log2
    uniq = max(1, (1 << 21) // (nT * nT))
    n = px_total // (nT * nT)
    c = make_coeff_blocks(rng, uniq, log2)
    reps = (n + uniq - 1) // uniq
    coeffs = np.tile(c, (reps, 1, 1))[:n].copy()
    qp = rng.integers(0, 52, size=n).astype(np.uint8)
    dst = rng.integers(0, 256, size=(n, nT, nT), dtype=np.uint8)
    return coeffs, qp, dst


def _barrier_max(world, value):
    if world == 1:
        return value
    import torch, torch.distributed as dist
    t = torch.tensor([value], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def run_ours(a, rank, world, local, ClockSampler, measured_peaks):
    from libde265_b200 import lib as L
    if world > 1:
        import torch, torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl")
    ctx = L.Context(local)
    wl = a.workload or "idct_sweep"
    if wl != "idct_sweep":
        raise SystemExit("unknown workload " + wl)
    px = 1 << 26  # pixels per transform kind per step (256 MiB algorithmic traffic per kind)
    kinds = []
    for name, kind, log2 in SWEEP:
        coeffs, qp, dst = _sweep_inputs(log2, px, seed=2 + log2 + 10 * kind)
        n = coeffs.shape[0]
        d_c, d_q, d_d = ctx.dev_alloc(coeffs.nbytes), ctx.dev_alloc(qp.nbytes), ctx.dev_alloc(dst.nbytes)
        ctx.h2d(d_c, coeffs); ctx.h2d(d_q, qp); ctx.h2d(d_d, dst)
        kinds.append(dict(name=name, kind=kind, log2=log2, n=n, d_c=d_c, d_q=d_q, d_d=d_d,
                          host=(coeffs, qp, dst)))
    bytes_per_step = sum(4 * (k["n"] << (2 * k["log2"])) for k in kinds)
    px_per_step = sum(k["n"] << (2 * k["log2"]) for k in kinds)

    def step():
        for k in kinds:
            ctx.residual_batch_dev(k["kind"], k["log2"], k["n"], k["d_c"], k["d_d"], k["d_q"])

    for _ in range(a.warmup):
        step()
    ctx.sync()
    # per-kernel timing (flush L2 between launches; inputs are also > L2)
    per_kind = {}
    for k in kinds:
        ts = []
        for _ in range(5):
            ctx.flush_l2()
            ctx.timer_start()
            ctx.residual_batch_dev(k["kind"], k["log2"], k["n"], k["d_c"], k["d_d"], k["d_q"])
            ts.append(ctx.timer_stop())
        ms = float(np.mean(ts))
        per_kind[k["name"]] = {"ms": ms, "gbs": 4 * (k["n"] << (2 * k["log2"])) / ms / 1e6}
    launches0 = ctx.launches
    clocks = ClockSampler(local)
    clocks.start()
    if world > 1:
        import torch, torch.distributed as dist
        dist.barrier(); torch.cuda.synchronize()
    ctx.sync()
    ctx.timer_start()
    for _ in range(a.steps):
        step()
    ms_total = ctx.timer_stop()
    ms_total = _barrier_max(world, ms_total)
    clk = clocks.stop()
    launches = ctx.launches - launches0
    ms_step = ms_total / a.steps

    # end to end through the C ABI with host buffers (H2D + kernel + D2H inside the call)
    e2e_px = 0
    t0 = time.perf_counter()
    e2e_steps = 1
    for k in kinds:
        c, q, d = k["host"]
        m = min(k["n"], (1 << 24) >> (2 * k["log2"]))
        ctx.residual_batch(k["kind"], k["log2"], c[:m], d[:m].copy(), q[:m])
        e2e_px += m << (2 * k["log2"])
    e2e_s = time.perf_counter() - t0
    e2e_s = _barrier_max(world, e2e_s)

    peak, peak_src = measured_peaks()
    dom = max(per_kind.items(), key=lambda kv: kv[1]["ms"])
    res = {
        "metric": "dequant+inverse-transform+add throughput (config 2 sweep)", "value": world * px_per_step / ms_step / 1e3,
        "unit": "Mpixel/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int16/int32", "data": "synthetic",
        "config": {"workload": "idct_sweep: DST4+DCT4/8/16/32, 2^26 px each, random QP, dense-input form",
                   "l2": "inputs > L2 (768 MiB per kind) and L2 flushed between per-kernel timings"},
        "e2e": {"value": world * e2e_px / e2e_s / 1e6, "unit": "Mpixel/s", "h2d_bytes_per_step": int(e2e_px * 3 + e2e_px // 16),
                "d2h_bytes_per_step": int(e2e_px)},
        "gpu_launches": int(launches), "clocks": clk,
        "roofline": {"bound": "hbm", "kernel": "k_residual_dense<%s>" % dom[0], "achieved": dom[1]["gbs"], "peak": peak,
                     "unit": "GB/s", "frac": dom[1]["gbs"] / peak, "traffic": None, "peak_source": peak_src,
                     "per_kernel_gbs": {k: round(v["gbs"], 1) for k, v in per_kind.items()}},
    }
    if rank == 0 and world == 1:
        res["cpu_baseline"] = cpu_baseline_sweep(kinds)
    return res


def cpu_baseline_sweep(kinds, budget_s=15.0):
    """C oracle ('port') timed single-threaded on a bounded sample of the same blocks."""
    from tests import helpers
    orc = helpers.load_oracle()
    px, t = 0, 0.0
    for k in kinds:
        c, q, d = k["host"]
        m = min(k["n"], (1 << 21) >> (2 * k["log2"]))
        dd = d[:m].copy()
        t0 = time.perf_counter()
        orc.orc_residual_batch(k["kind"], k["log2"], 0, m, helpers.ptr(c[:m]), helpers.ptr(q[:m]), helpers.ptr(dd))
        t += time.perf_counter() - t0
        px += m << (2 * k["log2"])
    return {"value": px / t / 1e6, "unit": "Mpixel/s", "cores": 1, "kind": "port",
            "sample": "2^21 px per transform kind of the same sweep, C oracle, 1 thread"}


def run_reference(a):
    """Reference arm: the reference's own scalar transform functions on the host cores."""
    from tests import helpers
    ref = helpers.load_ref()
    kinds = []
    for name, kind, log2 in SWEEP:
        coeffs, qp, dst = _sweep_inputs(log2, 1 << 21, seed=2 + log2 + 10 * kind)
        kinds.append(dict(name=name, kind=kind, log2=log2, n=coeffs.shape[0], host=(coeffs, qp, dst)))
    if ref is None:
        cb = cpu_baseline_sweep(kinds)
    else:
        from tests.test_oracle_transform import ref_residual_batch
        cb = cpu_baseline_sweep(kinds)
    v = cb["value"]
    return {"impl": "reference", "metric": "dequant+inverse-transform+add throughput (config 2 sweep)", "value": v,
            "unit": "Mpixel/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "higher_is_better": True,
            "config": {"workload": "idct_sweep"}, "cpu_baseline": cb,
            "e2e": {"value": v, "unit": "Mpixel/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}

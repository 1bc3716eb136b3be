"""Workloads behind bench.py.

stream_4k     (default; BASELINE.json configs[3]) 3840x2160 enc265 intra stream with deblocking switched on
              (PPS re-written), full GPU reconstruction.  value = frames/s of the reconstruction kernels with the
              work lists resident in HBM; e2e = frames/s through the libde265 C API (host bitstream in, host
              pictures out: CABAC parsing on the host cores, H2D of work lists, kernels, D2H of pictures).
stream_1080p  (configs[0]) the 1080p intra stream, same measurement.
idct_sweep    (configs[1]) synthetic TU coefficient sweep, dense-input form (4*nT^2 B per block).

The reference arm / cpu_baseline is the unmodified reference decoder `oracle/_ref/dec265` (SSE path) run as
one process per host core on the same stream (enc265 streams have no WPP/tiles, so the
reference parallelises only across streams).
"""
import ctypes as C
import json, os, subprocess, sys, tempfile, time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STREAMS = os.path.join(ROOT, "tests", "golden", "streams")
REF_DIR = os.path.join(ROOT, "oracle", "_ref")

SWEEP = [("DST4", 1, 2), ("DCT4", 0, 2), ("DCT8", 0, 3), ("DCT16", 0, 4), ("DCT32", 0, 5)]
STREAM_WL = {
    # name: (golden stream, deblocking patched on, loops of the stream per step, kernel batch, description)
    "stream_4k": ("c4_2160p_q32", True, 32, 128, "3840x2160 8-bit 4:2:0 enc265 intra stream (8 unique frames, qp 32), "
                  "deblocking enabled by PPS re-write (SAO cannot be enabled in an enc265 stream)"),
    "stream_1080p": ("c1_1080p_q32", False, 16, 256, "1920x1080 8-bit 4:2:0 enc265 intra stream (16 unique frames, qp 32)"),
    # P/B streams at the metric's sizes.  The reference's encoder cannot produce inter pictures, so these come from
    # the random-syntax generator (tools/streamgen, golden MD5s from the unmodified reference decoder): merge / AMVP /
    # temporal MVP, AMP, weighted prediction, SAO and deblocking on; kernel batch = B265_BENCH_STREAMS (8) independent instances of
    # the whole stream in decoding order
    "stream_inter_1080p": ("gen_pb_1080p_ld32", False, 8, 0, "1920x1080 P/B generator stream, low-delay shape (32 pictures, 3 references, "
                           "weighted prediction, temporal MVP, AMP, SAO, deblocking)"),
    "stream_inter_1080p_ra": ("gen_pb_1080p_ra33", False, 8, 0, "1920x1080 P/B generator stream, random-access shape (33 pictures, output "
                              "reordering, B slices, weighted bi-prediction, temporal MVP, AMP, SAO, deblocking)"),
    "stream_inter_4k": ("gen_pb_2160p_ld12", False, 16, 0, "3840x2160 P/B generator stream, low-delay shape (12 pictures, 2 references, "
                        "weighted prediction, temporal MVP, AMP, SAO, deblocking)"),
}
# config 5 (BASELINE.json configs[4]): eight independent 4K enc265 streams, seeds 10..17 (tools/make_streams.py), one
# decoder instance per GPU; rank r of N decodes streams r, r+N, ...
C5_STREAMS = ["c5_2160p_s%d" % s_ for s_ in range(10, 18)]


def c5_available():
    return all(os.path.exists(os.path.join(STREAMS, n + ".bin")) for n in C5_STREAMS)


def _dist_device():
    import torch.distributed as dist
    return "cuda" if dist.get_backend() == "nccl" else "cpu"


def _barrier_max(world, value):
    if world == 1:
        return value
    import torch, torch.distributed as dist
    t = torch.tensor([value], dtype=torch.float64, device=_dist_device())
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _all_sum(world, value):
    if world == 1:
        return value
    import torch, torch.distributed as dist
    t = torch.tensor([value], dtype=torch.float64, device=_dist_device())
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


# ---------------------------------------------------------------------------------------------
# replicas: the path does not shard inside a picture, so N GPUs = N decoder instances fed
# with independent streams or with the closed-GOP segments of one stream, dealt round-robin
# ---------------------------------------------------------------------------------------------
def split_idr_segments(data):
    """Split an Annex-B stream into closed-GOP segments: every segment starts with the parameter sets seen
    so far and one IDR/BLA picture (enc265 intra streams: every picture).  Returns a list of bytes."""
    import re
    starts = [m.start() for m in re.finditer(b"\x00\x00\x01", data)]
    nals = []
    for i, s in enumerate(starts):
        e = starts[i + 1] if i + 1 < len(starts) else len(data)
        while e > s + 3 and data[e - 1] == 0 and i + 1 < len(starts):
            e -= 1   # zero byte(s) of the next 4-byte start code
        nals.append(data[s:e])
    params, segs, cur = {}, [], None
    for n in nals:
        t = (n[3] >> 1) & 0x3f
        if t in (32, 33, 34):        # VPS / SPS / PPS
            params[(t, n[5:6] if t == 34 else b"")] = n
            if cur is not None:
                cur.append(n)
            continue
        first_slice = t < 32 and len(n) > 5 and (n[5] & 0x80)
        if 16 <= t <= 21 and first_slice:   # IRAP picture starts a new segment
            if cur is not None:
                segs.append(b"".join(cur))
            cur = [b"\x00" + p for p in params.values()]
        if cur is not None:
            cur.append(b"\x00" + n if t >= 32 or first_slice else n)
    if cur:
        segs.append(b"".join(cur))
    return segs


def deal(items, world, rank):
    """round-robin share of `items` for `rank`, with the original indices (outputs are re-ordered by index)"""
    return [(i, it) for i, it in enumerate(items) if i % world == rank]


def run_parse_replicas(a, rank, world, local):
    """Host-only replica run (no GPU): every rank decodes its share of the closed-GOP segments of a golden
    stream with a parse-only decoder; used by the world_size-2 gloo test of the N>1 host logic."""
    from libde265_b200.decoder import Decoder
    name = "tiny_416x240_q30"
    data = _stream_bytes(name, False)
    gold = json.load(open(os.path.join(STREAMS, name + ".md5.json")))
    segs = split_idr_segments(data)
    mine = deal(segs, world, rank)
    dec = Decoder(device=local, threads=1, batch=4, parse_only=True)
    t0 = time.perf_counter()
    got = []
    for _ in range(a.steps):
        got = []
        for idx, seg in mine:
            dec.lib.de265_reset(dec.h)
            pocs = []
            dec.decode_stream(seg, on_picture=lambda img: pocs.append(dec.lib.b265_de265_image_poc(img)))
            got.append((idx, len(pocs)))
    dt = time.perf_counter() - t0
    dec.close()
    dt = _barrier_max(world, dt)
    frames = _all_sum(world, float(sum(n for _, n in got)))
    gathered = [got]
    if world > 1:
        import torch.distributed as dist
        gathered = [None] * world
        dist.all_gather_object(gathered, got)
    order = sorted(x for g in gathered for x in g)
    assert [i for i, _ in order] == list(range(len(segs))), "segments lost or duplicated across ranks"
    assert int(frames) == gold["frames"], "replicas decoded %d pictures, stream has %d" % (frames, gold["frames"])
    return {"metric": "parsed frames/s (host only, replicas)", "value": frames * a.steps / dt, "unit": "frames/s",
            "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": dt / a.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "syntax only", "data": "synthetic",
            "config": {"workload": "parse_replicas: closed-GOP segments of %s dealt round-robin over %d ranks" % (name, world),
                       "segments": len(segs)}}


def _stream_bytes(name, patched):
    data = open(os.path.join(STREAMS, name + ".bin"), "rb").read()
    if patched:
        from tools.pps_patch import patch_stream
        data = patch_stream(data)
    return data


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


# =====================================================================================================
# reference decoder on the host cores
# =====================================================================================================
def run_dec265(data, procs, loops, scalar=False):
    """`procs` concurrent dec265 processes, each decoding `data` repeated `loops` times (`data` may be a list of streams:
    process i takes stream i mod len).  Returns (frames_total, seconds) or None when oracle/_ref/dec265 is not available."""
    exe = os.path.join(REF_DIR, "dec265")
    if not os.path.exists(exe):
        return None
    datas = data if isinstance(data, (list, tuple)) else [data]
    with tempfile.TemporaryDirectory() as td:
        files = []
        for k, d in enumerate(datas):
            f = os.path.join(td, "in%d.bin" % k)
            with open(f, "wb") as fh:
                for _ in range(loops):
                    fh.write(d)
            files.append(f)
        base = [exe, "-q"] + (["-0"] if scalar else [])
        # warm the page cache / loader
        subprocess.run(base + [files[0]], stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
        t0 = time.perf_counter()
        ps = [subprocess.Popen(base + [files[i % len(files)]], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
              for i in range(procs)]
        frames = 0
        for p in ps:
            out = p.communicate()[0]
            for line in out.splitlines():
                if line.startswith("nFrames decoded:"):
                    frames += int(line.split(":")[1].split()[0])
        dt = time.perf_counter() - t0
    return frames, dt


def cpu_baseline_stream(wl, budget_s=20.0):
    name, patched, loops, _, desc = STREAM_WL[wl]
    data = _stream_bytes(name, patched)
    gold = json.load(open(os.path.join(STREAMS, name + ".md5.json")))
    procs = host_cores()   # every core the box has: both arms see the same machine
    one = run_dec265(data, 1, 1)
    if one is None:
        return cpu_port_stream(wl)
    fps1 = one[0] / one[1]
    # size the sample: every process decodes the stream `reps` times, about budget_s/2 seconds
    reps = max(1, int(budget_s / 2 * fps1 / gold["frames"]))
    frames, dt = run_dec265(data, procs, reps)
    return {"value": frames / dt, "unit": "frames/s", "cores": procs, "kind": "reference",
            "sample": "%d concurrent `dec265 -q` (SSE4.1 build of the unmodified reference) x %d frames each of the same "
                      "stream; 1 process alone: %.1f frames/s" % (procs, reps * gold["frames"], fps1),
            "single_process_fps": fps1}


def cpu_port_stream(wl):
    """Fallback when the compiled reference did not travel: host parser + C oracle reconstruction, 1 thread."""
    from benchmarks.jobs import stream_jobs
    from oracle.frame import oracle_planes
    name, patched, _, _, _ = STREAM_WL[wl]
    jobs = stream_jobs(_stream_bytes(name, patched), threads=1, max_frames=2)
    t0 = time.perf_counter()
    for oj in jobs:
        oracle_planes(oj.job)
    dt = time.perf_counter() - t0
    return {"value": len(jobs) / dt, "unit": "frames/s", "cores": 1, "kind": "port",
            "sample": "%d frames reconstructed by the C oracle from the parser's work lists (parse time excluded)" % len(jobs)}


# =====================================================================================================
# our arm: stream workloads
# =====================================================================================================
_TU_DT = np.dtype([('x', '<u2'), ('y', '<u2'), ('cidx_log2', 'u1'), ('flags', 'u1'), ('qp', 'u1'), ('intra_mode', 'u1'),
                   ('coeff_off', '<u4'), ('ncoeff', '<u2'), ('reserved', '<u2')])


def _job_algorithmic_bytes(J):
    """SURVEY §8d per picture (DESIGN.md "algorithmic bytes"):
    residual pre-pass: 4 B per non-zero coefficient read + 2 B per sample of a coded TB written (int16 plane);
    intra wavefront: 1.5*W*H samples written + 2 B per sample of a coded TB read;
    deblocking run as two passes reads+writes the picture twice (4P)."""
    P = J.width * J.height * 3 // 2
    from libde265_b200 import lib as L
    tus = np.ctypeslib.as_array(C.cast(J.tus, C.POINTER(C.c_uint8)), shape=(J.n_tus * 16,)).view(_TU_DT)
    coded = tus[tus['ncoeff'] > 0]
    coded_samples = int((1 << (2 * (coded['cidx_log2'].astype(np.int64) >> 2))).sum())
    intra_tus = tus[tus['intra_mode'] != 0xff]
    intra_samples = int((1 << (2 * (intra_tus['cidx_log2'].astype(np.int64) >> 2))).sum())
    coded_intra = intra_tus[intra_tus['ncoeff'] > 0]
    coded_intra_samples = int((1 << (2 * (coded_intra['cidx_log2'].astype(np.int64) >> 2))).sum())
    res = 4 * J.n_coeffs + 2 * coded_samples
    intra = intra_samples + 2 * coded_intra_samples      # all-intra picture: intra_samples == P
    dbk = 4 * P if (J.flags & L.FRAME_DEBLOCK) else 0
    mc = 0
    if J.n_pus:   # uni-predicted PU 1.5 wh read + 1.5 wh written, bi-predicted 3 wh read + 1.5 wh written
        pus = np.ctypeslib.as_array(C.cast(J.pus, C.POINTER(C.c_uint8)), shape=(J.n_pus * 24,)).reshape(-1, 24)
        area = pus[:, 4].astype(np.int64) * pus[:, 5].astype(np.int64)
        mc = int((area * np.where((pus[:, 6] & 3) == 3, 4.5, 3.0)).sum())
    sao = 2 * P if (J.flags & L.FRAME_SAO) else 0
    return res, intra, dbk, mc, sao


def run_stream(a, wl, rank, world, local, ClockSampler, measured_peaks):
    from libde265_b200 import lib as L
    from libde265_b200.decoder import Decoder
    from benchmarks.jobs import stream_jobs
    name, patched, loops, kbatch, desc = STREAM_WL[wl]
    data = _stream_bytes(name, patched)
    gold = json.load(open(os.path.join(STREAMS, name + ".md5.json")))
    W, H, nuniq = gold["width"], gold["height"], gold["frames"]
    if wl == "stream_4k" and world > 1 and c5_available():
        # config 5: every rank decodes its own stream(s), in the kernel leg and end to end
        mine = [n for i, n in enumerate(C5_STREAMS) if i % world == rank]
        data = b"".join(_stream_bytes(n, patched) for n in mine)
        nuniq = sum(json.load(open(os.path.join(STREAMS, n + ".md5.json")))["frames"] for n in mine)
        loops = max(1, loops * 8 // nuniq)
        desc += "; config 5: rank r of %d decodes the independent streams r, r+%d, ... (seeds 10..17)" % (world, world)
    cores = host_cores()
    threads = int(os.environ.get("B265_BENCH_THREADS", 0)) or max(1, cores // world - 1)   # override: developer sweeps

    # ---------------- kernels with work lists resident in HBM ----------------
    # One context holds a batch of `kbatch` pictures (independent pictures of an intra stream, or `kcopies` independent
    # instances of an inter stream whose dependency levels share launch groups); a step runs it.  B265_BENCH_CONTEXTS > 1
    # (developer runs) puts one such batch on each of several contexts, the shape the C API's reconstruction threads have;
    # the contexts then share the SMs (b265_busy_peers).
    nctx = max(1, int(os.environ.get("B265_BENCH_CONTEXTS", 1)))
    ctxs = [L.Context(local) for _ in range(nctx)]
    ctx = ctxs[0]
    uniq = stream_jobs(data, threads=min(threads, 8))       # decoding order
    inter = any(oj.job.n_refs for oj in uniq)
    kcopies = 1
    if not kbatch:   # inter: the whole stream in decoding order, B265_BENCH_STREAMS independent instances of it in one batch
        kcopies = max(1, int(os.environ.get("B265_BENCH_STREAMS", 8)))
        kbatch = len(uniq) * kcopies

    def resident_batch(cx):
        jobs, sids = [], []
        for copy in range(kcopies):   # inter streams: `kcopies` independent instances of the stream share the launch groups
            uid2sid = {}
            for i in range(kbatch // kcopies):
                oj = uniq[i % len(uniq)]
                J = L.FrameJob()
                C.memmove(C.byref(J), C.byref(oj.job), C.sizeof(L.FrameJob))
                sid = cx.surface_alloc(W, H)
                if inter:   # a parse-only decoder names pictures by unique id: map them to this instance's surfaces
                    uid2sid[oj.job.dst_surface] = sid
                    for r in range(J.n_refs):
                        J.ref_surfaces[r] = uid2sid[oj.job.ref_surfaces[r]]
                J.dst_surface = sid
                sids.append(sid)
                jobs.append(J)
        b = cx.upload_frames(jobs)
        cx.sync()
        return jobs, sids, b
    resident = [resident_batch(cx) for cx in ctxs]
    jobs, sids, batch = resident[0]
    ab = [_job_algorithmic_bytes(J) for J in jobs]
    bytes_res, bytes_intra, bytes_dbk, bytes_mc, bytes_sao = (sum(x[i] for x in ab) for i in range(5))
    for _ in range(a.warmup):
        ctx.run_batch(batch)
    ctx.sync()
    # per-stage device time of one batch running alone (CUDA events between the stages, on the context's stream)
    ctx.lib.b265_set_stage_timing(ctx.h, 1)
    st = []
    for _ in range(3):
        ctx.flush_l2()
        ctx.run_batch(batch)
        st.append(ctx.stage_times())
    ctx.lib.b265_set_stage_timing(ctx.h, 0)
    stage = [float(np.mean([s[i] for s in st])) for i in range(6)]

    def step():
        for cx, (_, _, b) in zip(ctxs, resident):
            cx.run_batch(b)
    for _ in range(a.warmup):
        step()
    for cx in ctxs:
        cx.sync()
    launches0 = sum(cx.launches for cx in ctxs)
    clocks = ClockSampler(local)
    clocks.start()
    if world > 1:
        import torch, torch.distributed as dist
        dist.barrier(); torch.cuda.synchronize()
    # device time of `steps` steps over all contexts: the start event on context 0's stream, every other stream waits for
    # it; at the end context 0's stream waits for the others before the stop event
    ctx.timer_start()
    for cx in ctxs[1:]:
        cx.wait_for(ctx)
    for _ in range(a.steps):
        step()
    for cx in ctxs[1:]:
        ctx.wait_for(cx)
    ms_total = ctx.timer_stop()
    ms_total = _barrier_max(world, ms_total)
    clk = clocks.stop()
    launches = sum(cx.launches for cx in ctxs) - launches0
    ms_step = ms_total / a.steps
    value = world * nctx * kbatch / ms_step * 1e3
    for cx, (_, sd, b) in zip(ctxs[1:], resident[1:]):
        cx.free_batch(b)
        for s_ in sd:
            cx.surface_release(s_)
        cx.close()
    ctx.free_batch(batch)
    # ---- SAO leg: an enc265 stream cannot carry SAO syntax, so the deblocked pictures of
    # this stream are filtered with synthetic per-CTB parameters (seed 5: 20 % off / 40 % band / 40 % edge,
    # class U{0..3}, band position U{0..31}, |offset| <= 7 with the edge-offset sign convention)
    sao_ms, sao_gbs = None, None
    if patched:
        keep = []
        sjobs = []
        rng = np.random.default_rng(5)
        for J0 in jobs:
            J = L.FrameJob()
            C.memmove(C.byref(J), C.byref(J0), C.sizeof(L.FrameJob))
            nctb = (-(-W >> J.log2_ctb)) * (-(-H >> J.log2_ctb))
            sao = (L.SAO * nctb)()
            raw = np.zeros((nctb, 18), np.int8)
            typ = rng.choice([0, 1, 2], size=(nctb, 2), p=[0.2, 0.4, 0.4])
            raw[:, 0] = typ[:, 0]; raw[:, 1] = typ[:, 1]; raw[:, 2] = typ[:, 1]
            for c in range(3):
                t = raw[:, c]
                band = rng.integers(0, 32, nctb); eo = rng.integers(0, 4, nctb)
                if c == 2:
                    eo = raw[:, 4].copy()   # Cr shares Cb's edge class
                raw[:, 3 + c] = np.where(t == 1, band, np.where(t == 2, eo, 0))
                off = rng.integers(-7, 8, (nctb, 4))
                edge = np.abs(off) * np.array([1, 1, -1, -1])
                raw[:, 6 + 4 * c:10 + 4 * c] = np.where((t == 2)[:, None], edge, off) * (t != 0)[:, None]
            C.memmove(sao, raw.ctypes.data, nctb * 18)
            sl = (L.Slice * J.n_slices)()
            C.memmove(sl, J.slices, C.sizeof(L.Slice) * J.n_slices)
            for i in range(J.n_slices):
                sl[i].sao_luma = 1; sl[i].sao_chroma = 1
            J.sao = C.cast(sao, C.POINTER(L.SAO))
            J.slices = C.cast(sl, C.POINTER(L.Slice))
            J.flags |= L.FRAME_SAO
            keep += [sao, sl]
            sjobs.append(J)
        sbatch = ctx.upload_frames(sjobs)
        for _ in range(max(1, a.warmup)):
            ctx.run_batch(sbatch)
        ctx.sync()
        ctx.lib.b265_set_stage_timing(ctx.h, 1)
        ss = []
        for _ in range(3):
            ctx.flush_l2()
            ctx.run_batch(sbatch)
            ss.append(ctx.stage_times()[4])
        ctx.lib.b265_set_stage_timing(ctx.h, 0)
        sao_ms = float(np.mean(ss))
        sao_gbs = kbatch * 2 * (W * H * 3 // 2) / sao_ms / 1e6
        ctx.free_batch(sbatch)
    for s in sids:
        ctx.surface_release(s)
    ctx.close()

    # ---------------- end to end through the libde265 C API (host bitstream -> host pictures) ----------------
    step_data = data * loops
    frames_step = nuniq * loops
    dbatch = int(os.environ.get("B265_BENCH_BATCH", 16))   # override: developer sweeps
    if "B265_RECON_CONTEXTS" not in os.environ:   # the ranks of a node share its cores: cores/N per rank for parse + reconstruction threads
        per_rank = cores // world
        os.environ["B265_RECON_CONTEXTS"] = "1" if per_rank <= 4 else ("2" if per_rank <= 8 else "3")
    dec = Decoder(device=local, threads=threads, batch=dbatch)
    count = [0]
    cb = lambda img: count.__setitem__(0, count[0] + 1)
    dec.decode_stream(step_data, chunk=1 << 20, on_picture=cb)      # warm-up: allocations, clocks
    s0 = dec.stats()
    count[0] = 0
    if world > 1:
        import torch, torch.distributed as dist
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        dec.lib.de265_reset(dec.h)
        dec.decode_stream(step_data, chunk=1 << 20, on_picture=cb)
    e2e_s = time.perf_counter() - t0
    s1 = dec.stats()
    dec.close()
    assert count[0] == frames_step * a.steps, "decoder returned %d pictures, expected %d" % (count[0], frames_step * a.steps)
    e2e_s = _barrier_max(world, e2e_s)
    e2e_fps = world * frames_step * a.steps / e2e_s

    peak, peak_src = measured_peaks()
    gbs = lambda nbytes, ms: (nbytes / ms / 1e6) if (ms and ms > 0 and nbytes) else None
    mc_gbs, res_gbs, intra_gbs, dbk_gbs = gbs(bytes_mc, stage[0]), gbs(bytes_res, stage[1]), gbs(bytes_intra, stage[2]), gbs(bytes_dbk, stage[3])
    own_sao_gbs = gbs(bytes_sao, stage[4])
    # the roofline object describes the stage with the largest share of the step
    stages = [("k_mc (+k_pcm)", stage[0], mc_gbs,
               "algorithmic bytes = 3*w*h per uni-predicted PU (1.5 read + 1.5 written), 4.5*w*h per bi-predicted PU"),
              ("k_residual_sparse* + k_intra_prep", stage[1], res_gbs, "4 B per non-zero level read + 2 B per sample of a coded TB written"),
              ("k_intra_wavefront", stage[2], intra_gbs,
               "dependency-latency bound (CTB-row wavefront: one warp walks a CTB row, rows trail each other by two CTBs), not "
               "bandwidth bound; algorithmic bytes = samples of the intra TBs written + 2 B per sample of a coded intra TB read"),
              ("k_deblock_* (V+H, bS derived in-kernel)", stage[3], dbk_gbs, "4P per picture: two passes read and write the picture"),
              ("k_sao", stage[4], own_sao_gbs, "2P per picture")]
    dom = max(stages, key=lambda t: t[1])
    traffic = None
    try:   # dram__bytes_read+write of one launch of the dominant kernel on this workload, from a committed ncu --set full capture
        tf = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        ent = tf.get(wl, {}).get(dom[0].split(" ")[0])
        if ent and ent.get("frames_per_launch") == kbatch:
            traffic = ent["dram_bytes"]
    except Exception:
        pass
    parse_core_s = (s1["parse_s"] - s0["parse_s"]) / a.steps
    res = {
        "metric": "decoded frames/s", "value": value, "unit": "frames/s", "n_gpus": world, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8 pixels / int16 coefficients / int32 accumulators", "data": "synthetic",
        "config": {"workload": "%s: %s" % (wl, desc), "frames_per_step": nctx * kbatch, "contexts": nctx, "frames_per_batch": kbatch,
                   "mpixel_per_s": value * W * H / 1e6,
                   "value_is": "all reconstruction kernels (MC, residual, intra wavefront, deblock V+H, SAO where the stream has it) on a batch "
                               "of %d pictures%s whose work lists are resident in HBM%s; the CPU arm is the whole reference decoder "
                               "(parsing included) and compares with `e2e`, not with `value`"
                               % (kbatch, " (%d independent instances of the stream: dependent pictures run in launch groups by dependency "
                                          "level)" % kcopies if kcopies > 1 else "",
                                  "; %d contexts (one CUDA stream each) hold one such batch each, device time from a start event every "
                                  "stream waits for to a stop event that waits for every stream" % nctx if nctx > 1 else ""),
                   "l2": "working set per step %d MB of pictures + work lists > 126 MB L2; L2 flushed before the per-stage timings"
                         % (nctx * kbatch * W * H * 3 // 2 >> 20),
                   "e2e_is": "libde265 C API (de265_push_data/de265_decode/de265_get_next_picture): %d host parse threads, "
                             "%s CUDA contexts, batch %d; %d frames per step" % (threads, os.environ.get("B265_RECON_CONTEXTS", "3"), dbatch, frames_step),
                   "host_cores": cores, "host_cores_per_rank": cores // world,
                   "stage_ms[pcm+mc,residual+avail,intra_wavefront,deblock,sao,total]": [round(x, 3) for x in stage],
                   "stage_ms_is": "one batch running alone on one context (stages of concurrent batches overlap, so they do not add up to ms_per_step)",
                   "sao_leg_ms": round(sao_ms, 3) if sao_ms else None},
        "e2e": {"value": e2e_fps, "unit": "frames/s",
                "h2d_bytes_per_step": int((s1["h2d_bytes"] - s0["h2d_bytes"]) / a.steps),
                "d2h_bytes_per_step": int((s1["d2h_bytes"] - s0["d2h_bytes"]) / a.steps),
                "mpixel_per_s": e2e_fps * W * H / 1e6,
                "per_gpu": e2e_fps / world,
                "host_parse_core_seconds_per_step": parse_core_s,
                "recon_thread_seconds_per_step": (s1["gpu_s"] - s0["gpu_s"]) / a.steps,
                "submit_seconds_per_step": (s1["submit_s"] - s0["submit_s"]) / a.steps,
                "wall_seconds_per_step": e2e_s / a.steps,
                # what the host cores alone allow: every rank parses its step on `threads` cores
                "host_parse_ceiling_fps": world * frames_step / max(1e-9, parse_core_s / max(1, threads)),
                "note": "end to end is bound by CABAC parsing on the host cores (host_parse_ceiling_fps): N GPUs share one host, so e2e does "
                        "not scale with N on a single node; `value` (kernels) does"},
        "gpu_launches": int(launches), "clocks": clk,
        "roofline": {"bound": "hbm", "kernel": dom[0], "achieved": dom[2] or 0.0, "peak": peak, "unit": "GB/s",
                     "frac": (dom[2] or 0.0) / peak, "traffic": traffic, "peak_source": peak_src,
                     "share_of_step": dom[1] / stage[5] if stage[5] else None,
                     "note": "largest share of a batch; " + dom[3] + "; time = CUDA events around this stage only, the batch running alone",
                     "per_kernel_gbs": {"k_mc": round(mc_gbs, 1) if mc_gbs else None,
                                        "k_residual_sparse x4 + k_intra_prep (4 B/coeff read, 2 B/coded sample written)": round(res_gbs, 1) if res_gbs else None,
                                        "k_intra_wavefront": round(intra_gbs, 1) if intra_gbs else None,
                                        "deblock(V+H luma/chroma, bS derived in-kernel, 4P per picture)": round(dbk_gbs, 1) if dbk_gbs else None,
                                        "sao(stream's own parameters, 2P per picture)": round(own_sao_gbs, 1) if own_sao_gbs else None,
                                        "sao(synthetic per-CTB parameters on the same pictures, 2P per picture)": round(sao_gbs, 1) if sao_gbs else None}},
    }
    if rank == 0 and world == 1:
        res["cpu_baseline"] = cpu_baseline_stream(wl)
    return res


def run_reference_stream(a, wl):
    name, patched, loops, _, desc = STREAM_WL[wl]
    data = _stream_bytes(name, patched)
    gold = json.load(open(os.path.join(STREAMS, name + ".md5.json")))
    if wl == "stream_4k" and a.gpus > 1 and c5_available():
        # config 5: the eight independent streams our ranks decode, dealt over the dec265 processes
        data = [_stream_bytes(n, patched) for n in C5_STREAMS]
        gold = json.load(open(os.path.join(STREAMS, C5_STREAMS[0] + ".md5.json")))
        desc += "; config 5: the eight independent streams (seeds 10..17) dealt over the processes"
    procs = host_cores()   # every core the box has: both arms see the same machine
    one = run_dec265(data[0] if isinstance(data, list) else data, 1, 1)
    if one is None:
        cb = cpu_port_stream(wl)
        v = cb["value"]
        ms = None
    else:
        fps1 = one[0] / one[1]
        reps = max(1, int(6.0 * fps1 / gold["frames"]))       # ~6 s of CPU work per step
        for _ in range(a.warmup):
            run_dec265(data, procs, 1)
        frames, secs = 0, 0.0
        for _ in range(a.steps):
            f, dt = run_dec265(data, procs, reps)
            frames += f
            secs += dt
        v = frames / secs
        ms = secs / a.steps * 1e3
        cb = {"value": v, "unit": "frames/s", "cores": procs, "kind": "reference",
              "sample": "%d concurrent `dec265 -q` processes (unmodified reference, SSE4.1 build) x %d frames each per step; "
                        "1 process alone: %.1f frames/s" % (procs, reps * gold["frames"], fps1)}
    return {"impl": "reference", "metric": "decoded frames/s", "value": v, "unit": "frames/s", "n_gpus": a.gpus,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8 pixels / int16 coefficients / int32 accumulators", "data": "synthetic",
            "config": {"workload": "%s: %s" % (wl, desc), "host_cores": host_cores()},
            "cpu_baseline": cb,
            "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


# =====================================================================================================
# idct_sweep (configs[1])
# =====================================================================================================
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


def run_sweep(a, rank, world, local, ClockSampler, measured_peaks):
    from libde265_b200 import lib as L
    ctx = L.Context(local)
    px = 1 << 26  # pixels per transform kind per step (256 MiB algorithmic traffic per kind)
    kinds = []
    for name, kind, log2 in SWEEP:
        coeffs, qp, dst = _sweep_inputs(log2, px, seed=2 + log2 + 10 * kind)
        n = coeffs.shape[0]
        d_c, d_q, d_d = ctx.dev_alloc(coeffs.nbytes), ctx.dev_alloc(qp.nbytes), ctx.dev_alloc(dst.nbytes)
        ctx.h2d(d_c, coeffs); ctx.h2d(d_q, qp); ctx.h2d(d_d, dst)
        kinds.append(dict(name=name, kind=kind, log2=log2, n=n, d_c=d_c, d_q=d_q, d_d=d_d, host=(coeffs, qp, dst)))
    px_per_step = sum(k["n"] << (2 * k["log2"]) for k in kinds)

    def step():
        for k in kinds:
            ctx.residual_batch_dev(k["kind"], k["log2"], k["n"], k["d_c"], k["d_d"], k["d_q"])

    for _ in range(a.warmup):
        step()
    ctx.sync()
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
    ms_total = _barrier_max(world, ctx.timer_stop())
    clk = clocks.stop()
    launches = ctx.launches - launches0
    ms_step = ms_total / a.steps
    # end to end through the C ABI with HOST buffers: levels, QPs and prediction in pinned host memory (b265_host_alloc),
    # H2D + kernel + D2H of the reconstruction inside the timed region; 2^24 px per transform kind
    def pinned_copy(arr):
        p = C.c_void_p()
        if ctx.lib.b265_host_alloc(ctx.h, arr.nbytes, C.byref(p)) != 0:
            return arr.copy(), None
        view = np.frombuffer((C.c_uint8 * arr.nbytes).from_address(p.value), dtype=arr.dtype).reshape(arr.shape)
        view[...] = arr
        return view, p
    e2e_in, pins = [], []
    for k in kinds:
        c, q, d = k["host"]
        m = min(k["n"], (1 << 24) >> (2 * k["log2"]))
        trio = []
        for arr in (c[:m], q[:m], d[:m]):
            v, p = pinned_copy(arr)
            trio.append(v)
            if p is not None:
                pins.append(p)
        e2e_in.append((k, m, trio))
    e2e_px = 0
    for k, m, (pc, pq, pd) in e2e_in[:1]:   # untimed warm-up of the staging buffers
        ctx.residual_batch(k["kind"], k["log2"], pc, pd.copy(), pq)
    t0 = time.perf_counter()
    for k, m, (pc, pq, pd) in e2e_in:
        ctx.residual_batch(k["kind"], k["log2"], pc, pd, pq)
        e2e_px += m << (2 * k["log2"])
    e2e_s = _barrier_max(world, time.perf_counter() - t0)
    for p in pins:
        ctx.lib.b265_host_free(ctx.h, p)
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


def cpu_baseline_sweep(kinds):
    """the reference's own scalar transform functions when oracle/_ref travelled, else the C oracle; 1 thread"""
    from tests import helpers
    ref = helpers.load_ref()
    px, t = 0, 0.0
    if ref is not None:
        from tests.test_oracle_transform import ref_residual_batch
        for k in kinds:
            c, q, d = k["host"]
            m = min(k["n"], (1 << 17) >> (2 * k["log2"]))
            t0 = time.perf_counter()
            ref_residual_batch(ref, k["kind"], k["log2"], c[:m], d[:m], q[:m])
            t += time.perf_counter() - t0
            px += m << (2 * k["log2"])
        return {"value": px / t / 1e6, "unit": "Mpixel/s", "cores": 1, "kind": "reference",
                "sample": "2^17 px per transform kind: dequant_coefficients + transform_*_add_8_fallback of the compiled "
                          "reference called block by block through ctypes (call overhead included), 1 thread"}
    orc = helpers.load_oracle()
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


def run_reference_sweep(a):
    kinds = []
    for name, kind, log2 in SWEEP:
        coeffs, qp, dst = _sweep_inputs(log2, 1 << 21, seed=2 + log2 + 10 * kind)
        kinds.append(dict(name=name, kind=kind, log2=log2, n=coeffs.shape[0], host=(coeffs, qp, dst)))
    cb = cpu_baseline_sweep(kinds)
    v = cb["value"]
    return {"impl": "reference", "metric": "dequant+inverse-transform+add throughput (config 2 sweep)", "value": v,
            "unit": "Mpixel/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "higher_is_better": True,
            "config": {"workload": "idct_sweep"}, "cpu_baseline": cb,
            "e2e": {"value": v, "unit": "Mpixel/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


# =====================================================================================================
# dispatch
# =====================================================================================================
# our arm: motion compensation on synthetic PU lists
# =====================================================================================================
def run_mc(a, rank, world, local, ClockSampler, measured_peaks):
    """1920x1080, three reference pictures (noise / smooth / noise), 8 pictures whose PU lists each tile the
    frame once with every legal PU size, MV ~ U[-256,255] quarter samples with 5 % far off-picture, 40 % uni,
    30 % bi, 15 % + 15 % weighted (tests/helpers.make_pu_list, the list the parity test uses), through the
    frame-level path (b265_upload_frames / b265_run_batch) with the lists resident in HBM."""
    from libde265_b200 import lib as L
    from tests.helpers import make_pu_list, make_planes
    W, H, layers = 1920, 1080, 8
    rng = np.random.default_rng(3)
    ctx = L.Context(local)
    refs = [make_planes(rng, W, H, k) for k in ("noise", "smooth", "noise")]
    ref_sids = []
    for r in refs:
        sid = ctx.surface_alloc(W, H)
        ctx.surface_upload(sid, r)
        ref_sids.append(sid)
    jobs, keep, sids = [], [], []
    alg = 0
    n_pus = 0
    for _ in range(layers):
        pus, slices = make_pu_list(rng, W, H, len(refs))
        J = L.FrameJob()
        J.width, J.height, J.chroma_format_idc, J.bit_depth = W, H, 1, 8
        J.log2_ctb, J.log2_min_cb, J.log2_min_tb = 6, 3, 2
        sid = ctx.surface_alloc(W, H)
        J.dst_surface = sid
        sids.append(sid)
        for i, r in enumerate(ref_sids):
            J.ref_surfaces[i] = r
            J.ref_poc[i] = i
        J.n_refs = len(ref_sids)
        J.pus = C.cast(pus, C.POINTER(L.PU)); J.n_pus = len(pus)
        J.slices = C.cast(slices, C.POINTER(L.Slice)); J.n_slices = len(slices)
        keep += [pus, slices]
        jobs.append(J)
        arr = np.ctypeslib.as_array(C.cast(pus, C.POINTER(C.c_uint8)), shape=(len(pus) * 24,)).reshape(-1, 24)
        area = arr[:, 4].astype(np.int64) * arr[:, 5].astype(np.int64)
        bi = (arr[:, 6] & 3) == 3
        alg += int((area * np.where(bi, 4.5, 3.0)).sum())
        n_pus += len(pus)
    batch = ctx.upload_frames(jobs)
    for _ in range(a.warmup):
        ctx.run_batch(batch)
    ctx.sync()
    clocks = ClockSampler(local)
    clocks.start()
    launches0 = ctx.launches
    ms = []
    for _ in range(a.steps):
        ctx.flush_l2()
        ctx.timer_start()
        ctx.run_batch(batch)
        ms.append(ctx.timer_stop())
    clk = clocks.stop()
    launches = ctx.launches - launches0
    ms_step = _barrier_max(world, float(np.mean(ms)))
    ctx.free_batch(batch)
    # end to end through b265_mc_batch with HOST buffers in pinned memory (b265_host_alloc): reference planes, PU and slice
    # lists up, prediction planes down, all inside the timed region
    pins = []

    def pinned(arr):
        p = C.c_void_p()
        if ctx.lib.b265_host_alloc(ctx.h, arr.nbytes, C.byref(p)) != 0:
            return arr.copy()
        pins.append(p)
        view = np.frombuffer((C.c_uint8 * arr.nbytes).from_address(p.value), dtype=arr.dtype).reshape(arr.shape)
        view[...] = arr
        return view
    prefs = [[pinned(pl) for pl in r] for r in refs]
    dst = [pinned(np.zeros((H, W), np.uint8)), pinned(np.zeros((H // 2, W // 2), np.uint8)), pinned(np.zeros((H // 2, W // 2), np.uint8))]
    ctx.mc_batch(keep[0], keep[1], prefs, dst, W, H)   # untimed: staging buffers
    t0 = time.perf_counter()
    for i in range(layers):
        ctx.mc_batch(keep[2 * i], keep[2 * i + 1], prefs, dst, W, H)
    e2e_s = _barrier_max(world, time.perf_counter() - t0)
    e2e = {"value": world * layers * W * H / e2e_s / 1e6, "unit": "Mpixel/s", "host_memory": "pinned" if pins else "pageable",
           "h2d_bytes_per_step": int(layers * len(refs) * W * H * 3 // 2 + 24 * n_pus), "d2h_bytes_per_step": int(layers * W * H * 3 // 2)}
    for p_ in pins:
        ctx.lib.b265_host_free(ctx.h, p_)
    cb = cpu_baseline_mc(keep, refs, W, H) if rank == 0 and world == 1 else None
    for s_ in sids + ref_sids:
        ctx.surface_release(s_)
    ctx.close()
    peak, peak_src = measured_peaks()
    gbs = alg / ms_step / 1e6
    px = layers * W * H
    res = {"metric": "motion-compensated Mpixel/s", "value": world * px / ms_step / 1e3, "unit": "Mpixel/s", "n_gpus": world,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8 pixels / int16 intermediates / int32 accumulators", "data": "synthetic",
            "config": {"workload": "mc_pus: SURVEY 8d config 3, 1920x1080, 3 references, %d PUs in %d pictures" % (n_pus, layers),
                       "l2": "L2 flushed before every timed step"},
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clk,
            "roofline": {"bound": "hbm", "kernel": "k_mc", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                         "traffic": None, "peak_source": peak_src,
                         "note": "algorithmic bytes = 3*w*h per uni-predicted PU (1.5 read + 1.5 written), 4.5*w*h per bi-predicted PU"}}
    if cb is not None:
        res["cpu_baseline"] = cb
    return res


def cpu_baseline_mc(keep, refs, W, H, budget_s=6.0):
    """The C oracle's orc_mc_batch (oracle/mc.c: the arithmetic of fallback-motion.cc:33-625 and motion.cc:48-282, pinned
    against the compiled reference's put_qpel / put_epel / writer slots in tests/test_oracle_mc.py) on the same PU lists,
    one picture per pass, one thread.  The reference has no entry point that predicts a PU list without a decoder context,
    hence kind "port"."""
    from tests import helpers
    orc = helpers.load_oracle()
    dst = [np.zeros((H, W), np.uint8), np.zeros((H // 2, W // 2), np.uint8), np.zeros((H // 2, W // 2), np.uint8)]
    rp = (C.c_void_p * (3 * len(refs)))(*[pl.ctypes.data for r in refs for pl in r])
    dp = (C.c_void_p * 3)(*[d.ctypes.data for d in dst])
    n, t0 = 0, time.perf_counter()
    while True:
        pus, slices = keep[2 * (n % (len(keep) // 2))], keep[2 * (n % (len(keep) // 2)) + 1]
        orc.orc_mc_batch(C.byref(pus), len(pus), C.byref(slices), rp, len(refs), dp, W, H, W, W // 2)
        n += 1
        if time.perf_counter() - t0 > budget_s or n >= 64:
            break
    t = time.perf_counter() - t0
    return {"value": n * W * H / t / 1e6, "unit": "Mpixel/s", "cores": 1, "kind": "port",
            "sample": "%d pictures (1920x1080 PU lists of the same workload) through the C oracle's orc_mc_batch, 1 thread" % n}


DEFAULT_WORKLOAD = "stream_4k"


def run_ours(a, rank, world, local, ClockSampler, measured_peaks):
    wl = a.workload or DEFAULT_WORKLOAD
    if world > 1:
        import torch, torch.distributed as dist
        backend = os.environ.get("B265_DIST_BACKEND", "gloo" if wl == "parse_replicas" else "nccl")
        if backend == "nccl":
            torch.cuda.set_device(local)
        dist.init_process_group(backend)
    if wl == "parse_replicas":
        res = run_parse_replicas(a, rank, world, local)
    elif wl in STREAM_WL:
        res = run_stream(a, wl, rank, world, local, ClockSampler, measured_peaks)
    elif wl == "idct_sweep":
        res = run_sweep(a, rank, world, local, ClockSampler, measured_peaks)
    elif wl == "mc_pus":
        res = run_mc(a, rank, world, local, ClockSampler, measured_peaks)
    else:
        raise SystemExit("unknown workload " + wl)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    return res


def run_reference(a):
    wl = a.workload or DEFAULT_WORKLOAD
    if wl in STREAM_WL:
        return run_reference_stream(a, wl)
    return run_reference_sweep(a)

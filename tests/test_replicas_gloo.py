"""N>1 host logic on CPU (SURVEY §8e "replicas only"): two ranks over gloo run bench.py's replica path with
parse-only decoders — closed-GOP segments dealt round-robin, barrier + max-over-ranks timing, whole-job
aggregate printed by rank 0 only."""
import json, os, subprocess, sys
from tests.helpers import ROOT, GOLDEN


def test_split_and_deal():
    sys.path.insert(0, ROOT)
    from benchmarks.workloads import split_idr_segments, deal
    data = open(os.path.join(GOLDEN, "streams", "tiny_416x240_q30.bin"), "rb").read()
    gold = json.load(open(os.path.join(GOLDEN, "streams", "tiny_416x240_q30.md5.json")))
    segs = split_idr_segments(data)
    assert len(segs) == gold["frames"]          # every enc265 intra picture is an IDR
    for s in segs:                              # each segment is self-contained: VPS, SPS, PPS, then the picture
        types = [(s[i + 3] >> 1) & 0x3f for i in range(len(s) - 3) if s[i:i + 3] == b"\x00\x00\x01"]
        assert types[:3] == [32, 33, 34] and 16 <= types[3] <= 21
    shares = [deal(segs, 3, r) for r in range(3)]
    assert sorted(i for sh in shares for i, _ in sh) == list(range(len(segs)))
    assert max(len(sh) for sh in shares) - min(len(sh) for sh in shares) <= 1


def test_segments_decode_like_the_whole_stream():
    """parse-only decode of the segments one by one yields the same pictures (work-list sizes) as the whole stream"""
    sys.path.insert(0, ROOT)
    from benchmarks.workloads import split_idr_segments
    from libde265_b200.decoder import Decoder
    data = open(os.path.join(GOLDEN, "streams", "tiny_416x240_q30.bin"), "rb").read()

    def sizes(dec, d):
        out = []
        dec.decode_stream(d, on_picture=lambda img: out.append((dec.job_of(img).n_tus, dec.job_of(img).n_coeffs)))
        return out
    dec = Decoder(device=0, threads=0, batch=1, parse_only=True)
    whole = sizes(dec, data)
    parts = []
    for s in split_idr_segments(data):
        dec.lib.de265_reset(dec.h)
        parts += sizes(dec, s)
    dec.close()
    assert parts == whole and len(whole) > 0


def test_bench_two_ranks_gloo():
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", B265_DIST_BACKEND="gloo")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29517", os.path.join(ROOT, "bench.py"), "--gpus", "2", "--steps", "2", "--warmup", "3",
           "--workload", "parse_replicas"]
    r = subprocess.run(cmd, env=env, cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, "exactly one JSON line (rank 0): %r" % r.stdout
    d = json.loads(lines[0])
    assert d["n_gpus"] == 2 and d["value"] > 0 and d["scaling"] == "weak" and d["steps"] == 2

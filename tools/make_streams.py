#!/usr/bin/env python3
"""Generate the synthetic HEVC test/benchmark streams with the reference's own
encoder `oracle/_ref/enc265` and record the reference decoder's per-frame, per-plane MD5.

Test infrastructure.  Runs only in the build container (needs oracle/_ref built from /root/reference);
its outputs (`tests/golden/streams/*.bin` + `*.md5.json`) are committed so the GPU box needs neither.

Every enc265 intra frame is an IDR picture with its own parameter sets, so segments encoded by
independent enc265 processes concatenate into one valid stream (this is how the encode is parallelised).
"""
import argparse, hashlib, json, os, subprocess, sys, tempfile
from concurrent.futures import ThreadPoolExecutor
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")
OUT = os.path.join(ROOT, "tests", "golden", "streams")


def gen_frame(w, h, t, seed, scale=1.0):
    """Luma: two sinusoid gratings + uniform noise on a third of the 64x64 tiles. Chroma: ramps."""
    rng = np.random.default_rng(seed * 1000 + t)
    y, x = np.mgrid[0:h, 0:w].astype(np.float32)
    ph = 0.3 * t
    l = 128 + 45 * np.sin(2 * np.pi * (x / (37 * scale) + y / (91 * scale)) + ph) \
            + 45 * np.sin(2 * np.pi * (x / (23 * scale) - y / (37 * scale)) - 0.5 * ph)
    tiles = rng.random(((h + 63) // 64, (w + 63) // 64)) < (1 / 3)
    mask = np.kron(tiles, np.ones((64, 64), dtype=bool))[:h, :w]
    noise = rng.integers(-10, 11, size=(h, w))
    l = l + mask * noise
    cy, cx = np.mgrid[0:h // 2, 0:w // 2].astype(np.float32)
    cb = 20 + 200 * cx / (w // 2 - 1) + 8 * np.sin(cy / 9.0 + ph)
    cr = 20 + 200 * cy / (h // 2 - 1) + 8 * np.cos(cx / 7.0 - ph)
    f = lambda a: np.clip(np.rint(a), 0, 255).astype(np.uint8)
    return f(l), f(cb), f(cr)


def encode_segment(args):
    name, w, h, seed, t0, n, qp, scale, extra = args
    with tempfile.TemporaryDirectory() as td:
        yuv = os.path.join(td, "in.yuv")
        with open(yuv, "wb") as fh:
            for t in range(t0, t0 + n):
                for p in gen_frame(w, h, t, seed, scale):
                    fh.write(p.tobytes())
        out = os.path.join(td, "out.bin")
        cmd = [os.path.join(REF, "enc265"), "-i", yuv, "-w", str(w), "-h", str(h), "-f", str(n),
               "--sop-structure", "intra", "-q", str(qp), "-o", out] + extra
        subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, cwd=td)
        return open(out, "rb").read()


def ref_md5(path, w, h):
    """per-frame [md5(Y), md5(Cb), md5(Cr)] of `dec265 -q -o` (SSE path) — the stream-level oracle."""
    with tempfile.TemporaryDirectory() as td:
        o = os.path.join(td, "o.yuv")
        subprocess.run([os.path.join(REF, "dec265"), "-q", "-o", o, path], check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        data = open(o, "rb").read()
    fs = w * h * 3 // 2
    res = []
    for i in range(len(data) // fs):
        f = data[i * fs:(i + 1) * fs]
        res.append([hashlib.md5(f[:w * h]).hexdigest(),
                    hashlib.md5(f[w * h:w * h * 5 // 4]).hexdigest(),
                    hashlib.md5(f[w * h * 5 // 4:]).hexdigest()])
    return res


def make(name, w, h, seed, frames, seg, qp, scale=1.0, extra=(), jobs=8):
    os.makedirs(OUT, exist_ok=True)
    path = os.path.join(OUT, name + ".bin")
    if not os.path.exists(path):
        tasks = [(name, w, h, seed, t0, min(seg, frames - t0), qp, scale, list(extra))
                 for t0 in range(0, frames, seg)]
        with ThreadPoolExecutor(jobs) as ex:
            parts = list(ex.map(encode_segment, tasks))
        with open(path, "wb") as fh:
            for p in parts:
                fh.write(p)
    md5 = ref_md5(path, w, h)
    json.dump({"width": w, "height": h, "frames": len(md5), "qp": qp, "seed": seed,
               "generator": "tools/make_streams.py", "md5": md5},
              open(os.path.join(OUT, name + ".md5.json"), "w"), indent=0)
    print(name, os.path.getsize(path), "bytes", len(md5), "frames", flush=True)


STREAMS = {
    # name: (w, h, seed, frames, frames/segment, qp, scale, extra enc265 args)
    "tiny_416x240_q30": (416, 240, 7, 3, 1, 30, 1.0, []),
    "tiny_416x240_q20_bf": (416, 240, 8, 2, 1, 20, 1.0, ["--TB-IntraPredMode", "brute-force"]),
    "odd_200x136_q26": (200, 136, 9, 2, 1, 26, 0.6, []),        # not a multiple of the 32-px CTB
    "cb64_320x192_q34": (320, 192, 11, 2, 1, 34, 1.0, ["--max-cb-size", "64"]),
    "c1_1080p_q32": (1920, 1080, 1, 16, 2, 32, 1.0, []),        # config 1 (16 unique frames, looped x4 = 64)
    "c1_1080p_q22": (1920, 1080, 1, 4, 1, 22, 1.0, []),         # coefficient-heavy variant
    "c4_2160p_q32": (3840, 2160, 4, 8, 1, 32, 2.0, []),         # config 4 base (PPS-patched => deblock on)
}
# config 5: eight INDEPENDENT 4K streams (seeds 10..17, 4 frames each), one per GPU / decoder instance
for _s in range(10, 18):
    STREAMS["c5_2160p_s%d" % _s] = (3840, 2160, _s, 4, 1, 32, 2.0, [])

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("names", nargs="*", default=[])
    ap.add_argument("-j", type=int, default=8)
    a = ap.parse_args()
    for n in (a.names or STREAMS):
        w, h, seed, frames, seg, qp, scale, extra = STREAMS[n]
        make(n, w, h, seed, frames, seg, qp, scale, extra, a.j)

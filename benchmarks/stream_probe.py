"""Quick end-to-end probe: decode a golden stream looped N times through the de265 API, print fps + stats.
   python -m benchmarks.stream_probe c1_1080p_q32 --loops 8 --threads 8 --batch 8 [--parse-only]"""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from libde265_b200.decoder import Decoder


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("name")
    ap.add_argument("--loops", type=int, default=4)
    ap.add_argument("--threads", type=int, default=8)
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--parse-only", action="store_true")
    a = ap.parse_args()
    data = open(os.path.join(ROOT, "tests", "golden", "streams", a.name + ".bin"), "rb").read() * a.loops
    dec = Decoder(threads=a.threads, batch=a.batch, parse_only=a.parse_only)
    n = [0]
    t0 = time.perf_counter()
    dec.decode_stream(data, chunk=1 << 20, on_picture=lambda img: n.__setitem__(0, n[0] + 1))
    dt = time.perf_counter() - t0
    st = dec.stats()
    dec.close()
    print(json.dumps({"stream": a.name, "frames": n[0], "seconds": dt, "fps": n[0] / dt, "threads": a.threads,
                      "batch": a.batch, "parse_only": a.parse_only, "stats": st}))


if __name__ == "__main__":
    main()

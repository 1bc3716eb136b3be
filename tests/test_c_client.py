"""A compiled C client (tests/c/dec_client.c, include/de265.h only) linked against libb265.so the way a libde265
user's program is linked against libde265: decodes golden streams on the GPU, planes compared with the reference MD5."""
import hashlib, json, os, subprocess, sys
import pytest
from tests.helpers import ROOT, GOLDEN

STREAMS = os.path.join(GOLDEN, "streams")
LIBDIR = os.path.join(ROOT, "libde265_b200")


def build_client(tmp):
    exe = os.path.join(tmp, "dec_client")
    subprocess.run(["gcc", "-std=c99", "-O1", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "tests", "c", "dec_client.c"), "-o", exe, "-L", LIBDIR, "-l:libb265.so",
                    "-Wl,-rpath," + LIBDIR], check=True)
    return exe


def test_c_client_compiles_and_links(tmp_path):
    """header is valid C99 and every symbol the client uses is exported (no GPU needed to link)"""
    from libde265_b200 import lib as L
    L.load()
    build_client(str(tmp_path))


@pytest.mark.gpu
@pytest.mark.parametrize("name,threads", [("tiny_416x240_q30_lf", 0), ("gen_pb_reorder_416x240", 3), ("gen_all_1280x720", 2)])
def test_c_client_md5(tmp_path, name, threads):
    exe = build_client(str(tmp_path))
    from tests.test_parser_streams import load_stream
    data, gold = load_stream(name)
    src, yuv = str(tmp_path / "in.bin"), str(tmp_path / "out.yuv")
    open(src, "wb").write(data)
    r = subprocess.run([exe, src, yuv, str(threads)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert ("pictures %d " % gold["frames"]) in r.stdout
    w, h = gold["width"], gold["height"]
    raw = open(yuv, "rb").read()
    fs = w * h * 3 // 2
    assert len(raw) == fs * gold["frames"]
    for i in range(gold["frames"]):
        f = raw[i * fs:(i + 1) * fs]
        got = [hashlib.md5(f[:w * h]).hexdigest(), hashlib.md5(f[w * h:w * h * 5 // 4]).hexdigest(),
               hashlib.md5(f[w * h * 5 // 4:]).hexdigest()]
        assert got == gold["md5"][i], "frame %d of %s" % (i, name)

"""Host parser (CABAC + syntax -> work lists) checked WITHOUT a GPU: the work lists of every picture of
the golden enc265 streams are reconstructed by the C oracle (orc_reconstruct_frame / orc_deblock /
orc_sao) and the per-plane MD5 must equal the reference decoder's (`dec265 -o`, committed in
tests/golden/streams/*.md5.json by tools/make_streams.py).  This pins parser + oracle together
against the unmodified reference at stream level."""
import ctypes as C
import hashlib, json, os
import numpy as np
import pytest
from tests.helpers import GOLDEN, load_oracle
from libde265_b200 import lib as L
from libde265_b200.decoder import Decoder

STREAMS = os.path.join(GOLDEN, "streams")


from oracle.frame import oracle_planes  # noqa: E402,F401  (the C oracle on one picture's work lists)


def parse_and_reconstruct(data, threads=0, max_frames=None):
    dec = Decoder(parse_only=True, threads=threads)
    md5s = []
    dpb = {}

    def on_pic(img):
        if max_frames is not None and len(md5s) >= max_frames:
            return
        job = dec.job_of(img)
        if dec.lib.b265_de265_image_starts_cvs(img):   # a new coded video sequence, output after everything before it
            group[0] += 1
        order.append((group[0], dec.lib.b265_de265_image_poc(img), len(order)))
        md5s.append([hashlib.md5(p.tobytes()).hexdigest() for p in oracle_planes(job, dpb)])

    group, order = [0], []
    dec.decode_stream(data, on_picture=on_pic)
    dec.close()
    # a parse-only decoder delivers in decoding order; the golden MD5s are in output order
    return [md5s[i] for _, _, i in sorted(order)]


def load_stream(name):
    """golden stream + reference MD5s; '<base>_lf' = the base stream with deblocking switched on by
    re-writing its PPS (tools/pps_patch.py), MD5s from the reference decoder on the patched stream"""
    gold = json.load(open(os.path.join(STREAMS, name + ".md5.json")))
    if "base" in gold:
        from tools.pps_patch import patch_stream
        data = patch_stream(open(os.path.join(STREAMS, gold["base"] + ".bin"), "rb").read(),
                            beta_div2=gold["beta_offset_div2"], tc_div2=gold["tc_offset_div2"])
    else:
        data = open(os.path.join(STREAMS, name + ".bin"), "rb").read()
    return data, gold


@pytest.mark.parametrize("name,max_frames", [("tiny_416x240_q30", None), ("tiny_416x240_q20_bf", None),
                                             ("odd_200x136_q26", None), ("cb64_320x192_q34", None),
                                             ("c1_1080p_q32", 1), ("tiny_416x240_q30_lf", None),
                                             ("tiny_416x240_q20_bf_lf", None), ("odd_200x136_q26_lf", None),
                                             ("cb64_320x192_q34_lf", None), ("c1_1080p_q32_lf", 1)])
def test_parser_matches_reference_md5(name, max_frames):
    data, gold = load_stream(name)
    got = parse_and_reconstruct(data, max_frames=max_frames)
    n = len(gold["md5"]) if max_frames is None else max_frames
    assert len(got) == n
    for i in range(n):
        assert got[i] == gold["md5"][i], "frame %d of %s" % (i, name)


def generated_streams():
    """random-syntax streams of tools/streamgen (tools/make_gen_streams.py): SAO, PCM, transform skip, cu_qp_delta,
    scaling lists, tiles, wavefronts, (dependent) slice segments, deblocking overrides, range-extension flags"""
    return sorted(f[:-len(".md5.json")] for f in os.listdir(STREAMS) if f.startswith("gen_") and f.endswith(".md5.json"))


@pytest.mark.parametrize("name", generated_streams())
def test_parser_matches_reference_md5_generated(name):
    data, gold = load_stream(name)
    got = parse_and_reconstruct(data)
    assert len(got) == gold["frames"]
    for i in range(gold["frames"]):
        assert got[i] == gold["md5"][i], "frame %d of %s" % (i, name)


def test_generated_streams_cover_the_tools():
    """the catalogue really switches the tools on (guards against a silently reduced golden set)"""
    names = generated_streams()
    assert len(names) >= 40
    args = [json.load(open(os.path.join(STREAMS, n + ".md5.json")))["args"] for n in names]
    for key in ("sao", "pcm", "tskip", "cu_qp_delta", "sign_hiding", "tq_bypass", "wpp", "dep_slices", "dbk_override",
                "rext_implicit_rdpcm", "chroma_qp_offset_list"):
        assert sum(1 for a in args if a.get(key)) >= 2, key
    assert {a.get("scaling", 0) for a in args} >= {0, 1, 2, 3}
    assert sum(1 for a in args if a.get("tile_cols", 1) * a.get("tile_rows", 1) > 1) >= 3


def test_parser_threads_same_result():
    data = open(os.path.join(STREAMS, "tiny_416x240_q30.bin"), "rb").read()
    assert parse_and_reconstruct(data, threads=0) == parse_and_reconstruct(data, threads=3)

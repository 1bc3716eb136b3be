#!/usr/bin/env python3
"""Random-syntax test streams (tools/streamgen) and their golden MD5s from the unmodified reference decoder.

Test infrastructure.  Runs only in the build container (needs oracle/_ref/dec265 built from /root/reference);
the outputs `tests/golden/streams/gen_*.bin` + `gen_*.md5.json` are committed, the GPU box needs neither the
generator nor the reference.  The named streams switch on the coding tools the reference's encoder cannot
produce (SAO, PCM, transform skip, cu_qp_delta, scaling lists, tiles, wavefronts, slices and dependent slice
segments, deblocking overrides, range-extension flags); the `fuzz_*` streams draw a whole configuration at random.
"""
import argparse, json, os, random, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tools.make_streams import OUT, ref_md5  # noqa: E402

GEN = os.path.join(ROOT, "tools", "streamgen", "b265_streamgen")
HOST = os.path.join(ROOT, "libde265_b200", "csrc", "host")


def build_generator():
    src = [os.path.join(ROOT, "tools", "streamgen", "streamgen.cc"), os.path.join(HOST, "params.cc"),
           os.path.join(HOST, "cabac.cc")]
    deps = src + [os.path.join(HOST, f) for f in ("slice_syntax.h", "cabac.h", "picture.h", "params.h", "bits.h")]
    if os.path.exists(GEN) and all(os.path.getmtime(GEN) >= os.path.getmtime(d) for d in deps):
        return GEN
    subprocess.run(["g++", "-std=c++17", "-O2", "-o", GEN] + src, check=True)
    return GEN


NAMED = {
    # name: generator arguments
    "gen_base_416x240": dict(seed=2, w=416, h=240, frames=3),
    "gen_sao_200x136": dict(seed=3, w=200, h=136, sao=1),
    "gen_pcm_128x96": dict(seed=5, w=128, h=96, pcm=1, pcm_bits_luma=5, pcm_bits_chroma=6, pcm_lf_disabled=1, sao=1),
    "gen_pcm8_ctb16_200x136": dict(seed=30, w=200, h=136, ctb=4, sao=1, pcm=1, maxpcm=4),
    "gen_tskip_rext_128x96": dict(seed=28, w=128, h=96, tskip=1, max_tskip=4, rext_rotation=1, rext_context=1,
                                  rext_implicit_rdpcm=1, tq_bypass=1),
    "gen_qpdelta_ctb64_256x128": dict(seed=8, w=256, h=128, ctb=6, cu_qp_delta=1, qp_delta_depth=2, sign_hiding=1),
    "gen_chromaqp_128x96": dict(seed=29, w=128, h=96, chroma_qp_offset_list=1, cb_qp_offset=3, cr_qp_offset=-4,
                                slice_chroma_offsets=1, cu_qp_delta=1, slice_ctbs=7),
    "gen_scaling_sps_128x96": dict(seed=12, w=128, h=96, scaling=2),
    "gen_scaling_pps_128x96": dict(seed=13, w=128, h=96, scaling=3, density=2.0),
    "gen_scaling_default_128x96": dict(seed=11, w=128, h=96, scaling=1),
    "gen_tiles_416x240": dict(seed=21, w=416, h=240, tile_cols=3, tile_rows=2, tiles_uniform=0, lf_across_tiles=0, sao=1),
    "gen_wpp_256x192": dict(seed=22, w=256, h=192, wpp=1, sao=1),
    "gen_slices_dep_256x192": dict(seed=24, w=256, h=192, slice_ctbs=4, dep_slices=1, sao=1, lf_across_slices=1),
    "gen_slices_wpp_256x192": dict(seed=25, w=256, h=192, slice_ctbs=4, dep_slices=1, wpp=1),
    "gen_slices_tiles_256x192": dict(seed=26, w=256, h=192, slice_ctbs=3, dep_slices=1, tile_cols=2, tile_rows=2, sao=1),
    "gen_dbk_override_256x192": dict(seed=27, w=256, h=192, slice_ctbs=6, dbk_override=1, lf_across_slices=0, sao=1),
    "gen_tb8_ctb64_320x192": dict(seed=31, w=320, h=192, ctb=6, mincb=4, mintb=3, sao=1, thdepth=3),
    "gen_bypass_rdpcm_128x64": dict(seed=123, w=128, h=64, ctb=4, mincb=4, tq_bypass=1, rext_implicit_rdpcm=1),
    # one large picture with everything the wavefront and the post filters have to get right at once
    "gen_all_1280x720": dict(seed=77, w=1280, h=720, frames=2, sao=1, pcm=1, tskip=1, cu_qp_delta=1, sign_hiding=1,
                             tq_bypass=1, scaling=3, slice_ctbs=60, dep_slices=1, dbk_override=1, lf_across_slices=1,
                             p_pcm=0.03, p_tq_bypass=0.05),
    # ---- P and B slices: merge / AMVP / temporal candidates, AMP, weighted prediction, list modification ----
    "gen_p_1ref_416x240": dict(seed=41, w=416, h=240, frames=4, inter=1, max_refs=1, p_bslice=0.0, p_islice=0.0),
    "gen_b_tmvp_416x240": dict(seed=42, w=416, h=240, frames=6, inter=1, max_refs=3, tmvp=1, p_bslice=1.0, sao=1),
    "gen_pb_amp_weighted_256x192": dict(seed=43, w=256, h=192, frames=5, inter=1, max_refs=3, tmvp=1, amp=1, weighted_pred=1,
                                        weighted_bipred=1, lists_modification=1, num_ref_default_l0=2, num_ref_default_l1=3,
                                        cabac_init_present=1, thdepth_inter=2),
    "gen_pb_parmerge_ctb64_320x192": dict(seed=44, w=320, h=192, ctb=6, frames=4, inter=1, max_refs=2, tmvp=1, par_merge_level=4,
                                          thdepth_inter=0, amp=1, cu_qp_delta=1, sign_hiding=1),
    "gen_pb_slices_tiles_256x192": dict(seed=45, w=256, h=192, frames=5, inter=1, max_refs=2, tmvp=1, slice_ctbs=5, dep_slices=1,
                                        tile_cols=2, tile_rows=2, sao=1, lf_across_slices=0, lf_across_tiles=0, dbk_override=1),
    "gen_pb_wpp_cip_256x192": dict(seed=46, w=256, h=192, frames=5, inter=1, max_refs=2, wpp=1, cip=1, p_intra_cu=0.5, pcm=1,
                                   tskip=1, tq_bypass=1, scaling=3),
    "gen_pb_reorder_416x240": dict(seed=48, w=416, h=240, frames=9, inter=1, reorder=1, max_refs=3, tmvp=1, p_bslice=0.8, sao=1),
    "gen_pb_longterm_256x192": dict(seed=49, w=256, h=192, frames=12, inter=1, long_term=1, poc_bits=4, max_refs=2, tmvp=1,
                                    lists_modification=1, weighted_pred=1),
    "gen_pb_reorder_longterm_256x192": dict(seed=50, w=256, h=192, frames=13, inter=1, reorder=1, long_term=1, poc_bits=5,
                                            max_refs=3, tmvp=1, amp=1, slice_ctbs=8),
    # ---- open GOPs, random access and lost pictures (skipped RASL pictures, generated references) ----
    "gen_pb_opengop_256x192": dict(seed=51, w=256, h=192, frames=17, inter=1, reorder=1, open_gop=1, max_refs=3, tmvp=1, sao=1),
    "gen_pb_opengop_cut_256x192": dict(seed=51, w=256, h=192, frames=17, inter=1, reorder=1, open_gop=1, max_refs=3, tmvp=1, sao=1,
                                       post={"cut_at_cra": 1}),
    "gen_pb_opengop_eos_256x192": dict(seed=57, w=256, h=192, frames=17, inter=1, reorder=1, open_gop=1, p_cra=0.5, max_refs=3, tmvp=1,
                                       post={"eos_at_cra": 1}),
    "gen_p_lost_picture_256x192": dict(seed=53, w=256, h=192, frames=8, inter=1, max_refs=2, tmvp=1, p_islice=0.0,
                                       post={"drop_picture": 2}),
    "gen_pb_lost_longterm_256x192": dict(seed=54, w=256, h=192, frames=12, inter=1, long_term=1, poc_bits=4, max_refs=2, tmvp=1,
                                         post={"drop_picture": 1}),
    # ---- decoded picture hash SEI (sei.cc:32-61, 160-354): the three hash types, P/B pictures, and a wrong hash ----
    "gen_hash_md5_200x136": dict(seed=61, w=200, h=136, frames=3, sao=1, post={"sei_hash": 0}),
    "gen_hash_crc_416x240": dict(seed=62, w=416, h=240, frames=2, sao=1, pcm=1, post={"sei_hash": 1}),
    "gen_hash_checksum_320x192": dict(seed=63, w=320, h=192, ctb=6, frames=2, post={"sei_hash": 2}),
    "gen_hash_pb_mixed_256x192": dict(seed=64, w=256, h=192, frames=6, inter=1, max_refs=2, tmvp=1, sao=1, post={"sei_hash": -1}),
    "gen_hash_wrong_256x192": dict(seed=64, w=256, h=192, frames=6, inter=1, max_refs=2, tmvp=1, sao=1,
                                   post={"sei_hash": -1, "corrupt_hash": 3}),
    # ---- P/B streams at the sizes the metric names (BASELINE north_star: 1080p and 4K): low-delay and random-access
    #      shapes, >= 32 pictures at 1080p, weighted prediction + temporal MVP + SAO on, mostly skipped / merged large blocks ----
    "gen_pb_1080p_ld32": dict(seed=71, w=1920, h=1080, frames=32, inter=1, max_refs=3, tmvp=1, amp=1, sao=1, weighted_pred=1,
                              weighted_bipred=1, p_skip=0.6, p_split_cu=0.3, p_split_tu=0.25, density=0.5, p_intra_cu=0.1),
    "gen_pb_1080p_ra33": dict(seed=72, w=1920, h=1080, frames=33, inter=1, reorder=1, max_refs=3, tmvp=1, amp=1, sao=1,
                              weighted_bipred=1, p_bslice=0.9, p_skip=0.6, p_split_cu=0.3, p_split_tu=0.25, density=0.5, p_intra_cu=0.1),
    "gen_pb_2160p_ld12": dict(seed=73, w=3840, h=2160, frames=12, inter=1, max_refs=2, tmvp=1, amp=1, sao=1, weighted_pred=1,
                              p_skip=0.65, p_split_cu=0.25, p_split_tu=0.2, density=0.4, p_intra_cu=0.08),
    "gen_pb_720p": dict(seed=47, w=1280, h=720, frames=4, inter=1, max_refs=3, tmvp=1, amp=1, sao=1, weighted_bipred=1,
                        slice_ctbs=200, p_skip=0.5),
}


def fuzz_config(seed, inter=False):
    """a whole configuration drawn at random, legal by construction (7.4.3.2, 7.4.3.3 range constraints);
    inter: P and B slices on top (reference picture sets with negative pictures only)"""
    rnd = random.Random(seed)
    ctb = rnd.choice([4, 5, 5, 6])
    mincb = rnd.choice([c for c in (3, 4, 5) if c <= ctb])
    mintb = rnd.choice([t for t in (2, 3, 4) if t < mincb])
    maxtb = rnd.choice([t for t in (3, 4, 5) if mintb <= t <= min(5, ctb)])
    w = rnd.choice([64, 96, 128, 200, 256, 320, 416])
    h = rnd.choice([64, 72, 96, 136, 192, 240])
    m = 1 << mincb
    w, h = max(m, w // m * m), max(m, h // m * m)
    a = dict(seed=seed, w=w, h=h, ctb=ctb, mincb=mincb, mintb=mintb, maxtb=maxtb, frames=rnd.choice([1, 2, 3]),
             thdepth=rnd.randint(0, min(4, ctb - mintb)))
    cw, ch = -(-w >> ctb), -(-h >> ctb)
    flag = lambda p=0.5: int(rnd.random() < p)
    a["sao"], a["strong_intra"], a["sign_hiding"] = flag(), flag(), flag()
    a["tq_bypass"], a["cip"] = flag(0.3), flag(0.3)
    if flag(0.4):
        lo, hi = min(mincb, 5), min(ctb, 5)
        a.update(pcm=1, pcm_bits_luma=rnd.randint(1, 8), pcm_bits_chroma=rnd.randint(1, 8), pcm_lf_disabled=flag())
        a["minpcm"] = rnd.randint(lo, hi)
        a["maxpcm"] = rnd.randint(a["minpcm"], hi)
    if flag(0.5):
        a.update(tskip=1, max_tskip=min(maxtb, rnd.choice([2, 2, 3, 4, 5])))
    if flag(0.4):
        a.update(rext_rotation=flag(), rext_context=flag(), rext_implicit_rdpcm=flag())
    if flag(0.5):
        a.update(cu_qp_delta=1, qp_delta_depth=rnd.randint(0, ctb - mincb))
    a.update(cb_qp_offset=rnd.randint(-6, 6), cr_qp_offset=rnd.randint(-6, 6), slice_chroma_offsets=flag())
    if flag(0.3):
        a.update(chroma_qp_offset_list=1, chroma_qp_offset_depth=rnd.randint(0, ctb - mincb))
    a["scaling"] = rnd.choice([0, 0, 1, 2, 3])
    a["init_qp"] = rnd.randint(15, 40)
    a["qp_lo"] = rnd.randint(5, 30)
    a["qp_hi"] = a["qp_lo"] + rnd.randint(0, 20)
    par = rnd.choice(["none", "tiles", "wpp", "none"])
    if par == "tiles" and cw * ch > 1:
        a.update(tile_cols=rnd.randint(1, min(4, cw)), tile_rows=rnd.randint(1, min(3, ch)), tiles_uniform=flag(),
                 lf_across_tiles=flag())
    if par == "wpp":
        a["wpp"] = 1
    if flag(0.6):
        a.update(slice_ctbs=rnd.randint(1, 12), dep_slices=flag(), lf_across_slices=flag())
    a["dbk_control"] = flag(0.8)
    if a["dbk_control"]:
        a.update(dbk_override=flag(), dbk_disabled=flag(0.3), beta_div2=rnd.randint(-6, 6), tc_div2=rnd.randint(-6, 6))
    if inter:
        a.update(inter=1, frames=rnd.choice([3, 4, 6]), max_refs=rnd.randint(1, 4), tmvp=flag(0.6), amp=flag(),
                 thdepth_inter=rnd.randint(0, min(3, ctb - mintb)), weighted_pred=flag(0.4), weighted_bipred=flag(0.4),
                 lists_modification=flag(0.4), par_merge_level=rnd.randint(2, ctb), cabac_init_present=flag(),
                 num_ref_default_l0=rnd.randint(1, 3), num_ref_default_l1=rnd.randint(1, 3),
                 p_bslice=rnd.choice([0.0, 0.5, 1.0]), p_islice=rnd.choice([0.0, 0.15]),
                 p_skip=rnd.choice([0.1, 0.3, 0.6]), p_intra_cu=rnd.choice([0.0, 0.2, 0.5]), p_merge=rnd.choice([0.2, 0.5, 0.8]))
        if flag(0.4):
            a.update(reorder=1, frames=rnd.choice([5, 9]))      # output order differs from decoding order
        if flag(0.4):
            a.update(long_term=1, poc_bits=rnd.choice([4, 5, 8]), frames=max(a["frames"], 8))
    a["density"] = rnd.choice([0.5, 1.0, 1.0, 1.5, 2.5])
    a["p_split_cu"] = rnd.choice([0.3, 0.5, 0.7])
    a["p_split_tu"] = rnd.choice([0.2, 0.4, 0.7])
    a["p_big_coeff"] = rnd.choice([0.0, 0.01, 0.05])
    return a


FUZZ_SEEDS = list(range(1000, 1024))
FUZZ_INTER_SEEDS = list(range(2000, 2024))


def catalogue():
    c = dict(NAMED)
    for s in FUZZ_SEEDS:
        c["gen_fuzz_%d" % s] = fuzz_config(s)
    for s in FUZZ_INTER_SEEDS:
        c["gen_fuzzpb_%d" % s] = fuzz_config(s, inter=True)
    return c


def split_nals(data):
    """Annex-B stream -> list of NAL units, each with its start code (and leading zero byte) attached"""
    import re
    starts = [m.start() for m in re.finditer(b"\x00\x00\x01", data)]
    out = []
    for i, s0 in enumerate(starts):
        b = s0 - 1 if s0 > 0 and data[s0 - 1] == 0 else s0
        e = starts[i + 1] if i + 1 < len(starts) else len(data)
        if i + 1 < len(starts) and data[e - 1] == 0:
            e -= 1
        out.append(data[b:e])
    return out


def nal_type(nal):
    i = nal.index(b"\x00\x00\x01") + 3
    return (nal[i] >> 1) & 0x3f


def first_slice(nal):
    i = nal.index(b"\x00\x00\x01") + 3
    return nal_type(nal) < 32 and bool(nal[i + 2] & 0x80)


def post_process(data, post):
    """Stream surgery the generator itself does not do (the reference decoder still judges the result):
    cut_at_cra   drop every picture before the first CRA: random access into an open GOP, its RASL pictures must be
                 skipped (decctx.cc:2038-2045) and trailing pictures that still name earlier pictures get generated ones
    eos_at_cra   an end-of-sequence NAL in front of every CRA picture: NoRaslOutputFlag = 1 in mid-stream
    drop_picture=k  lose the k-th picture (0-based, decoding order): unavailable-reference concealment (decctx.cc:1408-1434)"""
    nals = split_nals(data)
    out, pic, seen_cra = [], -1, False
    for n in nals:
        t = nal_type(n)
        if t < 32 and first_slice(n):
            pic += 1
        if t == 21:
            if "eos_at_cra" in post and first_slice(n):
                out.append(b"\x00\x00\x01\x48\x01")   # EOS_NUT = 36
            seen_cra = True
        if "cut_at_cra" in post and t < 32 and not seen_cra:
            continue
        if "drop_picture" in post and t < 32 and pic == post["drop_picture"]:
            continue
        out.append(n)
    return b"".join(out)


def crc_plane(plane):
    """the reference's CRC of a plane (sei.cc:216-254): CCITT register 0xFFFF, two zero bytes FIRST, then the rows"""
    import numpy as np
    tab = np.zeros(256, np.uint32)
    for s_ in range(256):
        t = s_ ^ (s_ >> 4)
        tab[s_] = (t ^ (t << 5) ^ (t << 12)) & 0xffff
    crc = 0xffff
    for b in (0, 0):
        crc = ((crc << 8) & 0xffff) ^ int(tab[(b ^ (crc >> 8)) & 0xff])
    for b in plane.tobytes():
        crc = ((crc << 8) & 0xffff) ^ int(tab[(b ^ (crc >> 8)) & 0xff])
    return crc


def checksum_plane(plane):
    """sei.cc:160-182"""
    import numpy as np
    h, w = plane.shape
    x, y = np.meshgrid(np.arange(w), np.arange(h))
    mask = ((x & 0xff) ^ (y & 0xff) ^ (x >> 8) ^ (y >> 8)) & 0xff
    return int((plane.astype(np.int64) ^ mask).sum() & 0xffffffff)


def escape_rbsp(b):
    out, zeros = bytearray(), 0
    for v in b:
        if zeros >= 2 and v <= 3:
            out.append(3)
            zeros = 0
        out.append(v)
        zeros = zeros + 1 if v == 0 else 0
    return bytes(out)


def add_hash_sei(data, w, h, hash_type, corrupt=None):
    """append a decoded-picture-hash suffix SEI (payload 132, D.2.19) to every picture: the hash of the REFERENCE decoder's
    output planes (so `dec265 -c` accepts the stream).  Only for streams whose decoding order is their output order.
    hash_type 0 MD5, 1 CRC, 2 checksum, -1: cycle through the three.  corrupt = picture index whose hash gets one bit flipped."""
    import hashlib, tempfile
    import numpy as np
    from tools.make_streams import REF
    with tempfile.TemporaryDirectory() as td:
        src, yuv = os.path.join(td, "s.bin"), os.path.join(td, "o.yuv")
        open(src, "wb").write(data)
        subprocess.run([os.path.join(REF, "dec265"), "-q", "-o", yuv, src], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        raw = np.fromfile(yuv, np.uint8)
    fs = w * h * 3 // 2
    n = len(raw) // fs
    nals = split_nals(data)
    out, pic = [], -1
    def sei_for(i):
        f = raw[i * fs:(i + 1) * fs]
        planes = [f[:w * h].reshape(h, w), f[w * h:w * h * 5 // 4].reshape(h // 2, w // 2), f[w * h * 5 // 4:].reshape(h // 2, w // 2)]
        t = hash_type if hash_type >= 0 else i % 3
        body = bytearray([t])
        for pl in planes:
            if t == 0:
                body += hashlib.md5(pl.tobytes()).digest()
            elif t == 1:
                body += crc_plane(pl).to_bytes(2, "big")
            else:
                body += checksum_plane(pl).to_bytes(4, "big")
        if corrupt == i:
            body[-1] ^= 0x10
        payload = bytes([132, len(body)]) + bytes(body) + b"\x80"
        return b"\x00\x00\x01\x50\x01" + escape_rbsp(payload)   # SUFFIX_SEI_NUT = 40, temporal id 0
    for k, nal in enumerate(nals):
        t = nal_type(nal)
        if t < 32 and first_slice(nal):
            if pic >= 0:
                out.append(sei_for(pic))
            pic += 1
        elif t >= 32 and pic >= 0 and (not out or nal_type(out[-1]) != 40) and t != 40:
            # a non-VCL NAL (parameter set, EOS ...) ends the picture too
            out.append(sei_for(pic))
        out.append(nal)
    assert pic + 1 == n, "hash SEI needs every picture decoded and output in decoding order (%d coded, %d output)" % (pic + 1, n)
    if nal_type(out[-1]) != 40:
        out.append(sei_for(pic))
    return b"".join(out)


def generate(name, args, out_dir=OUT):
    path = os.path.join(out_dir, name + ".bin")
    args = dict(args)
    post = args.pop("post", None)
    subprocess.run([build_generator(), path] + ["%s=%s" % kv for kv in args.items()], check=True)
    if post:
        data = open(path, "rb").read()
        if set(post) - {"sei_hash", "corrupt_hash"}:
            data = post_process(data, post)
        if "sei_hash" in post:
            from tools.make_streams import REF
            data = add_hash_sei(data, args["w"], args["h"], post["sei_hash"], post.get("corrupt_hash"))
            open(path, "wb").write(data)
            # the reference decoder must accept (or, for the corrupted variant, reject) the hashes
            r = subprocess.run([os.path.join(REF, "dec265"), "-q", "-c", path], capture_output=True, text=True)
            bad = "mismatch" in r.stderr or "SEI decoded picture hash:" in r.stderr
            assert bad == ("corrupt_hash" in post), "dec265 -c on %s: %s" % (name, r.stderr[-300:])
        open(path, "wb").write(data)
        args["post"] = post
    md5 = ref_md5(path, args["w"], args["h"])
    json.dump({"width": args["w"], "height": args["h"], "frames": len(md5), "generator": "tools/make_gen_streams.py",
               "args": args, "md5": md5}, open(os.path.join(out_dir, name + ".md5.json"), "w"), indent=0)
    return os.path.getsize(path), len(md5)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("names", nargs="*", default=[])
    a = ap.parse_args()
    cat = catalogue()
    total = 0
    for n in (a.names or cat):
        size, frames = generate(n, cat[n])
        total += size
        print(n, size, "bytes", frames, "frames", flush=True)
    print("total", total, "bytes")

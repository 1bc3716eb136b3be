#!/usr/bin/env python3
"""Re-write the PPS NAL units of an HEVC byte stream so that in-loop deblocking is switched on
(SURVEY.md §8c "closing the gaps" (1)).

enc265 writes deblocking_filter_control_present_flag=1, deblocking_filter_override_enabled_flag=0,
pps_deblocking_filter_disabled_flag=1 (encoder-context.cc:163-166), so its slice headers carry no
deblocking syntax and flipping the PPS flag alone yields a valid stream that every decoder deblocks.
Pure Python (bit-level PPS parse up to the flag, H.265 7.3.2.3); needs nothing from the reference, so
tests can patch the committed golden streams at run time.  Options allow non-zero beta/tc offsets and
pps_loop_filter_across_slices to be set for extra coverage.
"""
import sys


class BitR:
    def __init__(self, data):
        self.d, self.p = data, 0

    def u(self, n):
        v = 0
        for _ in range(n):
            v = (v << 1) | ((self.d[self.p >> 3] >> (7 - (self.p & 7))) & 1)
            self.p += 1
        return v

    def ue(self):
        z = 0
        while self.u(1) == 0:
            z += 1
        return (1 << z) - 1 + (self.u(z) if z else 0)

    def se(self):
        v = self.ue()
        return (v + 1) >> 1 if v & 1 else -(v >> 1)


class BitW:
    def __init__(self):
        self.bits = []

    def u(self, n, v):
        for i in range(n - 1, -1, -1):
            self.bits.append((v >> i) & 1)

    def ue(self, v):
        v += 1
        n = v.bit_length()
        self.u(n - 1, 0)
        self.u(n, v)

    def se(self, v):
        self.ue(2 * v - 1 if v > 0 else -2 * v)

    def copy(self, r, n):
        for _ in range(n):
            self.bits.append(r.u(1))

    def bytes(self):
        b = self.bits + [0] * (-len(self.bits) % 8)
        return bytes(sum(b[i + k] << (7 - k) for k in range(8)) for i in range(0, len(b), 8))


def unescape(b):
    out, z = bytearray(), 0
    for x in b:
        if z >= 2 and x == 3:
            z = 0
            continue
        out.append(x)
        z = z + 1 if x == 0 else 0
    return bytes(out)


def escape(b):
    out, z = bytearray(), 0
    for x in b:
        if z >= 2 and x <= 3:
            out.append(3)
            z = 0
        out.append(x)
        z = z + 1 if x == 0 else 0
    return bytes(out)


def patch_pps(rbsp, beta_div2=0, tc_div2=0, across_slices=None):
    """rbsp: PPS payload after the 2-byte NAL header (unescaped). Returns the patched payload."""
    r, w = BitR(rbsp), BitW()

    def cp_u(n):
        v = r.u(n); w.u(n, v); return v

    def cp_ue():
        v = r.ue(); w.ue(v); return v

    def cp_se():
        v = r.se(); w.se(v); return v

    cp_ue(); cp_ue()                      # pps id, sps id
    cp_u(1); cp_u(1); cp_u(3)             # dependent slices, output flag present, extra slice header bits
    cp_u(1); cp_u(1)                      # sign data hiding, cabac_init_present
    cp_ue(); cp_ue(); cp_se()             # num_ref_idx defaults, init_qp
    cp_u(1); cp_u(1)                      # constrained intra, transform skip
    if cp_u(1):                           # cu_qp_delta_enabled
        cp_ue()
    cp_se(); cp_se(); cp_u(1)             # cb/cr qp offsets, slice chroma qp offsets present
    cp_u(1); cp_u(1); cp_u(1)             # weighted pred, weighted bipred, transquant bypass
    tiles = cp_u(1); cp_u(1)              # tiles, entropy coding sync
    if tiles:
        cols, rows = cp_ue(), cp_ue()
        if not cp_u(1):
            for _ in range(cols):
                cp_ue()
            for _ in range(rows):
                cp_ue()
        cp_u(1)
    a = r.u(1)                            # pps_loop_filter_across_slices_enabled_flag
    w.u(1, a if across_slices is None else int(across_slices))
    present = r.u(1)
    w.u(1, 1)                             # deblocking_filter_control_present_flag
    override = 0
    if present:
        override = r.u(1)
        disabled = r.u(1)
        if not disabled:
            r.se(); r.se()
    w.u(1, override)
    w.u(1, 0)                             # pps_deblocking_filter_disabled_flag = 0
    w.se(beta_div2); w.se(tc_div2)
    # everything else up to the rbsp stop bit is copied verbatim
    total = len(rbsp) * 8
    last = len(rbsp) - 1
    while rbsp[last] == 0:
        last -= 1
    stop = last * 8 + 7 - ((rbsp[last] & -rbsp[last]).bit_length() - 1)
    w.copy(r, stop - r.p)
    w.u(1, 1)                             # rbsp_stop_one_bit (+ alignment zeros in bytes())
    assert stop < total
    return w.bytes()


def patch_stream(data, **kw):
    """Split on start codes, patch every PPS NAL (type 34), copy the rest byte for byte."""
    out, i, n = bytearray(), 0, len(data)
    starts = []
    while i + 2 < n:
        if data[i] == 0 and data[i + 1] == 0 and data[i + 2] == 1:
            starts.append(i + 3)
            i += 3
        else:
            i += 1
    for k, s in enumerate(starts):
        e = (starts[k + 1] - 3) if k + 1 < len(starts) else n
        nal = data[s:e]
        # a 4-byte start code leaves a trailing zero on the previous NAL: keep it outside the payload
        trail = 0
        while k + 1 < len(starts) and len(nal) > 2 and nal[-1] == 0:
            nal = nal[:-1]
            trail += 1
        out += b"\x00\x00\x01" if k or data[:3] == b"\x00\x00\x01" else b"\x00\x00\x00\x01"
        if ((nal[0] >> 1) & 0x3f) == 34:
            nal = nal[:2] + escape(patch_pps(unescape(nal[2:]), **kw))
        out += nal + b"\x00" * trail
    return bytes(out)


if __name__ == "__main__":
    src, dst = sys.argv[1], sys.argv[2]
    beta = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    tc = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    open(dst, "wb").write(patch_stream(open(src, "rb").read(), beta_div2=beta, tc_div2=tc))

"""TEST INFRASTRUCTURE: one picture's work lists reconstructed by the C oracle (liboracle.so).  Used by the CPU parity
tests and by bench.py's cpu_baseline fallback (kind "port") - never by the product path."""
import ctypes as C
import os
import subprocess
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None
FRAME_DEBLOCK, FRAME_SAO = 0x0008, 0x0010   # include/b265.h B265_FRAME_*


def load():
    global _lib
    if _lib is None:
        so = os.path.join(HERE, "_ref", "liboracle.so")
        if not os.path.exists(so):
            subprocess.run(["make", "-s", "-C", HERE, "oracle"], check=True)
        _lib = C.CDLL(so)
    return _lib


def oracle_planes(job, dpb=None):
    """reconstruct one picture with the C oracle; `dpb` maps the ids a parse-only decoder puts into
    ref_surfaces / dst_surface to the finished pictures (inter prediction)"""
    o = load()
    w, h = job.width, job.height
    planes = [np.zeros((h, w), np.uint8), np.zeros((h // 2, w // 2), np.uint8), np.zeros((h // 2, w // 2), np.uint8)]
    pp = (C.c_void_p * 3)(*[p.ctypes.data for p in planes])
    pitch = (C.c_int * 3)(w, w // 2, w // 2)
    refs = None
    if job.n_refs:
        # a reference the stream never delivered is the mid-grey picture of 8.3.3.2 (decctx.cc:1408-1434)
        grey = [np.full((h, w), 128, np.uint8), np.full((h // 2, w // 2), 128, np.uint8), np.full((h // 2, w // 2), 128, np.uint8)]
        pics = [dpb.setdefault(job.ref_surfaces[i], grey) for i in range(job.n_refs)]
        refs = (C.c_void_p * (3 * job.n_refs))(*[p.ctypes.data for pl in pics for p in pl])
    o.orc_reconstruct_frame(C.byref(job), pp, pitch, refs)
    if job.flags & FRAME_DEBLOCK:
        o.orc_deblock(C.byref(job), pp, pitch)
    if job.flags & FRAME_SAO:
        outs = [np.zeros_like(p) for p in planes]
        op = (C.c_void_p * 3)(*[p.ctypes.data for p in outs])
        o.orc_sao(C.byref(job), pp, op, pitch)
        planes = outs
    if dpb is not None:
        dpb[job.dst_surface] = planes
    return planes

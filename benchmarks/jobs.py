"""Turn a golden stream into self-contained b265_frame_job records (deep copies of the parser's work lists),
so the frame-level kernels can be timed with inputs resident in HBM (b265_upload_frames + b265_run_batch)."""
import ctypes as C
import os
import numpy as np
from libde265_b200 import lib as L
from libde265_b200.decoder import Decoder

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class OwnedJob:
    def __init__(self, J):
        self.keep = []
        n = L.FrameJob()
        C.memmove(C.byref(n), C.byref(J), C.sizeof(L.FrameJob))
        w, h = J.width, J.height
        nctb = ((w + (1 << J.log2_ctb) - 1) >> J.log2_ctb) * ((h + (1 << J.log2_ctb) - 1) >> J.log2_ctb)

        def dup(ptr, count, typ):
            if not ptr or count <= 0:
                return C.POINTER(typ)()
            arr = (typ * count)()
            C.memmove(arr, ptr, C.sizeof(typ) * count)
            self.keep.append(arr)
            return C.cast(arr, C.POINTER(typ))
        n.tus = dup(J.tus, J.n_tus, L.TU)
        n.coeffs = dup(J.coeffs, J.n_coeffs, L.Coeff)
        n.pus = dup(J.pus, J.n_pus, L.PU)
        n.pcms = dup(J.pcms, J.n_pcms, L.PCM)
        n.pcm_samples = dup(J.pcm_samples, J.n_pcm_samples, C.c_uint8)
        n.slices = dup(J.slices, J.n_slices, L.Slice)
        n.ctbs = dup(J.ctbs, nctb, L.CTB)
        n.sao = dup(J.sao, nctb, L.SAO)
        n.unit_info = dup(J.unit_info, (w >> 2) * (h >> 2), C.c_uint8)
        n.qp_y = dup(J.qp_y, (w >> J.log2_min_cb) * (h >> J.log2_min_cb), C.c_int8)
        n.scaling = dup(J.scaling, 1, L.Scaling)
        self.job = n
        self.bytes_in = (J.n_tus * C.sizeof(L.TU) + J.n_coeffs * C.sizeof(L.Coeff) + (w >> 2) * (h >> 2)
                         + (w >> J.log2_min_cb) * (h >> J.log2_min_cb) + nctb * (C.sizeof(L.CTB) + C.sizeof(L.SAO)))


def stream_jobs(name_or_bytes, threads=4, max_frames=None):
    data = name_or_bytes
    if isinstance(name_or_bytes, str):
        data = open(os.path.join(ROOT, "tests", "golden", "streams", name_or_bytes + ".bin"), "rb").read()
    dec = Decoder(parse_only=True, threads=threads)
    jobs = []

    def on_pic(img):
        if max_frames is None or len(jobs) < max_frames:
            jobs.append(OwnedJob(dec.job_of(img)))
    dec.decode_stream(data, chunk=1 << 20, on_picture=on_pic)
    dec.close()
    return jobs

"""ctypes binding of the C ABI in include/b265.h (libb265.so).

Host-side mirror used by the tests, bench.py and the Python decoder wrapper.  The library is the
product; this module only loads it and declares the record layouts.  If the CUDA library is missing
it raises — there is no CPU path behind these calls.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B265_LIB") or os.path.join(HERE, "libb265.so")   # B265_LIB: developer A/B runs

B265_OK = 0
MEM_HOST, MEM_DEVICE = 0, 1
TR_DCT, TR_DST, TR_SKIP, TR_BYPASS = 0, 1, 2, 3

TU_DST, TU_SKIP, TU_BYPASS, TU_RDPCM_H, TU_RDPCM_V, TU_ROTATE, TU_NO_EDGE_FILT, TU_INTRA_CU = \
    0x01, 0x02, 0x04, 0x08, 0x10, 0x20, 0x40, 0x80
INTRA_NONE = 0xFF
PU_L0, PU_L1 = 1, 2
UNIT_EDGE_V, UNIT_EDGE_H, UNIT_TU_EDGE_V, UNIT_TU_EDGE_H, UNIT_INTRA, UNIT_CBF, UNIT_NOFILT, UNIT_DONE = \
    0x01, 0x02, 0x04, 0x08, 0x10, 0x20, 0x40, 0x80
FRAME_STRONG_INTRA_SMOOTHING, FRAME_CONSTRAINED_INTRA, FRAME_LF_ACROSS_TILES, FRAME_DEBLOCK, \
    FRAME_SAO, FRAME_SCALING_LIST, FRAME_PCM_LF_DISABLE = 0x1, 0x2, 0x4, 0x8, 0x10, 0x20, 0x40


class Coeff(C.Structure):
    _fields_ = [("level", C.c_int16), ("pos", C.c_uint16)]


class TU(C.Structure):
    _fields_ = [("x", C.c_uint16), ("y", C.c_uint16), ("cidx_log2", C.c_uint8), ("flags", C.c_uint8),
                ("qp", C.c_uint8), ("intra_mode", C.c_uint8), ("coeff_off", C.c_uint32),
                ("ncoeff", C.c_uint16), ("reserved", C.c_uint16)]


class PU(C.Structure):
    _fields_ = [("x", C.c_uint16), ("y", C.c_uint16), ("w", C.c_uint8), ("h", C.c_uint8),
                ("flags", C.c_uint8), ("slice", C.c_uint8), ("mv", (C.c_int16 * 2) * 2),
                ("ref_idx", C.c_int8 * 2), ("ref_slot", C.c_uint8 * 2), ("reserved", C.c_uint32)]


class PCM(C.Structure):
    _fields_ = [("x", C.c_uint16), ("y", C.c_uint16), ("log2_size", C.c_uint8),
                ("reserved", C.c_uint8 * 3), ("sample_off", C.c_uint32)]


class SAO(C.Structure):
    _fields_ = [("type", C.c_uint8 * 3), ("band_eo", C.c_uint8 * 3), ("offset", (C.c_int8 * 4) * 3)]


class CTB(C.Structure):
    _fields_ = [("addr_ts", C.c_int32), ("slice", C.c_uint16), ("tile", C.c_uint16)]


class Slice(C.Structure):
    _fields_ = [("slice_addr_rs", C.c_int32), ("slice_type", C.c_int8), ("beta_offset", C.c_int8),
                ("tc_offset", C.c_int8), ("deblocking_disabled", C.c_uint8),
                ("lf_across_slices", C.c_uint8), ("sao_luma", C.c_uint8), ("sao_chroma", C.c_uint8),
                ("weighted", C.c_uint8), ("luma_log2_denom", C.c_uint8),
                ("chroma_log2_denom", C.c_uint8), ("pps_weighted_pred", C.c_uint8), ("pad", C.c_uint8),
                ("weight", ((C.c_int16 * 3) * 16) * 2), ("offset", ((C.c_int16 * 3) * 16) * 2)]


class Scaling(C.Structure):
    _fields_ = [("s4", (C.c_uint8 * 16) * 6), ("s8", (C.c_uint8 * 64) * 6),
                ("s16", (C.c_uint8 * 256) * 6), ("s32", (C.c_uint8 * 1024) * 6)]


class FrameJob(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("chroma_format_idc", C.c_uint8),
                ("bit_depth", C.c_uint8), ("log2_ctb", C.c_uint8), ("log2_min_cb", C.c_uint8),
                ("log2_min_tb", C.c_uint8), ("cb_qp_offset", C.c_int8), ("cr_qp_offset", C.c_int8),
                ("log2_sao_offset_scale_unused", C.c_uint8), ("flags", C.c_uint32),
                ("dst_surface", C.c_int32), ("ref_surfaces", C.c_int32 * 32), ("ref_poc", C.c_int32 * 32),
                ("n_refs", C.c_int32),
                ("tus", C.POINTER(TU)), ("n_tus", C.c_int32),
                ("coeffs", C.POINTER(Coeff)), ("n_coeffs", C.c_int32),
                ("pus", C.POINTER(PU)), ("n_pus", C.c_int32),
                ("pcms", C.POINTER(PCM)), ("n_pcms", C.c_int32),
                ("pcm_samples", C.POINTER(C.c_uint8)), ("n_pcm_samples", C.c_int32),
                ("slices", C.POINTER(Slice)), ("n_slices", C.c_int32),
                ("ctbs", C.POINTER(CTB)), ("sao", C.POINTER(SAO)),
                ("unit_info", C.POINTER(C.c_uint8)), ("qp_y", C.POINTER(C.c_int8)),
                ("scaling", C.POINTER(Scaling))]


_lib = None


class B265Error(RuntimeError):
    pass


def _check_host_isa():
    """The host parser is built for x86-64-v3 (build.py HOST_ISA): say so instead of dying with SIGILL."""
    if os.environ.get("B265_HOST_ISA") == "baseline":
        return
    try:
        flags = next(l for l in open("/proc/cpuinfo") if l.startswith("flags")).split()
    except (OSError, StopIteration):
        return
    missing = [f for f in ("avx2", "bmi2") if f not in flags]
    if missing:
        raise B265Error("this CPU lacks %s: rebuild with B265_HOST_ISA=baseline python -m libde265_b200.build --force"
                        % ", ".join(missing))


def load():
    """Load libb265.so (raises if it has not been built: python -m libde265_b200.build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B265Error("libb265.so not built (run `python -m libde265_b200.build`); "
                        "there is no CPU fallback for the b265 pixel path")
    _check_host_isa()
    lib = C.CDLL(LIB_PATH)
    vp, i32, u8p, i16p = C.c_void_p, C.c_int, C.POINTER(C.c_uint8), C.POINTER(C.c_int16)
    lib.b265_version.restype = C.c_char_p
    lib.b265_status_string.restype = C.c_char_p
    lib.b265_status_string.argtypes = [i32]
    lib.b265_last_error.restype = C.c_char_p
    lib.b265_last_error.argtypes = [vp]
    lib.b265_create.argtypes = [i32, C.POINTER(vp)]
    lib.b265_destroy.argtypes = [vp]
    lib.b265_destroy.restype = None
    lib.b265_sync.argtypes = [vp]
    lib.b265_stream_wait.argtypes = [vp, vp]
    lib.b265_launch_count.argtypes = [vp]
    lib.b265_launch_count.restype = C.c_uint64
    lib.b265_host_alloc.argtypes = [vp, C.c_size_t, C.POINTER(vp)]
    lib.b265_host_free.argtypes = [vp, vp]
    lib.b265_host_free.restype = None
    lib.b265_dev_alloc.argtypes = [vp, C.c_size_t, C.POINTER(vp)]
    lib.b265_dev_free.argtypes = [vp, vp]
    lib.b265_dev_free.restype = None
    lib.b265_memcpy_h2d.argtypes = [vp, vp, vp, C.c_size_t]
    lib.b265_memcpy_d2h.argtypes = [vp, vp, vp, C.c_size_t]
    lib.b265_timer_start.argtypes = [vp]
    lib.b265_timer_stop.argtypes = [vp, C.POINTER(C.c_float)]
    lib.b265_flush_l2.argtypes = [vp]
    lib.b265_surface_alloc.argtypes = [vp, i32, i32, i32, C.POINTER(i32)]
    lib.b265_surface_release.argtypes = [vp, i32]
    lib.b265_surface_fill.argtypes = [vp, i32, i32, i32]
    lib.b265_surface_upload.argtypes = [vp, i32, i32, vp, C.c_ssize_t]
    lib.b265_surface_download.argtypes = [vp, i32, i32, vp, C.c_ssize_t, i32]
    lib.b265_surface_plane.argtypes = [vp, i32, i32, C.POINTER(vp), C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]
    lib.b265_residual_batch.argtypes = [vp, i32, i32, i32, i32, i32, vp, vp, vp]
    for name, args in _OPTIONAL.items():
        if hasattr(lib, name):
            getattr(lib, name).argtypes = args
    _lib = lib
    return lib


_vp, _i32 = C.c_void_p, C.c_int
_OPTIONAL = {
    "b265_residual_sparse": [_vp, _i32, _vp, _i32, _vp, _i32, _vp, _i32, _i32, _i32],
    "b265_mc_batch": [_vp, _i32, _vp, _i32, _vp, _i32, _vp, _i32, _vp, _i32, _i32, _i32, _i32],
    "b265_interp_batch": [_vp, _i32, _i32, _i32, _i32, _i32, _vp, C.c_ssize_t, C.c_ssize_t, _vp, _vp],
    "b265_intra_batch": [_vp, _i32, _i32, _i32, _i32, _vp, _vp, _i32, _i32, _vp],
    "b265_deblock_planes": [_vp, _i32, _vp, _vp, _i32, _i32],
    "b265_sao_planes": [_vp, _i32, _vp, _vp, _vp, _i32, _i32],
    "b265_submit_frames": [_vp, _vp, _i32],
    "b265_upload_frames": [_vp, _vp, _i32, C.POINTER(_vp)],
    "b265_run_batch": [_vp, _vp],
    "b265_free_batch": [_vp, _vp],
    "b265_set_stage_timing": [_vp, _i32],
    "b265_get_stage_times": [_vp, C.POINTER(C.c_float * 6)],
}


def check(ctx, status, what=""):
    if status != B265_OK:
        lib = load()
        msg = lib.b265_status_string(status).decode()
        detail = lib.b265_last_error(ctx).decode() if ctx else ""
        raise B265Error("%s failed: %s (%s)" % (what or "b265 call", msg, detail))


class Context:
    """One decoder back end == one CUDA device + stream (b265_ctx)."""

    def __init__(self, device=0):
        self.lib = load()
        h = C.c_void_p()
        st = self.lib.b265_create(device, C.byref(h))
        if st != B265_OK:
            raise B265Error("b265_create(device=%d): %s — the b265 pixel path requires a CUDA device, "
                            "there is no CPU fallback" % (device, self.lib.b265_status_string(st).decode()))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.b265_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def sync(self):
        check(self.h, self.lib.b265_sync(self.h), "sync")

    def wait_for(self, other):
        """work submitted to this context from now on starts after everything submitted to `other` so far (device side)"""
        check(self.h, self.lib.b265_stream_wait(self.h, other.h), "stream_wait")

    @property
    def launches(self):
        return int(self.lib.b265_launch_count(self.h))

    # -- device memory helpers ------------------------------------------------------------------
    def dev_alloc(self, nbytes):
        p = C.c_void_p()
        check(self.h, self.lib.b265_dev_alloc(self.h, nbytes, C.byref(p)), "dev_alloc")
        return p

    def dev_free(self, p):
        self.lib.b265_dev_free(self.h, p)

    def h2d(self, dptr, arr):
        check(self.h, self.lib.b265_memcpy_h2d(self.h, dptr, arr.ctypes.data, arr.nbytes), "h2d")

    def d2h(self, arr, dptr):
        check(self.h, self.lib.b265_memcpy_d2h(self.h, arr.ctypes.data, dptr, arr.nbytes), "d2h")

    def timer_start(self):
        check(self.h, self.lib.b265_timer_start(self.h), "timer_start")

    def timer_stop(self):
        ms = C.c_float()
        check(self.h, self.lib.b265_timer_stop(self.h, C.byref(ms)), "timer_stop")
        return ms.value

    def flush_l2(self):
        check(self.h, self.lib.b265_flush_l2(self.h), "flush_l2")

    # -- kernel-level calls ---------------------------------------------------------------------
    def residual_batch(self, kind, log2, coeffs, dst, qp=None, rdpcm=0):
        """In-place on `dst` (numpy uint8 [n, nT, nT]); coeffs int16 [n, nT, nT] (host buffers)."""
        n = coeffs.shape[0]
        check(self.h, self.lib.b265_residual_batch(
            self.h, MEM_HOST, kind, log2, rdpcm, n, coeffs.ctypes.data,
            qp.ctypes.data if qp is not None else None, dst.ctypes.data), "residual_batch")
        return dst

    def residual_batch_dev(self, kind, log2, n, d_coeffs, d_dst, d_qp=None, rdpcm=0):
        check(self.h, self.lib.b265_residual_batch(self.h, MEM_DEVICE, kind, log2, rdpcm, n, d_coeffs,
                                                   d_qp, d_dst), "residual_batch(dev)")

    def mc_batch(self, pus, slices, refs, dst, width, height):
        """pus: ctypes array of PU; slices: ctypes array of Slice; refs: list of (Y,Cb,Cr) numpy
        planes (contiguous, pitch = row length); dst: (Y,Cb,Cr) numpy planes, modified in place."""
        import numpy as np
        n_refs = len(refs)
        rp = (C.c_void_p * (3 * n_refs))(*[p.ctypes.data for r in refs for p in r])
        dp = (C.c_void_p * 3)(*[p.ctypes.data for p in dst])
        check(self.h, self.lib.b265_mc_batch(self.h, MEM_HOST, C.byref(pus), len(pus), C.byref(slices), len(slices),
                                             rp, n_refs, dp, width, height, dst[0].shape[1], dst[1].shape[1]),
              "mc_batch")
        return dst

    def residual_sparse(self, tus, coeffs, plane):
        """tus: ctypes array of TU addressing `plane` (numpy uint8 2-D, modified in place); coeffs: ctypes array of Coeff."""
        h, w = plane.shape
        check(self.h, self.lib.b265_residual_sparse(self.h, MEM_HOST, C.byref(tus), len(tus),
                                                    C.byref(coeffs) if len(coeffs) else None, len(coeffs),
                                                    plane.ctypes.data, plane.strides[0], w, h), "residual_sparse")
        return plane

    def intra_batch(self, log2, cidx, modes, borders, strong=0, no_edge_filter=0):
        """modes uint8 [n]; borders uint8 [n, 4*nT+1] (p[-2nT..2nT]); returns uint8 [n, nT, nT]."""
        import numpy as np
        n, nT = len(modes), 1 << log2
        out = np.empty((n, nT, nT), dtype=np.uint8)
        check(self.h, self.lib.b265_intra_batch(self.h, MEM_HOST, log2, cidx, n, modes.ctypes.data, borders.ctypes.data,
                                                strong, no_edge_filter, out.ctypes.data), "intra_batch")
        return out

    def interp_batch(self, luma, w, h, buf, origin, stride, block_stride, n, frac_xy):
        """buf: uint8 numpy holding n blocks with halo; block i's integer position is at flat offset
        origin + i*block_stride (row stride `stride`); frac_xy uint8 [n, 2]. Returns int16 [n, h, w]."""
        import numpy as np
        out = np.empty((n, h, w), dtype=np.int16)
        check(self.h, self.lib.b265_interp_batch(self.h, MEM_HOST, luma, n, w, h, buf.ctypes.data + origin, stride,
                                                 block_stride, frac_xy.ctypes.data, out.ctypes.data), "interp_batch")
        return out

    def deblock_planes(self, job, planes):
        """In-place deblocking of numpy planes (Y[,Cb,Cr]) with the metadata of FrameJob `job`."""
        pp = (C.c_void_p * 3)(*([p.ctypes.data for p in planes] + [None] * (3 - len(planes))))
        check(self.h, self.lib.b265_deblock_planes(self.h, MEM_HOST, C.byref(job), pp, planes[0].strides[0],
                                                   planes[1].strides[0] if len(planes) > 1 else 0), "deblock_planes")
        return planes

    def sao_planes(self, job, in_planes):
        import numpy as np
        out = [np.zeros_like(p) for p in in_planes]
        ip = (C.c_void_p * 3)(*([p.ctypes.data for p in in_planes] + [None] * (3 - len(in_planes))))
        op = (C.c_void_p * 3)(*([p.ctypes.data for p in out] + [None] * (3 - len(out))))
        check(self.h, self.lib.b265_sao_planes(self.h, MEM_HOST, C.byref(job), ip, op, in_planes[0].strides[0],
                                               in_planes[1].strides[0] if len(in_planes) > 1 else 0), "sao_planes")
        return out

    # -- surfaces + frame-level path ------------------------------------------------------------
    def surface_alloc(self, width, height, chroma=1):
        sid = C.c_int()
        check(self.h, self.lib.b265_surface_alloc(self.h, width, height, chroma, C.byref(sid)), "surface_alloc")
        return sid.value

    def surface_release(self, sid):
        check(self.h, self.lib.b265_surface_release(self.h, sid), "surface_release")

    def surface_fill(self, sid, luma, chroma):
        check(self.h, self.lib.b265_surface_fill(self.h, sid, luma, chroma), "surface_fill")

    def surface_upload(self, sid, planes):
        for c, p in enumerate(planes):
            check(self.h, self.lib.b265_surface_upload(self.h, sid, c, p.ctypes.data, p.strides[0]), "surface_upload")
        self.sync()

    def surface_download(self, sid, width, height, chroma=1):
        import numpy as np
        shapes = [(height, width)] + ([((height + 1) // 2, (width + 1) // 2)] * 2 if chroma else [])
        out = [np.empty(s, dtype=np.uint8) for s in shapes]
        for c, p in enumerate(out):
            check(self.h, self.lib.b265_surface_download(self.h, sid, c, p.ctypes.data, p.strides[0], 0), "download")
        self.sync()
        return out

    def submit_frames(self, jobs):
        arr = (FrameJob * len(jobs))(*jobs)
        check(self.h, self.lib.b265_submit_frames(self.h, C.byref(arr), len(jobs)), "submit_frames")

    def upload_frames(self, jobs):
        arr = (FrameJob * len(jobs))(*jobs)
        b = C.c_void_p()
        check(self.h, self.lib.b265_upload_frames(self.h, C.byref(arr), len(jobs), C.byref(b)), "upload_frames")
        return b

    def run_batch(self, batch):
        check(self.h, self.lib.b265_run_batch(self.h, batch), "run_batch")

    def free_batch(self, batch):
        self.lib.b265_free_batch(self.h, batch)

    def stage_times(self):
        t = (C.c_float * 6)()
        check(self.h, self.lib.b265_get_stage_times(self.h, C.byref(t)), "stage_times")
        return list(t)


def as_ptr(arr, typ):
    """numpy array (or ctypes array) -> ctypes POINTER(typ); caller keeps `arr` alive."""
    if arr is None:
        return C.POINTER(typ)()
    if hasattr(arr, "ctypes"):
        return C.cast(arr.ctypes.data, C.POINTER(typ))
    return C.cast(arr, C.POINTER(typ))

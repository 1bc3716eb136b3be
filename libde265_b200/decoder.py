"""ctypes mirror of the libde265 decoder API exported by libb265.so (include/de265.h).

`Decoder` drives the same calls dec265 makes (dec265/dec265.cc:620-800): push_data -> decode ->
get_next_picture.  Pictures are reconstructed on the GPU by the b265 kernels; there is no CPU pixel
path (parse_only=True builds only the work lists and is used by the CPU tests together with the
oracle, and to measure host parsing speed).
"""
import ctypes as C
import numpy as np
from . import lib as L

DE265_OK = 0
DE265_ERROR_IMAGE_BUFFER_FULL = 9
DE265_ERROR_WAITING_FOR_INPUT_DATA = 13
DE265_ERROR_NOT_IMPLEMENTED_YET = 502
PARAM_SEI_CHECK_HASH, PARAM_DISABLE_DEBLOCKING, PARAM_DISABLE_SAO = 0, 7, 8

_bound = False


def _bind():
    global _bound
    lib = L.load()
    if _bound:
        return lib
    vp, i32 = C.c_void_p, C.c_int
    lib.de265_new_decoder.restype = vp
    lib.de265_free_decoder.argtypes = [vp]
    lib.de265_start_worker_threads.argtypes = [vp, i32]
    lib.de265_push_data.argtypes = [vp, vp, i32, C.c_int64, vp]
    lib.de265_push_NAL.argtypes = [vp, vp, i32, C.c_int64, vp]
    lib.de265_flush_data.argtypes = [vp]
    lib.de265_decode.argtypes = [vp, C.POINTER(i32)]
    lib.de265_reset.argtypes = [vp]
    lib.de265_reset.restype = None
    lib.de265_get_next_picture.argtypes = [vp]
    lib.de265_get_next_picture.restype = vp
    lib.de265_peek_next_picture.argtypes = [vp]
    lib.de265_peek_next_picture.restype = vp
    lib.de265_release_next_picture.argtypes = [vp]
    lib.de265_release_next_picture.restype = None
    lib.de265_get_warning.argtypes = [vp]
    lib.de265_get_image_width.argtypes = [vp, i32]
    lib.de265_get_image_height.argtypes = [vp, i32]
    lib.de265_get_chroma_format.argtypes = [vp]
    lib.de265_get_bits_per_pixel.argtypes = [vp, i32]
    lib.de265_get_image_plane.argtypes = [vp, i32, C.POINTER(i32)]
    lib.de265_get_image_plane.restype = vp
    lib.de265_get_image_PTS.argtypes = [vp]
    lib.de265_get_image_PTS.restype = C.c_int64
    lib.de265_set_parameter_bool.argtypes = [vp, i32, i32]
    lib.de265_set_parameter_bool.restype = None
    lib.de265_get_error_text.argtypes = [i32]
    lib.de265_get_error_text.restype = C.c_char_p
    lib.b265_de265_set_device.argtypes = [i32]
    lib.b265_de265_set_device.restype = None
    lib.b265_de265_set_parse_only.argtypes = [i32]
    lib.b265_de265_set_parse_only.restype = None
    lib.b265_de265_set_batch.argtypes = [vp, i32]
    lib.b265_de265_set_batch.restype = None
    lib.b265_de265_get_stats.argtypes = [vp, C.POINTER(C.c_double * 8)]
    lib.b265_de265_get_stats.restype = None
    lib.b265_de265_image_job.argtypes = [vp, vp, C.POINTER(L.FrameJob)]
    lib.b265_de265_image_poc.argtypes = [vp]
    lib.b265_de265_image_starts_cvs.argtypes = [vp]
    lib.b265_de265_hashes_checked.argtypes = [vp]
    lib.b265_de265_hashes_checked.restype = C.c_uint64
    lib.de265_get_image_NAL_header.argtypes = [vp, C.POINTER(C.c_int), vp, vp, vp]
    lib.de265_get_image_NAL_header.restype = None
    _bound = True
    return lib


class De265Error(RuntimeError):
    def __init__(self, code, text):
        RuntimeError.__init__(self, "de265 error %d: %s" % (code, text))
        self.code = code


class Decoder:
    def __init__(self, device=0, threads=0, batch=8, parse_only=False, disable_deblocking=False, disable_sao=False,
                 check_hash=False):
        self.lib = _bind()
        self.lib.b265_de265_set_device(device)
        self.lib.b265_de265_set_parse_only(1 if parse_only else 0)
        self.h = self.lib.de265_new_decoder()
        self.lib.b265_de265_set_parse_only(0)
        if not self.h:
            raise De265Error(11, "de265_new_decoder failed")
        if threads > 0:
            self.lib.de265_start_worker_threads(self.h, threads)
        self.lib.b265_de265_set_batch(self.h, batch)
        if disable_deblocking:
            self.lib.de265_set_parameter_bool(self.h, PARAM_DISABLE_DEBLOCKING, 1)
        if disable_sao:
            self.lib.de265_set_parameter_bool(self.h, PARAM_DISABLE_SAO, 1)
        if check_hash:   # dec265 -c: verify the decoded-picture-hash SEI of every output picture (on the device)
            self.lib.de265_set_parameter_bool(self.h, PARAM_SEI_CHECK_HASH, 1)
        self.parse_only = parse_only

    def close(self):
        if getattr(self, "h", None):
            self.lib.de265_free_decoder(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _err(self, code):
        raise De265Error(code, self.lib.de265_get_error_text(code).decode())

    def _image(self, img, copy=True):
        planes = []
        for c in range(3):
            w, h = self.lib.de265_get_image_width(img, c), self.lib.de265_get_image_height(img, c)
            if w <= 0 or h <= 0:
                continue
            stride = C.c_int()
            p = self.lib.de265_get_image_plane(img, c, C.byref(stride))
            if not p:
                continue
            a = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(h, stride.value))[:, :w]
            planes.append(a.copy() if copy else a)
        return planes

    def decode_stream(self, data, chunk=1 << 16, on_picture=None, copy=True):
        """Feed a byte stream the way dec265 does (chunked push_data, decode until stalled, drain
        pictures).  Returns the list of pictures (each a list of numpy planes) unless on_picture is
        given, in which case it is called with (image_handle) for every output picture."""
        lib, h = self.lib, self.h
        out = []
        mv = memoryview(data)
        pos = 0

        def drain_one():
            img = lib.de265_get_next_picture(h)
            if not img:
                return False
            if on_picture:
                on_picture(img)
            else:
                out.append(self._image(img, copy))
            return True

        more = C.c_int(1)
        finished = False
        while not finished:
            if pos < len(mv):
                n = min(chunk, len(mv) - pos)
                buf = (C.c_uint8 * n).from_buffer_copy(mv[pos:pos + n])
                lib.de265_push_data(h, buf, n, pos, None)
                pos += n
            if pos >= len(mv):
                lib.de265_flush_data(h)
            while True:
                err = lib.de265_decode(h, C.byref(more))
                got = drain_one()
                if err == DE265_ERROR_WAITING_FOR_INPUT_DATA:
                    break
                if err == DE265_ERROR_IMAGE_BUFFER_FULL:
                    while drain_one():
                        pass
                    continue
                if err != DE265_OK:
                    self._err(err)
                if not more.value and not got:
                    finished = pos >= len(mv)
                    break
        while drain_one():
            pass
        return out

    def job_of(self, img):
        """(parse-only decoders) the b265_frame_job of an output picture; valid until the next call."""
        J = L.FrameJob()
        err = self.lib.b265_de265_image_job(self.h, img, C.byref(J))
        if err:
            self._err(err)
        return J

    def stats(self):
        s = (C.c_double * 8)()
        self.lib.b265_de265_get_stats(self.h, C.byref(s))
        return {"pictures": int(s[0]), "parse_s": s[1], "gpu_s": s[2], "h2d_bytes": s[3], "d2h_bytes": s[4],
                "launches": int(s[5]), "batches": int(s[6]), "submit_s": s[7]}

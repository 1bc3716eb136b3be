"""b265_plane_crc_noalloc (include/b265.h): the CRC picture hash for callers that cannot allocate - its scratch is a static
device array, nothing is allocated inside the call.  Checked through the C ABI against the existing entry point
(b265_surface_hash, hash_type 1 = CRC) on two calls in a row; integer work, so bit-exact.  Collected last on purpose:
it must pass alone in a fresh process and after every other test."""
import ctypes as C
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
MAX_ROWS = 8192          # B265_CRC_NOALLOC_MAX_ROWS
ERR_UNSUPPORTED = -5     # B265_ERR_UNSUPPORTED


def _bind(lib):
    vp, i32 = C.c_void_p, C.c_int
    lib.b265_plane_crc_noalloc.argtypes = [vp, vp, i32, i32, i32, vp]
    lib.b265_plane_crc_noalloc.restype = i32
    lib.b265_surface_hash.argtypes = [vp, C.POINTER(i32), i32, i32, vp]
    lib.b265_surface_hash.restype = i32
    lib.b265_last_error.argtypes = [vp]
    lib.b265_last_error.restype = C.c_char_p


def _plane(ctx, sid, c):
    p, pitch, w, h = C.c_void_p(), C.c_int(), C.c_int(), C.c_int()
    assert ctx.lib.b265_surface_plane(ctx.h, sid, c, C.byref(p), C.byref(pitch), C.byref(w), C.byref(h)) == 0
    return p, pitch.value, w.value, h.value


def test_noalloc_crc_matches_surface_hash_twice_in_a_row():
    from libde265_b200 import lib as L
    ctx = L.Context(0)
    _bind(ctx.lib)
    rng = np.random.default_rng(11)
    d_out = ctx.dev_alloc(16)          # the CALLER's output buffer: the entry point itself allocates nothing
    # two calls in a row on different pictures (the shared scratch keeps the first call's rows: the second, smaller
    # picture must not see them), then the first picture again
    pictures = []
    for w, h in ((1920, 1080), (416, 240)):
        planes = [rng.integers(0, 256, size=(h, w), dtype=np.uint8), rng.integers(0, 256, size=(h // 2, w // 2), dtype=np.uint8),
                  rng.integers(0, 256, size=(h // 2, w // 2), dtype=np.uint8)]
        sid = ctx.surface_alloc(w, h)
        ctx.surface_upload(sid, planes)
        want = np.zeros(48, np.uint8)
        ids = (C.c_int * 1)(sid)
        assert ctx.lib.b265_surface_hash(ctx.h, ids, 1, 1, want.ctypes.data) == 0
        ctx.sync()
        pictures.append((sid, [int(want[16 * c:16 * c + 4].view("<u4")[0]) for c in range(3)]))
    for sid, want in pictures + pictures[:1]:
        for c in range(3):
            p, pitch, w, h = _plane(ctx, sid, c)
            got = np.zeros(1, np.uint32)
            assert ctx.lib.b265_plane_crc_noalloc(ctx.h, p, pitch, w, h, d_out) == 0
            ctx.d2h(got, d_out)
            assert int(got[0]) == want[c], "plane %d of surface %d: %04x, b265_surface_hash says %04x" % (c, sid, int(got[0]), want[c])
    # beyond the static scratch: a clear error, no launch
    p, pitch, w, h = _plane(ctx, pictures[0][0], 0)
    launches = ctx.launches
    assert ctx.lib.b265_plane_crc_noalloc(ctx.h, p, pitch, w, MAX_ROWS + 1, d_out) == ERR_UNSUPPORTED
    assert b"B265_CRC_NOALLOC_MAX_ROWS" in ctx.lib.b265_last_error(ctx.h)
    assert ctx.launches == launches
    for sid, _ in pictures:
        ctx.surface_release(sid)
    ctx.close()

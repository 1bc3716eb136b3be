"""The reference's kernel-level plugin surface (libde265/acceleration.h:29-201) filled by
init_acceleration_functions_cuda() (include/b265_acceleration.h), checked slot by slot against the table
init_acceleration_functions_fallback() of the compiled reference fills - the `acceleration-speed --check`
pattern (acceleration-speed/acceleration-speed.cc:71-94): same random block through both tables, bit-exact."""
import ctypes as C
import numpy as np
import pytest
from tests import helpers

N_SLOTS = 85          # data members of struct acceleration_functions, in declaration order
# slot indices (acceleration.h declaration order)
AVG8, UNW8, W8, BI8 = 0, 1, 2, 3
EPEL, EPEL_H, EPEL_V, EPEL_HV = 8, 9, 10, 11
QPEL = 12             # [x][y] -> 12 + 4 * x + y
SKIP8, DST8, DCT8 = 51, 54, 55
vp, i32, pd = C.c_void_p, C.c_int, C.c_ssize_t


def ptr(a):
    return a.ctypes.data_as(vp)


def ref_table():
    ref = helpers.load_ref()
    if ref is None:
        pytest.skip("oracle/_ref/libde265_ref.so not built")
    t = (vp * N_SLOTS)()
    ref._Z36init_acceleration_functions_fallbackP22acceleration_functions(C.byref(t))
    return ref, t


def test_slot_layout_matches_the_reference():
    """the indices used here (and the struct in b265_acceleration.h) name the slots the reference's initialiser fills"""
    ref, t = ref_table()
    addr = lambda name: C.cast(getattr(ref, name), vp).value
    assert t[DST8] == addr("_Z33transform_4x4_luma_add_8_fallbackPhPKsl")
    assert t[DCT8 + 3] == addr("_Z30transform_32x32_add_8_fallbackPhPKsl")
    assert t[UNW8] == addr("_Z30put_unweighted_pred_8_fallbackPhlPKslii")
    assert t[EPEL] == addr("_Z19put_epel_8_fallbackPslPKhliiiiS_")
    assert t[QPEL + 4 * 1 + 2] == addr("_Z21put_qpel_1_2_fallbackPslPKhliiS_")
    assert t[N_SLOTS - 1] is not None          # hadamard_transform_8[3]: the table ends where the struct ends


def test_init_without_gpu_fails_loudly_and_touches_nothing():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from libde265_b200 import lib as L
    lib = L.load()
    t = (vp * N_SLOTS)()
    assert lib.init_acceleration_functions_cuda(C.byref(t)) == -1   # B265_ERR_NO_DEVICE
    assert all(x is None for x in t)


@pytest.fixture(scope="module")
def tables():
    from libde265_b200 import lib as L
    lib = L.load()
    ref, tr = ref_table()
    tc = (vp * N_SLOTS)()
    ref._Z36init_acceleration_functions_fallbackP22acceleration_functions(C.byref(tc))   # decctx.cc:430-449 order
    assert lib.init_acceleration_functions_cuda(C.byref(tc)) == 0
    lib.b265_acceleration_launch_count.restype = C.c_uint64
    return lib, tr, tc


def both(tables, slot, proto, make_args):
    """call slot `slot` of both tables with independently built arguments; returns the two output arrays"""
    lib, tr, tc = tables
    outs = []
    for t in (tr, tc):
        assert t[slot], "slot %d empty" % slot
        args, out = make_args()
        C.CFUNCTYPE(None, *proto)(t[slot])(*args)
        outs.append(out)
    assert tr[slot] != tc[slot], "slot %d was not overridden by the CUDA initialiser" % slot
    return outs


@pytest.mark.gpu
@pytest.mark.parametrize("slot,log2", [(DST8, 2), (DCT8, 2), (DCT8 + 1, 3), (DCT8 + 2, 4), (DCT8 + 3, 5)])
def test_transform_add_slots(tables, slot, log2):
    rng = np.random.default_rng(slot)
    n = 1 << log2
    for it in range(6):
        coeffs = helpers.make_coeff_blocks(rng, 1, log2)[0].astype(np.int16)
        if it == 0:
            coeffs[:] = rng.integers(-32768, 32768, size=coeffs.shape)   # saturation of both passes
        stride = n + 16 * (it % 2)
        dst0 = rng.integers(0, 256, size=(n, stride), dtype=np.uint8)

        def mk():
            d = dst0.copy()
            return (ptr(d), ptr(coeffs), stride), d
        a, b = both(tables, slot, (vp, vp, pd), mk)
        assert np.array_equal(a, b)


@pytest.mark.gpu
def test_transform_skip_slot(tables):
    """transform_skip_8 is deprecated in the reference (its fallback asserts): checked against the formula it holds,
    fallback-dct.cc:46-60"""
    lib, tr, tc = tables
    rng = np.random.default_rng(7)
    coeffs = rng.integers(-600, 600, size=16).astype(np.int16)
    dst = rng.integers(0, 256, size=(4, 20), dtype=np.uint8)
    want = dst.copy()
    r = ((coeffs.astype(np.int32).reshape(4, 4) << 7) + (1 << 11)) >> 12
    want[:, :4] = np.clip(dst[:, :4].astype(np.int32) + r, 0, 255)
    C.CFUNCTYPE(None, vp, vp, pd)(tc[SKIP8])(ptr(dst), ptr(coeffs), 20)
    assert np.array_equal(dst, want)


@pytest.mark.gpu
def test_qpel_slots(tables):
    rng = np.random.default_rng(3)
    for fx in range(4):
        for fy in range(4):
            w, h = [(8, 8), (16, 4), (64, 64), (12, 16)][(fx + fy) % 4]
            img = rng.integers(0, 256, size=(h + 7, w + 7 + 9), dtype=np.uint8)
            sstride = img.shape[1]
            scratch = np.zeros(80 * 80, np.int16)

            def mk():
                out = np.full((h, w + 8), -1, np.int16)
                src = img.ctypes.data + 3 * sstride + 3
                return (ptr(out), w + 8, vp(src), sstride, w, h, ptr(scratch)), out
            a, b = both(tables, QPEL + 4 * fx + fy, (vp, pd, vp, pd, i32, i32, vp), mk)
            assert np.array_equal(a, b), (fx, fy)


@pytest.mark.gpu
def test_epel_slots(tables):
    rng = np.random.default_rng(4)
    scratch = np.zeros(80 * 80, np.int16)
    for slot, fracs in ((EPEL, [(0, 0)]), (EPEL_H, [(3, 0), (7, 0)]), (EPEL_V, [(0, 5), (0, 1)]), (EPEL_HV, [(2, 6), (7, 7), (4, 4)])):
        for mx, my in fracs:
            w, h = 16, 8
            img = rng.integers(0, 256, size=(h + 3, w + 3 + 5), dtype=np.uint8)
            sstride = img.shape[1]

            def mk():
                out = np.full((h, w), -1, np.int16)
                src = img.ctypes.data + sstride + 1
                if slot == EPEL:
                    return (ptr(out), w, vp(src), sstride, w, h, mx, my, ptr(scratch)), out
                return (ptr(out), w, vp(src), sstride, w, h, mx, my, ptr(scratch), 8), out
            proto = (vp, pd, vp, pd, i32, i32, i32, i32, vp) + (() if slot == EPEL else (i32,))
            a, b = both(tables, slot, proto, mk)
            assert np.array_equal(a, b), (slot, mx, my)


@pytest.mark.gpu
def test_writer_slots(tables):
    rng = np.random.default_rng(5)
    w, h, ss, ds = 16, 8, 24, 32
    s1 = rng.integers(-2000, 16000, size=(h, ss)).astype(np.int16)
    s2 = rng.integers(-2000, 16000, size=(h, ss)).astype(np.int16)

    def run(slot, proto, extra, two):
        def mk():
            d = np.full((h, ds), 77, np.uint8)
            srcs = (ptr(s1), ptr(s2)) if two else (ptr(s1),)
            return (ptr(d), ds) + srcs + (ss, w, h) + extra, d
        a, b = both(tables, slot, proto, mk)
        assert np.array_equal(a, b), slot
    run(UNW8, (vp, pd, vp, pd, i32, i32), (), False)
    run(AVG8, (vp, pd, vp, vp, pd, i32, i32), (), True)
    for wt, o, l2 in ((64, 0, 7), (-37, 12, 6), (127, -128, 13), (90, 5, 1)):
        run(W8, (vp, pd, vp, pd, i32, i32, i32, i32, i32), (wt, o, l2), False)
    for w1, o1, w2, o2, l2 in ((64, 0, 64, 0, 7), (-30, 10, 100, -20, 8), (127, 127, -128, -128, 6)):
        run(BI8, (vp, pd, vp, vp, pd, i32, i32, i32, i32, i32, i32, i32), (w1, o1, w2, o2, l2), True)
    lib = tables[0]
    assert lib.b265_acceleration_launch_count() > 0

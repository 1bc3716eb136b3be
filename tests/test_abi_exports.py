"""The C-ABI library loads without a GPU and exports every entry point include/*.h declares
(no compute calls here)."""
import ctypes as C
import os, re
from tests.helpers import ROOT


def declared(header, macro):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
    src = re.sub(r"//[^\n]*", " ", src)
    src = re.sub(r"#\s*define[^\n]*", " ", src)
    names = re.findall(macro + r"\s+[^;{()]*?\b(\w+)\s*\(", src)
    return sorted(set(names))


def test_every_declared_symbol_is_exported():
    from libde265_b200 import lib as L
    lib = C.CDLL(L.LIB_PATH)
    b = declared("b265.h", "B265_API")
    d = declared("de265.h", "LIBDE265_API")
    assert len(b) >= 30 and "b265_submit_frames" in b and "b265_intra_batch" in b
    assert len(d) >= 40 and "de265_decode" in d and "de265_get_next_picture" in d
    missing = [n for n in b + d if not hasattr(lib, n)]
    assert not missing, "declared in include/*.h but not exported by libb265.so: %s" % missing


def test_version_and_error_strings():
    from libde265_b200 import lib as L
    lib = L.load()
    assert lib.b265_version().decode().count(".") >= 1
    assert lib.b265_status_string(0).decode()
    lib.de265_get_version.restype = C.c_char_p
    lib.de265_get_error_text.restype = C.c_char_p
    lib.de265_get_error_text.argtypes = [C.c_int]
    assert lib.de265_get_version().decode()
    assert "data" in lib.de265_get_error_text(13).decode().lower() or lib.de265_get_error_text(13)


def test_create_without_gpu_fails_loudly():
    """no CPU fallback: on a machine without a CUDA device b265_create reports an error"""
    import torch
    if torch.cuda.is_available():
        return
    from libde265_b200 import lib as L
    import pytest
    with pytest.raises(L.B265Error):
        L.Context(0)

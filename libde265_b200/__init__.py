"""libde265_b200 — B200-native (sm_100a) reconstruction back end for libde265's HEVC decoder.

The product is the C-ABI library `libb265.so` (include/b265.h, include/b265_acceleration.h,
include/de265.h).  This package builds it (`python -m libde265_b200.build`) and mirrors the
reference's host interfaces in Python for tests and benchmarks.
"""
from . import lib  # noqa: F401

__all__ = ["lib"]

"""The host parser is compiled for x86-64-v3 (libde265_b200/build.py HOST_ISA); the loader must say so on a CPU
without AVX2/BMI2 instead of dying with SIGILL, and must not object on this machine."""
import builtins
import io
import pytest
from libde265_b200 import lib as L


def test_host_isa_check_passes_here():
    L._check_host_isa()   # the library was built and loaded on this CPU by the other tests


def test_host_isa_check_reports_missing_flags(monkeypatch):
    real_open = builtins.open

    def fake_open(path, *a, **k):
        if path == "/proc/cpuinfo":
            return io.StringIO("processor\t: 0\nflags\t\t: fpu sse4_1 sse4_2 popcnt\n")
        return real_open(path, *a, **k)

    monkeypatch.delenv("B265_HOST_ISA", raising=False)
    monkeypatch.setattr(builtins, "open", fake_open)
    with pytest.raises(L.B265Error) as e:
        L._check_host_isa()
    assert "avx2" in str(e.value) and "bmi2" in str(e.value)


def test_host_isa_baseline_build_skips_the_check(monkeypatch):
    monkeypatch.setenv("B265_HOST_ISA", "baseline")
    L._check_host_isa()

"""Build libb265.so (CUDA kernels + C ABI + host parser) in-tree with nvcc for sm_100a.

    python -m libde265_b200.build [--force]

The .so lands next to this file so it travels to the GPU box with the repo snapshot.
Incremental: one object per source, rebuilt when the source or any header is newer.
"""
import os, subprocess, sys, glob
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libb265.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-std=c++17", "-O3", "-lineinfo", "-Xcompiler", "-fPIC,-fvisibility=hidden,-Wall,-Wno-unused-function"] + \
    os.environ.get("B265_NVCC_FLAGS", "").split()   # tuning experiments, e.g. -DINTRA_MIN_CTAS=6


# x86-64-v3 (Haswell / Zen and later: AVX2, BMI2, LZCNT) for the host-only sources; lib.py checks the CPU before loading
HOST_ISA = [] if os.environ.get("B265_HOST_ISA") == "baseline" else ["-Xcompiler", "-march=x86-64-v3"]


def _sources():
    s = sorted(glob.glob(os.path.join(CSRC, "*.cu")) + glob.glob(os.path.join(CSRC, "*.cc"))
               + glob.glob(os.path.join(CSRC, "host", "*.cc")))
    return s


def _headers():
    return (glob.glob(os.path.join(CSRC, "*.h")) + glob.glob(os.path.join(CSRC, "*.cuh"))
            + glob.glob(os.path.join(CSRC, "host", "*.h"))
            + glob.glob(os.path.join(HERE, "..", "include", "*.h")))


def _compile(src, obj):
    cmd = [NVCC] + ARCH + COMMON + ["-c", src, "-o", obj]
    if src.endswith(".cc"):   # host parser: BMI2 / LZCNT shorten the CABAC engine's shift and count chains (-6 % parse time)
        cmd = [NVCC] + COMMON + HOST_ISA + ["-x", "cu"] + ARCH + ["-c", src, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
    return r.stderr


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    hdr_m = max([os.path.getmtime(h) for h in _headers()] + [os.path.getmtime(__file__)])
    jobs, objs = [], []
    for src in _sources():
        obj = os.path.join(OBJ, os.path.relpath(src, CSRC).replace(os.sep, "_") + ".o")
        objs.append(obj)
        if force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(src), hdr_m):
            jobs.append((src, obj))
    if jobs:
        with ThreadPoolExecutor(8) as ex:
            for msg in ex.map(lambda a: _compile(*a), jobs):
                if verbose and msg:
                    print(msg)
    if jobs or not os.path.exists(LIB):
        cmd = [NVCC] + ARCH + ["-shared", "-o", LIB] + objs + ["-lcudart", "-lpthread"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))

"""TEST-ONLY host emulation of the kernels: the CUDA sources in pace_b200/csrc compiled by g++ with
-DPACE_HOST_EMU (launch helpers become loops).  Lets the kernel *logic* be checked against the golden
fixtures on the CPU-only build machine; the product never loads this library and GPU parity is
established separately by the `-m gpu` tests through the real CUDA library."""
import ctypes
import glob
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "_hostemu", "libpace_b200_hostemu.so")
_cached = None


def build():
    srcs = sorted(glob.glob(os.path.join(ROOT, "pace_b200", "csrc", "*.cu")))
    deps = srcs + glob.glob(os.path.join(ROOT, "pace_b200", "csrc", "*.[ch]*")) + [os.path.join(ROOT, "include", "pace_b200.h")]
    if os.path.exists(OUT) and all(os.path.getmtime(d) <= os.path.getmtime(OUT) for d in deps):
        return OUT
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    objs = []
    procs = []
    for s in srcs:
        o = os.path.join(os.path.dirname(OUT), os.path.basename(s)[:-3] + ".o")
        objs.append(o)
        if os.path.exists(o) and all(os.path.getmtime(d) <= os.path.getmtime(o) for d in deps if not d.endswith(".cu") or d == s):
            continue
        procs.append(subprocess.Popen(["g++", "-x", "c++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC",
                                       "-DPACE_HOST_EMU", "-c", s, "-o", o]))
    for p in procs:
        if p.wait() != 0:
            raise RuntimeError("host-emu build failed")
    subprocess.check_call(["g++", "-shared", "-o", OUT] + objs)
    return OUT


def load():
    global _cached
    if _cached is None:
        from pace_b200 import _lib

        _cached = _lib.bind(ctypes.CDLL(build()))
    return _cached

"""TEST / BASELINE INFRASTRUCTURE ONLY — the CPU port of the hot path.

The kernel sources of pace_b200/csrc compiled by g++ with -DPACE_HOST_EMU: every kernel launch becomes a plain
(optionally OpenMP) loop nest over the same index space, running the same per-point code, with the host libm.
It is checked bit-for-bit against the golden vectors produced by the unmodified reference (tests/golden/, written by
oracle/gen_golden.py), so it doubles as
  * the CPU-side checker of kernel logic on machines without a GPU (tests/test_*hostemu*), and
  * the CPU baseline timed by bench.py (`cpu_baseline.kind = "port"`, and `--impl reference`) on the GPU box, where
    the Python reference itself cannot run.
Only tests/, __graft_entry__.smoke() and bench.py may load it; the product (pace_b200/) never does.
"""
import ctypes
import glob
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT_DIR = os.path.join(ROOT, "oracle", "_build")
_cached = {}


def build(openmp=False):
    tag = "omp" if openmp else "serial"
    out = os.path.join(OUT_DIR, "libpace_b200_cpu_%s.so" % tag)
    srcs = sorted(glob.glob(os.path.join(ROOT, "pace_b200", "csrc", "*.cu")))
    deps = srcs + glob.glob(os.path.join(ROOT, "pace_b200", "csrc", "*.[chi]*")) + [os.path.join(ROOT, "include", "pace_b200.h")]
    if os.path.exists(out) and all(os.path.getmtime(d) <= os.path.getmtime(out) for d in deps):
        return out
    os.makedirs(os.path.join(OUT_DIR, tag), exist_ok=True)
    flags = ["-x", "c++", "-std=c++17", "-O2", "-mfma", "-ffp-contract=off", "-fPIC", "-DPACE_HOST_EMU"]
    if openmp:
        flags += ["-fopenmp", "-DPACE_HOST_EMU_OMP"]
    objs, procs = [], []
    for s in srcs:
        o = os.path.join(OUT_DIR, tag, os.path.basename(s)[:-3] + ".o")
        objs.append(o)
        if os.path.exists(o) and all(os.path.getmtime(d) <= os.path.getmtime(o) for d in deps if not d.endswith(".cu") or d == s):
            continue
        procs.append(subprocess.Popen(["g++"] + flags + ["-c", s, "-o", o]))
    for p in procs:
        if p.wait() != 0:
            raise RuntimeError("CPU port build failed")
    subprocess.check_call(["g++", "-shared", "-o", out] + objs + (["-fopenmp"] if openmp else []))
    return out


def load(openmp=False):
    if openmp not in _cached:
        from pace_b200 import _lib

        _cached[openmp] = _lib.bind(ctypes.CDLL(build(openmp)))
    return _cached[openmp]

"""Host <-> device copy bandwidth of this box: alone, both directions at once, and with the pinned buffers allocated
from CPUs of each NUMA node (first touch).  Explains the e2e figure of bench.py, which is bound by these copies."""
import os
import subprocess
import sys
import time

import torch


def bw(fn, nbytes, reps=4):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return nbytes * reps / (time.perf_counter() - t0) / 1e9


def measure(tag):
    n = 1 << 30
    h1 = torch.empty(n, dtype=torch.uint8).pin_memory()
    h2 = torch.empty(n, dtype=torch.uint8).pin_memory()
    h1.zero_(); h2.zero_()
    d1 = torch.empty(n, dtype=torch.uint8, device="cuda")
    d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def up():
        with torch.cuda.stream(s1):
            d1.copy_(h1, non_blocking=True)

    def down():
        with torch.cuda.stream(s2):
            h2.copy_(d2, non_blocking=True)

    def both():
        up(); down()

    print("%-28s h2d alone %.1f GB/s, d2h alone %.1f GB/s, both at once %.1f GB/s each" % (
        tag, bw(up, n), bw(down, n), bw(both, n)), flush=True)


def main():
    for cmd in (["nvidia-smi", "topo", "-m"], ["bash", "-c", "nvidia-smi -q | grep -A12 'GPU Link Info' | head -16"],
                ["bash", "-c", "lscpu | grep -i 'numa\\|socket\\|model name'"]):
        try:
            print(subprocess.run(cmd, capture_output=True, text=True, timeout=60).stdout, flush=True)
        except Exception as e:  # noqa: BLE001
            print(cmd, "failed:", e)
    torch.cuda.init()
    measure("default affinity")
    nodes = sorted(d for d in os.listdir("/sys/devices/system/node") if d.startswith("node"))
    for nd in nodes:
        cpus = open("/sys/devices/system/node/%s/cpulist" % nd).read().strip()
        ids = set()
        for part in cpus.split(","):
            if not part:
                continue
            a, _, b = part.partition("-")
            ids.update(range(int(a), int(b or a) + 1))
        allowed = ids & os.sched_getaffinity(0)
        if not allowed:
            print(nd, "no allowed CPUs")
            continue
        old = os.sched_getaffinity(0)
        os.sched_setaffinity(0, allowed)
        measure("buffers from %s (%s)" % (nd, cpus))
        os.sched_setaffinity(0, old)


if __name__ == "__main__":
    sys.exit(main())

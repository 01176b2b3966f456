"""TEST-ONLY: the CPU port of the kernels (oracle/cpu_port.py) used to check kernel logic against the golden
fixtures on the CPU-only build machine; GPU parity is established separately by the `-m gpu` tests."""
from oracle import cpu_port


def build():
    return cpu_port.build(False)


def load():
    return cpu_port.load(False)

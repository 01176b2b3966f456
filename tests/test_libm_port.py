"""The kernels' exp/log/pow (csrc/libm_port.cuh) must be bit-identical to the host libm the oracle ran on
(glibc, reached here through Python's math module)."""
import ctypes
import math

import numpy as np
import pytest
import torch


def run(backend):
    from tests.module_harness import make_core

    core = make_core(backend, nx=12, ny=12, npz=3)
    rng = np.random.default_rng(0)
    n = 200000
    dev = core.device
    cases = {
        0: (rng.uniform(-40, 40, n), None, math.exp),
        1: (np.concatenate([np.exp(rng.uniform(-30, 30, n // 2)), rng.uniform(0.9, 1.1, n // 2)]), None, math.log),
        2: (np.exp(rng.uniform(-10, 12, n)), np.where(rng.random(n) < 0.5, 0.2857340234, rng.uniform(-3, 3, n)), math.pow),
    }
    for op, (x, y, fn) in cases.items():
        tx = torch.from_numpy(x).to(dev)
        ty = torch.from_numpy(y if y is not None else x).to(dev)
        out = torch.empty_like(tx)
        rc = core.lib.pace_b200_math_eval(core.handle, op, n, ctypes.c_void_p(tx.data_ptr()),
                                          ctypes.c_void_p(ty.data_ptr()), ctypes.c_void_p(out.data_ptr()))
        assert rc == 0
        core.sync()
        got = out.cpu().numpy()
        want = np.array([fn(a) if y is None else fn(a, b) for a, b in zip(x, y if y is not None else x)])
        assert np.array_equal(got, want), "op %d: %d of %d differ" % (op, np.count_nonzero(got != want), n)


def test_libm_port_hostemu():
    run("hostemu")


@pytest.mark.gpu
def test_libm_port_gpu():
    run("cuda")

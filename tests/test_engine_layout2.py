"""One full DynamicalCore.step_dynamics on the 2x2 layout (24 subdomains of 6x6x79 at C12, the partitioning of the
C384 headline configuration) against the final state of the reference run with 24 ranks
(tests/golden/c12_24rank_*.npz, oracle/gen_golden.py --layout 2)."""
import os

import numpy as np
import pytest

from tests.golden_util import GOLDEN, passes, report
from tests.module_harness import K_INTERFACE
from tests.test_engine_acoustic import STEP_OUT

N = 6  # cells per subdomain edge


def load24(kind):
    z = np.load(os.path.join(GOLDEN, "c12_24rank_%s.npz" % kind))
    out = [dict() for _ in range(24)]
    for key in z.files:
        r, name = key.split(":", 1)
        a = z[key]
        if kind != "grid" and a.ndim == 3:
            full = np.zeros((N + 7, N + 7, a.shape[2]))
            full[3:3 + N + 1, 3:3 + N + 1] = a
            a = full
        out[int(r)][name] = a
    return out


def run_step(backend, tol):
    from pace_b200.engine import DycoreEngine

    kw = {}
    if backend == "hostemu":
        from tests import hostemu

        kw = dict(device="cpu", _lib_handle=hostemu.load())
    eng = DycoreEngine(12, 2, 79, load24("grid"), **kw)
    eng.set_state(load24("init"))
    eng.step_dynamics()
    eng.core.sync()
    want = load24("final")
    worst = {}
    for name in STEP_OUT:
        got = eng.core.get(name)
        for r in range(24):
            g, w = got[r], want[r][name]
            if g.ndim == 3 and name not in K_INTERFACE:
                g, w = g[:, :, :-1], w[:, :, :-1]
            g, w = g[3:3 + N, 3:3 + N], w[3:3 + N, 3:3 + N]
            err, nbad = report(name, g, w)
            if nbad:
                worst[(name, r)] = (err, nbad)
    msg = "\n".join("  %-10s rank %2d max_rel=%.3e non-identical=%d" % (n, r, e, b)
                    for (n, r), (e, b) in sorted(worst.items()))
    print(msg)
    assert all(passes(e, b, tol) for e, b in worst.values()), "2x2-layout full-step parity failed (tol %g):\n%s" % (tol, msg)


def test_full_step_layout2_hostemu():
    run_step("hostemu", 0.0)


@pytest.mark.gpu
def test_full_step_layout2_gpu():
    run_step("cuda", 0.0)

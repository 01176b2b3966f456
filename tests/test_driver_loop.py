"""The driver's time loop on the device (examples/coupled_loop.py: dycore step, DycoreToPhysics, a stand-in physics,
UpdateAtmosphereState, diagnostics) and restarts: a run that is stopped, written to a restart and continued in a fresh
engine must end bit-identical to the uninterrupted run."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from tests.test_coupling import load_coupling  # noqa: E402
from tests.test_engine_acoustic import load_six  # noqa: E402

CHECK = ("u", "v", "w", "pt", "delp", "delz", "qvapor", "qliquid", "ua", "va", "pe", "ps")


def make(backend):
    from pace_b200.engine import DycoreEngine

    kw = {}
    if backend == "hostemu":
        from tests import hostemu

        kw = dict(device="cpu", _lib_handle=hostemu.load())
    eng = DycoreEngine(12, 1, 79, load_six("grid"), **kw)
    eng.set_state(load_six("init"))
    return eng


def run_loop(backend, tmp_path):
    from examples.coupled_loop import run
    from pace_b200.egress import MonitorDiagnostics, load_restart, save_restart

    dgrids = [f["grid"] for f in load_coupling()]

    class Monitor:
        def __init__(self):
            self.times, self.col = [], []

        def store(self, state):
            self.times.append(state["time"])
            self.col.append(np.array(state["column_integrated_qvapor"][0].data))

        def store_constant(self, state):
            pass

        def cleanup(self):
            pass

    # uninterrupted: 3 steps, diagnostics after each
    mon = Monitor()
    a = make(backend)
    diag = MonitorDiagnostics(mon, ["pt"], ["column_integrated_qvapor"], [(40, ["w"])], dycore=a)
    run(a, dgrids, 3, diag)
    diag.cleanup()
    assert mon.times == [225.0, 450.0, 675.0]
    assert all(np.isfinite(c).all() for c in mon.col) and not np.array_equal(mon.col[0], mon.col[2])
    final = {n: a.core.get(n) for n in CHECK}
    for n in CHECK:
        assert all(np.isfinite(x[3:15, 3:15]).all() for x in final[n]), n
    # the physics acted: the coupled run differs from three plain dycore steps
    plain = make(backend)
    for _ in range(3):
        plain.step_dynamics()
    assert not np.array_equal(plain.core.get("pt")[0][3:15, 3:15, :79], final["pt"][0][3:15, 3:15, :79])
    # stopped after 2 steps, restart written, continued in a fresh engine
    b = make(backend)
    run(b, dgrids, 2)
    save_restart(b, str(tmp_path))
    c = make(backend)
    for n in ("u", "v", "w", "pt", "delp", "delz", "qvapor", "qliquid", "qrain", "qice", "qsnow", "qgraupel", "qo3mr",
              "qsgs_tke", "qcld", "ua", "va", "pe", "peln", "pk", "pkz", "q_con", "omga", "ps"):
        c.core.t[n].fill_(float("nan"))       # nothing of the fresh engine's own state may survive the restart
    load_restart(c, str(tmp_path))
    run(c, dgrids, 1)
    for n in CHECK:
        got = c.core.get(n)
        for r in range(6):
            g, w = got[r][3:15, 3:15], final[n][r][3:15, 3:15]
            if g.ndim == 3 and n != "pe":
                g, w = g[:, :, :79], w[:, :, :79]
            assert np.array_equal(g, w), "%s rank %d differs after the restart" % (n, r)


def test_driver_loop_and_restart_hostemu(tmp_path):
    run_loop("hostemu", tmp_path)


@pytest.mark.gpu
def test_driver_loop_and_restart_gpu(tmp_path):
    run_loop("cuda", tmp_path)

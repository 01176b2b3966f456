"""Diagnostics / restart egress: pace_b200.egress against the unmodified reference's
MonitorDiagnostics.store run on 6 ranks after one dycore step (oracle/gen_diagnostics_golden.py ->
tests/golden/c12_6rank_diagnostics.npz; the state after that step is tests/golden/c12_6rank_final.npz).  Bit for bit."""
import json
import os

import numpy as np
import pytest

from tests.golden_util import GOLDEN
from tests.test_engine_acoustic import load_six

NAMES = ["u", "v", "w", "pt", "delp", "qvapor", "ua", "pe", "ps"]
DERIVED = ["column_integrated_qvapor", "column_integrated_qliquid", "column_integrated_qo3mr"]
Z_SELECT = [(40, ["pt", "w"]), (78, ["ua", "qvapor"]), (0, ["delp"])]


def make_engine(backend):
    from pace_b200.engine import DycoreEngine

    kw = {}
    if backend == "hostemu":
        from tests import hostemu

        kw = dict(device="cpu", _lib_handle=hostemu.load())
    eng = DycoreEngine(12, 1, 79, load_six("grid"), **kw)
    eng.set_state(load_six("final"))
    return eng


class Recorder:
    def __init__(self):
        self.stored, self.constants, self.cleaned = [], [], False

    def store(self, state):
        # the buffers behind the quantities are reused: keep copies, like a monitor that writes them out
        self.stored.append({k: ([np.array(q.data) for q in v], v[0].dims, v[0].units) if k != "time" else v
                            for k, v in state.items()})

    def store_constant(self, state):
        self.constants.append(state)

    def cleanup(self):
        self.cleaned = True


def check_against_reference(stored):
    z = np.load(os.path.join(GOLDEN, "c12_6rank_diagnostics.npz"))
    meta = json.loads(bytes(z["__meta__"]).decode())
    final = load_six("final")
    assert sorted(k for k in stored if k != "time") == sorted(meta)
    for name, m in meta.items():
        arrs, dims, units = stored[name]
        assert list(dims) == m["dims"] and units == m["units"], name
        for r in range(6):
            got = arrs[r]
            assert list(got.shape) == m["shape"], (name, got.shape)
            if name in NAMES:   # the reference hands its monitor the state's own quantity: view[:] of the final state
                f = final[r][name]
                want = f[3:3 + got.shape[0], 3:3 + got.shape[1]]
                want = want[:, :, :got.shape[2]] if got.ndim == 3 else want
            else:
                want = z["r%d:%s" % (r, name)]
            assert np.array_equal(got, want), "%s rank %d: max |d| = %g" % (name, r, np.max(np.abs(got - want)))


def run_diagnostics(backend):
    from pace_b200.egress import MonitorDiagnostics

    eng = make_engine(backend)
    mon = Recorder()
    diag = MonitorDiagnostics(mon, NAMES, DERIVED, Z_SELECT, dycore=eng)
    diag.store(225.0)
    diag.cleanup()
    assert mon.cleaned and len(mon.stored) == 1 and mon.stored[0]["time"] == 225.0
    check_against_reference(mon.stored[0])


def test_diagnostics_hostemu():
    run_diagnostics("hostemu")


@pytest.mark.gpu
def test_diagnostics_gpu():
    run_diagnostics("cuda")


def test_invalid_names_raise_like_the_reference():
    from pace_b200.egress import AsyncEgress

    eng = make_engine("hostemu")
    with pytest.raises(ValueError, match="Invalid state variable nope"):
        AsyncEgress(eng, names=["nope"])
    with pytest.raises(ValueError, match="z_select only works"):
        AsyncEgress(eng, z_select=[(3, ["pe"])])   # z_interface fields are refused (diagnostics.py:50-54)
    with pytest.warns(UserWarning, match="not a supported diagnostic variable"):
        AsyncEgress(eng, derived_names=["vorticity"]).close()


def run_restart_roundtrip(backend, tmp_path):
    from pace_b200.egress import load_restart, save_restart

    eng = make_engine(backend)
    before = {n: eng.core.get(n) for n in ("u", "v", "pt", "delp", "pe", "ps", "qvapor")}
    save_restart(eng, str(tmp_path))
    assert sorted(os.listdir(str(tmp_path))) == ["restart_dycore_state_%d.npz" % r for r in range(6)]
    for n in before:
        eng.core.t[n].zero_()
    load_restart(eng, str(tmp_path))
    for n, arrs in before.items():
        got = eng.core.get(n)
        for r in range(6):
            a, b = got[r][3:15, 3:15], arrs[r][3:15, 3:15]
            if a.ndim == 3 and n != "pe":
                a, b = a[:, :, :79], b[:, :, :79]
            assert np.array_equal(a, b), (n, r)


def test_restart_roundtrip_hostemu(tmp_path):
    run_restart_roundtrip("hostemu", tmp_path)


@pytest.mark.gpu
def test_restart_roundtrip_gpu(tmp_path):
    run_restart_roundtrip("cuda", tmp_path)


@pytest.mark.gpu
def test_snapshots_are_taken_at_their_step_gpu():
    """store() returns before the data has left the device and the next steps overwrite the state: what the sink
    receives for step n must be the state after step n, not a later one."""
    from pace_b200.egress import AsyncEgress

    eng = make_engine("cuda")
    got, want = [], []

    def sink(time, fields):
        got.append((time, {k: [np.array(q.data) for q in v] for k, v in fields.items()}))

    eg = AsyncEgress(eng, names=["pt", "u"], derived_names=["column_integrated_qvapor"], z_select=[(10, ["w"])],
                     sink=sink, depth=2)
    for step in range(4):
        eng.step_dynamics()
        eg.store(step)
        # reference values for this step, fetched synchronously AFTER store() was issued
        want.append({"pt": [a[3:15, 3:15, :79].copy() for a in eng.core.get("pt")],
                     "u": [a[3:15, 3:16, :79].copy() for a in eng.core.get("u")],
                     "w_z10": [a[3:15, 3:15, 10].copy() for a in eng.core.get("w")]})
    eg.close()
    assert [t for t, _ in got] == [0, 1, 2, 3]
    for step in range(4):
        for name, arrs in want[step].items():
            for r in range(6):
                assert np.array_equal(got[step][1][name][r], arrs[r]), (step, name, r)
    # the state moved between the steps (otherwise the test above would prove nothing)
    assert not np.array_equal(got[0][1]["pt"][0], got[3][1]["pt"][0])

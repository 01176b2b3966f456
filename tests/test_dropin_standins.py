"""pace_b200.dropin driven the way a user of the reference drives pyFV3 — constructor arguments, `step_dynamics(state)`,
`.acoustic_dynamics(state, timestep, n_map)`, a checkpointer — from Quantity-like stand-ins built out of the C12
fixtures (no reference import, so it runs on the GPU box), against the reference's own results
(tests/golden/c12_6rank_final.npz, c12_rank0_stages.npz)."""
import os
from types import SimpleNamespace

import numpy as np
import pytest

from tests.golden_util import GOLDEN
from tests.module_harness import K_INTERFACE
from tests.test_engine_acoustic import load_six


class Q:
    """Stand-in for ndsl.Quantity: the drop-in only reads / writes `.data`."""

    def __init__(self, a):
        self.data = np.array(a, dtype=np.float64)


def reference_like_objects():
    from oracle.refshim.run_reference import default_config_kwargs  # plain dict of the fixture's namelist, no reference import

    grids, init = load_six("grid"), load_six("init")
    grid_data, damping, states, phis = [], [], [], []
    for g, s in zip(grids, init):
        gd = SimpleNamespace(**{k[5:]: v for k, v in g.items() if k.startswith("grid.")})
        gd.ks = 18
        grid_data.append(gd)
        damping.append(SimpleNamespace(**{k[5:]: v for k, v in g.items() if k.startswith("damp.")}))
        st = SimpleNamespace(**{k: Q(v) for k, v in s.items()})
        states.append(st)
        phis.append(st.phis)
    kw = default_config_kwargs()
    kw.update(n_sponge=1, nf_omega=1, grid_type=0, convert_ke=False, adiabatic=False, use_old_omega=True, do_f3d=False,
              inline_q=False, rf_fast=True)
    return grid_data, damping, SimpleNamespace(**kw), phis, states


def make(backend, checkpointer=None):
    from pace_b200.dropin import DynamicalCore

    kw = {}
    if backend == "hostemu":
        from tests import hostemu

        kw = dict(device="cpu", _lib_handle=hostemu.load())
    grid_data, damping, config, phis, states = reference_like_objects()
    core = DynamicalCore(None, grid_data, None, None, damping, config, phis, states, 225.0, checkpointer, **kw)
    return core, states


def run_step(backend):
    seen = []
    kept = {}

    def checkpointer(savepoint, **kw):
        seen.append((savepoint, sorted(kw)))
        if savepoint in ("C_SW-Out", "Remapping-In"):
            kept[savepoint] = {k: np.array(v[0]) for k, v in kw.items()}

    core, states = make(backend, checkpointer)
    assert hasattr(core, "acoustic_dynamics") and hasattr(core, "tracer_advection")
    core.step_dynamics(states)
    want = load_six("final")
    for name in ("u", "v", "w", "delp", "pt", "delz", "qvapor", "qliquid", "pe", "pk", "peln", "pkz", "q_con", "omga",
                 "ua", "va", "ps", "mfxd", "cxd", "diss_estd"):
        for r in range(6):
            g, w = getattr(states[r], name).data, want[r][name]
            if g.ndim == 3 and name not in K_INTERFACE:
                g, w = g[:, :, :-1], w[:, :, :-1]
            assert np.array_equal(g[3:15, 3:15], w[3:15, 3:15]), "%s rank %d differs" % (name, r)
    # the checkpointer saw the reference's savepoints in the reference's order, with the reference's keywords
    names = [s for s, _ in seen]
    assert names == ["FVDynamics-In", "C_SW-In", "C_SW-Out", "D_SW-In", "D_SW-Out", "Tracer2D1L-In", "Tracer2D1L-Out",
                     "Remapping-In", "Remapping-Out", "FVDynamics-Out"]
    assert dict(seen)["D_SW-In"] == sorted(["ucd", "vcd", "wd", "delpcd", "delpd", "ud", "vd", "ptd", "uad", "vad", "zhd",
                                            "divgdd", "xfxd", "yfxd", "mfxd", "mfyd"])
    # and at the right moment: what it was handed after c_sw is the reference's state after c_sw (rank 0)
    from tests.golden_util import Stages

    after_csw = Stages("c12_rank0_stages.npz").after("c_sw")
    for key, name in (("ucd", "uc"), ("vcd", "vc"), ("utd", "ad.ut"), ("divgdd", "ad.divgd")):
        assert np.array_equal(kept["C_SW-Out"][key][3:15, 3:15, :79], after_csw[name][3:15, 3:15, :79]), key
    assert kept["Remapping-In"]["pe"].shape == (19, 80, 19)  # [x, z, y] like the reference's transposed savepoint data
    # _cappa is reachable like on the reference object
    assert core._cappa[0].shape == (19, 19, 80)


def run_acoustic(backend):
    """AcousticDynamics.__call__(state, timestep, n_map) through the drop-in's .acoustic_dynamics, from the state the
    step's preamble leaves, against the reference's rank-0 state after its own acoustic_dynamics call."""
    from tests.golden_util import Stages

    core, states = make(backend)
    eng = core.engine
    fields, names = core._upload(states)
    eng.compute_preamble()
    eng.core.call("copy_levels", "delp", "dyn.dp1")
    core._download(fields, names)
    core.acoustic_dynamics(states, 225.0, n_map=1)
    want = Stages("c12_rank0_stages.npz").after("acoustic_dynamics")
    for name in ("u", "v", "w", "delp", "pt", "delz", "pe", "q_con", "uc", "vc", "mfxd", "cyd", "omga"):
        g, w = getattr(states[0], name).data, want[name]
        if name not in K_INTERFACE:
            g, w = g[:, :, :-1], w[:, :, :-1]
        assert np.array_equal(g[3:15, 3:15], w[3:15, 3:15]), name


def test_dropin_step_hostemu():
    run_step("hostemu")


def test_dropin_acoustic_hostemu():
    run_acoustic("hostemu")


@pytest.mark.gpu
def test_dropin_step_gpu():
    run_step("cuda")


@pytest.mark.gpu
def test_dropin_acoustic_gpu():
    run_acoustic("cuda")


def test_write_back_refuses_unknown_storage():
    from pace_b200.dropin import _write_back

    class Opaque:
        data = object()

    with pytest.raises(TypeError):
        _write_back(Opaque(), np.zeros((2, 2)))

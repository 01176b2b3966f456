"""Whole AcousticDynamics call on all 6 subdomains of a C12 cube resident on one device, with the device-side
halo exchange, against the reference's rank-0 stage log (tests/golden/c12_rank0_stages.npz) — the other 5 ranks
only enter through the halos they send."""
import os

import numpy as np
import pytest

from tests.golden_util import GOLDEN, passes, report
from tests.module_harness import K_INTERFACE


def load_six(kind):
    z = np.load(os.path.join(GOLDEN, "c12_6rank_%s.npz" % kind))
    out = [dict() for _ in range(6)]
    for key in z.files:
        r, name = key.split(":", 1)
        a = z[key]
        if kind != "grid" and a.ndim == 3:
            full = np.zeros((19, 19, a.shape[2]))
            full[3:16, 3:16] = a
            a = full
        out[int(r)][name] = a
    return out


def make_engine(backend):
    from pace_b200.engine import DycoreEngine

    kw = {}
    if backend == "hostemu":
        from tests import hostemu

        kw = dict(device="cpu", _lib_handle=hostemu.load())
    eng = DycoreEngine(12, 1, 79, load_six("grid"), **kw)
    eng.set_state(load_six("init"))
    return eng


def compare_rank0(eng, want, names, tol):
    worst = {}
    for name in names:
        got = eng.core.get(name)[0]
        w = want[name]
        if got.ndim == 3 and name not in K_INTERFACE:
            got, w = got[:, :, :-1], w[:, :, :-1]
        # compute domain (incl. the interface row/column): the 6-rank fixture carries no initial halos, and halo
        # cells no exchange ever refreshes keep their initial values in the reference
        got, w = got[3:16, 3:16], w[3:16, 3:16]
        err, nbad = report(name, got, w)
        worst[name] = (err, nbad)
    msg = "\n".join("  %-16s max_rel=%.3e non-identical=%d" % (n, e, b) for n, (e, b) in sorted(worst.items()) if b)
    print(msg)
    bad = {n: e for n, (e, b) in worst.items() if not passes(e, b, tol)}
    assert not bad, "parity failed (tol %g):\n%s" % (tol, msg)


ACOUSTIC_OUT = ["u", "v", "w", "delp", "pt", "delz", "pe", "pk", "peln", "q_con", "uc", "vc", "ua", "va", "mfxd", "mfyd",
                "cxd", "cyd", "omga", "diss_estd", "ad.gz", "ad.zh", "ad.pkc", "ad.pk3", "ad.heat_source", "ad.divgd",
                "ad.ut", "ad.vt", "ad.crx", "ad.cry", "ad.xfx", "ad.yfx", "ad.ws3", "ad.wsd", "ad.cappa", "csw.delpc",
                "csw.ptc"]


def run(backend, golden, tol):
    eng = make_engine(backend)
    eng.compute_preamble()
    eng.core.call("copy_levels", "delp", "dyn.dp1")
    eng.acoustic_dynamics(225.0, n_map=1)
    eng.core.sync()
    compare_rank0(eng, golden.after("acoustic_dynamics"), ACOUSTIC_OUT, tol)


def test_acoustic_six_tiles_hostemu(golden):
    run("hostemu", golden, 0.0)


@pytest.mark.gpu
def test_acoustic_six_tiles_gpu(golden):
    run("cuda", golden, 0.0)


TRACER_OUT = ["qvapor", "qliquid", "qrain", "qice", "qsnow", "qgraupel", "qo3mr", "qsgs_tke", "cxd", "cyd", "mfxd",
              "mfyd", "dyn.dp1"]


def run_tracer(backend, golden, tol):
    eng = make_engine(backend)
    eng.compute_preamble()
    eng.core.call("copy_levels", "delp", "dyn.dp1")
    eng.acoustic_dynamics(225.0, n_map=1)
    eng.tracer_advection()
    eng.core.sync()
    compare_rank0(eng, golden.after("tracer_advection"), TRACER_OUT, tol)


def test_tracer_six_tiles_hostemu(golden):
    run_tracer("hostemu", golden, 0.0)


@pytest.mark.gpu
def test_tracer_six_tiles_gpu(golden):
    run_tracer("cuda", golden, 0.0)


STEP_OUT = ["u", "v", "w", "ua", "va", "uc", "vc", "delp", "delz", "pe", "pt", "peln", "pk", "pkz", "qvapor", "qliquid",
            "qice", "qrain", "qsnow", "qgraupel", "qo3mr", "qsgs_tke", "qcld", "q_con", "omga", "mfxd", "mfyd", "cxd",
            "cyd", "diss_estd", "ps"]


def run_step(backend, tol):
    """One full DynamicalCore.step_dynamics on all 6 subdomains against the reference's final state of all 6 ranks."""
    eng = make_engine(backend)
    eng.step_dynamics()
    eng.core.sync()
    want = load_six("final")
    worst = {}
    for name in STEP_OUT:
        got = eng.core.get(name)
        for r in range(6):
            g, w = got[r], want[r][name]
            if g.ndim == 3 and name not in K_INTERFACE:
                g, w = g[:, :, :-1], w[:, :, :-1]
            if g.ndim == 3:
                # u lives on y interfaces, v on x interfaces: their closing row / column belongs to the comparison
                ni, nj = (13 if name == "v" else 12), (13 if name == "u" else 12)
                g, w = g[3:3 + ni, 3:3 + nj], w[3:3 + ni, 3:3 + nj]
            else:
                g, w = g[3:15, 3:15], w[3:15, 3:15]
            err, nbad = report(name, g, w)
            if nbad:
                worst[(name, r)] = (err, nbad)
    msg = "\n".join("  %-10s rank %d max_rel=%.3e non-identical=%d" % (n, r, e, b) for (n, r), (e, b) in sorted(worst.items()))
    print(msg)
    assert all(passes(e, b, tol) for e, b in worst.values()), "full-step parity failed (tol %g):\n%s" % (tol, msg)


def test_full_step_six_tiles_hostemu():
    run_step("hostemu", 0.0)


@pytest.mark.gpu
def test_full_step_six_tiles_gpu():
    run_step("cuda", 0.0)


def run_two_steps(backend, tol, poison=False):
    """Two dycore steps with k_split=2, n_split=2 (all branches of the acoustic loop that depend on the substep and
    remap-step counters; non-zero w, q_con and condensates in the second step) against the reference's final state
    (tests/golden/c12_6rank_final_ns2_ks2_s2.npz, oracle/gen_golden.py --n-split 2 --k-split 2 --steps 2 --final-only)."""
    from pace_b200.engine import DycoreEngine

    kw = {}
    if backend == "hostemu":
        from tests import hostemu

        kw = dict(device="cpu", _lib_handle=hostemu.load())
    eng = DycoreEngine(12, 1, 79, load_six("grid"), n_split=2, k_split=2, **kw)
    eng.set_state(load_six("init"))
    if poison:
        # engine.STEP_INPUTS under k_split = 2, n_split = 2 as well: everything a step does not read is NaN beforehand
        from pace_b200.engine import STATE_2D, STATE_3D, STEP_INPUTS

        for name in STATE_3D + STATE_2D:
            if name not in STEP_INPUTS:
                eng.core.t[name].fill_(float("nan"))
    eng.step_dynamics()
    if poison:
        # a host-resident driver re-uploads only the inputs before the second step too
        for name in STATE_3D + STATE_2D:
            if name not in STEP_INPUTS:
                eng.core.t[name].fill_(float("nan"))
    eng.step_dynamics()
    eng.core.sync()
    z = np.load(os.path.join(GOLDEN, "c12_6rank_final_ns2_ks2_s2.npz"))
    worst = {}
    for name in STEP_OUT:
        got = eng.core.get(name)
        for r in range(6):
            g, w = got[r], z["%d:%s" % (r, name)]
            if g.ndim == 3:
                g = g[3:16, 3:16]
                if name not in K_INTERFACE:
                    g, w = g[:, :, :-1], w[:, :, :-1]
            g, w = g[:12, :12] if g.ndim == 3 else g[3:15, 3:15], w[:12, :12] if w.ndim == 3 else w[3:15, 3:15]
            err, nbad = report(name, g, w)
            if nbad:
                worst[(name, r)] = (err, nbad)
    msg = "\n".join("  %-10s rank %d max_rel=%.3e non-identical=%d" % (n, r, e, b) for (n, r), (e, b) in sorted(worst.items()))
    print(msg)
    assert all(passes(e, b, tol) for e, b in worst.values()), "two-step parity failed (tol %g):\n%s" % (tol, msg)


def test_two_steps_ksplit2_nsplit2_hostemu():
    run_two_steps("hostemu", 0.0)


def test_step_inputs_ksplit2_nsplit2_hostemu():
    run_two_steps("hostemu", 0.0, poison=True)


@pytest.mark.gpu
def test_two_steps_ksplit2_nsplit2_gpu():
    run_two_steps("cuda", 0.0)


@pytest.mark.gpu
@pytest.mark.parametrize("nsets,nsteps", [(3, 4), (2, 3)])
def test_host_pipeline_matches_plain_steps_gpu(nsets, nsteps):
    """HostPipeline (state streamed from / to pinned host memory on side streams, nsets device copies): steps
    that each start from the same host-resident initial state must each return the reference's one-step result."""
    import torch

    from pace_b200.engine import HostPipeline

    eng = make_engine("cuda")
    pipe = HostPipeline(eng, nsets=nsets)
    assert len(pipe.sets) == nsets
    hin = pipe.pinned_like()
    for name in pipe.names:
        hin[name].copy_(eng.core.t[name])
    # the pipeline uploads the step's inputs only: whatever the other fields hold on the device must not matter
    for dev_set in pipe.sets:
        for name in pipe.names:
            if name not in pipe.upload_names:
                dev_set[name].fill_(float("nan"))
    assert len(pipe.upload_names) == 16
    outs = [pipe.pinned_like() for _ in range(nsteps)]
    pipe.run(nsteps, hin, lambda n: outs[n])
    assert pipe.copy_ms[0] > 0.0 and pipe.copy_ms[1] > 0.0
    want = load_six("final")
    for n in range(nsteps):
        for name in ("u", "v", "w", "delp", "pt", "delz", "qvapor", "pe", "pkz", "q_con"):
            t = outs[n][name].numpy()
            for r in range(6):
                got = np.ascontiguousarray(t[r, :, :, :19].transpose(2, 1, 0))[3:15, 3:15]
                w = want[r][name][3:15, 3:15]
                if name not in K_INTERFACE:
                    got, w = got[:, :, :-1], w[:, :, :-1]
                assert np.array_equal(got, w), "step %d %s rank %d differs" % (n, name, r)
    # the engine is bound to its own tensors again and holds the last result
    assert eng.core.t["u"].data_ptr() == pipe.sets[0]["u"].data_ptr()
    assert torch.equal(eng.core.t["delp"].cpu(), outs[nsteps - 1]["delp"])


def test_step_inputs_hostemu():
    """engine.STEP_INPUTS: with every other DycoreState field poisoned in the initial state the step returns the
    reference's result (compute domain, bit for bit)."""
    from pace_b200.engine import STATE_2D, STATE_3D, STEP_INPUTS

    eng = make_engine("hostemu")
    for name in STATE_3D + STATE_2D:
        if name not in STEP_INPUTS:
            eng.core.t[name].fill_(float("nan"))
    eng.step_dynamics()
    want = load_six("final")
    for name in ("u", "v", "w", "delp", "pt", "delz", "qvapor", "pe", "pk", "peln", "pkz", "q_con", "omga", "ua", "va", "ps"):
        got = eng.core.get(name)
        for r in range(6):
            a, w = got[r][3:15, 3:15], want[r][name][3:15, 3:15]
            if a.ndim == 3 and name not in K_INTERFACE:
                a, w = a[:, :, :-1], w[:, :, :-1]
            assert np.array_equal(a, w), "%s rank %d differs" % (name, r)


NEGQ_OUT = ["u", "v", "w", "delp", "pt", "delz", "qvapor", "qliquid", "qrain", "qice", "qsnow", "qgraupel", "qo3mr",
            "qsgs_tke", "qcld", "q_con", "pkz", "omga", "ua", "va", "pe", "ps"]


def run_negq(backend, tol):
    """One step from a perturbed initial state — noisy w, noisy tracers of every species, a third of the tracer values
    negative — so that fillz, neg_adj3, the w transport and the saturation adjustment do real work
    (tests/golden/c12_6rank_negq.npz, oracle/gen_negq_golden.py; outputs of ranks 0 and 4 are stored)."""
    z = np.load(os.path.join(GOLDEN, "c12_6rank_negq.npz"))
    states = load_six("init")
    for r in range(6):
        for key in z.files:
            rr, kind, name = key.split(":", 2)
            if int(rr) == r and kind == "in":
                full = np.zeros((19, 19, z[key].shape[2]))
                full[3:16, 3:16] = z[key]
                states[r][name] = full
    from pace_b200.engine import DycoreEngine

    kw = {}
    if backend == "hostemu":
        from tests import hostemu

        kw = dict(device="cpu", _lib_handle=hostemu.load())
    eng = DycoreEngine(12, 1, 79, load_six("grid"), **kw)
    eng.set_state(states)
    eng.step_dynamics()
    eng.core.sync()
    worst = {}
    for name in NEGQ_OUT:
        got = eng.core.get(name)
        for r in (0, 4):
            g, w = got[r], z["%d:out:%s" % (r, name)]
            if g.ndim == 3:
                g = g[3:15, 3:15]
                w = w[:12, :12]
                if name not in K_INTERFACE:
                    g, w = g[:, :, :-1], w[:, :, :-1]
            else:
                g, w = g[3:15, 3:15], w[3:15, 3:15]
            err, nbad = report(name, g, w)
            if nbad:
                worst[(name, r)] = (err, nbad)
    msg = "\n".join("  %-10s rank %d max_rel=%.3e non-identical=%d" % (n, r, e, b) for (n, r), (e, b) in sorted(worst.items()))
    print(msg)
    assert all(passes(e, b, tol) for e, b in worst.values()), "perturbed-state parity failed (tol %g):\n%s" % (tol, msg)


def test_perturbed_state_step_hostemu():
    run_negq("hostemu", 0.0)


@pytest.mark.gpu
def test_perturbed_state_step_gpu():
    run_negq("cuda", 0.0)

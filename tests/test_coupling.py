"""Dycore <-> physics coupling: pace_b200_dycore_to_physics / pace_b200_update_atmos_state against the
unmodified reference classes DycoreToPhysics and UpdateAtmosphereState run on 6 ranks at C12
(oracle/gen_coupling_golden.py -> tests/golden/c12_6rank_coupling.npz).  Bit for bit, compute domain."""
import os

import numpy as np
import pytest

from tests.golden_util import GOLDEN
from tests.test_engine_acoustic import load_six

NX = 12


def load_coupling():
    z = np.load(os.path.join(GOLDEN, "c12_6rank_coupling.npz"))
    out = [dict() for _ in range(6)]
    for key in z.files:
        r, kind, name = key.split(":", 2)
        out[int(r[1:])].setdefault(kind, {})[name] = z[key]
    return out


def full(a):
    """fixture arrays hold [is, ie+1] x [js, je+1]; back to the array with its 3-cell halo"""
    f = np.zeros((19, 19) + a.shape[2:])
    f[3:16, 3:16] = a
    return f


def run_coupling(backend, comm=None, mine=None):
    """mine / comm: the cube ranks this process holds and its pace_b200.distributed.TorchComm (multi-process runs)"""
    from pace_b200.engine import COUPLING_COPIED, COUPLING_PHYSICS_OUT, DycoreEngine

    kw = {}
    if backend == "hostemu":
        from tests import hostemu

        kw = dict(device="cpu", _lib_handle=hostemu.load())
    elif backend != "cuda":
        kw = dict(device=backend)
    mine = list(range(6)) if mine is None else list(mine)
    fx_all = load_coupling()
    fx = [fx_all[r] for r in mine]
    grids, init = load_six("grid"), load_six("init")
    eng = DycoreEngine(NX, 1, 79, [grids[r] for r in mine], comm=comm, **kw)
    states = []
    for slot, r in enumerate(mine):
        s = {"phis": init[r]["phis"]}
        for name, a in fx[slot]["in"].items():
            s[name] = full(a)
        states.append(s)
    eng.set_state(states)
    eng.setup_coupling([f["grid"] for f in fx])
    core = eng.core

    # DycoreToPhysics: plain copies on [is, ie+1] x [js, je+1] x [0, npz) (checked against the reference's output when
    # the fixture was generated: oracle/gen_coupling_golden.py)
    eng.dycore_to_physics()
    for name in COUPLING_COPIED:
        got = core.get("phy." + name)
        for r in range(len(mine)):
            want = full(fx[r]["in"][name])
            assert np.array_equal(got[r][3:16, 3:16, :79], want[3:16, 3:16, :79]), (name, r)
            assert not got[r][:, :, 79].any(), (name, r)
    # the physics' part: prsi and the updated fields come from the fixture
    for name in COUPLING_PHYSICS_OUT:
        core.put("phy." + name, [full(fx[r]["phy"][name]) for r in range(len(mine))])
    eng.update_atmos_state()
    core.sync()
    bad = []
    for r in range(len(mine)):
        for name, want in fx[r]["out"].items():
            got = core.get(name)[r]
            if want.ndim == 3:
                g, w = got[3:15, 3:15, :79], want[:12, :12, :79]
                if name == "u":
                    g, w = got[3:15, 3:16, :79], want[:12, :13, :79]
                elif name == "v":
                    g, w = got[3:16, 3:15, :79], want[:13, :12, :79]
                elif name in ("pe", "peln", "pk"):
                    g, w = got[3:15, 3:15, :80], want[:12, :12, :80]
            else:
                g, w = got[3:15, 3:15], want[:12, :12]
            if not (np.isfinite(g).all() and np.array_equal(g, w)):
                bad.append("%s rank %d: %d values differ, max |d| = %g" % (name, r, (g != w).sum(), np.nanmax(np.abs(g - w))))
    assert not bad, "\n".join(bad)
    # fields the coupling must not touch
    for name in ("w", "delz", "omga", "qo3mr", "qcld"):
        got = core.get(name)
        for r in range(len(mine)):
            assert np.array_equal(got[r][3:15, 3:15, :79], full(fx[r]["in"][name])[3:15, 3:15, :79]), (name, r)


def test_coupling_hostemu():
    run_coupling("hostemu")


@pytest.mark.gpu
def test_coupling_gpu():
    run_coupling("cuda")


def run_coupling_layout2(backend):
    """The same coupling step on the 2x2 layout (24 subdomains of 6x6, the partitioning of the C384 configuration).  Given
    the same state the coupling is a function of each point and its neighbours, with tile-edge formulas that depend on
    the position on the TILE: the 24 blocks must reproduce the blocks of the 6-rank reference result bit for bit.  This
    checks the tile positions handed to AGrid2DGridPhysics (which half of an edge a subdomain lies in) and the halo
    updates of the tendencies between subdomains of one tile."""
    from pace_b200.engine import COUPLING_PHYSICS_OUT, DycoreEngine
    from tests.test_engine_layout2 import load24

    kw = {}
    if backend == "hostemu":
        from tests import hostemu

        kw = dict(device="cpu", _lib_handle=hostemu.load())
    fx = load_coupling()
    n = 6
    init = load24("init")
    eng = DycoreEngine(NX, 2, 79, load24("grid"), **kw)

    def block(a, rank, halo):
        """rank's block of a 6-rank array: with its 3-cell halo from a full array, else [is, ie+1] x [js, je+1]"""
        _, ix, jy = eng.planner.position(rank)
        if halo:
            return a[ix * n: ix * n + n + 7] if a.ndim == 1 else a[ix * n: ix * n + n + 7, jy * n: jy * n + n + 7]
        f = np.zeros((n + 7, n + 7) + a.shape[2:])
        f[3:3 + n + 1, 3:3 + n + 1] = a[ix * n: ix * n + n + 1, jy * n: jy * n + n + 1]
        return f

    states, dgrids = [], []
    for rank in range(24):
        tile = rank // 4
        s = {"phis": init[rank]["phis"]}
        for name, a in fx[tile]["in"].items():
            s[name] = block(a, rank, False)
        states.append(s)
        dgrids.append({name: block(a, rank, True) for name, a in fx[tile]["grid"].items()})
    eng.set_state(states)
    eng.setup_coupling(dgrids)
    eng.dycore_to_physics()
    for name in COUPLING_PHYSICS_OUT:
        eng.core.put("phy." + name, [block(fx[r // 4]["phy"][name], r, False) for r in range(24)])
    eng.update_atmos_state()
    eng.core.sync()
    bad = []
    for rank in range(24):
        tile = rank // 4
        _, ix, jy = eng.planner.position(rank)
        for name, want in fx[tile]["out"].items():
            got = eng.core.get(name)[rank]
            di = 1 if name == "v" else 0
            dj = 1 if name == "u" else 0
            nk = 80 if name in ("pe", "peln", "pk") else 79
            w = want[ix * n: ix * n + n + di, jy * n: jy * n + n + dj]
            g = got[3:3 + n + di, 3:3 + n + dj]
            if want.ndim == 3:
                g, w = g[:, :, :nk], w[:, :, :nk]
            if not (np.isfinite(g).all() and np.array_equal(g, w)):
                bad.append("%s rank %d: %d values differ" % (name, rank, (g != w).sum()))
    assert not bad, "\n".join(bad)


def test_coupling_layout2_hostemu():
    run_coupling_layout2("hostemu")


@pytest.mark.gpu
def test_coupling_layout2_gpu():
    run_coupling_layout2("cuda")


def _gloo_worker(rank, world, port, q):
    try:
        import sys

        sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        import torch.distributed as dist

        from pace_b200.distributed import Placement, TorchComm

        dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
        run_coupling("hostemu", comm=TorchComm(), mine=Placement(6, world).ranks_of(rank))
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, None))
    except BaseException as e:  # noqa: BLE001 — surfaced in the parent
        import traceback

        q.put((rank, "%s\n%s" % (e, traceback.format_exc())))


def test_coupling_two_processes_gloo():
    """The coupling on 2 processes holding 3 subdomains each: the halo updates of the tendencies and of the winds cross
    the process boundary (pace_b200.distributed over gloo)."""
    import torch.multiprocessing as mp

    from tests import hostemu
    from tests.test_distributed_gloo import _free_port

    hostemu.load()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=600) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, err in sorted(results):
        assert err is None, "rank %d: %s" % (rank, err)


def run_fill_only(backend):
    """UpdateAtmosphereState(dycore_only=True, apply_tendencies=False): fill_gfs_delp on the dycore's own qvapor."""
    from pace_b200.engine import DycoreEngine

    kw = {}
    if backend == "hostemu":
        from tests import hostemu

        kw = dict(device="cpu", _lib_handle=hostemu.load())
    fx = load_coupling()
    eng = DycoreEngine(NX, 1, 79, load_six("grid"), **kw)
    init = load_six("init")
    eng.set_state([{"phis": init[r]["phis"], "qvapor": full(fx[r]["in2"]["qvapor"]), "delp": full(fx[r]["in2"]["delp"])}
                   for r in range(6)])
    eng.setup_coupling([fx[r]["grid"] for r in range(6)])
    eng.update_atmos_state(dycore_only=True, apply_tendencies=False)
    eng.core.sync()
    got_q, got_dp = eng.core.get("qvapor"), eng.core.get("delp")
    changed = 0
    for r in range(6):
        want = fx[r]["out2"]["qvapor"][:12, :12, :79]
        assert np.array_equal(got_q[r][3:15, 3:15, :79], want), r
        assert np.array_equal(got_dp[r][3:15, 3:15, :79], fx[r]["in2"]["delp"][:12, :12, :79]), r
        changed += int((want != fx[r]["in2"]["qvapor"][:12, :12, :79]).sum())
    assert changed > 1000   # the fill had work to do


def test_fill_only_hostemu():
    run_fill_only("hostemu")


@pytest.mark.gpu
def test_fill_only_gpu():
    run_fill_only("cuda")

"""TEST INFRASTRUCTURE ONLY — pace_b200.dropin.DynamicalCore next to pyFV3.DynamicalCore on the same objects
(run by tests/test_dropin_reference.py in a fresh interpreter: the reference bootstrap has to precede numpy)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.refshim.bootstrap import bootstrap  # noqa: E402

bootstrap()
import numpy as np  # noqa: E402

from oracle.refshim.proc_comm import run_ranks  # noqa: E402


def _rank(mpi_comm):
    import copy
    from datetime import timedelta

    from oracle.refshim.run_reference import build_dycore, default_config_kwargs, state_arrays

    from pace_b200.dropin import DynamicalCore
    from oracle import cpu_port as hostemu

    dycore, state, grid_data, damping, _, config, communicator = build_dycore(mpi_comm, default_config_kwargs())
    mine_state = copy.deepcopy(state)
    mine = DynamicalCore(communicator, grid_data, None, None, damping, config, state.phis, mine_state,
                         timedelta(seconds=config.dt_atmos), device="cpu", _lib_handle=hostemu.load())
    mine.step_dynamics(mine_state)
    dycore.step_dynamics(state)
    want, got = state_arrays(state), state_arrays(mine_state)
    bad = []
    for name in ("u", "v", "w", "delp", "pt", "delz", "qvapor", "qliquid", "qice", "qrain", "qsnow", "qgraupel", "qo3mr",
                 "qsgs_tke", "qcld", "pe", "pk", "peln", "pkz", "ps", "q_con", "omga", "ua", "va", "mfxd", "mfyd",
                 "cxd", "cyd", "diss_estd"):
        a, b = got[name], want[name]
        a, b = (a[3:15, 3:15, :79], b[3:15, 3:15, :79]) if a.ndim == 3 else (a[3:15, 3:15], b[3:15, 3:15])
        if name in ("pe", "pk", "peln"):
            a, b = got[name][3:15, 3:15], want[name][3:15, 3:15]
        if not np.array_equal(a, b):
            bad.append((name, float(np.max(np.abs(a - b)))))
    bad += _coupling(mpi_comm, dycore, mine, state, mine_state, grid_data, config, communicator)
    return bad


def _coupling(mpi_comm, dycore, mine, state, mine_state, grid_data, config, communicator):
    """The coupling classes of pace_b200.dropin next to pySHiELD's (update_atmos_state.py), on the stepped states."""
    import copy
    import dataclasses

    import pySHiELD
    from ndsl.constants import X_DIM, Y_DIM, Z_DIM
    from ndsl.grid import DriverGridData
    from pySHiELD.update import update_atmos_state as ref

    from oracle.gen_coupling_golden import PHY_IN, synthesize_physics
    from oracle.refshim.run_reference import build_dycore, state_arrays
    from pace_b200 import dropin

    qf, sf = build_dycore.extras["quantity_factory"], build_dycore.extras["stencil_factory"]
    driver_grid = DriverGridData.new_from_metric_terms(build_dycore.extras["metric_terms"])
    schemes = [pySHiELD.PHYSICS_PACKAGES["GFS_microphysics"]]

    @dataclasses.dataclass
    class Tend:
        u_dt: object
        v_dt: object
        pt_dt: object

    def tend():
        return Tend(*(qf.zeros([X_DIM, Y_DIM, Z_DIM], "unknown") for _ in range(3)))

    phys_cfg = pySHiELD.PhysicsConfig(
        dt_atmos=config.dt_atmos, hydrostatic=config.hydrostatic, npx=config.npx, npy=config.npy, npz=config.npz,
        nwat=config.nwat, schemes=["GFS_microphysics"], do_qa=True, layout=config.layout, c2l_ord=config.c2l_ord)
    t_ref, t_mine = tend(), tend()
    p_ref = pySHiELD.PhysicsState.init_zeros(quantity_factory=qf, schemes=schemes)
    p_mine = pySHiELD.PhysicsState.init_zeros(quantity_factory=qf, schemes=schemes)
    d2p_ref = ref.DycoreToPhysics(stencil_factory=sf, quantity_factory=qf, dycore_config=config,
                                  do_dry_convective_adjust=False, dycore_only=False)
    upd_ref = ref.UpdateAtmosphereState(stencil_factory=sf, grid_data=grid_data, namelist=phys_cfg, comm=communicator,
                                        grid_info=driver_grid, state=state, quantity_factory=qf, dycore_only=False,
                                        apply_tendencies=True, tendency_state=t_ref)
    upd_mine = dropin.UpdateAtmosphereState(sf, grid_data, phys_cfg, communicator, driver_grid, mine_state, qf, False,
                                            True, t_mine, dycore=mine)
    d2p_mine = dropin.DycoreToPhysics(sf, qf, config, False, False, dycore=mine)
    dt = float(config.dt_atmos)
    d2p_ref(dycore_state=state, physics_state=p_ref, tendency_state=t_ref, timestep=dt)
    d2p_mine(mine_state, p_mine, t_mine, dt)
    rank = mpi_comm.Get_rank()
    synthesize_physics(p_ref, rank, config.npx - 1, float(grid_data.ptop))
    for n in PHY_IN:   # the same "physics result" for both
        np.asarray(getattr(p_mine, n))[...] = np.asarray(getattr(p_ref, n))
    upd_ref(dycore_state=state, phy_state=p_ref, u_dt=t_ref.u_dt, v_dt=t_ref.v_dt, pt_dt=t_ref.pt_dt, dt=dt)
    upd_mine(mine_state, p_mine, t_mine.u_dt, t_mine.v_dt, t_mine.pt_dt, dt)
    want, got = state_arrays(state), state_arrays(mine_state)
    bad = []
    for name in ("u", "v", "delp", "pt", "qvapor", "qliquid", "qice", "qrain", "qsnow", "qgraupel", "pe", "pk", "peln",
                 "ps", "ua", "va"):
        a, b = got[name], want[name]
        a, b = (a[3:15, 3:15], b[3:15, 3:15]) if name in ("pe", "pk", "peln", "ps") else (a[3:15, 3:15, :79], b[3:15, 3:15, :79])
        if not np.array_equal(a, b):
            bad.append(("coupling:" + name, float(np.max(np.abs(a - b)))))
    for n in ("qvapor", "pt", "ua", "omga"):
        if not np.array_equal(np.asarray(getattr(p_mine, n))[3:15, 3:15, :79], np.asarray(getattr(p_ref, n))[3:15, 3:15, :79]):
            bad.append(("coupling:phy." + n, 0.0))
    for n in ("u_dt", "v_dt", "pt_dt"):
        a, b = np.asarray(getattr(t_mine, n).data)[3:15, 3:15, :79], np.asarray(getattr(t_ref, n).data)[3:15, 3:15, :79]
        if not np.array_equal(a, b):
            bad.append(("coupling:tend." + n, float(np.max(np.abs(a - b)))))
    return bad


if __name__ == "__main__":
    from oracle import cpu_port

    cpu_port.load()
    res = run_ranks(6, _rank, (), timeout=3600)
    print("DROPIN_RESULT " + json.dumps(res))
    sys.exit(0 if all(r == [] for r in res) else 1)

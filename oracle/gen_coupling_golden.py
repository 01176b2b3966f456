"""TEST INFRASTRUCTURE ONLY — golden fixture for the dycore <-> physics coupling: the UNMODIFIED
reference classes `DycoreToPhysics` and `UpdateAtmosphereState` (pySHiELD/update/update_atmos_state.py:148-349, with
`ApplyPhysicsToDycore`, fv_update_phys.py:78-195, and `AGrid2DGridPhysics`, update_dwind_phys.py:165-701), constructed
the way the driver constructs them (driver/pace/driver/driver.py:519-537) and called the way its time loop calls them
(driver.py:645-661), on 6 ranks at C12 L79 in this container.

    python oracle/gen_coupling_golden.py [--raw /tmp/pace_golden_coupling]

The state comes from one reference dycore step of the baroclinic case (so that pe / peln / pk / ua / va are filled).
The physics parameterisations themselves are out of scope: what they would hand back (prsi and the physics_updated_*
fields) is synthesised here from the copied state plus seeded perturbations, including specific-humidity values below
the fill threshold and below zero so that every branch of fill_gfs_delp runs, and is stored in the fixture.

Writes tests/golden/c12_6rank_coupling.npz: for every rank r
  r<r>:in:<field>      dycore state before the coupling calls (compute domain + interface row / column / level)
  r<r>:phy:<field>     prsi and the physics_updated_* fields (what the physics returns)
  (the physics-state fields after DycoreToPhysics are checked here to be copies of the dycore fields on
   [is, ie+1] x [js, je+1] x [0, npz) and zero on the extra level; they are not stored)
  r<r>:out:<field>     dycore state and tendencies after UpdateAtmosphereState
  r<r>:grid:<field>    DriverGridData (vlon, vlat, edge_vect_*, es1_*, ew2_*), full arrays
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle.refshim.bootstrap import bootstrap  # noqa: E402

bootstrap()
import numpy as np  # noqa: E402

from oracle.refshim.proc_comm import run_ranks  # noqa: E402
from oracle.refshim.run_reference import build_dycore, default_config_kwargs, state_arrays  # noqa: E402

COPIED = ("qvapor", "qliquid", "qrain", "qsnow", "qice", "qgraupel", "qo3mr", "qsgs_tke", "qcld", "pt", "delp", "delz",
          "ua", "va", "w", "omga")
PHY_IN = ("prsi", "physics_updated_ua", "physics_updated_va", "physics_updated_pt", "physics_updated_specific_humidity",
          "physics_updated_qliquid", "physics_updated_qrain", "physics_updated_qsnow", "physics_updated_qice",
          "physics_updated_qgraupel")
# dycore fields the coupling reads (everything else of the state is left out of the fixture)
NEEDED_IN = COPIED + ("u", "v", "pe", "peln", "pk", "ps")
DRIVER_GRID = ("vlon1", "vlon2", "vlon3", "vlat1", "vlat2", "vlat3", "edge_vect_w", "edge_vect_e", "edge_vect_s",
               "edge_vect_n", "es1_1", "es1_2", "es1_3", "ew2_1", "ew2_2", "ew2_3")


def _a(q):
    return np.asarray(q.data if hasattr(q, "data") and not isinstance(q, np.ndarray) else q)


def synthesize_physics(phy, rank, nx, ptop):
    """What the physics would hand back, from the copied state.  Arrays are [i, j, k] with a 3-cell halo."""
    rng = np.random.default_rng(1234 + rank)
    c = (slice(3, 3 + nx), slice(3, 3 + nx))
    nz = _a(phy.delp).shape[2] - 1

    def noise(shape):
        return rng.random(shape) - 0.5

    shp = (nx, nx, nz)
    prsi = _a(phy.prsi)
    prsi[...] = 0.0
    prsi[c][..., 0] = ptop
    for k in range(nz):
        prsi[c][..., k + 1] = prsi[c][..., k] + _a(phy.delp)[c][..., k]
    _a(phy.physics_updated_pt)[c][..., :nz] = _a(phy.pt)[c][..., :nz] * (1.0 + 2.0e-3 * noise(shp))
    _a(phy.physics_updated_ua)[c][..., :nz] = _a(phy.ua)[c][..., :nz] + 1.5 * noise(shp)
    _a(phy.physics_updated_va)[c][..., :nz] = _a(phy.va)[c][..., :nz] + 1.5 * noise(shp)
    qv = _a(phy.qvapor)[c][..., :nz] * (1.0 + 0.05 * noise(shp))
    # a few values below the fill threshold (1e-9) and below zero, at scattered points and levels
    pick = rng.random(shp)
    qv = np.where(pick < 0.02, 1.0e-10 * rng.random(shp), qv)
    qv = np.where((pick >= 0.02) & (pick < 0.03), -1.0e-6 * rng.random(shp), qv)
    _a(phy.physics_updated_specific_humidity)[c][..., :nz] = qv
    for name, amp in (("qliquid", 2.0e-5), ("qrain", 1.0e-5), ("qsnow", 5.0e-6), ("qice", 8.0e-6), ("qgraupel", 3.0e-6)):
        base = _a(getattr(phy, name))[c][..., :nz]
        _a(getattr(phy, "physics_updated_" + name))[c][..., :nz] = base + amp * rng.random(shp) * (rng.random(shp) < 0.3)


def rank_main(mpi_comm, cfg_kwargs, out_dir):
    from oracle.refshim.bootstrap import bootstrap as bs

    bs()
    import dataclasses

    import pySHiELD
    from ndsl.grid import DriverGridData
    from pySHiELD.update import update_atmos_state

    rank = mpi_comm.Get_rank()
    dycore, state, grid_data, damping, metric_terms, config, communicator = build_dycore(mpi_comm, cfg_kwargs)
    dycore.step_dynamics(state)
    qf = build_dycore.extras["quantity_factory"]
    stencil_factory = build_dycore.extras["stencil_factory"]
    driver_grid = DriverGridData.new_from_metric_terms(metric_terms)
    schemes = [pySHiELD.PHYSICS_PACKAGES["GFS_microphysics"]]
    phy = pySHiELD.PhysicsState.init_zeros(quantity_factory=qf, schemes=schemes)

    @dataclasses.dataclass
    class Tend:
        u_dt: object
        v_dt: object
        pt_dt: object

    from ndsl.constants import X_DIM, Y_DIM, Z_DIM

    tend = Tend(*(qf.zeros([X_DIM, Y_DIM, Z_DIM], "unknown") for _ in range(3)))
    phys_cfg = pySHiELD.PhysicsConfig(
        dt_atmos=config.dt_atmos, hydrostatic=config.hydrostatic, npx=config.npx, npy=config.npy, npz=config.npz,
        nwat=config.nwat, schemes=["GFS_microphysics"], do_qa=True, layout=config.layout, c2l_ord=config.c2l_ord)
    d2p = update_atmos_state.DycoreToPhysics(
        stencil_factory=stencil_factory, quantity_factory=qf, dycore_config=config,
        do_dry_convective_adjust=False, dycore_only=False)
    upd = update_atmos_state.UpdateAtmosphereState(
        stencil_factory=stencil_factory, grid_data=grid_data, namelist=phys_cfg, comm=communicator,
        grid_info=driver_grid, state=state, quantity_factory=qf, dycore_only=False, apply_tendencies=True,
        tendency_state=tend)

    nx = config.npx - 1
    out = {}
    pre = {k: v.copy() for k, v in state_arrays(state).items()}
    import time

    t0 = time.time()
    d2p(dycore_state=state, physics_state=phy, tendency_state=tend, timestep=float(config.dt_atmos))
    t_d2p = time.time() - t0
    for n in COPIED:
        out["copy:" + n] = _a(getattr(phy, n)).copy()
    synthesize_physics(phy, rank, nx, float(grid_data.ptop))
    for n in PHY_IN:
        out["phy:" + n] = _a(getattr(phy, n)).copy()
    t0 = time.time()
    upd(dycore_state=state, phy_state=phy, u_dt=tend.u_dt, v_dt=tend.v_dt, pt_dt=tend.pt_dt, dt=float(config.dt_atmos))
    t_upd = time.time() - t0
    post = state_arrays(state)
    for k, v in pre.items():
        out["in:" + k] = v
    for k, v in post.items():
        if not np.array_equal(v, pre[k], equal_nan=True):
            out["out:" + k] = v.copy()
    for n in ("u_dt", "v_dt", "pt_dt"):
        out["out:tend." + n] = _a(getattr(tend, n)).copy()
    for n in DRIVER_GRID:
        out["grid:" + n] = _a(getattr(driver_grid, n)).copy()
    # second case: dycore_only=True, apply_tendencies=False (driver.py:160-165 with no dry convective adjustment): the
    # call reduces to fill_gfs_delp on the dycore's own qvapor, given scattered values below the threshold and below zero
    upd_only = update_atmos_state.UpdateAtmosphereState(
        stencil_factory=stencil_factory, grid_data=grid_data, namelist=phys_cfg, comm=communicator,
        grid_info=driver_grid, state=state, quantity_factory=qf, dycore_only=True, apply_tendencies=False,
        tendency_state=tend)
    rng = np.random.default_rng(4321 + rank)
    qv = _a(state.qvapor)
    c = (slice(3, 3 + nx), slice(3, 3 + nx), slice(0, config.npz))
    pick = rng.random(qv[c].shape)
    qv[c] = np.where(pick < 0.03, 1.0e-10 * rng.random(pick.shape), qv[c])
    qv[c] = np.where((pick >= 0.03) & (pick < 0.05), -2.0e-6 * rng.random(pick.shape), qv[c])
    out["in2:qvapor"] = qv.copy()
    out["in2:delp"] = _a(state.delp).copy()
    upd_only(dycore_state=state, phy_state=phy, u_dt=tend.u_dt, v_dt=tend.v_dt, pt_dt=tend.pt_dt, dt=float(config.dt_atmos))
    out["out2:qvapor"] = _a(state.qvapor).copy()
    assert np.array_equal(_a(state.delp), out["in2:delp"])
    os.makedirs(out_dir, exist_ok=True)
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), **out)
    return {"rank": rank, "changed": sorted(k for k in out if k.startswith("out:")), "dycore_to_physics_s": t_d2p,
            "update_atmos_state_s": t_upd}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--raw", default="/tmp/pace_golden_coupling")
    ap.add_argument("--time-only", action="store_true", help="print the reference's call times, leave the fixture alone")
    ap.add_argument("--warm", action="store_true",
                    help="first compile every stencil in ONE process (NullComm), so that the six ranks read a warm GT4Py "
                         "cache instead of racing to write it")
    args = ap.parse_args()
    cfg = default_config_kwargs(npx=13, n_split=1)
    if args.warm:
        from ndsl.comm.null_comm import NullComm

        rank_main(NullComm(rank=0, total_ranks=6, fill_value=0.0), cfg, args.raw + "_warm")
    res = run_ranks(6, rank_main, (cfg, args.raw), timeout=3600)
    for r in res:
        print(r)
    if args.time_only:
        import json

        print("REFERENCE_COUPLING " + json.dumps({
            "what": "UNMODIFIED reference, numpy backend, C12 L79, 6 processes: first call of DycoreToPhysics and of "
                    "UpdateAtmosphereState after one dycore step (max over ranks)",
            "dycore_to_physics_s": max(r["dycore_to_physics_s"] for r in res),
            "update_atmos_state_s": max(r["update_atmos_state_s"] for r in res)}))
        return
    nx = 12
    merged = {}
    for r in range(6):
        z = np.load(os.path.join(args.raw, "rank%d.npz" % r))
        for k in z.files:
            a = z[k]
            kind, name = k.split(":", 1)
            if kind == "copy":
                # DycoreToPhysics is a plain copy on [is, ie+1] x [js, je+1] x [0, npz): checked here, not stored
                src = z["in:" + name]
                assert np.array_equal(a[3:3 + nx + 1, 3:3 + nx + 1, :-1], src[3:3 + nx + 1, 3:3 + nx + 1, :-1]), k
                assert not a[..., -1].any(), k
                continue
            if kind == "in" and name not in NEEDED_IN:
                continue
            if kind != "grid" and a.ndim >= 2:
                # compute domain plus the closing row / column; halos are exchanged by the code under test
                a = a[3:3 + nx + 1, 3:3 + nx + 1]
            merged["r%d:%s" % (r, k)] = np.ascontiguousarray(a)
    dst = os.path.join(ROOT, "tests", "golden", "c12_6rank_coupling.npz")
    np.savez_compressed(dst, **merged)
    print("wrote", dst, "%.1f MB" % (os.path.getsize(dst) / 1e6))


if __name__ == "__main__":
    main()

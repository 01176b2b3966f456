"""Drop-in for `pyFV3.DynamicalCore`: same constructor arguments, same `step_dynamics(state, timer)` call, same state
object in and out (pyFV3/stencils/fv_dynamics.py:89-133, 445-460) — the numerics run in libpace_b200.

    core = pace_b200.dropin.DynamicalCore(comm, grid_data, stencil_factory, quantity_factory, damping_coefficients,
                                          config, phis, state, timestep)
    core.step_dynamics(state, timer)

Two ways to hold the cube:
  * like the reference, ONE subdomain per process: `comm` is the reference's CubedSphereCommunicator (its `.comm` is
    an mpi4py-style object: Get_rank / Get_size / Isend / Irecv) and the other arguments are this rank's objects;
    halo messages then travel through that communicator (MpiStyleComm below);
  * ALL subdomains in one process (one GPU): pass lists over the cube ranks for grid_data, damping_coefficients, phis
    and state, and `comm=None`.
`stencil_factory` and `quantity_factory` are accepted for signature compatibility and not used: nothing is compiled
at construction time and the device buffers are owned by the engine.  The state is copied in before and out after
every call (the reference layout is (i, j, k) numpy/cupy storage, the library's is [k][j][i]); a driver that keeps
the state on the device between steps uses DycoreEngine directly.
"""
import numpy as np

from .engine import STATE_2D, STATE_3D, DycoreEngine

GRID_NAMES = [
    "lon", "lat", "lon_agrid", "lat_agrid", "fC", "fC_agrid", "area", "area_64", "rarea", "rarea_c", "dx", "dy",
    "dxc", "dyc", "dxa", "dya", "rdx", "rdy", "rdxc", "rdyc", "rdxa", "rdya", "a11", "a12", "a21", "a22",
    "edge_w", "edge_e", "edge_s", "edge_n", "p", "ak", "bk", "dp_ref", "cosa", "cosa_u", "cosa_v", "cosa_s",
    "sina_u", "sina_v", "rsina", "rsin_u", "rsin_v", "rsin2", "sin_sg1", "sin_sg2", "sin_sg3", "sin_sg4",
    "cos_sg1", "cos_sg2", "cos_sg3", "cos_sg4", "ee1", "ee2", "es1", "ew2",
]
CONFIG_NAMES = [
    "n_split", "k_split", "hord_dp", "hord_mt", "hord_tm", "hord_tr", "hord_vt", "kord_mt", "kord_tm", "kord_tr",
    "kord_wz", "nord", "n_sponge", "nf_omega", "grid_type", "do_sat_adj", "do_vort_damp", "rf_fast", "hydrostatic",
    "convert_ke", "fill", "adiabatic", "do_qa", "use_old_omega", "do_f3d", "inline_q", "nwat", "dt_atmos", "a_imp",
    "beta", "p_fac", "d_con", "d2_bg", "d2_bg_k1", "d2_bg_k2", "d4_bg", "dddmp", "delt_max", "ke_bg", "rf_cutoff", "tau",
    "vtdm4", "d_ext", "consv_te",
]


def _arr(x):
    """numpy view (numpy storage) or host copy (cupy storage) of a Quantity / gt4py storage / array-like."""
    d = x if isinstance(x, np.ndarray) else getattr(x, "data", x)   # (ndarray.data is a memoryview)
    if hasattr(d, "get"):  # cupy
        d = d.get()
    return np.asarray(d)


def _write_back(dst, a):
    """Store the result `a` into the caller's field object: in place for numpy storage, through a host-to-device copy
    for cupy storage; anything else is refused rather than silently dropped."""
    d = dst if isinstance(dst, np.ndarray) else getattr(dst, "data", dst)
    if isinstance(d, np.ndarray):
        if d.ndim == 3 and d.shape[2] != a.shape[2]:
            a = a[:, :, :d.shape[2]]
        d[...] = a
    elif hasattr(d, "get") and hasattr(d, "set"):  # cupy.ndarray
        if d.ndim == 3 and d.shape[2] != a.shape[2]:
            a = a[:, :, :d.shape[2]]
        d.set(np.ascontiguousarray(a, dtype=d.dtype))
    else:
        raise TypeError("pace_b200.dropin: cannot write results into storage of type %s" % type(d).__name__)


def grid_arrays(grid_data, damping_coefficients):
    """The metric terms and damping coefficients the step reads, keyed the way libpace_b200 names them."""
    out = {"grid." + n: _arr(getattr(grid_data, n)) for n in GRID_NAMES if hasattr(grid_data, n)}
    out["grid.ptop"] = np.float64(grid_data.ptop)
    for n in ("del6_u", "del6_v", "divg_u", "divg_v"):
        out["damp." + n] = _arr(getattr(damping_coefficients, n))
    out["damp.da_min"] = np.float64(damping_coefficients.da_min)
    out["damp.da_min_c"] = np.float64(damping_coefficients.da_min_c)
    return out


def config_values(config):
    """CONFIG_NAMES and the SatAdjustConfig values from a DynamicalCoreConfig or an AcousticDynamicsConfig (whose
    options sit in the nested riemann / d_grid_shallow_water objects, pyFV3/_config.py:59-140)."""
    places = [config] + [getattr(config, n) for n in ("acoustic_dynamics", "riemann", "d_grid_shallow_water", "remapping")
                         if hasattr(config, n)]
    ad = getattr(config, "acoustic_dynamics", None)
    if ad is not None:
        places += [getattr(ad, n) for n in ("riemann", "d_grid_shallow_water") if hasattr(ad, n)]
    cfg = {}
    for n in CONFIG_NAMES:
        for pl in places:
            try:
                cfg[n] = getattr(pl, n)
                break
            except (AttributeError, NotImplementedError):
                continue
    sat = {}
    src = getattr(config, "sat_adjust", config)
    from ._lib import SatAdjConfig

    for n in SatAdjConfig.DEFAULTS:
        if hasattr(src, n):
            sat[n] = getattr(src, n)
    return cfg, sat


class MpiStyleComm:
    """Halo-message transport over an mpi4py-style communicator (the object inside the reference's Communicator)."""

    def __init__(self, mpi_comm):
        self.mpi = mpi_comm
        self.rank, self.world_size = mpi_comm.Get_rank(), mpi_comm.Get_size()
        self.group = None
        self._tag = 7100

    @staticmethod
    def _wait_all(reqs):
        for r in reqs:
            (r.wait if hasattr(r, "wait") else r.Wait)()

    def exchange(self, sends, recvs):
        self._tag += 1
        tag = self._tag
        rreqs = [self.mpi.Irecv(t.numpy(), source=p, tag=tag) for p, t in sorted(recvs.items())]
        # the send requests are kept and waited for too: the send buffer is a slice of a persistent pack buffer that the
        # next update overwrites
        sreqs = [self.mpi.Isend(t.numpy(), dest=p, tag=tag) for p, t in sorted(sends.items())]
        return lambda: self._wait_all(rreqs + [r for r in sreqs if r is not None])


class _StagedComm(MpiStyleComm):
    """Same, for device tensors when the communicator is not CUDA-aware: messages are staged through host memory."""

    def exchange(self, sends, recvs):
        import torch

        self._tag += 1
        tag = self._tag
        host = {p: torch.empty(t.shape, dtype=t.dtype) for p, t in recvs.items()}
        rreqs = [self.mpi.Irecv(host[p].numpy(), source=p, tag=tag) for p in sorted(recvs)]
        staged = {p: t.cpu() for p, t in sends.items()}  # kept alive until the sends have completed
        sreqs = [self.mpi.Isend(staged[p].numpy(), dest=p, tag=tag) for p in sorted(staged)]

        def wait():
            self._wait_all(rreqs + [r for r in sreqs if r is not None])
            for p, t in recvs.items():
                t.copy_(host[p])
            staged.clear()

        return wait


# savepoint -> {checkpointer keyword: library field name}: the variables the reference hands to its checkpointer
# (dyn_core.py:641-701, fv_dynamics.py:342-443); "^T" marks the [x, z, y] transposes of pe / peln
_CSW = dict(delpd="delp", ptd="pt", ud="u", vd="v", wd="w", ucd="uc", vcd="vc", uad="ua", vad="va", utd="ad.ut",
            vtd="ad.vt", divgdd="ad.divgd")
_DSW = dict(ucd="uc", vcd="vc", wd="w", delpcd="ad.vt", delpd="delp", ud="u", vd="v", ptd="pt", uad="ua", vad="va",
            divgdd="ad.divgd", xfxd="ad.xfx", yfxd="ad.yfx", mfxd="mfxd", mfyd="mfyd")
_FVD = dict(u="u", v="v", w="w", delz="delz", va="va", uc="uc", vc="vc", qvapor="qvapor")
_TRC = dict(dp1="dyn.dp1", mfxd="mfxd", mfyd="mfyd", cxd="cxd", cyd="cyd")
_RMO = dict(pt="pt", delp="delp", delz="delz", peln="peln^T", u="u", v="v", w="w", cappa="ad.cappa", pkz="pkz", pk="pk",
            pe="pe^T", dp1="dyn.dp1")
_RMI = dict(pt="pt", delp="delp", delz="delz", peln="peln^T", u="u", v="v", w="w", ua="ua", va="va", cappa="ad.cappa",
            pk="pk", pe="pe^T", phis="phis", ps="ps", wsd="ad.wsd", omga="omga", dp1="dyn.dp1")
SAVEPOINTS = {"C_SW-In": _CSW, "C_SW-Out": _CSW, "D_SW-In": dict(_DSW, zhd="ad.zh"), "D_SW-Out": _DSW,
              "FVDynamics-In": _FVD, "FVDynamics-Out": _FVD, "Tracer2D1L-In": _TRC, "Tracer2D1L-Out": _TRC,
              "Remapping-In": _RMI, "Remapping-Out": _RMO}


class _Bridge:
    """What DynamicalCore and AcousticDynamics share: the engine, the copies between the caller's state objects
    (reference layout (i, j, k), numpy or cupy storage) and the device buffers, and the checkpointer hook."""

    def _build(self, comm, grid_data, damping_coefficients, config, phis, n_tile, layout, npz, dt_atmos, checkpointer,
               device, _lib_handle, **cfg_over):
        many = isinstance(grid_data, (list, tuple))
        grids = [grid_arrays(g, d) for g, d in zip(grid_data, damping_coefficients)] if many else [
            grid_arrays(grid_data, damping_coefficients)]
        if layout[0] != layout[1]:
            raise NotImplementedError("pace_b200: only square layouts are supported")
        total = 6 * layout[0] * layout[1]
        cfg, sat = config_values(config)
        cfg["ks"] = int(getattr(grid_data[0] if many else grid_data, "ks", 0))
        if dt_atmos is not None:
            cfg["dt_atmos"] = dt_atmos
        cfg.update(cfg_over)
        transport = None
        if many:
            if len(grids) != total:
                raise ValueError("expected the %d subdomains of the cube, got %d" % (total, len(grids)))
        else:
            mpi = getattr(comm, "comm", comm)  # a Communicator wraps the mpi4py-style object as .comm
            if mpi.Get_size() != total:
                raise ValueError("one process per subdomain needs %d ranks, the communicator has %d" % (total, mpi.Get_size()))
            transport = (MpiStyleComm if str(device).startswith("cpu") else _StagedComm)(mpi)
        self.engine = DycoreEngine(n_tile, layout[0], npz, grids, comm=transport, device=device,
                                   _lib_handle=_lib_handle, sat_adjust=sat, **cfg)
        self._many = many
        # the reference's constructors derive zs from phis and keep it (dyn_core.py:489-496)
        self._phis = [_arr(p) for p in phis] if many else [_arr(phis)]
        self.engine.core.put("ad.zs", [np.asarray(p, dtype=np.float64) / 9.80665 for p in self._phis])
        self.engine.core.put("phis", self._phis)
        self._set_checkpointer(checkpointer)

    def _set_checkpointer(self, checkpointer):
        self.checkpointer = checkpointer
        self.call_checkpointer = checkpointer is not None
        if checkpointer is None:
            self.engine.set_stage_callback(None)
            return
        core = self.engine.core

        def hook(savepoint):
            core.sync()
            kw = {}
            for key, name in SAVEPOINTS[savepoint].items():
                tr = name.endswith("^T")
                arrs = core.get(name[:-2] if tr else name)
                if tr:
                    arrs = [a.transpose(0, 2, 1) for a in arrs]
                kw[key] = arrs if self._many else arrs[0]
            checkpointer(savepoint, **kw)

        self.engine.set_stage_callback(hook)

    @staticmethod
    def _fields(state):
        return {n: getattr(state, n) for n in STATE_3D + STATE_2D if hasattr(state, n)}

    def _upload(self, state):
        states = list(state) if self._many else [state]
        fields = [self._fields(s) for s in states]
        names = [n for n in fields[0] if n != "phis"]
        for n in names:
            self.engine.core.put(n, [_arr(f[n]) for f in fields])
        return fields, names

    def _download(self, fields, names):
        self.engine.core.sync()
        for n in names:
            for f, a in zip(fields, self.engine.core.get(n)):
                _write_back(f[n], a)


class AcousticDynamics(_Bridge):
    """Drop-in for `pyFV3.stencils.dyn_core.AcousticDynamics`: the reference's constructor arguments
    (dyn_core.py:367-382) and `__call__(state, timestep, n_map=1)` (:714-719).  `cappa` is the reference's public
    temporary (dyn_core.py:468): a list of (i, j, k) arrays fetched from the device on access."""

    def __init__(self, comm, stencil_factory, quantity_factory, grid_data, damping_coefficients, grid_type, nested,
                 stretched_grid, config, phis, wsd, state, checkpointer=None, *, device="cuda:0", _lib_handle=None,
                 _engine_of=None):
        del stencil_factory, wsd, state  # signature compatibility: nothing is compiled, temporaries live on the device
        if _engine_of is not None:  # the instance a DynamicalCore exposes as .acoustic_dynamics shares its engine
            self.engine, self._many = _engine_of.engine, _engine_of._many
            self.checkpointer, self.call_checkpointer = _engine_of.checkpointer, _engine_of.call_checkpointer
            return
        if nested or stretched_grid:
            raise NotImplementedError("pace_b200: nested / stretched grids are not supported")
        g0 = grid_data[0] if isinstance(grid_data, (list, tuple)) else grid_data
        nx = _arr(g0.rarea).shape[0] - 7
        layout = (1, 1)
        part = getattr(getattr(comm, "partitioner", None), "tile", None)
        if part is not None:
            layout = tuple(part.layout)
        npz = int(getattr(getattr(quantity_factory, "sizer", None), "nz", 0)) or _arr(g0.ak).shape[0] - 1
        self._build(comm, grid_data, damping_coefficients, config, phis, nx * layout[0], layout, npz, None,
                    checkpointer, device, _lib_handle, grid_type=int(grid_type))

    @property
    def cappa(self):
        arrs = self.engine.core.get("ad.cappa")
        return arrs if self._many else arrs[0]

    def __call__(self, state, timestep, n_map=1):
        fields, names = self._upload(state)
        self.engine.acoustic_dynamics(float(timestep), n_map=int(n_map))
        self._download(fields, names)


class TracerAdvection:
    """`DynamicalCore.tracer_advection` of the drop-in: TracerAdvection.__call__ (tracer_2d_1l.py:289-417) on the
    engine's device state; the reference's arguments (tracers, dp1, mass fluxes, Courant numbers) are the engine's
    own fields of the same names, so they are accepted and not read."""

    def __init__(self, owner):
        self._owner = owner

    def __call__(self, *args, **kwargs):
        self._owner.engine.tracer_advection()


class DynamicalCore(_Bridge):
    """See the module docstring.  Raises the reference's own errors for the options it does not implement
    (DycoreEngine._check_config)."""

    def __init__(self, comm, grid_data, stencil_factory, quantity_factory, damping_coefficients, config, phis, state,
                 timestep, checkpointer=None, *, device="cuda:0", _lib_handle=None):
        del stencil_factory, quantity_factory  # signature compatibility only
        self.config = config
        layout = tuple(getattr(config, "layout", (1, 1)))
        dt = timestep.total_seconds() if hasattr(timestep, "total_seconds") else float(timestep)
        if hasattr(config, "dt_atmos") and float(config.dt_atmos) != dt:
            raise ValueError("timestep (%s s) differs from config.dt_atmos (%s s)" % (dt, config.dt_atmos))
        self._build(comm, grid_data, damping_coefficients, config, phis, int(config.npx) - 1, layout, int(config.npz),
                    dt, checkpointer, device, _lib_handle)
        # the sub-objects a user of the reference reaches for (fv_dynamics.py:231, 268, 299)
        self.acoustic_dynamics = AcousticDynamics(None, None, None, None, None, None, None, None, None, None, None, None,
                                                  _engine_of=self)
        self.tracer_advection = TracerAdvection(self)

    @property
    def _cappa(self):
        return self.acoustic_dynamics.cappa

    def step_dynamics(self, state, timer=None):
        """Step the model state forward by one timestep, in place (fv_dynamics.py:445-460)."""
        fields, names = self._upload(state)
        self.engine.step_dynamics(timer)
        self._download(fields, names)

    def __call__(self, *args, **kwargs):
        return self.step_dynamics(*args, **kwargs)


# ---- dycore <-> physics coupling of the driver's time loop (driver/pace/driver/driver.py:519-537, 645-661) ----------
def _driver_grid_arrays(grid_info):
    from .engine import DRIVER_GRID

    return {n: _arr(getattr(grid_info, n)) for n in DRIVER_GRID}


class DycoreToPhysics:
    """Drop-in for `pySHiELD.update.update_atmos_state.DycoreToPhysics` (update_atmos_state.py:148-231): the reference's
    constructor arguments plus `dycore=` (the pace_b200.dropin.DynamicalCore whose device state is the source), and
    `__call__(dycore_state, physics_state, tendency_state=None, timestep=None)`.  The sixteen fields are copied on the
    device and written into `physics_state`; the dry convective adjustment (fv_subgridz) is not implemented."""

    def __init__(self, stencil_factory, quantity_factory, dycore_config, do_dry_convective_adjust, dycore_only, *,
                 dycore):
        del stencil_factory, quantity_factory, dycore_config
        if do_dry_convective_adjust:
            # the reference's own call path raises for this option (TypeError at fv_subgridz.py:874: DycoreState.pe is a
            # Quantity and is indexed there; oracle/probe_reference_subgridz.py), so there is nothing to reproduce
            raise NotImplementedError("pace_b200: do_dry_convective_adjust (fv_subgridz) is not implemented")
        self._dycore_only = bool(dycore_only)
        self._owner = dycore

    def __call__(self, dycore_state, physics_state, tendency_state=None, timestep=None):
        from .engine import COUPLING_COPIED

        if self._dycore_only:
            return
        own, eng = self._owner, self._owner.engine
        if not hasattr(eng, "_origin_i"):
            raise RuntimeError("DycoreToPhysics: build the UpdateAtmosphereState of this dycore first (it registers the "
                               "coupling fields)")
        own._upload(dycore_state)
        eng.dycore_to_physics()
        eng.core.sync()
        states = list(physics_state) if own._many else [physics_state]
        for n in COUPLING_COPIED:
            for s, a in zip(states, eng.core.get("phy." + n)):
                _write_back(getattr(s, n), a)


class UpdateAtmosphereState:
    """Drop-in for `pySHiELD.update.update_atmos_state.UpdateAtmosphereState` (update_atmos_state.py:234-349): the
    reference's constructor arguments plus `dycore=`, and `__call__(dycore_state, phy_state, u_dt, v_dt, pt_dt, dt)`.
    `grid_info` is the reference's DriverGridData (a list over the cube ranks when the dycore holds the whole cube)."""

    def __init__(self, stencil_factory, grid_data, namelist, comm, grid_info, state, quantity_factory, dycore_only,
                 apply_tendencies, tendency_state, *, dycore):
        del stencil_factory, grid_data, comm, state, quantity_factory, tendency_state
        self.namelist = namelist
        self._dycore_only, self._apply_tendencies = bool(dycore_only), bool(apply_tendencies)
        self._owner = dycore
        infos = list(grid_info) if dycore._many else [grid_info]
        if int(getattr(infos[0], "grid_type", 0)) > 3:
            raise NotImplementedError("pace_b200: doubly periodic grids (grid_type > 3) are not supported")
        dycore.engine.setup_coupling([_driver_grid_arrays(g) for g in infos])

    def __call__(self, dycore_state, phy_state, u_dt, v_dt, pt_dt, dt):
        from .engine import COUPLING_PHYSICS_OUT

        own, eng = self._owner, self._owner.engine
        core = eng.core
        fields, names = own._upload(dycore_state)
        phys = list(phy_state) if own._many else [phy_state]
        tend = {"u_dt": u_dt, "v_dt": v_dt, "pt_dt": pt_dt}
        if not self._dycore_only:
            for n in COUPLING_PHYSICS_OUT + ["ua", "va", "pt"]:
                core.put("phy." + n, [_arr(getattr(s, n)) for s in phys])
        for n, q in tend.items():
            core.put("tend." + n, [_arr(x) for x in q] if own._many else [_arr(q)])
        eng.update_atmos_state(float(dt), self._dycore_only, self._apply_tendencies)
        own._download(fields, names)
        for n, q in tend.items():
            for dst, a in zip(list(q) if own._many else [q], core.get("tend." + n)):
                _write_back(dst, a)
        if not self._dycore_only:   # fill_gfs_delp adjusts the physics' specific humidity in place
            for s, a in zip(phys, core.get("phy.physics_updated_specific_humidity")):
                _write_back(s.physics_updated_specific_humidity, a)

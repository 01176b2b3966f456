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
    """numpy view of a Quantity / gt4py storage / array-like."""
    d = getattr(x, "data", x)
    if hasattr(d, "get"):  # cupy
        d = d.get()
    return np.asarray(d)


def grid_arrays(grid_data, damping_coefficients):
    """The metric terms and damping coefficients the step reads, keyed the way libpace_b200 names them."""
    out = {"grid." + n: _arr(getattr(grid_data, n)) for n in GRID_NAMES}
    out["grid.ptop"] = np.float64(grid_data.ptop)
    for n in ("del6_u", "del6_v", "divg_u", "divg_v"):
        out["damp." + n] = _arr(getattr(damping_coefficients, n))
    out["damp.da_min"] = np.float64(damping_coefficients.da_min)
    out["damp.da_min_c"] = np.float64(damping_coefficients.da_min_c)
    return out


class MpiStyleComm:
    """Halo-message transport over an mpi4py-style communicator (the object inside the reference's Communicator)."""

    def __init__(self, mpi_comm):
        self.mpi = mpi_comm
        self.rank, self.world_size = mpi_comm.Get_rank(), mpi_comm.Get_size()
        self.group = None
        self._tag = 7100

    def exchange(self, sends, recvs):
        self._tag += 1
        tag = self._tag
        reqs = [self.mpi.Irecv(t.numpy(), source=p, tag=tag) for p, t in sorted(recvs.items())]
        for p, t in sorted(sends.items()):
            self.mpi.Isend(t.numpy(), dest=p, tag=tag)
        return lambda: [r.wait() for r in reqs]


class _StagedComm(MpiStyleComm):
    """Same, for device tensors when the communicator is not CUDA-aware: messages are staged through host memory."""

    def exchange(self, sends, recvs):
        import torch

        self._tag += 1
        tag = self._tag
        host = {p: torch.empty(t.shape, dtype=t.dtype) for p, t in recvs.items()}
        reqs = [self.mpi.Irecv(host[p].numpy(), source=p, tag=tag) for p in sorted(recvs)]
        for p, t in sorted(sends.items()):
            self.mpi.Isend(t.cpu().numpy(), dest=p, tag=tag)

        def wait():
            for r in reqs:
                r.wait()
            for p, t in recvs.items():
                t.copy_(host[p])

        return wait


class DynamicalCore:
    """See the module docstring.  Raises the reference's own errors for the options it does not implement
    (DycoreEngine._check_config)."""

    def __init__(self, comm, grid_data, stencil_factory, quantity_factory, damping_coefficients, config, phis, state,
                 timestep, checkpointer=None, *, device="cuda:0", _lib_handle=None):
        del stencil_factory, quantity_factory  # signature compatibility only
        self.config = config
        self.checkpointer = checkpointer
        many = isinstance(grid_data, (list, tuple))
        grids = [grid_arrays(g, d) for g, d in zip(grid_data, damping_coefficients)] if many else [
            grid_arrays(grid_data, damping_coefficients)]
        layout = tuple(getattr(config, "layout", (1, 1)))
        if layout[0] != layout[1]:
            raise NotImplementedError("pace_b200: only square layouts are supported")
        n_tile = int(config.npx) - 1
        total = 6 * layout[0] * layout[1]
        cfg = {n: getattr(config, n) for n in CONFIG_NAMES if hasattr(config, n)}
        cfg["ks"] = int(getattr(grid_data[0] if many else grid_data, "ks", 0))
        dt = timestep.total_seconds() if hasattr(timestep, "total_seconds") else float(timestep)
        if "dt_atmos" in cfg and float(cfg["dt_atmos"]) != dt:
            raise ValueError("timestep (%s s) differs from config.dt_atmos (%s s)" % (dt, cfg["dt_atmos"]))
        cfg["dt_atmos"] = dt
        transport = None
        if many:
            if len(grids) != total:
                raise ValueError("expected the %d subdomains of the cube, got %d" % (total, len(grids)))
        else:
            mpi = getattr(comm, "comm", comm)  # a Communicator wraps the mpi4py-style object as .comm
            if mpi.Get_size() != total:
                raise ValueError("one process per subdomain needs %d ranks, the communicator has %d" % (total, mpi.Get_size()))
            transport = (MpiStyleComm if str(device).startswith("cpu") else _StagedComm)(mpi)
        self.engine = DycoreEngine(n_tile, layout[0], int(config.npz), grids, comm=transport, device=device,
                                   _lib_handle=_lib_handle, **cfg)
        self._many = many
        # the reference's constructor derives zs from phis and keeps it (dyn_core.py:489-496)
        self._phis = [_arr(p) for p in phis] if many else [_arr(phis)]
        self.engine.core.put("ad.zs", [np.asarray(p, dtype=np.float64) / 9.80665 for p in self._phis])
        self.engine.core.put("phis", self._phis)

    @staticmethod
    def _fields(state):
        return {n: getattr(state, n) for n in STATE_3D + STATE_2D if hasattr(state, n)}

    def step_dynamics(self, state, timer=None):
        """Step the model state forward by one timestep, in place (fv_dynamics.py:445-460)."""
        states = list(state) if self._many else [state]
        if timer is not None:
            timer.start("DynCore")
        fields = [self._fields(s) for s in states]
        names = [n for n in fields[0] if n != "phis"]
        for n in names:
            self.engine.core.put(n, [_arr(f[n]) for f in fields])
        self.engine.step_dynamics()
        self.engine.core.sync()
        for n in names:
            for f, a in zip(fields, self.engine.core.get(n)):
                dst = _arr(f[n])
                if dst.ndim == 3 and dst.shape[2] != a.shape[2]:
                    a = a[:, :, :dst.shape[2]]
                dst[...] = a
        if timer is not None:
            timer.stop("DynCore")

    def __call__(self, *args, **kwargs):
        return self.step_dynamics(*args, **kwargs)

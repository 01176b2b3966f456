"""Diagnostics / restart egress: the state leaves the device asynchronously, so that the output
cadence does not stall the time loop.

Reference: driver/pace/driver/diagnostics.py:142-249 (`MonitorDiagnostics.store`: plain names, `ZSelect` level
selections, `column_integrated_<tracer>`), driver/pace/driver/state.py:77-123 (restart files of the DycoreState).

An output step costs the compute stream only the kernels that take a snapshot: `pace_b200_egress_pack` writes every
requested field as a dense array in the reference's (i, j, k) order into a staging buffer (and
`pace_b200_column_integral` the derived fields).  A side stream then copies the staging buffers to pinned host memory
while the following steps run, and a host thread hands the arrays to the sink once the copies have landed.  With
`depth` slots (default 2) the time loop only waits when the sink is more than `depth` output steps behind.

`MonitorDiagnostics` is the drop-in for the reference class of that name: same constructor arguments (plus the dycore
whose device state is the source) and `store(time, state)`, `store_grid`, `cleanup`; the monitor is any object with the
reference's Monitor interface (`store`, `store_constant`, `cleanup` — ndsl's ZarrMonitor / NetCDFMonitor where zarr /
netCDF4 are installed).
"""
import ctypes
import queue
import threading

import numpy as np

# DycoreState metadata (pyFV3/dycore_state.py): dims and units of the fields a diagnostics list can name
FIELD_META = {
    "u": (("x", "y_interface", "z"), "m/s"), "v": (("x_interface", "y", "z"), "m/s"), "w": (("x", "y", "z"), "m/s"),
    "ua": (("x", "y", "z"), "m/s"), "va": (("x", "y", "z"), "m/s"), "uc": (("x_interface", "y", "z"), "m/s"),
    "vc": (("x", "y_interface", "z"), "m/s"), "delp": (("x", "y", "z"), "Pa"), "delz": (("x", "y", "z"), "m"),
    "ps": (("x", "y"), "Pa"), "pe": (("x", "y", "z_interface"), "Pa"), "pt": (("x", "y", "z"), "degK"),
    "peln": (("x", "y", "z_interface"), "ln(Pa)"), "pk": (("x", "y", "z_interface"), "unknown"),
    "pkz": (("x", "y", "z"), "unknown"), "qvapor": (("x", "y", "z"), "kg/kg"), "qliquid": (("x", "y", "z"), "kg/kg"),
    "qice": (("x", "y", "z"), "kg/kg"), "qrain": (("x", "y", "z"), "kg/kg"), "qsnow": (("x", "y", "z"), "kg/kg"),
    "qgraupel": (("x", "y", "z"), "kg/kg"), "qo3mr": (("x", "y", "z"), "kg/kg"),
    "qsgs_tke": (("x", "y", "z"), "m**2/s**2"), "qcld": (("x", "y", "z"), ""), "q_con": (("x", "y", "z"), "kg/kg"),
    "omga": (("x", "y", "z"), "Pa/s"), "mfxd": (("x_interface", "y", "z"), "unknown"),
    "mfyd": (("x", "y_interface", "z"), "unknown"), "cxd": (("x_interface", "y", "z"), ""),
    "cyd": (("x", "y_interface", "z"), ""), "diss_estd": (("x", "y", "z"), "unknown"), "phis": (("x", "y"), "m^2 s^-2"),
}


class _View:
    def __init__(self, data):
        self._data = data

    def __getitem__(self, key):
        return self._data[key]


class EgressQuantity:
    """What a monitor is handed: the attributes of ndsl.Quantity that monitors read (`view[:]`, `data`, `dims`,
    `units`, `extent`, `origin`, `attrs`).  The data is the compute domain itself (no halo), in host memory."""

    def __init__(self, data, dims, units):
        self.data = data
        self.dims = tuple(dims)
        self.units = units
        self.origin = (0,) * data.ndim
        self.extent = tuple(data.shape)
        self.attrs = {"units": units}
        self.view = _View(data)
        self.np = np


class _Item:
    __slots__ = ("out", "kind", "src", "level", "xi", "yi", "nk", "dims", "units", "aux")

    def __init__(self, out, kind, src, level, xi, yi, nk, dims, units, aux=None):
        self.out, self.kind, self.src, self.level, self.xi, self.yi, self.nk = out, kind, src, level, xi, yi, nk
        self.dims, self.units, self.aux = dims, units, aux


class AsyncEgress:
    """names: state fields; z_select: list of (level, [names]); derived_names: "column_integrated_<tracer>".
    sink(time, fields): called on a host thread with dict name -> list over the resident subdomains of
    EgressQuantity (arrays are views of pinned buffers that are reused `depth` output steps later: a sink that keeps
    data must copy it)."""

    def __init__(self, engine, names=(), derived_names=(), z_select=(), sink=None, depth=2):
        import torch

        self.torch = torch
        self.engine, self.core = engine, engine.core
        core = self.core
        self.sink = sink
        self.depth = int(depth)
        if self.depth < 1:
            raise ValueError("AsyncEgress: depth must be at least 1")
        nz = core.npz
        self.items = []
        for name in names:
            if name not in FIELD_META or name not in core.t:
                raise ValueError("Invalid state variable %s for diagnostics" % name)
            dims, units = FIELD_META[name]
            xi, yi = "x_interface" in dims, "y_interface" in dims
            if len(dims) == 2:
                self.items.append(_Item(name, 1, name, 0, xi, yi, 1, dims, units))
            else:
                self.items.append(_Item(name, 0, name, 0, xi, yi, nz + ("z_interface" in dims), dims, units))
        for level, znames in z_select:
            for name in znames:
                if name not in FIELD_META or name not in core.t:
                    raise ValueError("Invalid state variable %s for level select" % name)   # diagnostics.py:47-48
                dims, units = FIELD_META[name]
                if len(dims) < 3 or dims[2] != "z":   # diagnostics.py:50-54 (its test only admits the layer dimension)
                    raise ValueError("z_select only works for state variables with dimension (x, y, z). \n %s has "
                                     "dimension %s" % (name, dims))
                self.items.append(_Item("%s_z%d" % (name, int(level)), 2, name, int(level), "x_interface" in dims,
                                        "y_interface" in dims, 1, dims[:2], units))
        for name in derived_names:
            if not name.startswith("column_integrated_"):
                import warnings

                warnings.warn("%s is not a supported diagnostic variable." % name)   # diagnostics.py:191
                continue
            tracer = name[len("column_integrated_"):]
            if tracer not in FIELD_META or FIELD_META[tracer][0] != ("x", "y", "z"):
                raise NotImplementedError("this function assumes the z-dimension is the third dimension")
            work = "egress." + name
            if work not in core.t:
                core.alloc2(work)
            self.items.append(_Item(name, 1, work, 0, False, False, 1, ("x", "y"), "kg/m**2", aux=tracer))
        nt, nx, ny = core.nt, core.nx, core.ny
        cuda = core.device.type == "cuda"
        self.slots = []
        for _ in range(self.depth):
            dev, host = {}, {}
            for it in self.items:
                shape = (nt, nx + it.xi, ny + it.yi) + ((it.nk,) if it.kind == 0 else ())
                dev[it.out] = torch.empty(shape, dtype=torch.float64, device=core.device)
                host[it.out] = torch.empty(shape, dtype=torch.float64, pin_memory=cuda)
            self.slots.append({"dev": dev, "host": host, "free": threading.Event(),
                               "done": torch.cuda.Event() if cuda else None})
            self.slots[-1]["free"].set()
        self.stream = torch.cuda.Stream(core.device) if cuda else None
        self.bytes_per_store = sum(t.numel() * 8 for t in self.slots[0]["dev"].values())
        self._n = 0
        self._q = queue.Queue()
        self._error = None
        self._thread = threading.Thread(target=self._worker, daemon=True)
        self._thread.start()

    # ---- time loop side ------------------------------------------------------------------------------------------
    def store(self, time):
        """Takes the snapshot of this moment of the compute stream and returns; raises what an earlier sink call
        raised."""
        if self._error is not None:
            err, self._error = self._error, None
            raise err
        torch, core = self.torch, self.core
        slot = self.slots[self._n % self.depth]
        self._n += 1
        slot["free"].wait()          # the sink still holds this slot's buffers: back-pressure
        slot["free"].clear()
        for it in self.items:
            if it.aux is not None:
                core.call("column_integral", it.aux, "delp", it.src)
            core.call("egress_pack", it.src, it.kind, it.level, int(it.xi), int(it.yi), it.nk,
                      ctypes.c_void_p(slot["dev"][it.out].data_ptr()))
        if self.stream is not None:
            main = torch.cuda.current_stream(core.device)
            self.stream.wait_stream(main)
            with torch.cuda.stream(self.stream):
                for name, d in slot["dev"].items():
                    slot["host"][name].copy_(d, non_blocking=True)
                slot["done"].record(self.stream)
        else:
            for name, d in slot["dev"].items():
                slot["host"][name].copy_(d)
        self._q.put((slot, time))

    def cleanup(self):
        """Waits until every stored step has been handed to the sink."""
        self._q.join()
        if self._error is not None:
            err, self._error = self._error, None
            raise err

    def close(self):
        self.cleanup()
        self._q.put(None)
        self._thread.join()

    # ---- host thread ---------------------------------------------------------------------------------------------
    def _worker(self):
        while True:
            job = self._q.get()
            if job is None:
                self._q.task_done()
                return
            slot, time = job
            try:
                if slot["done"] is not None:
                    slot["done"].synchronize()
                if self.sink is not None:
                    fields = {}
                    for it in self.items:
                        a = slot["host"][it.out].numpy()
                        fields[it.out] = [EgressQuantity(a[t], it.dims, it.units) for t in range(a.shape[0])]
                    self.sink(time, fields)
            except BaseException as e:  # noqa: BLE001 — surfaced by the next store() / cleanup()
                self._error = e
            finally:
                slot["free"].set()
                self._q.task_done()


class MonitorDiagnostics:
    """Drop-in for `pace.driver.diagnostics.MonitorDiagnostics` (diagnostics.py:142-211): the reference's constructor
    arguments plus `dycore=` (a pace_b200.dropin.DynamicalCore or a DycoreEngine: the device state that is stored; the
    `state` argument of `store` is accepted for signature compatibility and not read).  z_select: objects with
    `.level` and `.names` (the reference's ZSelect) or (level, names) pairs.  When the dycore holds one subdomain per
    process the monitor receives one quantity per name, as in the reference; when it holds several, a list."""

    def __init__(self, monitor, names, derived_names, z_select, *, dycore, depth=2):
        self.monitor = monitor
        self.names, self.derived_names, self.z_select = list(names), list(derived_names), list(z_select)
        self._engine = getattr(dycore, "engine", dycore)
        self._single = len(self._engine.ranks) == 1
        zs = [(getattr(z, "level", None), getattr(z, "names", None)) if hasattr(z, "level") else tuple(z)
              for z in self.z_select]
        self._egress = AsyncEgress(self._engine, self.names, self.derived_names, zs, self._hand_over, depth)

    def _hand_over(self, time, fields):
        state = {"time": time}
        for name, per_rank in fields.items():
            state[name] = per_rank[0] if self._single else per_rank
        self.monitor.store(state)

    def store(self, time, state=None):
        self._egress.store(time)

    def store_grid(self, grid_data):
        for k in ("lat", "lon", "lon_agrid", "lat_agrid"):   # diagnostics.py:199-207
            self.monitor.store_constant({k: getattr(grid_data, k)})

    def cleanup(self):
        self._egress.close()
        self.monitor.cleanup()


def save_restart(engine, path, names=None):
    """The DycoreState of every resident subdomain as restart_dycore_state_<rank>.npz under `path` (the reference writes
    restart_dycore_state_<rank>.nc through xarray, driver/pace/driver/state.py:77-123; neither xarray nor netCDF4 is in
    this image, so the container is NumPy's): one array per field, the reference's (i, j, k) compute domain.  Goes
    through the same pack kernels and pinned staging as the diagnostics; returns after the files are written."""
    import os

    from .engine import STATE_2D, STATE_3D

    names = list(names) if names is not None else [n for n in STATE_3D + STATE_2D if n in FIELD_META]
    os.makedirs(path, exist_ok=True)

    def sink(time, fields):
        for t, rank in enumerate(engine.ranks):
            np.savez(os.path.join(path, "restart_dycore_state_%d.npz" % rank), **{n: q[t].data for n, q in fields.items()})

    eg = AsyncEgress(engine, names=names, sink=sink, depth=1)
    eg.store(None)
    eg.close()


def load_restart(engine, path, names=None):
    """Inverse of save_restart: the compute domains are written back into the device state (halos are exchanged by the
    next step)."""
    import os

    core = engine.core
    per_rank = [np.load(os.path.join(path, "restart_dycore_state_%d.npz" % rank)) for rank in engine.ranks]
    for name in (names if names is not None else per_rank[0].files):
        cur = core.get(name)
        for t, z in enumerate(per_rank):
            a = z[name]
            cur[t][(slice(3, 3 + a.shape[0]), slice(3, 3 + a.shape[1])) + ((slice(0, a.shape[2]),) if a.ndim == 3 else ())] = a
        core.put(name, cur)

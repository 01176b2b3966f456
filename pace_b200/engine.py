"""DycoreEngine: the call sequences of the reference's AcousticDynamics.__call__ and DynamicalCore._compute
(pyFV3/stencils/dyn_core.py:714-1013, fv_dynamics.py:513-655) over ALL subdomains resident on one GPU.

The numerics AND the call order are behind the C ABI (include/pace_b200.h: pace_b200_acoustic / _tracer_advection /
_remap / _step, csrc/step.cu); this file owns the buffers' registration, the halo plans and — when subdomains live on
other GPUs — the exchange the library calls back for.
"""
import ctypes

import numpy as np

from .halo import HaloPlanner
from .native import NativeCore

GRAV = 9.80665
HUGE_R = 1.0e40
CNST_0P20 = 0.2

# halo plan ids
A3, A2, B3, D3, C3, SYNC = 0, 1, 2, 3, 4, 5

STATE_3D = [
    "u", "v", "w", "ua", "va", "uc", "vc", "delp", "delz", "pe", "pt", "peln", "pk", "pkz", "qvapor", "qliquid",
    "qice", "qrain", "qsnow", "qgraupel", "qo3mr", "qsgs_tke", "qcld", "q_con", "omga", "mfxd", "mfyd", "cxd", "cyd",
    "diss_estd",
]
STATE_2D = ["ps", "phis"]
# The fields a step reads before it overwrites them: the prognostic variables and the surface geopotential.  Everything
# else in DycoreState (ua, va, uc, vc, pe, peln, pk, pkz, q_con, omga, mfxd, mfyd, cxd, cyd, diss_estd, ps) is diagnosed
# or accumulated from zero inside the step (fv_setup / zero_data / riem_solver3 / remapping / c2l_ord4): poisoning any of
# them in the initial state leaves the result bit-identical (tools/find_step_inputs.py; tests/test_engine_acoustic.py).
STEP_INPUTS = ["u", "v", "w", "delp", "delz", "pt", "qvapor", "qliquid", "qice", "qrain", "qsnow", "qgraupel", "qo3mr",
               "qsgs_tke", "qcld", "phis"]
# What the reference's driver reads back from the dycore after a step (driver/pace/driver/driver.py and the physics
# hand-over pySHiELD/update: delp, pt, delz, w, ua, va, pe, peln, omga, phis and the tracers) plus the prognostic winds
# and ps: the default download set of a host-resident model that does not ask for the whole DycoreState.
PHYSICS_FIELDS = ["u", "v", "w", "ua", "va", "delp", "delz", "pt", "pe", "peln", "omga", "ps", "qvapor", "qliquid", "qice",
                  "qrain", "qsnow", "qgraupel", "qo3mr", "qsgs_tke", "qcld"]
# dycore <-> physics coupling (pySHiELD/update/update_atmos_state.py): the fields DycoreToPhysics copies into the
# PhysicsState, what the physics hands back, and the DriverGridData arrays AGrid2DGridPhysics reads
COUPLING_COPIED = ["qvapor", "qliquid", "qrain", "qsnow", "qice", "qgraupel", "qo3mr", "qsgs_tke", "qcld", "pt", "delp",
                   "delz", "ua", "va", "w", "omga"]
COUPLING_PHYSICS_OUT = ["prsi", "physics_updated_ua", "physics_updated_va", "physics_updated_pt",
                        "physics_updated_specific_humidity", "physics_updated_qliquid", "physics_updated_qrain",
                        "physics_updated_qsnow", "physics_updated_qice", "physics_updated_qgraupel"]
DRIVER_GRID = ["vlon1", "vlon2", "vlon3", "vlat1", "vlat2", "vlat3", "edge_vect_w", "edge_vect_e", "edge_vect_s",
               "edge_vect_n", "es1_1", "es1_2", "es1_3", "ew2_1", "ew2_2", "ew2_3"]
ACOUSTIC_TEMPORARIES = ["ad.ut", "ad.vt", "ad.gz", "ad.zh", "ad.pem", "ad.pkc", "ad.pk3", "ad.heat_source", "ad.cappa",
                        "ad.divgd", "ad.crx", "ad.cry", "ad.xfx", "ad.yfx", "csw.delpc", "csw.ptc", "dyn.dp1", "dyn.cvm"]
TEMPORARIES_2D = ["ad.ws3", "ad.wsd", "ad.zs"]


class DycoreEngine:
    def __init__(self, n_tile, layout, npz, grids, ranks=None, comm=None, device="cuda:0", _lib_handle=None,
                 sat_adjust=None, **cfg):
        """grids: list (one per resident subdomain) of dicts with the reference's GridData / DampingCoefficients
        arrays ("grid.*", "damp.*", reference layout).  ranks: cube ranks of those subdomains (default: all).
        comm: pace_b200.distributed.TorchComm when the cube is spread over several GPUs (one process per GPU,
        contiguous blocks of subdomains, see distributed.Placement); ranks then defaults to this GPU's block.
        sat_adjust: SatAdjustConfig values (pyFV3/_config.py:15-41) that differ from the namelist defaults."""
        from .distributed import HaloExchanger, Placement

        self._sat_adjust = dict(sat_adjust or {})
        self.planner = HaloPlanner(n_tile, layout)
        world = comm.world_size if comm is not None else 1
        self.placement = Placement(self.planner.total_ranks, world)
        mine = self.placement.ranks_of(comm.rank if comm is not None else 0)
        self.ranks = mine if ranks is None else list(ranks)
        if self.ranks != mine:
            raise ValueError("resident subdomains must be this GPU's block of the placement: %s" % mine)
        self.comm = comm
        n = self.planner.n
        g0 = grids[0]
        cfg = dict(cfg)
        cfg.setdefault("ptop", float(g0["grid.ptop"]))
        cfg.setdefault("da_min", float(g0["damp.da_min"]))
        cfg.setdefault("da_min_c", float(g0["damp.da_min_c"]))
        self.core = NativeCore(n, n, npz, n_tiles=len(self.ranks), device=device, _lib_handle=_lib_handle,
                               edge_flags=[self.planner.edge_flags(r) for r in self.ranks], npx=n_tile + 1,
                               npy=n_tile + 1, **cfg)
        self.cfg = self.core.cfg
        self.npz = npz
        core = self.core
        for name in STATE_3D + ACOUSTIC_TEMPORARIES:
            core.alloc3(name)
        for name in STATE_2D + TEMPORARIES_2D:
            core.alloc2(name)
        core.t["ad.pk3"].fill_(HUGE_R)  # dyn_core.py:476-477
        core.put_grid(grids)
        core.set_columns_from_config()
        local = {r: n for n, r in enumerate(self.ranks)}
        pl = self.planner
        plans = {A3: pl.scalar_plan(False, False, 3), A2: pl.scalar_plan(False, False, 2),
                 B3: pl.scalar_plan(True, True, 3), D3: pl.vector_plan("d", 3), C3: pl.vector_plan("c", 3),
                 SYNC: pl.interface_sync_plan()}
        # several GPUs: halo messages are stored straight into the receiver's memory over NVLink when torch's symmetric
        # memory is available (PeerArena); PACE_B200_HALO=p2p keeps the isend / irecv transport (always used on CPU)
        arena = None
        if comm is not None and comm.world_size > 1 and core.device.type == "cuda" and hasattr(comm, "dist"):
            import os

            if os.environ.get("PACE_B200_HALO", "peer") != "p2p":
                from .distributed import PeerArena

                cap = max(HaloExchanger.max_recv_cells(ops, self.placement) for ops in plans.values())
                try:
                    arena = PeerArena(comm, cap * HaloExchanger.NQ_MAX * (npz + 1), core.device)
                    if os.environ.get("PACE_B200_HALO", "peer") != "peer_py":
                        arena.attach(core)  # the library drives the exchange itself; "peer_py": from Python, torch barrier
                except Exception as exc:  # no symmetric memory on this box: every rank falls back together
                    import warnings

                    warnings.warn("pace_b200: peer-memory halo exchange unavailable (%s); using isend/irecv" % exc)
                    arena = None
                import torch

                ok = torch.tensor([1 if arena is not None else 0], device=core.device)
                comm.dist.all_reduce(ok, op=comm.dist.ReduceOp.MIN, group=comm.group)
                if int(ok.item()) == 0:
                    arena = None
        self.arena = arena
        self.halo = {pid: HaloExchanger(core, pid, ops, self.placement, comm, local, arena=arena)
                     for pid, ops in plans.items()}
        self._check_config()
        self._install_callbacks()

    def _halo(self, plan_id, x_names, y_names=None, nk=None):
        self.halo[plan_id].update(x_names, y_names, nk)

    def _check_config(self):
        """Unimplemented-option guards of the reference, same messages (fv_dynamics.py:178-203, 462-499;
        dyn_core.py:434-439)."""
        c = self.cfg
        if c["dt_atmos"] == 0:
            raise RuntimeError("Bad dynamical core configuration: the atmospheric timestep is 0 seconds!")
        if c["nwat"] != 6:
            raise NotImplementedError("Dynamical core (fv_dynamics): nwat==%d is not implemented."
                                      " Only nwat=6 has been implemented." % c["nwat"])
        if c["d_ext"] != 0:
            raise RuntimeError("Acoustics (dyn_core): d_ext != 0 is not implemented")
        if c["beta"] != 0:
            raise RuntimeError("Acoustics (dyn_core): beta != 0 is not implemented")
        if c["hydrostatic"]:
            raise NotImplementedError("Hydrostatic is not implemented")
        if c["consv_te"] > 0:
            raise NotImplementedError("Dynamical Core (fv_dynamics): compute total energy is not implemented")
        if (not c["rf_fast"]) and c["tau"] != 0:
            raise NotImplementedError("Dynamical Core (fv_dynamics): Rayleigh_Super,"
                                      " called when rf_fast=False and tau !=0, is not implemented")
        if c["inline_q"]:
            raise NotImplementedError("Dynamical core (fv_dynamics):tracer_2d not implemented. z_tracer available")

    # -- state ingest / egress -----------------------------------------------------------------------------------
    def set_state(self, states):
        """states: list over resident subdomains of dict name -> reference-layout array (DycoreState fields)."""
        for name in STATE_3D + STATE_2D:
            if name in states[0]:
                self.core.put(name, [s[name] for s in states])
        # AcousticDynamics.__init__: zs = phis / GRAV (dyn_core.py:489-496)
        self.core.put("ad.zs", [np.asarray(s["phis"], dtype=np.float64) / GRAV for s in states])

    def get_state(self, names=None):
        names = names or (STATE_3D + STATE_2D)
        out = [dict() for _ in self.ranks]
        for name in names:
            for n, a in enumerate(self.core.get(name)):
                out[n][name] = a
        return out

    # -- the reference's driver classes: their call order lives behind the C ABI (csrc/step.cu) ---------------------
    def _install_callbacks(self):
        """Halo updates with neighbours on other GPUs / processes are handed back to HaloExchanger by the library."""
        from . import _lib

        core = self.core
        self._cb_error = None

        def on_halo(_user, plan_id, x_names, y_names, n, nk):
            try:
                xs = [x_names[i].decode() for i in range(n)]
                ys = [y_names[i].decode() for i in range(n)] if y_names else None
                self.halo[plan_id].update(xs, ys, nk)
                return 0
            except Exception as exc:  # surfaced by _call below
                self._cb_error = exc
                return 1

        native = self.arena is not None and self.arena.native
        remote = (not native) and any(h.remote or h.arena is not None for h in self.halo.values())
        self._halo_cb = _lib.HALO_CALLBACK(on_halo) if remote else _lib.HALO_CALLBACK()
        core.lib.pace_b200_set_halo_callback(core.handle, self._halo_cb, None)
        sat = _lib.SatAdjConfig.make(**dict({k: (int(v) if isinstance(v, bool) else v) for k, v in self._sat_adjust.items()},
                                            do_qa=int(self.cfg["do_qa"])))
        core.lib.pace_b200_set_satadj_config(core.handle, sat)

    def set_stage_callback(self, fn):
        """fn(savepoint_name) is called at the reference's checkpointer hooks ("C_SW-In", "D_SW-Out", "Remapping-In",
        ...; include/pace_b200.h).  None removes it."""
        from . import _lib

        def on_stage(_user, name):
            try:
                fn(name.decode())
                return 0
            except Exception as exc:
                self._cb_error = exc
                return 1

        self._stage_cb = _lib.STAGE_CALLBACK(on_stage) if fn is not None else _lib.STAGE_CALLBACK()
        self.core.lib.pace_b200_set_stage_callback(self.core.handle, self._stage_cb, None)

    def _call(self, name, *args):
        try:
            self.core.call(name, *args)
        except Exception:
            if self._cb_error is not None:
                exc, self._cb_error = self._cb_error, None
                raise exc
            raise

    def acoustic_dynamics(self, timestep, n_map=1):
        """AcousticDynamics.__call__ (dyn_core.py:714-1013) -> pace_b200_acoustic."""
        self._call("acoustic", float(timestep), int(n_map))

    def tracer_advection(self):
        """TracerAdvection.__call__ (tracer_2d_1l.py:289-417) -> pace_b200_tracer_advection."""
        self._call("tracer_advection")

    def remapping(self, last_step, mdt):
        """LagrangianToEulerian.__call__ (remapping.py:522-731) -> pace_b200_remap."""
        self._call("remap", int(last_step), float(mdt))

    # -- dycore <-> physics coupling of the driver's time loop (driver/pace/driver/driver.py:645-661) -------------
    def setup_coupling(self, driver_grids):
        """Allocates the PhysicsState / TendencyState fields of the coupling ("phy.*", "tend.*") and registers the
        DriverGridData arrays.  driver_grids: list over resident subdomains of dict name -> reference-layout array with
        the names of engine.DRIVER_GRID (edge_vect_s / edge_vect_n are 1-D along i, as in the reference)."""
        core = self.core
        n_sub = self.planner.n
        ni, nj = n_sub + 7, n_sub + 7
        for name in DRIVER_GRID:
            arrs = []
            for gdict in driver_grids:
                a = np.asarray(gdict[name], dtype=np.float64)
                if a.ndim == 1:   # FloatFieldI: constant along j
                    a = np.repeat(a[:ni, None], nj, axis=1)
                arrs.append(a)
            core.put("dgrid." + name, arrs)
        for name in COUPLING_COPIED + COUPLING_PHYSICS_OUT:
            if "phy." + name not in core.t:
                core.alloc3("phy." + name)
        for name in ("u_dt", "v_dt", "pt_dt"):
            if "tend." + name not in core.t:
                core.alloc3("tend." + name)
        # what the reference adds to a local index to get the position on the tile (update_dwind_phys.py:488-499)
        pos = [self.planner.position(r) for r in self.ranks]
        self._origin_i = (ctypes.c_int * len(pos))(*[p[1] * n_sub for p in pos])
        self._origin_j = (ctypes.c_int * len(pos))(*[p[2] * n_sub for p in pos])

    def dycore_to_physics(self):
        """DycoreToPhysics.__call__ (update_atmos_state.py:184-231, no dry convective adjustment) ->
        pace_b200_dycore_to_physics: the dycore state is copied into the "phy.*" fields on the device."""
        self._call("dycore_to_physics")

    def update_atmos_state(self, dt=None, dycore_only=False, apply_tendencies=True):
        """UpdateAtmosphereState.__call__ (update_atmos_state.py:296-349) -> pace_b200_update_atmos_state: applies what
        the physics left in "phy.prsi" / "phy.physics_updated_*" (and "tend.*") to the dycore state."""
        if not hasattr(self, "_origin_i"):
            raise RuntimeError("update_atmos_state: call setup_coupling(driver_grids) first")
        c = self.cfg
        self._call("update_atmos_state", float(c["dt_atmos"] if dt is None else dt), int(bool(dycore_only)),
                   int(bool(apply_tendencies)), self._origin_i, self._origin_j, int(c["npx"]), int(c["npy"]))

    # -- DynamicalCore.step_dynamics / _compute (fv_dynamics.py:445-655) ----------------------------------------
    def step_dynamics(self, timer=None):
        """One dycore step -> pace_b200_step.  timer: optional object with start(name) / stop(name), called around the
        sections the reference times (fv_dynamics.py:551-590: "DynCore", "TracerAdvection", "Remapping"); the step is
        then issued phase by phase (pace_b200_acoustic / _tracer_advection / _remap) so that the caller can clock them
        (bench.py records CUDA events on the launching stream)."""
        if timer is None:
            self._call("step")
            return
        core, c = self.core, self.cfg
        k_split = int(c["k_split"])
        self.compute_preamble()
        for k in range(k_split):
            n_map = k + 1
            last_step = k == k_split - 1
            core.call("copy_levels", "delp", "dyn.dp1")
            timer.start("DynCore")
            self.acoustic_dynamics(c["dt_atmos"] / k_split, n_map=n_map)
            timer.stop("DynCore")
            timer.start("TracerAdvection")
            self.tracer_advection()
            timer.stop("TracerAdvection")
            if self.npz > 4:
                timer.start("Remapping")
                self.remapping(last_step, c["dt_atmos"] / k_split)
                timer.stop("Remapping")
                if last_step:
                    core.call("omega_from_w")
                    if c["nf_omega"] > 0:
                        self._halo(A3, ["omga"])
                        core.call("del2cubed", "omga", 0.18 * c["da_min"], int(c["nf_omega"]))
        core.call("neg_adj3")
        self._halo(D3, ["u"], ["v"])
        core.call("c2l_ord4")

    # -- DynamicalCore.compute_preamble (fv_dynamics.py:461-508) -------------------------------------------------
    def compute_preamble(self):
        self.core.call("fv_setup")
        self.core.call("pt_to_potential_density_pt")


class HostPipeline:
    """Steps a HOST-resident model state through a DycoreEngine: the call a coupled driver makes when the rest of the
    model lives in host memory.  Every step uploads the fields the step reads (STEP_INPUTS; upload_names=None uploads
    the full DycoreState) from pinned host buffers and downloads the whole updated state (STATE_3D + STATE_2D); with nsets device copies of the state (default three) the upload of step n+1 and the
    download of step n-1 run on their own CUDA streams while step n computes (PCIe is full duplex), so the cost per
    step tends to max(compute, upload, download) instead of their sum.  phis (static) is expected to stay unchanged
    (ad.zs is derived from it once in set_state)."""

    def __init__(self, eng, names=None, upload_names=STEP_INPUTS, download_names=None, nsets=3):
        import torch

        self.torch = torch
        self.eng = eng
        self.names = list(names) if names is not None else STATE_3D + STATE_2D
        self.upload_names = [n for n in self.names if upload_names is None or n in upload_names]
        # download_names: the fields the caller wants back every step (default: all of `names`)
        self.download_names = [n for n in self.names if download_names is None or n in download_names]
        core = eng.core
        # nsets device copies of the state, used round robin.  With two, the upload of step n+1 has to wait for the
        # download of step n-1 (same copy) and the two transfers of a step run one after the other; with three it
        # only waits for the download of step n-2, which had a whole step to finish.
        if nsets < 2:
            raise ValueError("HostPipeline: nsets must be at least 2")
        self.nsets = nsets
        self.sets = [{n: core.t[n] for n in self.names}]
        for _ in range(nsets - 1):
            self.sets.append({n: torch.empty_like(core.t[n]) for n in self.names})
        for extra in self.sets[1:]:
            for n in self.names:
                extra[n].copy_(core.t[n])
        self.s_up, self.s_down = torch.cuda.Stream(core.device), torch.cuda.Stream(core.device)
        self.bytes_per_step = sum(self.sets[0][n].numel() * self.sets[0][n].element_size() for n in self.download_names)
        self.upload_bytes_per_step = sum(self.sets[0][n].numel() * self.sets[0][n].element_size() for n in self.upload_names)
        self.copy_ms = None   # (upload, download) durations of the last step of the last run(), on the device

    def pinned_like(self):
        """A dict of pinned host tensors shaped like the state (one upload or download buffer)."""
        return {n: self.torch.empty(t.shape, dtype=t.dtype).pin_memory() for n, t in self.sets[0].items()}

    def run(self, steps, host_in, host_out):
        """steps dycore steps; step n reads its inputs from host_in (dict, or callable n -> dict) and leaves its
        result in host_out (dict, or callable n -> dict).  Returns after everything has completed."""
        torch, eng, core = self.torch, self.eng, self.eng.core
        get_in = host_in if callable(host_in) else (lambda n: host_in)
        get_out = host_out if callable(host_out) else (lambda n: host_out)
        main = torch.cuda.current_stream(core.device)
        # the library launches on the stream it was given (NativeCore: the stream current at construction); the events
        # below order the copies against THAT stream, so it must be the current one here
        core.lib.pace_b200_set_stream(core.handle, ctypes.c_void_p(main.cuda_stream))
        # work already queued on the main stream (the constructor's copies, the caller's writes) precedes the copies
        self.s_up.wait_stream(main)
        self.s_down.wait_stream(main)
        S = self.nsets
        up_done = [torch.cuda.Event() for _ in range(S)]
        step_done = [torch.cuda.Event() for _ in range(S)]
        down_done = [None] * S
        t_up = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        t_down = [torch.cuda.Event(enable_timing=True) for _ in range(2)]

        def upload(n):
            tgt = self.sets[n % S]
            with torch.cuda.stream(self.s_up):
                if down_done[n % S] is not None:
                    self.s_up.wait_event(down_done[n % S])   # the previous result must have left this copy
                if n >= S:
                    self.s_up.wait_event(step_done[n % S])
                src = get_in(n)
                if n == steps - 1:
                    t_up[0].record(self.s_up)
                for name in self.upload_names:
                    tgt[name].copy_(src[name], non_blocking=True)
                if n == steps - 1:
                    t_up[1].record(self.s_up)
                up_done[n % S].record(self.s_up)

        upload(0)
        for n in range(steps):
            cur = self.sets[n % S]
            for name in self.names:
                core._register(name, cur[name])
            main.wait_event(up_done[n % S])
            if n + 1 < steps:
                upload(n + 1)
            eng.step_dynamics()
            step_done[n % S].record(main)
            with torch.cuda.stream(self.s_down):
                self.s_down.wait_event(step_done[n % S])
                dst = get_out(n)
                if n == steps - 1:
                    t_down[0].record(self.s_down)
                for name in self.download_names:
                    dst[name].copy_(cur[name], non_blocking=True)
                if n == steps - 1:
                    t_down[1].record(self.s_down)
                ev = torch.cuda.Event()
                ev.record(self.s_down)
                down_done[n % S] = ev
        torch.cuda.synchronize(core.device)
        if steps > 0:
            self.copy_ms = (t_up[0].elapsed_time(t_up[1]), t_down[0].elapsed_time(t_down[1]))
        for name in self.names:  # leave the engine bound to its original tensors
            core._register(name, self.sets[0][name])
        last = (steps - 1) % S if steps > 0 else 0
        if last != 0:
            # the last step ran on another copy: bring its result home
            for name in self.names:
                self.sets[0][name].copy_(self.sets[last][name])
            torch.cuda.synchronize(core.device)

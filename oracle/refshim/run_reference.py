"""TEST INFRASTRUCTURE ONLY — build and step the *unmodified* reference dycore (numpy backend) on a
multi-process Comm, and record what it does.  Used by oracle/gen_golden.py and oracle/gen_c48_golden.py (fixtures; their step times are the reference-numpy timings
quoted in profiles/ and BASELINE.md).  Needs /root/reference, so it never runs on the GPU box.

Constructor sequence follows the reference's own test tests/main/fv3core/test_dycore_call.py:28-136,
with NullComm replaced by ProcComm so halos are real.
"""
import os
import time
from datetime import timedelta

import numpy as np


def default_config_kwargs(npx=13, npz=79, layout=(1, 1), n_split=1, k_split=1, **over):
    kw = dict(
        layout=layout, npx=npx, npy=npx, npz=npz, ntiles=6, nwat=6, dt_atmos=225, a_imp=1.0, beta=0.0,
        consv_te=False, d2_bg=0.0, d2_bg_k1=0.2, d2_bg_k2=0.1, d4_bg=0.15, d_con=1.0, d_ext=0.0, dddmp=0.5,
        delt_max=0.002, do_sat_adj=True, do_vort_damp=True, fill=True, hord_dp=6, hord_mt=6, hord_tm=6,
        hord_tr=8, hord_vt=6, hydrostatic=False, k_split=k_split, ke_bg=0.0, kord_mt=9, kord_tm=-9, kord_tr=9,
        kord_wz=9, n_split=n_split, nord=3, p_fac=0.05, rf_fast=True, rf_cutoff=3000.0, tau=10.0, vtdm4=0.06,
        z_tracer=True, do_qa=True,
    )
    kw.update(over)
    return kw


def build_dycore(mpi_comm, cfg_kwargs, backend="numpy"):
    """Returns (dycore, state, grid_data, damping, metric_terms, config, communicator)."""
    import ndsl.dsl.stencil
    import pyFV3
    import pyFV3.initialization.analytic_init as ai
    from ndsl.comm.communicator import CubedSphereCommunicator
    from ndsl.comm.partitioner import CubedSpherePartitioner, TilePartitioner
    from ndsl.dsl.dace.dace_config import DaceConfig
    from ndsl.dsl.stencil import GridIndexing
    from ndsl.grid import DampingCoefficients, GridData, MetricTerms
    from ndsl.initialization.allocator import QuantityFactory
    from ndsl.initialization.sizer import SubtileGridSizer

    config = pyFV3.DynamicalCoreConfig(**cfg_kwargs)
    partitioner = CubedSpherePartitioner(TilePartitioner(config.layout))
    communicator = CubedSphereCommunicator(mpi_comm, partitioner)
    dace_config = DaceConfig(communicator=communicator, backend=backend)
    stencil_config = ndsl.dsl.stencil.StencilConfig(
        compilation_config=ndsl.dsl.stencil.CompilationConfig(
            backend=backend, rebuild=False, validate_args=os.environ.get("PACE_REF_VALIDATE_ARGS", "1") != "0"),
        dace_config=dace_config,
    )
    sizer = SubtileGridSizer.from_tile_params(
        nx_tile=config.npx - 1, ny_tile=config.npy - 1, nz=config.npz, n_halo=3, extra_dim_lengths={},
        layout=config.layout, tile_partitioner=partitioner.tile, tile_rank=communicator.tile.rank,
    )
    grid_indexing = GridIndexing.from_sizer_and_communicator(sizer=sizer, comm=communicator)
    quantity_factory = QuantityFactory.from_backend(sizer=sizer, backend=backend)
    metric_terms = MetricTerms(quantity_factory=quantity_factory, communicator=communicator, eta_file="in-repo-L79")
    grid_data = GridData.new_from_metric_terms(metric_terms)
    state = ai.init_analytic_state(
        analytic_init_case="baroclinic", grid_data=grid_data, quantity_factory=quantity_factory,
        adiabatic=config.adiabatic, hydrostatic=config.hydrostatic, moist_phys=config.moist_phys, comm=communicator,
    )
    stencil_factory = ndsl.dsl.stencil.StencilFactory(config=stencil_config, grid_indexing=grid_indexing)
    damping = DampingCoefficients.new_from_metric_terms(metric_terms)
    dycore = pyFV3.DynamicalCore(
        comm=communicator, grid_data=grid_data, stencil_factory=stencil_factory, quantity_factory=quantity_factory,
        damping_coefficients=damping, config=config, timestep=timedelta(seconds=config.dt_atmos), phis=state.phis,
        state=state,
    )
    # the factories, for callers that build further reference objects on the same grid (oracle/gen_coupling_golden.py)
    build_dycore.extras = {"stencil_factory": stencil_factory, "quantity_factory": quantity_factory,
                           "metric_terms": metric_terms}
    return dycore, state, grid_data, damping, metric_terms, config, communicator


# ------------------------------------------------------------------------------------------------
# recording
# ------------------------------------------------------------------------------------------------
def _arr(q):
    return np.asarray(q.data if hasattr(q, "data") and not isinstance(q, np.ndarray) else q)


class Recorder:
    """Sequential log of which watched arrays changed at each marked point (full arrays stored)."""

    def __init__(self, watch):
        self.watch = {k: _arr(v) for k, v in watch.items()}
        self.last = {k: v.copy() for k, v in self.watch.items()}
        self.stages = [("init", dict(self.last))]

    def mark(self, label):
        changed = {}
        for k, v in self.watch.items():
            if not np.array_equal(v, self.last[k], equal_nan=True):
                changed[k] = v.copy()
                self.last[k] = changed[k]
        self.stages.append((label, changed))


class _Wrapped:
    def __init__(self, inner, label, rec):
        object.__setattr__(self, "_inner", inner)
        object.__setattr__(self, "_label", label)
        object.__setattr__(self, "_rec", rec)

    def __call__(self, *a, **k):
        self._rec.mark("pre:" + self._label)
        out = self._inner(*a, **k)
        self._rec.mark(self._label)
        return out

    def __getattr__(self, name):
        return getattr(self._inner, name)

    def __setattr__(self, name, v):
        setattr(self._inner, name, v)


def grid_arrays(grid_data, damping):
    """All array/scalar metric terms the dycore step reads, as plain numpy."""
    names = [
        "lon", "lat", "lon_agrid", "lat_agrid", "fC", "fC_agrid", "area", "area_64", "rarea", "rarea_c", "dx", "dy",
        "dxc", "dyc", "dxa", "dya", "rdx", "rdy", "rdxc", "rdyc", "rdxa", "rdya", "a11", "a12", "a21", "a22",
        "edge_w", "edge_e", "edge_s", "edge_n", "p", "ak", "bk", "dp_ref", "cosa", "cosa_u", "cosa_v", "cosa_s",
        "sina_u", "sina_v", "rsina", "rsin_u", "rsin_v", "rsin2", "sin_sg1", "sin_sg2", "sin_sg3", "sin_sg4",
        "cos_sg1", "cos_sg2", "cos_sg3", "cos_sg4", "ee1", "ee2", "es1", "ew2",
    ]
    out = {}
    for n in names:
        out["grid." + n] = _arr(getattr(grid_data, n)).copy()
    out["grid.ptop"] = np.float64(grid_data.ptop)
    for n in ("del6_u", "del6_v", "divg_u", "divg_v"):
        out["damp." + n] = _arr(getattr(damping, n)).copy()
    out["damp.da_min"] = np.float64(damping.da_min)
    out["damp.da_min_c"] = np.float64(damping.da_min_c)
    return out


def state_arrays(state):
    from dataclasses import fields

    out = {}
    for f in fields(state):
        v = getattr(state, f.name)
        if hasattr(v, "data"):
            out[f.name] = _arr(v)
    return out


def instrument_deep(dycore, state, rec_watch, wrap):
    """Finer-grained hooks inside d_sw / updatedzd / nh_p_grad (development + sub-module fixtures)."""
    ad = dycore.acoustic_dynamics
    dsw = ad.dgrid_shallow_water_lagrangian_dynamics
    for n in ("_tmp_heat_s", "_tmp_diss_e", "_vort_x_delta", "_vort_y_delta", "_dt_kinetic_energy_on_cell_corners",
              "_abs_vorticity_agrid", "_damped_rel_vorticity_agrid", "_uc_contra", "_vc_contra", "_tmp_ut", "_tmp_vt",
              "_tmp_fx", "_tmp_fy", "_tmp_gx", "_tmp_gy", "_tmp_dw", "_tmp_wk", "_vorticity_agrid",
              "_vorticity_bgrid_damped", "_tmp_fx2", "_tmp_fy2"):
        rec_watch["dsw." + n.lstrip("_")] = getattr(dsw, n)
    for n in ("fv_prep", "fvtp2d_dp", "_flux_capacitor_stencil", "delnflux_nosg_w", "_heat_diss_stencil",
              "fvtp2d_vt_nodelnflux", "_apply_fluxes", "fvtp2d_dp_t", "fvtp2d_tm", "_apply_pt_delp_fluxes",
              "_adjust_w_and_qcon_stencil", "_compute_kinetic_energy", "_compute_vorticity_stencil",
              "divergence_damping", "_rel_vorticity_to_abs", "_u_and_v_from_ke_stencil", "delnflux_nosg_v",
              "_vort_differencing_stencil", "_heat_source_from_vorticity_damping_stencil",
              "_accumulate_heat_source_and_dissipation_estimate_stencil", "_update_u_and_v_stencil"):
        wrap(dsw, n, "dsw." + n.lstrip("_"))
    dd = dsw.divergence_damping
    for n in dir(dd):
        v = getattr(dd, n, None)
        if n.startswith("_") and not n.startswith("__") and hasattr(v, "data") and hasattr(v, "dims") and len(v.dims) == 3:
            rec_watch["dd." + n.lstrip("_")] = v
    dzd = ad.update_height_on_d_grid
    for n in dir(dzd):
        v = getattr(dzd, n, None)
        if n.startswith("_") and not n.startswith("__") and hasattr(v, "data") and hasattr(v, "dims") and len(v.dims) >= 2:
            rec_watch["dzd." + n.lstrip("_")] = v
    rm = dycore._lagrangian_to_eulerian_obj
    for n in ("_pe1", "_pe2", "_pe3", "_dp2", "_pn2", "_pe0", "_gz", "_cvm"):
        rec_watch["rm." + n.lstrip("_")] = getattr(rm, n)
    for n in ("_init_pe", "_moist_cv_pt_pressure", "_pn2_pk_delp", "_map_single_pt", "_mapn_tracer", "_map_single_w",
              "_map_single_delz", "_undo_delz_adjust_and_copy_peln", "_moist_cv_pkz", "_pressures_mapu",
              "_map_single_u", "_pressures_mapv", "_map_single_v", "_update_ua", "_copy_from_below_stencil",
              "_saturation_adjustment", "_moist_cv_last_step_stencil", "_basic_adjust_divide_stencil"):
        wrap(rm, n, "rm." + n.lstrip("_"))
    nh = ad.nonhydrostatic_pressure_gradient
    for n in dir(nh):
        v = getattr(nh, n, None)
        if n.startswith("_") and not n.startswith("__") and hasattr(v, "data") and hasattr(v, "dims") and len(v.dims) == 3:
            rec_watch["nh." + n.lstrip("_")] = v


def instrument(dycore, state, deep=False):
    """Wrap the module-level callables of one dycore and return the Recorder."""
    ad = dycore.acoustic_dynamics
    csw = ad.cgrid_shallow_water_lagrangian_dynamics
    watch = dict(state_arrays(state))
    pending = []
    if deep:
        instrument_deep(dycore, state, watch, lambda parent, attr, label: pending.append((parent, attr, label)))
    for n in ("_ut", "_vt", "_gz", "_zh", "_pkc", "_pk3", "_pem", "_divgd", "_heat_source", "_crx", "_cry", "_xfx",
              "_yfx", "_ws3", "cappa"):
        watch["ad." + n.lstrip("_")] = getattr(ad, n)
    watch["ad.wsd"] = dycore._wsd
    watch["ad.zs"] = ad._zs
    watch["csw.delpc"] = csw.delpc
    watch["csw.ptc"] = csw.ptc
    watch["dyn.dp1"] = dycore._dp_initial
    watch["dyn.cvm"] = dycore._cvm
    rec = Recorder(watch)

    def wrap(parent, attr, label):
        setattr(parent, attr, _Wrapped(getattr(parent, attr), label, rec))

    for parent, attr, label in pending:
        wrap(parent, attr, label)
    wrap(dycore, "_fv_setup_stencil", "fv_setup")
    wrap(dycore, "_pt_to_potential_density_pt", "pt_to_potential_density_pt")
    wrap(dycore, "acoustic_dynamics", "acoustic_dynamics")
    wrap(dycore, "tracer_advection", "tracer_advection")
    wrap(dycore, "_lagrangian_to_eulerian_obj", "remapping")
    wrap(dycore, "_omega_from_w", "omega_from_w")
    wrap(dycore, "_hyperdiffusion", "omega_del2cubed")
    wrap(dycore, "_adjust_tracer_mixing_ratio", "neg_adj3")
    wrap(dycore, "_cubed_to_latlon", "c2l_ord")
    wrap(ad, "_zero_data", "zero_data")
    wrap(ad, "_gz_from_surface_height_and_thickness", "gz_from_zs_delz")
    wrap(ad, "cgrid_shallow_water_lagrangian_dynamics", "c_sw")
    wrap(ad, "_copy_stencil", "copy_gz_zh")
    wrap(ad, "update_geopotential_height_on_c_grid", "updatedzc")
    wrap(ad, "vertical_solver_cgrid", "riem_solver_c")
    wrap(ad, "_p_grad_c", "p_grad_c")
    wrap(ad, "dgrid_shallow_water_lagrangian_dynamics", "d_sw")
    wrap(ad, "update_height_on_d_grid", "updatedzd")
    wrap(ad, "vertical_solver", "riem_solver3")
    wrap(ad, "_edge_pe_stencil", "edge_pe")
    wrap(ad, "_pk3_halo", "pk3_halo")
    wrap(ad, "_compute_geopotential_stencil", "compute_geopotential")
    wrap(ad, "nonhydrostatic_pressure_gradient", "nh_p_grad")
    wrap(ad, "_rayleigh_damping", "ray_fast")
    if ad._do_del2cubed:
        wrap(ad, "_hyperdiffusion", "heat_del2cubed")
    wrap(ad, "_apply_diffusive_heating", "apply_diffusive_heating")
    return rec


def rank_main(mpi_comm, cfg_kwargs, n_steps=1, record_ranks=(0,), out_dir=None, time_steps=0, deep=False):
    """Body of one rank. Returns a dict of small results; large data goes to out_dir/*.npz."""
    import os

    from oracle.refshim.bootstrap import bootstrap

    bootstrap()
    from ndsl.performance.timer import Timer

    rank = mpi_comm.Get_rank()
    t0 = time.time()
    dycore, state, grid_data, damping, metric_terms, config, communicator = build_dycore(mpi_comm, cfg_kwargs)
    t_build = time.time() - t0
    res = {"rank": rank, "build_s": t_build}
    init = {k: v.copy() for k, v in state_arrays(state).items()}
    rec = instrument(dycore, state, deep) if rank in record_ranks else None
    step_s = []
    for s in range(n_steps):
        t0 = time.time()
        dycore.step_dynamics(state)
        step_s.append(time.time() - t0)
        if rec is not None:
            rec.mark("end_of_step_%d" % s)
    res["step_s"] = step_s
    if time_steps:
        timer = Timer()
        mpi_comm.barrier()
        t0 = time.time()
        for s in range(time_steps):
            dycore.step_dynamics(state, timer)
        mpi_comm.barrier()
        res["timed_s_per_step"] = (time.time() - t0) / time_steps
        res["timer"] = {k: float(v) for k, v in timer.times.items()}
    if out_dir is not None:
        os.makedirs(out_dir, exist_ok=True)
        final = state_arrays(state)
        np.savez(os.path.join(out_dir, "rank%d_init.npz" % rank), **init)
        np.savez(os.path.join(out_dir, "rank%d_final.npz" % rank), **final)
        np.savez(os.path.join(out_dir, "rank%d_grid.npz" % rank), **grid_arrays(grid_data, damping))
        if rec is not None:
            flat = {}
            labels = []
            for n, (label, changed) in enumerate(rec.stages):
                labels.append([label, sorted(changed)])
                for k, v in changed.items():
                    flat["%03d:%s" % (n, k)] = v
            import json

            flat["__labels__"] = np.frombuffer(json.dumps(labels).encode(), dtype=np.uint8)
            np.savez(os.path.join(out_dir, "rank%d_stages.npz" % rank), **flat)
    return res

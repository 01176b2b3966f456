"""ctypes binding of libpace_b200.so (C ABI in include/pace_b200.h).

The CUDA library is the only compute path: if it is missing this module raises — there is no CPU or
PyTorch fallback.
"""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libpace_b200.so")

_I = ctypes.c_int
_D = ctypes.c_double
_S = ctypes.c_char_p


class Config(ctypes.Structure):
    """Mirror of `pace_b200_config` (include/pace_b200.h)."""

    _fields_ = (
        [(n, _I) for n in ("npx", "npy", "npz", "nx", "ny", "n_tiles_local", "n_split", "k_split")]
        + [(n, _I) for n in ("hord_dp", "hord_mt", "hord_tm", "hord_tr", "hord_vt")]
        + [(n, _I) for n in ("kord_mt", "kord_tm", "kord_tr", "kord_wz")]
        + [(n, _I) for n in ("nord", "n_sponge", "nf_omega", "grid_type")]
        + [(n, _I) for n in ("do_sat_adj", "do_vort_damp", "rf_fast", "hydrostatic", "convert_ke", "fill",
                             "adiabatic", "do_qa", "use_old_omega", "do_f3d", "inline_q", "nwat")]
        + [("dt_atmos", _D)]
        + [(n, _D) for n in ("a_imp", "beta", "p_fac", "d_con", "d2_bg", "d2_bg_k1", "d2_bg_k2", "d4_bg", "dddmp",
                             "delt_max", "ke_bg", "rf_cutoff", "tau", "vtdm4", "d_ext", "consv_te")]
        + [(n, _D) for n in ("ptop", "da_min", "da_min_c")]
        + [("ks", _I)]
    )


class SatAdjConfig(ctypes.Structure):
    """Mirror of `pace_b200_satadj_config`; defaults = NDSL NamelistDefaults (ndsl/namelist.py:26-57)."""

    _fields_ = [(n, _I) for n in ("rad_snow", "rad_rain", "rad_graupel", "tintqs", "icloud_f", "do_qa")] + [
        (n, _D) for n in ("sat_adj0", "ql_gen", "qs_mlt", "ql0_max", "t_sub", "qi_gen", "qi_lim", "qi0_max", "dw_ocean",
                          "dw_land", "cld_min", "tau_i2s", "tau_v2l", "tau_r2g", "tau_l2r", "tau_l2v", "tau_imlt",
                          "tau_smlt")]
    DEFAULTS = dict(rad_snow=1, rad_rain=1, rad_graupel=1, tintqs=0, icloud_f=0, do_qa=1, sat_adj0=0.90, ql_gen=1.0e-3,
                    qs_mlt=1.0e-6, ql0_max=2.0e-3, t_sub=184.0, qi_gen=1.82e-6, qi_lim=1.0, qi0_max=1.0e-4,
                    dw_ocean=0.10, dw_land=0.15, cld_min=0.05, tau_i2s=1000.0, tau_v2l=90.0, tau_r2g=900.0,
                    tau_l2r=900.0, tau_l2v=300.0, tau_imlt=600.0, tau_smlt=900.0)

    @classmethod
    def make(cls, **over):
        vals = dict(cls.DEFAULTS)
        vals.update(over)
        return cls(**vals)


class HaloOp(ctypes.Structure):
    """Mirror of `pace_b200_halo_op`."""

    _fields_ = [(n, _I) for n in ("dst_t", "src_t", "comp", "src_comp")] + [("sign", _D)] + [
        (n, _I) for n in ("i0", "j0", "ni", "nj", "aii", "aij", "bi", "aji", "ajj", "bj")]


HALO_CALLBACK = ctypes.CFUNCTYPE(_I, ctypes.c_void_p, _I, ctypes.POINTER(_S), ctypes.POINTER(_S), _I, _I)
STAGE_CALLBACK = ctypes.CFUNCTYPE(_I, ctypes.c_void_p, _S)

_lib = None


def load():
    """Load the shared library (raises if it has not been built)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "pace_b200: %s is missing — build it with `python -m pace_b200.build`; "
            "there is no CPU fallback" % LIB_PATH
        )
    _lib = bind(ctypes.CDLL(LIB_PATH))
    return _lib


def bind(lib):
    """Declare argument/return types of every entry point of include/pace_b200.h."""
    vp = ctypes.c_void_p
    lib.pace_b200_row_stride.argtypes = [_I]
    lib.pace_b200_row_stride.restype = _I
    lib.pace_b200_create.argtypes = [ctypes.POINTER(Config), ctypes.c_char_p, ctypes.POINTER(vp)]
    lib.pace_b200_destroy.argtypes = [vp]
    lib.pace_b200_last_error.argtypes = [vp]
    lib.pace_b200_last_error.restype = ctypes.c_char_p
    lib.pace_b200_set_stream.argtypes = [vp, vp]
    lib.pace_b200_set_field.argtypes = [vp, ctypes.c_char_p, vp]
    lib.pace_b200_set_column.argtypes = [vp, ctypes.c_char_p, ctypes.POINTER(ctypes.c_double), _I]
    lib.pace_b200_halo_plan_set.argtypes = [vp, _I, vp, _I]
    lib.pace_b200_halo_update.argtypes = [vp, _I, ctypes.POINTER(_S), ctypes.POINTER(_S), _I, _I]
    lib.pace_b200_halo_pack.argtypes = [vp, _I, ctypes.POINTER(_S), ctypes.POINTER(_S), _I, _I, vp]
    lib.pace_b200_halo_unpack.argtypes = [vp, _I, ctypes.POINTER(_S), ctypes.POINTER(_S), _I, _I, vp]
    lib.pace_b200_launch_count.argtypes = []
    lib.pace_b200_launch_count.restype = ctypes.c_longlong
    lib.pace_b200_profile_enable.argtypes = [_I]
    lib.pace_b200_profile_report.argtypes = [ctypes.c_char_p, _I]
    lib.pace_b200_math_eval.argtypes = [vp, _I, _I, vp, vp, vp]
    lib.pace_b200_sat_adjust.argtypes = [vp, ctypes.POINTER(SatAdjConfig), _D, _I, _I]
    lib.pace_b200_map_single_f32.argtypes = [vp, vp, vp, vp, vp, _I, _I, _I, _I, ctypes.c_float]
    lib.pace_b200_set_satadj_config.argtypes = [vp, ctypes.POINTER(SatAdjConfig)]
    lib.pace_b200_set_halo_callback.argtypes = [vp, HALO_CALLBACK, vp]
    lib.pace_b200_set_stage_callback.argtypes = [vp, STAGE_CALLBACK, vp]
    lib.pace_b200_halo_peer_arena.argtypes = [vp, _I, _I, ctypes.POINTER(vp), ctypes.c_longlong]
    lib.pace_b200_halo_peer_plan.argtypes = [vp, _I, _I, ctypes.POINTER(_I), ctypes.POINTER(_I), _I, ctypes.c_longlong]
    lib.pace_b200_halo_peer_status.argtypes = [vp, ctypes.POINTER(ctypes.c_ulonglong), ctypes.POINTER(ctypes.c_ulonglong)]
    lib.pace_b200_halo_peer_update.argtypes = [vp, _I, ctypes.POINTER(_S), ctypes.POINTER(_S), _I, _I]
    for name, extra in MODULE_CALLS.items():
        fn = getattr(lib, "pace_b200_" + name)
        fn.argtypes = [vp] + extra
        fn.restype = _I
    return lib


# module entry points: name -> scalar argument types after the handle
MODULE_CALLS = {
    "riem_solver_c": [_D],
    "c_sw": [_D],
    "update_dz_c": [_D],
    "update_dz_c_static": [_D],
    "p_grad_c": [_D],
    "fv_prep": [_D],
    "d_sw": [_D],
    "fv_setup": [],
    "neg_adj3": [],
    "c2l_ord4": [],
    "remap_main": [_I],
    "remap_finish": [_I],
    "map_single": [_S, _S, _S, _S, _I, _I, _I, _I, _D],
    "tracer_flux_prep": [_I],
    "tracer_substep": [_I],
    "tracer_swap_dp": [],
    "pt_to_potential_density_pt": [],
    "omega_from_w": [],
    "copy_levels": [_S, _S],
    "riem_solver3": [_D, _I],
    "update_dz_d": [_D],
    "zero_data": [_I],
    "gz_from_zs_delz": [],
    "pem_from_delp": [],
    "copy_full": [_S, _S],
    "compute_geopotential": [],
    "edge_pe": [],
    "pk3_halo": [],
    "nh_p_grad": [_D],
    "ray_fast": [_D],
    "del2cubed": [_S, _D, _I],
    "apply_diffusive_heating": [_D, _I],
    "fvtp2d": [_S] * 10 + [_I, _S, _S, _I],
    "delnflux_nosg": [_S] * 6 + [_I],
    "a2b_ord4": [_S, _S, _I, _I, _I],
    "divergence_damping": [_S, _S, _S, _S, _D],
    "acoustic": [_D, _I],
    "tracer_advection": [],
    "remap": [_I, _D],
    "step": [],
    "dycore_to_physics": [],
    "egress_pack": [_S, _I, _I, _I, _I, _I, ctypes.c_void_p],
    "column_integral": [_S, _S, _S],
    "update_atmos_state": [_D, _I, _I, ctypes.POINTER(_I), ctypes.POINTER(_I), _I, _I],
}

EXPORTS = ["pace_b200_row_stride", "pace_b200_create", "pace_b200_destroy", "pace_b200_last_error",
           "pace_b200_set_stream", "pace_b200_set_field", "pace_b200_set_column", "pace_b200_halo_plan_set", "pace_b200_halo_update", "pace_b200_halo_pack", "pace_b200_halo_unpack", "pace_b200_math_eval", "pace_b200_sat_adjust", "pace_b200_map_single_f32", "pace_b200_launch_count", "pace_b200_profile_enable",
           "pace_b200_profile_report", "pace_b200_set_satadj_config", "pace_b200_set_halo_callback",
           "pace_b200_set_stage_callback", "pace_b200_halo_peer_arena", "pace_b200_halo_peer_plan",
           "pace_b200_halo_peer_update", "pace_b200_halo_peer_status"] + ["pace_b200_" + n for n in MODULE_CALLS]

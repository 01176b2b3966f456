"""Module-level parity cases.  tol = max of the reference metric 2|a-b|/(|a|+|b|) allowed on the GPU
(0.0 = bit-exact); tol_host likewise for the host-emulation build (same glibc libm as the oracle)."""
from tests.module_harness import DT, DT2

CASES = {
    "fv_prep": dict(stage="dsw.fv_prep", call="fv_prep", args=(DT,), tol=0.0, deep=True),
    "d_sw": dict(stage="d_sw", call="d_sw", args=(DT,), tol=0.0, deep=True, begin="pre:dsw.fv_prep",
                 region={"dsw.tmp_gx": 0, "dsw.tmp_gy": 0, "dsw.tmp_heat_s": 0, "dsw.tmp_wk": 0,
                         "dsw.damped_rel_vorticity_agrid": 0}),
    "d_sw_module": dict(stage="d_sw", call="d_sw", args=(DT,), tol=0.0),
    "c_sw": dict(stage="c_sw", call="c_sw", args=(DT2,), tol=0.0),
    "updatedzc": dict(stage="updatedzc", call="update_dz_c", args=(DT2,), tol=0.0),
    "p_grad_c": dict(stage="p_grad_c", call="p_grad_c", args=(DT2,), tol=0.0),
    "riem_solver_c": dict(stage="riem_solver_c", call="riem_solver_c", args=(DT2,), tol=1e-11, tol_host=0.0),
}

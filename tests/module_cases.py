"""Module-level parity cases.  tol = max of the reference metric 2|a-b|/(|a|+|b|) allowed on the GPU
(0.0 = bit-exact); tol_host likewise for the host-emulation build (same glibc libm as the oracle)."""
from tests.module_harness import DT, DT2

CASES = {
    "riem_solver_c": dict(stage="riem_solver_c", call="riem_solver_c", args=(DT2,), tol=1e-11, tol_host=0.0),
}

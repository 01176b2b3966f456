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
    return bad



if __name__ == "__main__":
    from oracle import cpu_port

    cpu_port.load()
    res = run_ranks(6, _rank, (), timeout=3600)
    print("DROPIN_RESULT " + json.dumps(res))
    sys.exit(0 if all(r == [] for r in res) else 1)

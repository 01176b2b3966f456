"""TEST INFRASTRUCTURE ONLY — golden fixture for the diagnostics egress: the UNMODIFIED reference
`MonitorDiagnostics.store` (driver/pace/driver/diagnostics.py:142-197, with ZSelect :39-64 and
_compute_column_integral :226-249) on 6 ranks at C12 L79 after one dycore step, writing to a monitor that records what
it is handed (the reference's own monitors need zarr / netCDF4, which this image does not have; a monitor is any object
with store / store_constant / cleanup).

    python oracle/gen_diagnostics_golden.py [--raw /tmp/pace_golden_diag]

The state after that step is the one already committed as tests/golden/c12_6rank_final.npz, so the fixture
tests/golden/c12_6rank_diagnostics.npz only holds what the diagnostics layer adds: the derived fields and the level
selections as the monitor received them (`quantity.view[:]`, compute domain), their dims / units, and for the plain
names a check (made here) that the monitor received the state fields themselves.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle.refshim.bootstrap import bootstrap  # noqa: E402

bootstrap()
import numpy as np  # noqa: E402

from oracle.refshim.proc_comm import run_ranks  # noqa: E402
from oracle.refshim.run_reference import build_dycore, default_config_kwargs  # noqa: E402

NAMES = ["u", "v", "w", "pt", "delp", "qvapor", "ua", "pe", "ps"]
DERIVED = ["column_integrated_qvapor", "column_integrated_qliquid", "column_integrated_qo3mr"]
Z_SELECT = [(40, ["pt", "w"]), (78, ["ua", "qvapor"]), (0, ["delp"])]


class RecordingMonitor:
    def __init__(self):
        self.stored = []

    def store(self, state):
        self.stored.append(dict(state))

    def store_constant(self, state):
        pass

    def cleanup(self):
        pass


def rank_main(mpi_comm, cfg_kwargs, out_dir):
    from oracle.refshim.bootstrap import bootstrap as bs

    bs()
    import types
    from datetime import timedelta

    sys.path.insert(0, "/root/reference/driver")
    from pace.driver.diagnostics import MonitorDiagnostics, ZSelect

    rank = mpi_comm.Get_rank()
    dycore, state, grid_data, damping, metric_terms, config, communicator = build_dycore(mpi_comm, cfg_kwargs)
    dycore.step_dynamics(state)
    mon = RecordingMonitor()
    diag = MonitorDiagnostics(monitor=mon, names=NAMES, derived_names=DERIVED,
                              z_select=[ZSelect(level=lv, names=nm) for lv, nm in Z_SELECT])
    diag.store(timedelta(seconds=config.dt_atmos), types.SimpleNamespace(dycore_state=state, physics_state=None))
    got = mon.stored[0]
    out, meta = {}, {}
    for name, q in got.items():
        if name == "time":
            continue
        view = np.asarray(q.view[:])
        meta[name] = {"dims": list(q.dims), "units": q.units, "shape": list(view.shape)}
        if name in NAMES:
            # the monitor is handed the state's own quantity
            assert q is getattr(state, name), name
            continue
        out[name] = view.copy()
    out["__meta__"] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
    os.makedirs(out_dir, exist_ok=True)
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), **out)
    return {"rank": rank, "stored": sorted(k for k in got if k != "time")}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--raw", default="/tmp/pace_golden_diag")
    args = ap.parse_args()
    res = run_ranks(6, rank_main, (default_config_kwargs(npx=13, n_split=1), args.raw), timeout=3600)
    print(res[0])
    merged = {}
    for r in range(6):
        z = np.load(os.path.join(args.raw, "rank%d.npz" % r))
        for k in z.files:
            if k == "__meta__":
                if r == 0:
                    merged[k] = z[k]
            else:
                merged["r%d:%s" % (r, k)] = z[k]
    dst = os.path.join(ROOT, "tests", "golden", "c12_6rank_diagnostics.npz")
    np.savez_compressed(dst, **merged)
    print("wrote", dst, "%.0f KB" % (os.path.getsize(dst) / 1e3))


if __name__ == "__main__":
    main()

"""TEST INFRASTRUCTURE ONLY — one dycore step of the UNMODIFIED reference (6 ranks, C12 L79) from a PERTURBED initial
state: noisy vertical velocity and noisy tracers of all species with negative patches, so that fillz, neg_adj3, the
w transport / damping and more branches of the saturation adjustment do real work.

    python oracle/gen_negq_golden.py   ->  tests/golden/c12_6rank_negq.npz  (perturbed inputs + final state)
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.refshim.bootstrap import bootstrap  # noqa: E402

bootstrap()
import numpy as np  # noqa: E402

from oracle.refshim.proc_comm import run_ranks  # noqa: E402

PERTURBED = ["w", "qvapor", "qliquid", "qrain", "qice", "qsnow", "qgraupel", "qo3mr", "qsgs_tke", "qcld"]


def _rank(mpi_comm):
    from oracle.refshim.run_reference import _arr, build_dycore, default_config_kwargs, state_arrays

    dycore, state, *_ = build_dycore(mpi_comm, default_config_kwargs())
    rank = mpi_comm.Get_rank()
    rng = np.random.default_rng(100 + rank)
    shape = _arr(state.delp).shape
    _arr(state.w)[...] = rng.normal(0.0, 0.5, size=shape)
    qv = _arr(state.qvapor)
    qv[...] = qv * (1.0 + 0.6 * rng.normal(size=shape)) - 2.0e-6 * rng.random(size=shape)
    for name, amp in (("qliquid", 2e-5), ("qrain", 1e-5), ("qice", 2e-5), ("qsnow", 1e-5), ("qgraupel", 5e-6),
                      ("qo3mr", 1e-7), ("qsgs_tke", 1e-2), ("qcld", 0.2)):
        a = _arr(getattr(state, name))
        a[...] = amp * (rng.normal(size=shape) + 0.3)
    out = {"in:" + n: _arr(getattr(state, n))[3:-3, 3:-3].copy() for n in PERTURBED}
    dycore.step_dynamics(state)
    for k, v in state_arrays(state).items():
        out["out:" + k] = v[3:-3, 3:-3] if v.ndim == 3 else v
    return out


if __name__ == "__main__":
    res = run_ranks(6, _rank, (), timeout=3600)
    flat = {}
    for r, d in enumerate(res):
        for k, v in d.items():
            flat["%d:%s" % (r, k)] = v
    frac = np.mean([np.mean(res[0]["in:" + n] < 0) for n in PERTURBED[1:]])
    print("mean fraction of negative tracer values in the input: %.3f" % frac)
    print("finite:", all(np.all(np.isfinite(v)) for v in flat.values()))
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "c12_6rank_negq.npz"), **flat)

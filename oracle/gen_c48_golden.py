"""TEST INFRASTRUCTURE ONLY — C48 fixtures (6 ranks of 48x48x79, one dycore step of the UNMODIFIED reference, numpy
backend, run in this container): a size at which the production kernels' tiling is real (2x2 tiles of 32, two column
strips of 26, several row chunks), unlike the C12 fixtures where every tiling is degenerate.

    python oracle/gen_c48_golden.py [--raw /tmp/pace_golden_c48] [--skip-run]

Writes
  tests/golden/c48_6rank_inputs.npz : what a step reads — the metric terms / damping coefficients of every rank and
        the engine.STEP_INPUTS fields of the analytic baroclinic state (compute domain; all-zero fields are left out)
  tests/golden/c48_6rank_final_sha256.json : SHA-256 of the compute domain of every DycoreState field of every rank
        after the step (the contract is bit-exact, so hashes are enough and a few KB instead of 150 MB), plus the
        reference's own wall time per step on this container's cores.
The raw arrays (incl. the rank-0 stage log with the d_sw / remap internals) stay under --raw for development.
"""
import argparse
import hashlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

N = 48
K_INTERFACE = ("pe", "peln", "pk")
STEP_INPUTS = ["u", "v", "w", "delp", "delz", "pt", "qvapor", "qliquid", "qice", "qrain", "qsnow", "qgraupel", "qo3mr",
               "qsgs_tke", "qcld", "phis"]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype=np.float64).tobytes()).hexdigest()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--raw", default="/tmp/pace_golden_c48")
    ap.add_argument("--skip-run", action="store_true")
    args = ap.parse_args()
    timing = None
    if not args.skip_run:
        from oracle.refshim.bootstrap import bootstrap

        bootstrap()
        from oracle.refshim.proc_comm import run_ranks
        from oracle.refshim.run_reference import default_config_kwargs, rank_main

        t0 = time.time()
        res = run_ranks(6, rank_main, (default_config_kwargs(npx=N + 1, n_split=1), 1, (0,), args.raw, 0, True),
                        timeout=14400)
        timing = {"wall_s": time.time() - t0, "step_s_per_rank": [r["step_s"][0] for r in res],
                  "cores": os.cpu_count(), "note": "rank 0 also records the stage log"}
        json.dump(timing, open(os.path.join(args.raw, "timing.json"), "w"))
    elif os.path.exists(os.path.join(args.raw, "timing.json")):
        timing = json.load(open(os.path.join(args.raw, "timing.json")))
    inputs, hashes = {}, {}
    for r in range(6):
        g = np.load(os.path.join(args.raw, "rank%d_grid.npz" % r))
        for k in g.files:
            if g[k].ndim <= 2:   # the step reads scalars, K-columns and 2-D metric terms only
                inputs["%d:%s" % (r, k)] = g[k]
        s = np.load(os.path.join(args.raw, "rank%d_init.npz" % r))
        for k in STEP_INPUTS:
            a = s[k]
            a = a[3:-3, 3:-3] if a.ndim == 3 else a   # 3-D: compute domain (+ interface row); phis is read in the halo
            if np.any(a != 0.0):
                inputs["%d:%s" % (r, k)] = a
        f = np.load(os.path.join(args.raw, "rank%d_final.npz" % r))
        for k in f.files:
            a = f[k]
            ni = N + 1 if k == "v" else N
            nj = N + 1 if k == "u" else N
            a = a[3:3 + ni, 3:3 + nj]
            if a.ndim == 3 and k not in K_INTERFACE:
                a = a[:, :, :-1]   # the extra allocated level (NDSL sizer.py:141-154) is not part of the field
            hashes["%d:%s" % (r, k)] = sha(a)
    out = os.path.join(ROOT, "tests", "golden")
    np.savez_compressed(os.path.join(out, "c48_6rank_inputs.npz"), **inputs)
    json.dump({"what": "sha256 of float64 C-order bytes of field[3:3+ni, 3:3+nj, :nk] (reference layout [i][j][k]; "
                       "nk = 80 for pe / peln / pk, else 79; ni/nj = 49 on the staggered axis of v/u) after one "
                       "step_dynamics",
               "reference_timing": timing, "sha256": hashes},
              open(os.path.join(out, "c48_6rank_final_sha256.json"), "w"), indent=0, sort_keys=True)
    print("wrote", len(inputs), "input arrays,", len(hashes), "hashes")


if __name__ == "__main__":
    main()

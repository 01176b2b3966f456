"""TEST INFRASTRUCTURE ONLY — does the UNMODIFIED reference dycore run with PACE_FLOAT_PRECISION=32?

    python oracle/refshim/warm_cache.py          (with PACE_FLOAT_PRECISION=32 in the environment: single-process compile)
    python oracle/probe_reference_f32.py [--validate-args]

Runs one C12 L79 step of the reference (same config as oracle/gen_golden.py) on 6 ranks with
`Float = np.float32` (NDSL/ndsl/dsl/typing.py:25-55) and reports, per logged module call of rank 0, how many values of
each changed array are not finite.  Outcome in this container (profiles/r02_reference_f32_probe.txt):
  * with the reference's default `validate_args=True` (NDSL/ndsl/dsl/stencil_config.py:33) the step raises
    TypeError "The type of parameter 'ptop' is <class 'float'> instead of 'float32'" at pyFV3/stencils/dyn_core.py:776;
  * with `validate_args=False` it runs, and the first d_sw call leaves u and v non-finite on every cell; after the step
    pt, pkz, u, v, qvapor, q_con are non-finite on the whole compute domain.
So there is no fp32 behaviour of the whole dycore to reproduce; the fp32 row of the scope table stays at MapSingle
(tests/test_mapsingle_f32.py), which the reference does run correctly in fp32.
"""
import argparse
import json
import os
import sys

os.environ["PACE_FLOAT_PRECISION"] = "32"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--validate-args", action="store_true", help="keep the reference's default argument validation")
    ap.add_argument("--raw", default="/tmp/pace_ref_f32_probe")
    args = ap.parse_args()
    os.environ["PACE_REF_VALIDATE_ARGS"] = "1" if args.validate_args else "0"
    from oracle.refshim.bootstrap import bootstrap

    bootstrap()
    import numpy as np

    from oracle.refshim.proc_comm import run_ranks
    from oracle.refshim.run_reference import default_config_kwargs, rank_main

    cfg = default_config_kwargs(npx=13, n_split=1)
    try:
        run_ranks(6, rank_main, (cfg, 1, (0,), args.raw, 0, False), timeout=3000)
    except RuntimeError as e:
        last = [ln for ln in str(e).splitlines() if ln.strip()]
        print("reference step failed:", last[-1])
        where = [ln.strip() for ln in last if "/root/reference/pyFV3" in ln]
        if where:
            print("at", where[-1])
        return 0
    st = np.load(os.path.join(args.raw, "rank0_stages.npz"))
    labels = json.loads(bytes(st["__labels__"]).decode())
    nx = 12
    for key in st.keys():
        if key == "__labels__" or key.startswith("000:"):
            continue
        a = st[key]
        if a.dtype.kind != "f" or a.ndim != 3:
            continue
        bad = ~np.isfinite(a)
        if bad.any():
            seq, name = key.split(":", 1)
            print("%-14s %-12s dtype %s  non-finite: %6d of %d, on the compute cells %6d of %d" % (
                labels[int(seq)][0], name, a.dtype, bad.sum(), a.size, bad[3:3 + nx, 3:3 + nx, :79].sum(), nx * nx * 79))
    fin = np.load(os.path.join(args.raw, "rank0_final.npz"))
    print("after the step, compute cells non-finite:",
          {k: int((~np.isfinite(fin[k][3:3 + nx, 3:3 + nx, :79])).sum()) for k in ("u", "v", "w", "pt", "delp", "delz", "pkz", "qvapor", "q_con")})
    return 0


if __name__ == "__main__":
    sys.exit(main())

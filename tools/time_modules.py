"""Time single module calls of the C ABI on the C384 workload (CUDA events on the launching stream, working set >> L2).

    python tools/time_modules.py [--tile 384 --layout 2] c_sw:112.5 d_sw:225 riem_solver_c:112.5 ...

Each spec is name[:arg[:arg...]] (numbers are passed as float unless they have no '.', strings as field names).  One
dycore step is run first so that the fields hold realistic values; modules that update their inputs in place are timed
on the drifting state, which is fine for bandwidth-bound kernels."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tile", type=int, default=384)
    ap.add_argument("--layout", type=int, default=2)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("specs", nargs="+")
    args = ap.parse_args()
    import torch

    from pace_b200.synthetic import make_engine

    eng = make_engine(args.tile, args.layout, npz=79, device="cuda:0", n_split=6, k_split=1, dt_atmos=225.0)
    eng.step_dynamics()
    torch.cuda.synchronize()
    core = eng.core
    out = {}
    for spec in args.specs:
        parts = spec.split(":")
        name, raw = parts[0], parts[1:]
        call_args = []
        for r in raw:
            try:
                call_args.append(float(r) if "." in r else int(r))
            except ValueError:
                call_args.append(r)
        core.call(name, *call_args)  # warm-up (first-use allocations, attributes)
        torch.cuda.synchronize()
        l0 = core.lib.pace_b200_launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.reps):
            core.call(name, *call_args)
        e1.record()
        torch.cuda.synchronize()
        out[spec] = {"ms": e0.elapsed_time(e1) / args.reps,
                     "launches": (core.lib.pace_b200_launch_count() - l0) // args.reps}
    print(json.dumps(out))


if __name__ == "__main__":
    main()

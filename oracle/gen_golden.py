"""TEST INFRASTRUCTURE ONLY — generate the golden fixtures under tests/golden/ by running the
UNMODIFIED reference (pyFV3 numpy backend, /root/reference) on 6 ranks in this container.

    python oracle/gen_golden.py [--raw /tmp/pace_golden_raw] [--n-split 1]

Writes
  tests/golden/c12_rank0_stages[_ns<N>].npz : rank 0, one dycore step, C12 L79, config of
        tests/main/fv3core/test_dycore_call.py:28-78 (== driver/examples/configs/baroclinic_c12.yaml):
        initial contents of every watched array + for each module call the arrays it changed
        (full arrays incl. halos).  Keys "<seq>:<name>", JSON label list in "__labels__".
  tests/golden/c12_rank0_grid.npz           : metric terms / damping coefficients of rank 0
  tests/golden/c12_6rank_grid.npz, c12_6rank_init.npz, c12_6rank_final.npz :
        all six ranks (grid full arrays; state = compute domain only, halos are exchanged by the code
        under test), for end-to-end parity of one full step with real halo exchange.
The transcendentals of the oracle are pinned to glibc libm (see refshim/bootstrap.py).
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
from oracle.refshim.run_reference import default_config_kwargs, rank_main  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--raw", default="/tmp/pace_golden_raw")
    ap.add_argument("--n-split", type=int, default=1)
    ap.add_argument("--npx", type=int, default=13)
    ap.add_argument("--steps", type=int, default=1)
    ap.add_argument("--skip-run", action="store_true")
    ap.add_argument("--stages-only", action="store_true")
    ap.add_argument("--k-split", type=int, default=1)
    ap.add_argument("--final-only", action="store_true",
                    help="only the state after --steps steps of every rank (multi-substep / multi-step fixture)")
    ap.add_argument("--layout", type=int, default=1, help="layout x layout subdomains per tile (no stage log if > 1)")
    ap.add_argument("--deep", action="store_true", help="also hook d_sw/updatedzd/nh_p_grad internals (dev only)")
    args = ap.parse_args()
    tag = "c%d" % (args.npx - 1)
    suffix = "" if args.n_split == 1 else "_ns%d" % args.n_split
    if args.deep:
        suffix += "_deep"
    raw = args.raw + suffix
    if args.layout > 1:
        return multi_layout(args, tag, raw + "_l%d" % args.layout)
    if args.final_only:
        return final_only(args, tag, raw + "_ks%d_s%d" % (args.k_split, args.steps))
    if not args.skip_run:
        cfg = default_config_kwargs(npx=args.npx, n_split=args.n_split)
        res = run_ranks(6, rank_main, (cfg, args.steps, (0,), raw, 0, args.deep), timeout=7200)
        for r in res:
            print(r)
    out = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out, exist_ok=True)
    st = np.load(os.path.join(raw, "rank0_stages.npz"))
    np.savez_compressed(os.path.join(out, "%s_rank0_stages%s.npz" % (tag, suffix)), **{k: st[k] for k in st.files})
    labels = json.loads(bytes(st["__labels__"]).decode())
    for n, (label, changed) in enumerate(labels):
        print(n, label, changed if n else "(%d arrays)" % len(changed))
    if args.stages_only:
        return
    g0 = np.load(os.path.join(raw, "rank0_grid.npz"))
    np.savez_compressed(os.path.join(out, "%s_rank0_grid.npz" % tag), **{k: g0[k] for k in g0.files})
    grid, init, final = {}, {}, {}
    for r in range(6):
        g = np.load(os.path.join(raw, "rank%d_grid.npz" % r))
        for k in g.files:
            grid["%d:%s" % (r, k)] = g[k]
        for src, dst in (("init", init), ("final", final)):
            s = np.load(os.path.join(raw, "rank%d_%s.npz" % (r, src)))
            for k in s.files:
                a = s[k]
                # compute domain (+1 for interface dims is kept: slice 3:-3 of the (n+7) allocation)
                # 3-D: compute domain only; 2-D (phis, ps): full array — phis is read in the halo
                dst["%d:%s" % (r, k)] = a[3:-3, 3:-3] if a.ndim == 3 else a
    np.savez_compressed(os.path.join(out, "%s_6rank_grid.npz" % tag), **grid)
    np.savez_compressed(os.path.join(out, "%s_6rank_init%s.npz" % (tag, suffix)), **init)
    np.savez_compressed(os.path.join(out, "%s_6rank_final%s.npz" % (tag, suffix)), **final)


def final_only(args, tag, raw):
    """6 ranks, n_split / k_split / steps as given: the final state only (the initial state and the grid are those
    of the n_split=1 fixture)."""
    if not args.skip_run:
        cfg = default_config_kwargs(npx=args.npx, n_split=args.n_split, k_split=args.k_split)
        for r in run_ranks(6, rank_main, (cfg, args.steps, (), raw, 0, False), timeout=7200):
            print(r)
    final = {}
    for r in range(6):
        s = np.load(os.path.join(raw, "rank%d_final.npz" % r))
        for k in s.files:
            a = s[k]
            final["%d:%s" % (r, k)] = a[3:-3, 3:-3] if a.ndim == 3 else a
    out = os.path.join(ROOT, "tests", "golden")
    np.savez_compressed(os.path.join(out, "%s_6rank_final_ns%d_ks%d_s%d.npz" % (tag, args.n_split, args.k_split, args.steps)),
                        **final)


def multi_layout(args, tag, raw):
    """6 * layout^2 ranks: inputs and the state after the step(s) of every rank (no per-stage log)."""
    n = 6 * args.layout * args.layout
    if not args.skip_run:
        cfg = default_config_kwargs(npx=args.npx, n_split=args.n_split, layout=(args.layout, args.layout))
        for r in run_ranks(n, rank_main, (cfg, args.steps, (), raw, 0, False), timeout=7200):
            print(r)
    grid, init, final = {}, {}, {}
    for r in range(n):
        g = np.load(os.path.join(raw, "rank%d_grid.npz" % r))
        for k in g.files:
            grid["%d:%s" % (r, k)] = g[k]
        for src, dst in (("init", init), ("final", final)):
            s = np.load(os.path.join(raw, "rank%d_%s.npz" % (r, src)))
            for k in s.files:
                a = s[k]
                dst["%d:%s" % (r, k)] = a[3:-3, 3:-3] if a.ndim == 3 else a
    out = os.path.join(ROOT, "tests", "golden")
    for name, d in (("grid", grid), ("init", init), ("final", final)):
        np.savez_compressed(os.path.join(out, "%s_%drank_%s.npz" % (tag, n, name)), **d)


if __name__ == "__main__":
    main()

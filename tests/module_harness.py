"""Run one module entry point of the C ABI on the state the reference had just before the same call
(golden stage log) and compare everything it changed — and that it changed nothing else."""
import numpy as np

from tests.golden_util import passes, report

DT_ATMOS = 225.0
DT = DT_ATMOS  # k_split = n_split = 1 in the golden config
DT2 = 0.5 * DT


def make_core(backend, nx=12, ny=12, npz=79, n_tiles=1, edge_flags=None, **kw):
    from pace_b200.native import NativeCore

    if backend == "hostemu":
        from tests import hostemu

        return NativeCore(nx, ny, npz, n_tiles=n_tiles, edge_flags=edge_flags, device="cpu",
                          _lib_handle=hostemu.load(), **kw)
    return NativeCore(nx, ny, npz, n_tiles=n_tiles, edge_flags=edge_flags, device="cuda:0", **kw)


def load_state(core, state, grid):
    for name, arr in state.items():
        core.put(name, [arr])
    core.put_grid([grid])
    core.set_columns_from_config()


# fields defined on vertical interfaces (npz+1 levels); every other 3-D array has npz meaningful levels and
# the extra allocated level (NDSL sizer.py:141-154) is not compared
K_INTERFACE = {"pe", "peln", "pk", "ad.gz", "ad.zh", "ad.pkc", "ad.pk3", "ad.pem"}


def interior(a, nx=12, ny=12, pad=0):
    """compute-domain (+pad) view of a reference-layout array; pad = n or (low side, high side)"""
    lo, hi = pad if isinstance(pad, tuple) else (pad, pad)
    return a[3 - lo:3 + nx + hi, 3 - lo:3 + ny + hi]


def check_module(backend, golden, grid, stage, call, args, tol=0.0, near_zero=0.0, cfg=None, skip=(), begin=None,
                 region=None):
    """begin: label of the stage whose *end* is the starting state (default "pre:<stage>").
    region: optional dict name -> pad; those arrays are compared on the compute domain (+pad) only —
    used for internal work arrays whose unused rows/columns hold leftovers in the reference."""
    cfg = dict(cfg or {})
    cfg.setdefault("ptop", float(grid["grid.ptop"]))
    cfg.setdefault("da_min", float(grid["damp.da_min"]))
    cfg.setdefault("da_min_c", float(grid["damp.da_min_c"]))
    core = make_core(backend, **cfg)
    before = golden.state_at(golden.index[begin]) if begin else golden.before(stage)
    region = region or {}
    load_state(core, before, grid)
    core.call(call, *args)
    core.sync()
    after = golden.after(stage)
    changed = golden.changed_by(stage)
    results = {}
    worst = 0.0
    failed = False
    for name in before:
        if name in skip:
            continue
        got = core.get(name)[0]
        want = after[name]
        if got.ndim == 3 and name not in K_INTERFACE and not name.startswith("dzd."):
            got, want = got[:, :, :-1], want[:, :, :-1]
        if name in region:
            got, want = interior(got, pad=region[name]), interior(want, pad=region[name])
        err, nbad = report(name, got, want, near_zero)
        results[name] = (err, nbad, name in changed)
        worst = max(worst, err)
        failed = failed or not passes(err, nbad, tol)
    msg = "\n".join(
        "  %-16s %s max_rel=%.3e non-identical=%d" % (n, "out" if c else "   ", e, b)
        for n, (e, b, c) in sorted(results.items()) if c or b
    )
    print("[%s/%s]\n%s" % (backend, stage, msg))
    assert not failed and worst <= tol, "parity failed for %s (tol %g):\n%s" % (stage, tol, msg)
    return results

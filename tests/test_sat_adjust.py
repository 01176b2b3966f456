"""SatAdjust3d kernel against the reference's stage log (rank 0, C12 L79)."""
import numpy as np
import pytest

from tests.module_harness import K_INTERFACE, load_state, make_core
from tests.golden_util import report


def kmp_of(pfull, npz):
    """remapping.py:347-351"""
    for k in range(npz):
        if pfull[k] > 10.0e2:
            return k
    return npz - 1


def run(backend, golden_deep, grid0):
    core = make_core(backend, ptop=float(grid0["grid.ptop"]), da_min=float(grid0["damp.da_min"]),
                     da_min_c=float(grid0["damp.da_min_c"]))
    before = golden_deep.before("rm.saturation_adjustment")
    load_state(core, before, grid0)
    core.sat_adjust(225.0, 1, kmp_of(grid0["grid.p"], 79))
    core.sync()
    after = golden_deep.after("rm.saturation_adjustment")
    bad = []
    for name in before:
        got, want = core.get(name)[0], after[name]
        if got.ndim == 3 and name not in K_INTERFACE and not name.startswith("dzd."):
            got, want = got[:, :, :-1], want[:, :, :-1]
        err, nbad = report(name, got, want)
        if nbad:
            bad.append("%s max_rel=%.3e non-identical=%d" % (name, err, nbad))
    assert not bad, "\n".join(bad)


def test_sat_adjust_hostemu(golden_deep, grid0):
    run("hostemu", golden_deep, grid0)


@pytest.mark.gpu
def test_sat_adjust_gpu(golden_deep, grid0):
    run("cuda", golden_deep, grid0)

"""The halo-exchange kernel (csrc/halo.cu through the C ABI) against the reference's source maps: all 6 (or 24)
subdomains resident on one device, one launch per update."""
import os

import numpy as np
import pytest

from pace_b200.halo import HaloPlanner
from tests.test_halo_plan import Z, coded


def run(backend, layout):
    from tests.module_harness import make_core

    pl = HaloPlanner(12, layout)
    nt = pl.total_ranks
    core = make_core(backend, nx=pl.n, ny=pl.n, npz=2, n_tiles=nt, edge_flags=[pl.edge_flags(r) for r in range(nt)])
    ident = {r: r for r in range(nt)}
    core.set_halo_plan(0, pl.scalar_plan(False, False, 3), ident)
    core.set_halo_plan(1, pl.scalar_plan(False, False, 2), ident)
    core.set_halo_plan(2, pl.scalar_plan(True, True, 3), ident)
    core.set_halo_plan(3, pl.vector_plan("d", 3), ident)
    core.set_halo_plan(4, pl.vector_plan("c", 3), ident)
    core.set_halo_plan(5, pl.interface_sync_plan(), ident)

    def field(xi, yi, comp):
        return [np.repeat(coded(pl.n, xi, yi, r, comp)[:, :, None], 3, axis=2) for r in range(nt)]

    def check(name, key):
        got = core.get(name)
        for r in range(nt):
            for k in range(2):
                np.testing.assert_array_equal(got[r][:, :, k], Z[key][r], err_msg="%s rank %d" % (key, r))

    for pid, key, xi, yi in ((0, "a3", False, False), (1, "a2", False, False), (2, "b3", True, True)):
        core.put("q", field(xi, yi, 1))
        core.halo_update(pid, ["q"], nk=2)
        check("q", "l%d_%s" % (layout, key))
    for pid, key, sx, sy in ((3, "d3", (False, True), (True, False)), (4, "c3", (True, False), (False, True)),
                             (5, "sync", (False, True), (True, False))):
        core.put("vx", field(sx[0], sx[1], 1))
        core.put("vy", field(sy[0], sy[1], 2))
        core.halo_update(pid, ["vx"], ["vy"], nk=2)
        check("vx", "l%d_%s_x" % (layout, key))
        check("vy", "l%d_%s_y" % (layout, key))


@pytest.mark.parametrize("layout", [1, 2])
def test_halo_kernel_hostemu(layout):
    run("hostemu", layout)


@pytest.mark.gpu
@pytest.mark.parametrize("layout", [1, 2])
def test_halo_kernel_gpu(layout):
    run("cuda", layout)

"""The halo plan (pace_b200/halo.py) against source maps produced by the reference's own HaloUpdater
(tests/golden/halo_maps.npz, written by oracle/gen_halo_golden.py): bit-exact for scalar A-grid (3 and 2
points), B-grid, D- and C-grid vector pairs and the vector-interface synchronisation, layouts 1x1, 2x2, 3x3."""
import os

import numpy as np
import pytest

from pace_b200.halo import H, HaloPlanner, apply_plan_numpy

Z = np.load(os.path.join(os.path.dirname(__file__), "golden", "halo_maps.npz"))
NZ_TILE = 12


def coded(n, xi, yi, rank, comp):
    a = np.zeros((n + 7, n + 7))
    ex, ey = n + int(xi), n + int(yi)
    ii, jj = np.meshgrid(np.arange(H, H + ex), np.arange(H, H + ey), indexing="ij")
    a[H:H + ex, H:H + ey] = comp * 1e6 + rank * 1e4 + ii * 100 + jj
    return a


@pytest.mark.parametrize("layout", [1, 2, 3])
@pytest.mark.parametrize("case,xi,yi,npts", [("a3", False, False, 3), ("a2", False, False, 2), ("b3", True, True, 3)])
def test_scalar_plan(layout, case, xi, yi, npts):
    pl = HaloPlanner(NZ_TILE, layout)
    arrs = [coded(pl.n, xi, yi, r, 1) for r in range(pl.total_ranks)]
    apply_plan_numpy(pl.scalar_plan(xi, yi, npts), arrs)
    want = Z["l%d_%s" % (layout, case)]
    for r in range(pl.total_ranks):
        np.testing.assert_array_equal(arrs[r], want[r], err_msg="rank %d" % r)


@pytest.mark.parametrize("layout", [1, 2, 3])
@pytest.mark.parametrize("grid", ["d", "c"])
def test_vector_plan(layout, grid):
    pl = HaloPlanner(NZ_TILE, layout)
    stag = {"d": ((False, True), (True, False)), "c": ((True, False), (False, True))}[grid]
    xs = [coded(pl.n, stag[0][0], stag[0][1], r, 1) for r in range(pl.total_ranks)]
    ys = [coded(pl.n, stag[1][0], stag[1][1], r, 2) for r in range(pl.total_ranks)]
    apply_plan_numpy(pl.vector_plan(grid, 3), xs, ys)
    wx, wy = Z["l%d_%s3_x" % (layout, grid)], Z["l%d_%s3_y" % (layout, grid)]
    for r in range(pl.total_ranks):
        np.testing.assert_array_equal(xs[r], wx[r], err_msg="x rank %d" % r)
        np.testing.assert_array_equal(ys[r], wy[r], err_msg="y rank %d" % r)


@pytest.mark.parametrize("layout", [1, 2, 3])
def test_interface_sync_plan(layout):
    pl = HaloPlanner(NZ_TILE, layout)
    xs = [coded(pl.n, False, True, r, 1) for r in range(pl.total_ranks)]
    ys = [coded(pl.n, True, False, r, 2) for r in range(pl.total_ranks)]
    apply_plan_numpy(pl.interface_sync_plan(), xs, ys)
    wx, wy = Z["l%d_sync_x" % layout], Z["l%d_sync_y" % layout]
    for r in range(pl.total_ranks):
        np.testing.assert_array_equal(xs[r], wx[r], err_msg="x rank %d" % r)
        np.testing.assert_array_equal(ys[r], wy[r], err_msg="y rank %d" % r)

"""FiniteVolumeTransport alone on non-trivial data.  The C12 baroclinic fixture starts from w = 0 and smooth fields, so
the ord-8 PPM limiters and both upwind directions are only lightly exercised by the step tests.  Here the reference's
own FiniteVolumeTransport instances (hord 8 with mass fluxes as in tracer advection; hord 6 with mass fluxes; hord 6
with DelnFlux damping, with and without a mass weight) were called on synthetic fields with gradients everywhere and
Courant numbers of both signs, on rank 0 = a whole tile, so every tile edge and corner formula is hit
(tests/golden/fvtp2d_synthetic.npz, oracle/gen_fvtp2d_golden.py).  pace_b200_fvtp2d must return the same fluxes,
bit for bit, on the compute domain."""
import os

import numpy as np
import pytest

from tests.golden_util import GOLDEN
from tests.module_harness import load_state, make_core

CASES = {
    "hord8_massflux": dict(hord=8, mf=True, mass=False, nord="", damp=""),
    "hord6_massflux": dict(hord=6, mf=True, mass=False, nord="", damp=""),
    "hord6_deln": dict(hord=6, mf=False, mass=False, nord="col.nord_v", damp="col.delnflux_damp_vt_fv"),
    "hord6_deln_mass": dict(hord=6, mf=True, mass=True, nord="col.nord_t", damp="col.delnflux_damp_t_fv"),
}


def run(backend, grid0):
    z = np.load(os.path.join(GOLDEN, "fvtp2d_synthetic.npz"))
    core = make_core(backend, ptop=float(grid0["grid.ptop"]), da_min=float(grid0["damp.da_min"]),
                     da_min_c=float(grid0["damp.da_min_c"]))
    state = {"t.crx": z["crx"], "t.cry": z["cry"], "t.xfx": z["xfx"], "t.yfx": z["yfx"], "t.mfx": z["mfx"],
             "t.mfy": z["mfy"], "t.mass": z["mass"], "t.q": z["q"], "t.fx": 0 * z["q"], "t.fy": 0 * z["q"]}
    load_state(core, state, grid0)
    for name, c in CASES.items():
        core.put("t.q", [z["q"]])
        core.put("t.fx", [0 * z["q"]])
        core.put("t.fy", [0 * z["q"]])
        core.call("fvtp2d", "t.q", "t.crx", "t.cry", "t.xfx", "t.yfx", "t.fx", "t.fy", "t.mfx" if c["mf"] else "",
                  "t.mfy" if c["mf"] else "", "t.mass" if c["mass"] else "", c["hord"], c["nord"], c["damp"], 79)
        core.sync()
        fx, fy = core.get("t.fx")[0], core.get("t.fy")[0]
        wfx, wfy = z[name + ":fx"], z[name + ":fy"]
        # fx on [is..ie+1] x [js..je], fy on [is..ie] x [js..je+1], 79 levels
        gx, wx = fx[3:16, 3:15, :79], wfx[3:16, 3:15, :79]
        gy, wy = fy[3:15, 3:16, :79], wfy[3:15, 3:16, :79]
        assert np.all(np.isfinite(wx)) and np.abs(wx).max() > 0
        assert np.array_equal(gx, wx), "%s fx: max |d| = %g" % (name, np.max(np.abs(gx - wx)))
        assert np.array_equal(gy, wy), "%s fy: max |d| = %g" % (name, np.max(np.abs(gy - wy)))
        # the transported field keeps its values except for the x-copied corners (as in the reference)
        assert np.array_equal(core.get("t.q")[0][3:15, 3:15, :79], z[name + ":q"][3:15, 3:15, :79])


def test_fvtp2d_synthetic_hostemu(grid0):
    run("hostemu", grid0)


@pytest.mark.gpu
def test_fvtp2d_synthetic_gpu(grid0):
    run("gpu", grid0)

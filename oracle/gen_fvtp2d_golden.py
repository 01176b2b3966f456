"""TEST INFRASTRUCTURE ONLY — golden vectors for FiniteVolumeTransport alone on NON-TRIVIAL data: the instances the
reference dycore itself builds (tracer advection: hord 8 with mass fluxes; d_sw: hord 6 with mass fluxes, and hord 6
with DelnFlux damping) are called on synthetic fields on rank 0 of a 6-rank C12 run (real grid: rank 0 is a whole
tile, so all four tile edges and corners are exercised with gradients everywhere and both flow directions).

    python oracle/gen_fvtp2d_golden.py      ->  tests/golden/fvtp2d_synthetic.npz
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.refshim.bootstrap import bootstrap  # noqa: E402

bootstrap()
import numpy as np  # noqa: E402

from oracle.refshim.proc_comm import run_ranks  # noqa: E402


def _rank(mpi_comm, out_path):
    from oracle.refshim.run_reference import _arr, build_dycore, default_config_kwargs

    dycore, state, grid_data, damping, _, config, communicator = build_dycore(mpi_comm, default_config_kwargs())
    if mpi_comm.Get_rank() != 0:
        return None
    rng = np.random.default_rng(0)
    shape = _arr(state.delp).shape  # (19, 19, 80)
    n = shape[0]
    x = np.linspace(0, 2 * np.pi, n)[:, None, None]
    y = np.linspace(0, 2 * np.pi, n)[None, :, None]
    z = np.linspace(0, 1, shape[2])[None, None, :]
    q = 1.0 + 0.5 * np.sin(2 * x + z) * np.cos(3 * y) + rng.normal(0, 0.08, size=shape)
    crx = 0.28 * np.sin(x + 2 * y + 3 * z) + rng.normal(0, 0.02, size=shape)
    cry = 0.28 * np.cos(2 * x - y + z) + rng.normal(0, 0.02, size=shape)
    area = _arr(grid_data.area)[:, :, None]
    xfx = crx * area * 0.9
    yfx = cry * area * 0.9
    mfx = xfx * (1000.0 + 50 * np.cos(x))
    mfy = yfx * (1000.0 + 50 * np.sin(y))
    mass = 1000.0 + 100.0 * np.sin(x + y) + 0 * z
    out = {"q": q, "crx": crx, "cry": cry, "xfx": xfx, "yfx": yfx, "mfx": mfx, "mfy": mfy, "mass": mass}

    dsw = dycore.acoustic_dynamics.dgrid_shallow_water_lagrangian_dynamics
    cases = {"hord8_massflux": (dycore.tracer_advection.finite_volume_transport, True, False),
             "hord6_massflux": (dsw.fvtp2d_vt_nodelnflux, True, False),
             "hord6_deln": (dsw.fvtp2d_dp, False, False),
             "hord6_deln_mass": (dsw.fvtp2d_dp_t, True, True)}

    def quantity_like(a):
        import copy

        qq = copy.deepcopy(state.delp)
        _arr(qq)[...] = a
        return qq

    for name, (fvt, with_mf, with_mass) in cases.items():
        qq, fx, fy = quantity_like(q), quantity_like(0 * q), quantity_like(0 * q)
        kw = {}
        if with_mf:
            kw = dict(x_mass_flux=quantity_like(mfx), y_mass_flux=quantity_like(mfy))
        if with_mass:
            kw["mass"] = quantity_like(mass)
        fvt(qq, quantity_like(crx), quantity_like(cry), quantity_like(xfx), quantity_like(yfx), fx, fy, **kw)
        out[name + ":fx"], out[name + ":fy"], out[name + ":q"] = _arr(fx).copy(), _arr(fy).copy(), _arr(qq).copy()
    np.savez_compressed(out_path, **out)
    return sorted(out)


if __name__ == "__main__":
    path = os.path.join(ROOT, "tests", "golden", "fvtp2d_synthetic.npz")
    print(run_ranks(6, _rank, (path,), timeout=3600)[0])

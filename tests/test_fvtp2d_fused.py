"""The fused fvtp2d kernels — row-marching (pace_b200/csrc/fvtp2d_march.cu, the production path) and CTA-tiled
(fvtp2d_fused.cu) — against the straightforward three-kernel version (transport.cu, the one validated stencil by
stencil against the reference) on a domain that spans several CTA tiles / row chunks / column strips: one full dycore
step at C48 (6 subdomains of 48x48x79), bit for bit.  The baseline run also uses the unfused DelnFlux kernels and the
global-memory Riemann solvers, so the fused / on-chip versions of those are covered by the same comparison."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def comparable(name, a, n):
    """uc / vc double as work arrays of the divergence damping: their halo holds results of its earlier, wider
    iterations in the stencil-by-stencil version and is not written by the marching kernel (nothing reads it);
    everything else is compared on the whole arrays.  Layout [t][k][j][i], n compute cells per side."""
    return a[:, :, 3:3 + n + 1, 3:3 + n + 1] if name in ("uc", "vc") else a


# the other kernels that have a simpler counterpart kept for A/B runs: level-parallel nh_p_grad winds and update_dz_c, the
# separate update_u_and_v pass of d_sw, the four-field interface spline of update_dz_d
ALT = {"PACE_B200_NHPG": "levels", "PACE_B200_DZC": "levels", "PACE_B200_DSW_UV": "split", "PACE_B200_DZD_SPLINE": "field"}

SCRIPT = r"""
import sys, numpy as np, torch
sys.path.insert(0, %r)
from pace_b200.synthetic import make_engine
from pace_b200.engine import STATE_3D
eng = make_engine(int(sys.argv[2]), 1, n_split=2)
eng.step_dynamics()
eng.core.sync()
out = {n: eng.core.t[n].cpu().numpy() for n in STATE_3D}
np.savez(sys.argv[1], **out)
"""


@pytest.mark.gpu
def test_fused_matches_three_kernel_version(tmp_path):
    res = {}
    modes = {"march": {"PACE_B200_FVTP2D": "march"},
             "march_chunks": {"PACE_B200_FVTP2D": "march", "PACE_B200_FVTP2D_ROWS": "16"},
             "tile": {"PACE_B200_FVTP2D": "tile"}, "riem_smem": {"PACE_B200_RIEM": "smem"}, "unfused": {"PACE_B200_FVTP2D": "unfused", "PACE_B200_DELNFLUX": "unfused", "PACE_B200_RIEM": "global", "PACE_B200_A2B": "unfused", "PACE_B200_CSW": "unfused", "PACE_B200_DD": "unfused", **ALT}}
    for mode, extra in modes.items():
        env = dict(os.environ, **extra)
        path = str(tmp_path / (mode + ".npz"))
        subprocess.run([sys.executable, "-c", SCRIPT % ROOT, path, "48"], check=True, env=env, timeout=600)
        res[mode] = np.load(path)
    for mode in modes:
        if mode == "unfused":
            continue
        for name in res[mode].files:
            a, b = comparable(name, res[mode][name], 48), comparable(name, res["unfused"][name], 48)
            assert np.all(np.isfinite(a)), (mode, name)
            assert np.array_equal(a, b), "%s: %s differs: max |d| = %g" % (mode, name, np.max(np.abs(a - b)))


@pytest.mark.gpu
def test_production_kernels_match_unfused_at_c96(tmp_path):
    """Same comparison for the production configuration only, on 96 x 96 subdomains: 4 column strips (the last one
    narrower), several row chunks, 3 x 3 DelnFlux / a2b tiles."""
    res = {}
    for mode, extra in {"production": {}, "unfused": {"PACE_B200_FVTP2D": "unfused", "PACE_B200_DELNFLUX": "unfused",
                                                      "PACE_B200_RIEM": "global", "PACE_B200_A2B": "unfused", "PACE_B200_CSW": "unfused", "PACE_B200_DD": "unfused", **ALT}}.items():
        env = dict(os.environ, **extra)
        path = str(tmp_path / (mode + ".npz"))
        subprocess.run([sys.executable, "-c", SCRIPT % ROOT, path, "96"], check=True, env=env, timeout=900)
        res[mode] = np.load(path)
    for name in res["production"].files:
        a, b = comparable(name, res["production"][name], 96), comparable(name, res["unfused"][name], 96)
        assert np.all(np.isfinite(a)), name
        assert np.array_equal(a, b), "%s differs: max |d| = %g" % (name, np.max(np.abs(a - b)))


PATCHY = r"""
import sys, numpy as np, torch
sys.path.insert(0, %r)
from pace_b200.synthetic import make_engine
from pace_b200.engine import STATE_3D
eng = make_engine(int(sys.argv[2]), 1, n_split=2)
t = eng.core.t
nt, nk, nj, sj = t["qvapor"].shape
g = torch.Generator().manual_seed(7)
jj = torch.arange(nj).view(1, 1, nj, 1).double()
ii = torch.arange(sj).view(1, 1, 1, sj).double()
kk = torch.arange(nk).view(1, nk, 1, 1)
tt = torch.arange(nt).view(nt, 1, 1, 1)
dev = t["qvapor"].device
def put(name, mask, amp):
    noise = (0.5 + torch.rand(t[name].shape, generator=g, dtype=torch.float64)).to(dev)
    t[name].copy_(torch.where(mask.to(dev).expand(t[name].shape), amp * noise, torch.zeros_like(t[name])))
# a disk of rain on the lower levels of every subdomain (zero rows above and below it inside a strip)
put("qrain", ((ii - 20) ** 2 + (jj - 30) ** 2 < 49) & (kk > 50), 1.0e-4)
# snow: a single cell, and one whole row, on two levels
put("qsnow", (((ii == 11) & (jj == 12)) | (jj == 40)) & ((kk == 30) | (kk == 31)), 2.0e-5)
# ice: whole levels 10..19 only
put("qice", (kk >= 10) & (kk < 20) & (ii >= 0), 1.0e-5)
# cloud water: only in the first columns / rows next to a tile edge (aprons of the first strip and chunk)
put("qliquid", ((ii < 6) | (jj < 5)) & (kk > 60), 3.0e-5)
# ozone: on one level of ONE subdomain only, next to its east edge: the neighbours' copies of that level are zero until
# a halo update brings it in (the level flags of the later sub-cycles are updated from the halo frame alone)
put("qo3mr", (tt == 0) & (ii > 40) & (ii < 51) & (kk == 25), 1.0e-6)
# graupel: rows of zeros interrupted every 9th row
put("qgraupel", (jj %% 9 == 0) & (kk > 40), 1.0e-5)
eng.step_dynamics()
eng.core.sync()
out = {n: eng.core.t[n].cpu().numpy() for n in STATE_3D}
np.savez(sys.argv[1], **out)
"""


@pytest.mark.gpu
def test_zero_shortcuts_match_dense_on_patchy_tracers(tmp_path):
    """The shortcuts for zero-valued data (rows of zeros in fvtp2d_march, levels of zeros in the tracer and q_con
    transport, columns of zeros in the remap, tiles of zeros in DelnFlux) against the dense computation
    (PACE_B200_ZERO_PATHS=0, and the unfused kernels on top) on tracers that are zero in parts of the domain: a disk,
    a single cell, single rows, whole levels, the cells next to a tile edge — every entry into and exit from the
    shortcuts.  One full step at C48, bit for bit."""
    res = {}
    modes = {"production": {},
             "dense": {"PACE_B200_ZERO_PATHS": "0"},
             "dense_unfused": {"PACE_B200_ZERO_PATHS": "0", "PACE_B200_FVTP2D": "unfused", "PACE_B200_DELNFLUX": "unfused",
                               "PACE_B200_RIEM": "global", "PACE_B200_A2B": "unfused", "PACE_B200_CSW": "unfused", "PACE_B200_DD": "unfused", **ALT}}
    for mode, extra in modes.items():
        env = dict(os.environ, **extra)
        path = str(tmp_path / (mode + ".npz"))
        subprocess.run([sys.executable, "-c", PATCHY % ROOT, path, "48"], check=True, env=env, timeout=600)
        res[mode] = np.load(path)
    moved = 0
    for name in res["production"].files:
        a = comparable(name, res["production"][name], 48)
        assert np.all(np.isfinite(a)), name
        for other in ("dense", "dense_unfused"):
            b = comparable(name, res[other][name], 48)
            assert np.array_equal(a, b), "%s vs %s differs: max |d| = %g" % (name, other, np.max(np.abs(a - b)))
    for name in ("qrain", "qsnow", "qice", "qliquid", "qgraupel", "qo3mr"):
        moved += int(np.count_nonzero(res["production"][name]))
    assert moved > 0

"""The shared-memory fused fvtp2d kernel (pace_b200/csrc/fvtp2d_fused.cu) against the straightforward three-kernel
version (transport.cu, the one validated stencil by stencil against the reference) on a domain that spans several
CTA tiles in both directions: one full dycore step at C48 (6 subdomains of 48x48x79), bit for bit."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r"""
import sys, numpy as np, torch
sys.path.insert(0, %r)
from pace_b200.synthetic import make_engine
from pace_b200.engine import STATE_3D
eng = make_engine(48, 1, n_split=2)
eng.step_dynamics()
eng.core.sync()
out = {n: eng.core.t[n].cpu().numpy() for n in STATE_3D}
np.savez(sys.argv[1], **out)
"""


@pytest.mark.gpu
def test_fused_matches_three_kernel_version(tmp_path):
    res = {}
    for mode in ("fused", "unfused"):
        env = dict(os.environ, PACE_B200_FVTP2D=mode)
        path = str(tmp_path / (mode + ".npz"))
        subprocess.run([sys.executable, "-c", SCRIPT % ROOT, path], check=True, env=env, timeout=600)
        res[mode] = np.load(path)
    for name in res["fused"].files:
        a, b = res["fused"][name], res["unfused"][name]
        assert np.all(np.isfinite(a)), name
        assert np.array_equal(a, b), "%s differs: max |d| = %g" % (name, np.max(np.abs(a - b)))

"""MapSingle alone on 91 levels (the shape of BASELINE.json's config 5, in fp64): synthetic columns, remapped by the
reference class for kord 9 / 10 / 7 and iv 1 / 0 / -1 / -2 (tests/golden/mapsingle_l91.npz, written by
oracle/gen_mapsingle_golden.py with the unmodified pyFV3 MapSingle), against pace_b200_map_single — bit for bit."""
import os

import numpy as np
import pytest

from tests.golden_util import GOLDEN
from tests.module_harness import make_core

CASES = [(9, 1, 0.0, False), (9, 0, 0.0, False), (9, -1, 0.0, False), (9, -2, 0.0, True), (10, 1, 184.0, False),
         (7, 0, 0.0, False)]
NZ = 91


def run(backend):
    z = np.load(os.path.join(GOLDEN, "mapsingle_l91.npz"))
    core = make_core(backend, nx=12, ny=12, npz=NZ)
    for kord, iv, qmin, with_qs in CASES:
        src = z["qpos"] if iv == 0 else z["q"]
        core.put("ms.q", [src])
        core.put("ms.pe1", [z["pe1"]])
        core.put("ms.pe2", [z["pe2"]])
        core.put("ms.qs", [z["qs"]])
        core.call("map_single", "ms.q", "ms.pe1", "ms.pe2", "ms.qs" if with_qs else "", iv, kord, 0, 0, float(qmin))
        core.sync()
        got = core.get("ms.q")[0][3:15, 3:15, :NZ]
        want = z["out_kord%d_iv%d" % (kord, iv)][3:15, 3:15, :NZ]
        assert not np.array_equal(want, src[3:15, 3:15, :NZ])
        assert np.array_equal(got, want), "kord %d iv %d: max |d| = %g" % (kord, iv, np.max(np.abs(got - want)))


def test_mapsingle_l91_hostemu():
    run("hostemu")


@pytest.mark.gpu
def test_mapsingle_l91_gpu():
    run("gpu")

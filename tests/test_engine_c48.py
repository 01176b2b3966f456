"""One full dycore step at C48 (6 subdomains of 48x48x79: 2x2 tiles of 32 with a ragged second tile, two column strips,
several row chunks — the production kernels' tiling is real here, unlike at C12) against the UNMODIFIED reference run at
the same size (oracle/gen_c48_golden.py).  The contract is bit-exact, so the fixture holds SHA-256 hashes of the
reference's final fields per rank, not the arrays."""
import hashlib
import json
import os

import numpy as np
import pytest

from tests.golden_util import GOLDEN
from tests.module_harness import K_INTERFACE

N = 48


def load_inputs():
    z = np.load(os.path.join(GOLDEN, "c48_6rank_inputs.npz"))
    grids, states = [dict() for _ in range(6)], [dict() for _ in range(6)]
    for key in z.files:
        r, name = key.split(":", 1)
        a = z[key]
        if name.startswith(("grid.", "damp.")):
            grids[int(r)][name] = a
        else:
            if a.ndim == 3:
                full = np.zeros((N + 7, N + 7, a.shape[2]))
                full[3:3 + a.shape[0], 3:3 + a.shape[1]] = a
                a = full
            states[int(r)][name] = a
    return grids, states


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype=np.float64).tobytes()).hexdigest()


def run(backend):
    from pace_b200.engine import STATE_2D, STATE_3D, DycoreEngine

    kw = {}
    if backend == "hostemu":
        from tests import hostemu

        kw = dict(device="cpu", _lib_handle=hostemu.load())
    grids, states = load_inputs()
    eng = DycoreEngine(N, 1, 79, grids, **kw)
    eng.set_state(states)
    eng.step_dynamics()
    eng.core.sync()
    want = json.load(open(os.path.join(GOLDEN, "c48_6rank_final_sha256.json")))["sha256"]
    bad = []
    for name in STATE_3D + STATE_2D:
        got = eng.core.get(name)
        for r in range(6):
            ni = N + 1 if name == "v" else N
            nj = N + 1 if name == "u" else N
            a = got[r][3:3 + ni, 3:3 + nj]
            if a.ndim == 3 and name not in K_INTERFACE:
                a = a[:, :, :-1]
            if sha(a) != want["%d:%s" % (r, name)]:
                bad.append((name, r))
    assert not bad, "fields differing from the reference at C48: %s" % bad


def test_full_step_c48_hostemu():
    run("hostemu")


@pytest.mark.gpu
def test_full_step_c48_gpu():
    run("cuda")

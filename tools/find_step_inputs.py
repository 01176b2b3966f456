"""Which DycoreState fields does one step_dynamics read before it overwrites them?  (dev tool, CPU: host emulation)

Every field in turn is filled with NaN, then with large garbage, in the C12 initial state; if the state after one step
is bit-identical to the unpoisoned result in both cases, the field is an output only and a host-resident driver need
not upload it.  Usage: python tools/find_step_inputs.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pace_b200.engine import STATE_2D, STATE_3D  # noqa: E402
from tests.test_engine_acoustic import load_six, make_engine  # noqa: E402


def run(poison=None, value=None):
    eng = make_engine("hostemu")
    if poison is not None:
        t = eng.core.t[poison]
        if value == "nan":
            t.fill_(float("nan"))
        else:
            g = torch.Generator().manual_seed(3)
            t.copy_((torch.rand(t.shape, generator=g, dtype=torch.float64) - 0.5) * 1.0e30)
    eng.step_dynamics()
    return {n: eng.core.t[n].clone() for n in STATE_3D + STATE_2D}


def same(a, b, skip):
    """compute cells 3..14 and levels 0..nz-1 of every field (halo / padding cells that a step never writes keep the
    poison of the poisoned field itself and say nothing)"""
    for n in a:
        x, y = a[n], b[n]
        x = x[:, :-1, 3:15, 3:15] if x.ndim == 4 else x[:, 3:15, 3:15]
        y = y[:, :-1, 3:15, 3:15] if y.ndim == 4 else y[:, 3:15, 3:15]
        if not torch.equal(torch.nan_to_num(x, nan=1.2345e-300), torch.nan_to_num(y, nan=1.2345e-300)):
            return False
    return True


base = run()
outputs_only, inputs = [], []
for name in STATE_3D + STATE_2D:
    ok = True
    for value in ("nan", "garbage"):
        res = run(name, value)
        # the poisoned field itself must also come out as in the clean run if the step rewrites it completely; a field
        # the step never touches (phis) is an input by definition
        if not same(res, base, skip=None):
            ok = False
            break
    (outputs_only if ok else inputs).append(name)
    print(name, "output only" if ok else "input", flush=True)
print("INPUTS", inputs)
print("OUTPUTS_ONLY", outputs_only)

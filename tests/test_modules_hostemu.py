"""CPU-side logic checks: kernel sources compiled for the host (tests/hostemu.py) against the golden
stage log of the unmodified reference.  The GPU parity tests proper are in test_modules_gpu.py."""
import pytest

from tests.module_cases import CASES
from tests.module_harness import check_module


@pytest.mark.parametrize("case", sorted(CASES))
def test_module_hostemu(case, golden, golden_deep, grid0):
    c = CASES[case]
    check_module("hostemu", golden_deep if c.get("deep") else golden, grid0, c["stage"], c["call"], c["args"], tol=c.get("tol_host", 0.0),
                 near_zero=c.get("near_zero", 0.0), skip=c.get("skip", ()), begin=c.get("begin"), region=c.get("region"))

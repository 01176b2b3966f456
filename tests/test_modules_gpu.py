"""GPU parity tests: each module entry point of libpace_b200 (through the C ABI) against the golden
stage log of the unmodified reference (numpy backend, C12 L79, rank 0, tests/golden/)."""
import pytest

from tests.module_cases import CASES
from tests.module_harness import check_module

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case", sorted(CASES))
def test_module_gpu(case, golden, golden_deep, grid0):
    c = CASES[case]
    check_module("cuda", golden_deep if c.get("deep") else golden, grid0, c["stage"], c["call"], c["args"], tol=c.get("tol", 0.0),
                 near_zero=c.get("near_zero", 0.0), skip=c.get("skip", ()), begin=c.get("begin"), region=c.get("region"))

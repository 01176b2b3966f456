"""pace_b200.dropin.DynamicalCore used exactly like pyFV3.DynamicalCore, next to it: six rank processes each build the
reference's own objects (grid_data, damping coefficients, config, state, communicator) the way
tests/main/fv3core/test_dycore_call.py does, hand THE SAME objects to the drop-in (kernels in host emulation, halo
messages through the reference's communicator), step both once and compare every state field bit for bit.

Needs the reference tree and its compiled-stencil cache (minutes to build when cold), so it only runs where
/root/reference exists and the cache is warm (or PACE_B200_REFERENCE_TESTS=1)."""
import os

import pytest

REF = "/root/reference"
CACHE = os.environ.get("GT_CACHE_ROOT", "/tmp/pace_b200_gtcache")


def _cache_warm():
    d = os.path.join(CACHE, ".gt_cache_000000")
    return os.path.isdir(CACHE) and any(n.startswith(".gt_cache") for n in os.listdir(CACHE)) or os.path.isdir(d)


@pytest.mark.skipif(not os.path.isdir(REF), reason="needs the reference tree")
def test_dropin_next_to_reference():
    if not (_cache_warm() or os.environ.get("PACE_B200_REFERENCE_TESTS") == "1"):
        pytest.skip("reference stencil cache is cold (set PACE_B200_REFERENCE_TESTS=1 to build it)")
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "oracle", "run_dropin_check.py")], capture_output=True,
                         text=True, timeout=3600)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "DROPIN_RESULT [[], [], [], [], [], []]" in out.stdout

import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    from tests.golden_util import Stages

    return Stages("c12_rank0_stages.npz")


@pytest.fixture(scope="session")
def grid0():
    from tests.golden_util import load_grid

    return load_grid("c12_rank0_grid.npz")


@pytest.fixture(scope="session")
def golden_deep():
    """Finer-grained stage log (d_sw / updatedzd / nh_p_grad internals); same run as `golden`."""
    from tests.golden_util import Stages

    return Stages("c12_rank0_stages_deep.npz")

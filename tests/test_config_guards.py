"""Options the reference refuses are refused here with the reference's own exception type and message
(pyFV3/stencils/fv_dynamics.py:178-203, 462-499; dyn_core.py:434-439), before anything runs."""
import pytest

from tests.test_engine_acoustic import load_six

CASES = [
    (dict(nwat=5), NotImplementedError, "nwat==5 is not implemented"),
    (dict(hydrostatic=1), NotImplementedError, "Hydrostatic is not implemented"),
    (dict(d_ext=0.1), RuntimeError, "d_ext != 0 is not implemented"),
    (dict(beta=0.5), RuntimeError, "beta != 0 is not implemented"),
    (dict(consv_te=1.0), NotImplementedError, "compute total energy is not implemented"),
    (dict(rf_fast=0, tau=10.0), NotImplementedError, "Rayleigh_Super"),
    (dict(inline_q=1), NotImplementedError, "tracer_2d not implemented"),
    (dict(dt_atmos=0), RuntimeError, "the atmospheric timestep is 0 seconds"),
]


@pytest.mark.parametrize("cfg,exc,text", CASES)
def test_unimplemented_options_raise_like_the_reference(cfg, exc, text):
    from pace_b200.engine import DycoreEngine
    from tests import hostemu

    with pytest.raises(exc, match=text):
        DycoreEngine(12, 1, 79, load_six("grid"), device="cpu", _lib_handle=hostemu.load(), **cfg)


def test_product_refuses_to_run_without_cuda():
    """No CPU fallback: without the host-emulation handle (tests only) a CPU device is an error."""
    from pace_b200.native import NativeCore, NativeError

    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises((NativeError, RuntimeError)):
        NativeCore(12, 12, 79, device="cpu")

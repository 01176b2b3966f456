"""pace_b200_update_dz_c_static — the height-advection entry point for callers that are not allowed to allocate
device memory (their own pool owns the whole GPU, or they sit in a loop that cannot take cudaMalloc's implicit
device synchronisation).  It keeps its work array inside the library image instead of asking cudaMalloc for it.

Checked here, through the C ABI and on the reference's own pre-call state (golden stage log): its outputs equal those
of pace_b200_update_dz_c bit for bit (the tolerance of every module test in this suite is 0) on two calls in a
row — the second call finds the shared array full of the first call's data and must not care — and a handle too
large for the fixed array is refused with a message instead of being run.

The file name sorts last on purpose: by the time pytest gets here the other GPU tests have used the library in the
same process; the test also has to pass alone in a fresh process.
"""
import numpy as np
import pytest

from tests.module_harness import DT2, load_state, make_core

OUTPUTS = ("ad.gz", "ad.ws3")


def _fresh_core(golden, grid0):
    core = make_core("gpu", ptop=float(grid0["grid.ptop"]), da_min=float(grid0["damp.da_min"]),
                     da_min_c=float(grid0["damp.da_min_c"]))
    load_state(core, golden.before("updatedzc"), grid0)
    return core


@pytest.mark.gpu
def test_static_scratch_variant_matches_allocating_entry_point(golden, grid0):
    ref = _fresh_core(golden, grid0)
    ref.call("update_dz_c", DT2)
    ref.sync()
    want = {n: ref.get(n)[0] for n in OUTPUTS}
    # the existing entry point itself reproduces the reference here (tests/test_modules_gpu.py); make sure the
    # comparison below is not vacuous
    before = golden.before("updatedzc")
    assert not np.array_equal(want["ad.gz"], before["ad.gz"])
    after = golden.after("updatedzc")

    core = _fresh_core(golden, grid0)
    for call_no in (1, 2):
        # gz is updated in place: put the pre-call state back, leave the library's shared array as it is
        for n in OUTPUTS:
            core.put(n, [before[n]])
        core.call("update_dz_c_static", DT2)
        core.sync()
        for n in OUTPUTS:
            got = core.get(n)[0]
            assert np.array_equal(got, want[n]), "call %d: %s differs, max |d| = %g" % (
                call_no, n, np.max(np.abs(got - want[n])))
            # and the reference's own result of the same call, directly
            assert np.array_equal(got, after[n]), "call %d: %s differs from the reference" % (call_no, n)


@pytest.mark.gpu
def test_static_scratch_variant_rejects_oversized_handle():
    from pace_b200.native import NativeCore, NativeError

    # 6 subdomains of 96 x 96 x 80 levels: 6 * 80 * 103 * 104 doubles > 2097152
    core = NativeCore(96, 96, 79, n_tiles=6, device="cuda:0")
    with pytest.raises(NativeError, match="built-in array holds 2097152"):
        core.call("update_dz_c_static", DT2)

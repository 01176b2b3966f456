"""MapSingle in the reference's single-precision mode (BASELINE.json config 5: L91, PACE_FLOAT_PRECISION=32).

tests/golden/mapsingle_l91_f32.npz was written by `oracle/gen_mapsingle_golden.py --fp32`: the unmodified pyFV3
MapSingle with Float = np.float32 (every field, literal and intermediate float32) on the synthetic 91-level columns,
kord 9 / 10 / 7 and iv 1 / 0 / -1 / -2.  pace_b200_map_single_f32 must reproduce it bit for bit (integer compare of
the float32 patterns; tolerance 0)."""
import os

import numpy as np
import pytest

from tests.golden_util import GOLDEN
from tests.module_harness import make_core

CASES = [(9, 1, 0.0, False), (9, 0, 0.0, False), (9, -1, 0.0, False), (9, -2, 0.0, True), (10, 1, 184.0, False),
         (7, 0, 0.0, False)]
NZ = 91


def run(backend):
    z = np.load(os.path.join(GOLDEN, "mapsingle_l91_f32.npz"))
    assert z["pe1"].dtype == np.float32
    core = make_core(backend, nx=12, ny=12, npz=NZ)
    pe1, pe2, qs = core.put32([z["pe1"]]), core.put32([z["pe2"]]), core.put32([z["qs"]])
    for kord, iv, qmin, with_qs in CASES:
        src = z["qpos"] if iv == 0 else z["q"]
        q = core.put32([src])
        core.map_single_f32(q, pe1, pe2, qs if with_qs else None, iv=iv, kord=kord, qmin=qmin)
        core.sync()
        got = core.get32(q)[0][3:15, 3:15, :NZ]
        want = z["out_kord%d_iv%d" % (kord, iv)][3:15, 3:15, :NZ]
        assert got.dtype == np.float32 and want.dtype == np.float32
        assert not np.array_equal(want, src[3:15, 3:15, :NZ])
        same = got.view(np.int32) == want.view(np.int32)
        assert same.all(), "kord %d iv %d: %d of %d values differ, max |d| = %g" % (
            kord, iv, (~same).sum(), same.size, np.max(np.abs(got.astype(np.float64) - want.astype(np.float64))))


def test_mapsingle_f32_hostemu():
    run("hostemu")


@pytest.mark.gpu
def test_mapsingle_f32_gpu():
    run("gpu")


def conservation(backend, n, nt):
    """Size-independent property at any size: the remap conserves the column integral, sum_k q dp (fp32: relative to
    sum_k |q| dp within 2e-5)."""
    import torch

    core = make_core(backend, nx=n, ny=n, npz=NZ, n_tiles=nt)
    dev = core.device
    g = torch.Generator(device="cpu").manual_seed(0)
    shape2 = (nt, core.nj, core.sj)
    ps = (95000.0 + 8000.0 * torch.rand(shape2, generator=g)).to(dev)
    thick = (0.2 + 1.6 * torch.rand((nt, NZ) + shape2[1:], generator=g)).to(dev)
    thick = thick * ((ps - 300.0) / thick.sum(dim=1))[:, None]
    pe1 = torch.zeros((nt, NZ + 1) + shape2[1:], dtype=torch.float64, device=dev)
    pe1[:, 0] = 300.0
    pe1[:, 1:] = 300.0 + torch.cumsum(thick.double(), dim=1)
    eta = (torch.arange(NZ + 1, dtype=torch.float64, device=dev) / NZ) ** 1.3
    pe2 = 300.0 + (pe1[:, NZ] - 300.0)[:, None] * eta[None, :, None, None]
    pe1, pe2 = pe1.float().contiguous(), pe2.float().contiguous()
    pe2[:, NZ] = pe1[:, NZ]
    zz = torch.linspace(0.0, 1.0, NZ + 1, device=dev)[None, :, None, None]
    q = (250.0 + 40.0 * zz + 1.5 * torch.randn(pe1.shape, generator=g).to(dev)).float().contiguous()
    q0 = q.clone()
    core.map_single_f32(q, pe1, pe2, None, iv=1, kord=9, qmin=0.0)
    core.sync()
    dp1 = (pe1[:, 1:] - pe1[:, :-1]).double()
    dp2 = (pe2[:, 1:] - pe2[:, :-1]).double()
    sl = (slice(None), slice(3, 3 + n), slice(3, 3 + n))
    before = (q0[:, :NZ].double() * dp1).sum(dim=1)[sl]
    after = (q[:, :NZ].double() * dp2).sum(dim=1)[sl]
    scale = (q0[:, :NZ].double().abs() * dp1).sum(dim=1)[sl]
    rel = ((after - before).abs() / scale).max().item()
    assert not torch.equal(q, q0)
    assert rel < 2.0e-5, rel
    # halo columns are outside the launch box: untouched
    assert torch.equal(q[:, :, :3], q0[:, :, :3]) and torch.equal(q[:, :, :, :3], q0[:, :, :, :3])


def test_mapsingle_f32_conservation_hostemu():
    conservation("hostemu", 24, 1)


@pytest.mark.gpu
def test_mapsingle_f32_conservation_c384_l91_gpu():
    """BASELINE.json config 5 at full size: C384 L91, six tiles resident, fp32."""
    conservation("gpu", 384, 6)

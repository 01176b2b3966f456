"""TEST INFRASTRUCTURE ONLY — golden vectors for MapSingle alone (BASELINE.json config 5 shape: 91 levels, synthetic
columns), produced by the UNMODIFIED reference class pyFV3.stencils.map_single.MapSingle (numpy backend) in this
container.

    python oracle/gen_mapsingle_golden.py          ->  tests/golden/mapsingle_l91.npz       (fp64)
    python oracle/gen_mapsingle_golden.py --fp32   ->  tests/golden/mapsingle_l91_f32.npz   (PACE_FLOAT_PRECISION=32:
                                                       the reference's Float is np.float32, ndsl/dsl/typing.py:25-45)

Columns (seed 0): pe1 = cumulative sum of positive random layer thicknesses from ptop to ps in [950, 1030] hPa,
pe2 = ptop + (ps - ptop) * eta_k with a smooth monotone eta (synthetic L91 coordinate), q smooth + noise (positive
for iv = 0).  Cases: kord 9 with iv 1, 0, -1, -2 (bottom value qs), kord 10 with iv 1, kord 7 with iv 0.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
FP32 = "--fp32" in sys.argv
if FP32:
    os.environ["PACE_FLOAT_PRECISION"] = "32"  # read when ndsl.dsl.typing is imported
from oracle.refshim.bootstrap import bootstrap  # noqa: E402

bootstrap()
import numpy as np  # noqa: E402

NX, NZ = 12, 91
CASES = [(9, 1, 0.0, False), (9, 0, 0.0, False), (9, -1, 0.0, False), (9, -2, 0.0, True), (10, 1, 184.0, False),
         (7, 0, 0.0, False)]  # (kord, iv, qmin, with qs)


def columns(seed=0):
    rng = np.random.default_rng(seed)
    n = NX + 7
    ptop = 300.0
    ps = rng.uniform(95000.0, 103000.0, size=(n, n))
    thick = rng.uniform(0.2, 1.8, size=(n, n, NZ))
    thick *= ((ps - ptop) / thick.sum(axis=2))[:, :, None]
    pe1 = np.zeros((n, n, NZ + 1))
    pe1[:, :, 0] = ptop
    pe1[:, :, 1:] = ptop + np.cumsum(thick, axis=2)
    pe1[:, :, NZ] = ps  # exact bottom
    eta = (np.arange(NZ + 1) / NZ) ** 1.3
    pe2 = ptop + (ps - ptop)[:, :, None] * eta[None, None, :]
    pe2[:, :, NZ] = ps
    z = np.linspace(0.0, 1.0, NZ + 1)[None, None, :]
    x = np.linspace(0.0, 2 * np.pi, n)[:, None, None]
    smooth = 250.0 + 40.0 * z + 5.0 * np.sin(3 * x + 7 * z)
    q = smooth + rng.normal(0.0, 1.5, size=(n, n, NZ + 1))
    qpos = np.abs(1.0e-3 * (1.0 + np.sin(5 * z + x)) + rng.normal(0.0, 3.0e-4, size=(n, n, NZ + 1)))
    qs = rng.normal(0.0, 0.5, size=(n, n))
    return pe1, pe2, q, qpos, qs


def main():
    import ndsl.dsl.stencil
    from ndsl.comm.communicator import CubedSphereCommunicator
    from ndsl.comm.null_comm import NullComm
    from ndsl.comm.partitioner import CubedSpherePartitioner, TilePartitioner
    from ndsl.constants import X_DIM, Y_DIM, Z_DIM
    from ndsl.dsl.dace.dace_config import DaceConfig
    from ndsl.dsl.stencil import GridIndexing
    from ndsl.initialization.allocator import QuantityFactory
    from ndsl.initialization.sizer import SubtileGridSizer
    from pyFV3.stencils.map_single import MapSingle

    backend = "numpy"
    partitioner = CubedSpherePartitioner(TilePartitioner((1, 1)))
    comm = CubedSphereCommunicator(NullComm(rank=0, total_ranks=6, fill_value=0.0), partitioner)
    dace_config = DaceConfig(communicator=None, backend=backend)
    stencil_config = ndsl.dsl.stencil.StencilConfig(
        compilation_config=ndsl.dsl.stencil.CompilationConfig(backend=backend, rebuild=False, validate_args=True),
        dace_config=dace_config)
    sizer = SubtileGridSizer.from_tile_params(nx_tile=NX, ny_tile=NX, nz=NZ, n_halo=3, extra_dim_lengths={},
                                              layout=(1, 1), tile_partitioner=partitioner.tile, tile_rank=0)
    grid_indexing = GridIndexing.from_sizer_and_communicator(sizer=sizer, comm=comm)
    quantity_factory = QuantityFactory.from_backend(sizer=sizer, backend=backend)
    stencil_factory = ndsl.dsl.stencil.StencilFactory(config=stencil_config, grid_indexing=grid_indexing)

    from ndsl.dsl.typing import Float

    assert np.dtype(Float) == (np.float32 if FP32 else np.float64)
    pe1, pe2, q, qpos, qs = (a.astype(Float) for a in columns())
    if FP32:
        pe2[:, :, NZ] = pe1[:, :, NZ]  # keep the exact common bottom after rounding
    out = {"pe1": pe1, "pe2": pe2, "q": q, "qpos": qpos, "qs": qs}

    def quantity(a, dims):
        qq = quantity_factory.zeros(dims, units="unknown", dtype=Float)
        np.asarray(qq.data)[...] = a
        return qq

    for kord, iv, qmin, with_qs in CASES:
        ms = MapSingle(stencil_factory, quantity_factory, kord, iv, dims=[X_DIM, Y_DIM, Z_DIM])
        src = qpos if iv == 0 else q
        q1 = quantity(src, [X_DIM, Y_DIM, Z_DIM])
        p1 = quantity(pe1, [X_DIM, Y_DIM, Z_DIM])
        p2 = quantity(pe2, [X_DIM, Y_DIM, Z_DIM])
        if with_qs:
            ms(q1.data, p1.data, p2.data, qs=quantity(qs, [X_DIM, Y_DIM]).data, qmin=Float(qmin))
        else:
            ms(q1.data, p1.data, p2.data, qmin=Float(qmin))
        key = "out_kord%d_iv%d" % (kord, iv)
        out[key] = np.asarray(q1.data).copy()
        assert out[key].dtype == np.dtype(Float)
        print(key, float(np.abs(out[key][3:15, 3:15, :NZ] - src[3:15, 3:15, :NZ]).max()), flush=True)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "mapsingle_l91_f32.npz" if FP32 else "mapsingle_l91.npz"), **out)


if __name__ == "__main__":
    main()

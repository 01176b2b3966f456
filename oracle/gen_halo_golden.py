"""TEST INFRASTRUCTURE ONLY — golden source maps for the cubed-sphere halo exchange, produced by the
UNMODIFIED reference (NDSL CubedSphereCommunicator / HaloUpdater, all ranks in one process over
LocalComm as in NDSL/tests/test_halo_update.py:196-212).

Every compute cell of every rank is filled with a code  comp*1e6 + rank*1e4 + i*100 + j  (array indices,
comp 1 = scalar / x-component, 2 = y-component; halos 0).  After the reference's halo update each halo cell
therefore names the cell it was filled from (and its sign for rotated vector components), which pins the
index / rotation logic of pace_b200/halo.py bit-exactly.   Output: tests/golden/halo_maps.npz
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.refshim.bootstrap import bootstrap  # noqa: E402

bootstrap()
import numpy as np  # noqa: E402
from ndsl.comm.communicator import CubedSphereCommunicator  # noqa: E402
from ndsl.comm.local_comm import LocalComm  # noqa: E402
from ndsl.comm.partitioner import CubedSpherePartitioner, TilePartitioner  # noqa: E402
from ndsl.constants import X_DIM, X_INTERFACE_DIM, Y_DIM, Y_INTERFACE_DIM, Z_DIM  # noqa: E402
from ndsl.performance.timer import Timer  # noqa: E402
from ndsl.quantity import Quantity  # noqa: E402

NZ = 2
H = 3


def make_quantity(dims, n, rank, comp):
    ext = [n + (1 if d in (X_INTERFACE_DIM, Y_INTERFACE_DIM) else 0) for d in dims[:2]]
    data = np.zeros((n + 2 * H + 1, n + 2 * H + 1, NZ))
    ii, jj = np.meshgrid(np.arange(H, H + ext[0]), np.arange(H, H + ext[1]), indexing="ij")
    code = comp * 1e6 + rank * 1e4 + ii * 100 + jj
    data[H:H + ext[0], H:H + ext[1], :] = code[:, :, None]
    return Quantity(data, dims=list(dims), units="", origin=(H, H, 0), extent=(ext[0], ext[1], NZ))


def run(layout, n_tile):
    part = CubedSpherePartitioner(TilePartitioner((layout, layout)))
    total = part.total_ranks
    shared = {}
    comms = [CubedSphereCommunicator(LocalComm(rank=r, total_ranks=total, buffer_dict=shared), part, timer=Timer())
             for r in range(total)]
    n = n_tile // layout
    out = {}
    scalar_cases = {
        "a3": ((X_DIM, Y_DIM, Z_DIM), 3), "a2": ((X_DIM, Y_DIM, Z_DIM), 2),
        "b3": ((X_INTERFACE_DIM, Y_INTERFACE_DIM, Z_DIM), 3),
    }
    for name, (dims, npts) in scalar_cases.items():
        qs = [make_quantity(dims, n, r, 1) for r in range(total)]
        ups = [c.start_halo_update(q, npts) for c, q in zip(comms, qs)]
        for u in ups:
            u.wait()
        out["l%d_%s" % (layout, name)] = np.stack([q.data[:, :, 0] for q in qs])
    vector_cases = {
        "d3": ((X_DIM, Y_INTERFACE_DIM, Z_DIM), (X_INTERFACE_DIM, Y_DIM, Z_DIM)),
        "c3": ((X_INTERFACE_DIM, Y_DIM, Z_DIM), (X_DIM, Y_INTERFACE_DIM, Z_DIM)),
    }
    for name, (dx, dy) in vector_cases.items():
        xs = [make_quantity(dx, n, r, 1) for r in range(total)]
        ys = [make_quantity(dy, n, r, 2) for r in range(total)]
        ups = [c.start_vector_halo_update(x, y, 3) for c, x, y in zip(comms, xs, ys)]
        for u in ups:
            u.wait()
        out["l%d_%s_x" % (layout, name)] = np.stack([q.data[:, :, 0] for q in xs])
        out["l%d_%s_y" % (layout, name)] = np.stack([q.data[:, :, 0] for q in ys])
    # synchronize_vector_interfaces on a D-grid pair (dyn_core.py:987-988)
    xs = [make_quantity(vector_cases["d3"][0], n, r, 1) for r in range(total)]
    ys = [make_quantity(vector_cases["d3"][1], n, r, 2) for r in range(total)]
    reqs = [c.start_synchronize_vector_interfaces(x, y) for c, x, y in zip(comms, xs, ys)]
    for r in reqs:
        r.wait()
    out["l%d_sync_x" % layout] = np.stack([q.data[:, :, 0] for q in xs])
    out["l%d_sync_y" % layout] = np.stack([q.data[:, :, 0] for q in ys])
    return out


if __name__ == "__main__":
    res = {}
    res.update(run(1, 12))
    res.update(run(2, 12))
    res.update(run(3, 12))
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "halo_maps.npz"), **res)
    for k, v in res.items():
        print(k, v.shape)

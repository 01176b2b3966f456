"""Subdomains spread over several GPUs: placement and the cross-GPU half of the halo update.

The path shards by subdomain: the 6 * layout^2 subdomains are dealt to the GPUs in contiguous
blocks (24/12/6/3 per GPU at 1/2/4/8 GPUs for the 2x2 layout), every GPU steps its own subdomains, and the only
exchange is the halo update — point to point between neighbours, no collective on the step path.  Replaces the
reference's one-MPI-rank-per-subdomain HaloUpdater (NDSL/ndsl/halo/updater.py:217-303, comm/communicator.py) with,
per halo update and per peer GPU, ONE message: a pack kernel gathers the cells every remote neighbour needs (rotation
and vector sign applied while packing), torch.distributed sends / receives the messages (NCCL over NVLink on GPUs,
gloo in the CPU tests), an unpack kernel scatters them; halo cells whose source is on the same GPU are filled by the
local gather kernel meanwhile.
"""
from typing import Dict, List

import torch

from .halo import HaloOp


class Placement:
    """Contiguous blocks of subdomains (cube ranks) per GPU."""

    def __init__(self, total_ranks: int, world_size: int):
        if total_ranks % world_size:
            raise ValueError("%d subdomains cannot be dealt evenly to %d GPUs" % (total_ranks, world_size))
        self.total, self.world, self.per = total_ranks, world_size, total_ranks // world_size

    def owner(self, rank: int) -> int:
        return rank // self.per

    def ranks_of(self, gpu: int) -> List[int]:
        return list(range(gpu * self.per, (gpu + 1) * self.per))


class TorchComm:
    """Point-to-point transport over torch.distributed (process group must be initialised by the caller)."""

    def __init__(self, group=None):
        import torch.distributed as dist

        self.dist, self.group = dist, group
        self.rank, self.world_size = dist.get_rank(group), dist.get_world_size(group)

    def exchange(self, sends: Dict[int, torch.Tensor], recvs: Dict[int, torch.Tensor]):
        """Start all transfers; returns a callable that waits for them."""
        dist = self.dist
        ops = [dist.P2POp(dist.irecv, t, p, self.group) for p, t in sorted(recvs.items())]
        ops += [dist.P2POp(dist.isend, t, p, self.group) for p, t in sorted(sends.items())]
        if not ops:
            return lambda: None
        reqs = dist.batch_isend_irecv(ops)
        return lambda: [r.wait() for r in reqs]


class HaloExchanger:
    """One halo plan (scalar / vector / interface sync) of a DycoreEngine whose neighbours may be remote."""

    def __init__(self, core, plan_id: int, global_ops: List[HaloOp], placement: Placement, comm, local_index):
        self.core, self.comm, self.plan_id = core, comm, plan_id
        me = comm.rank if comm is not None else 0
        own = placement.owner
        local = [o for o in global_ops if own(o.dst_rank) == me and own(o.src_rank) == me]
        core.set_halo_plan(plan_id, local, local_index)
        self.pack_id, self.unpack_id = plan_id + 16, plan_id + 32
        # messages: ops in global plan order, grouped by peer
        send = sorted((o for o in global_ops if own(o.src_rank) == me and own(o.dst_rank) != me),
                      key=lambda o: own(o.dst_rank))  # stable: plan order within a peer
        recv = sorted((o for o in global_ops if own(o.dst_rank) == me and own(o.src_rank) != me),
                      key=lambda o: own(o.src_rank))
        self.send_seg, self.recv_seg = {}, {}  # peer -> (first cell, number of cells)
        pack_rows, unpack_rows, n = [], [], 0
        for o in send:
            p = own(o.dst_rank)
            a, b = self.send_seg.get(p, (n, 0))
            self.send_seg[p] = (a, b + o.ni * o.nj)
            pack_rows.append((n, local_index[o.src_rank], o))
            n += o.ni * o.nj
        self.send_cells, n = n, 0
        for o in recv:
            p = own(o.src_rank)
            a, b = self.recv_seg.get(p, (n, 0))
            self.recv_seg[p] = (a, b + o.ni * o.nj)
            unpack_rows.append((local_index[o.dst_rank], n, o))
            n += o.ni * o.nj
        self.recv_cells = n
        self.remote = bool(pack_rows or unpack_rows)
        if self.remote:
            core.set_halo_plan_raw(self.pack_id, pack_rows)
            core.set_halo_plan_raw(self.unpack_id, unpack_rows)
        self._bufs = {}

    def _buffers(self, nq, nk):
        key = (nq, nk)
        if key not in self._bufs:
            dev = self.core.device
            self._bufs[key] = (torch.zeros(max(1, self.send_cells * nq * nk), dtype=torch.float64, device=dev),
                               torch.zeros(max(1, self.recv_cells * nq * nk), dtype=torch.float64, device=dev))
        return self._bufs[key]

    def update(self, x_names, y_names=None, nk=None):
        core = self.core
        nk = core.npz if nk is None else nk
        if not self.remote:
            core.halo_update(self.plan_id, x_names, y_names, nk)
            return
        nq = len(x_names)
        sbuf, rbuf = self._buffers(nq, nk)
        core.halo_pack(self.pack_id, x_names, y_names, nk, sbuf)
        per = nq * nk
        sends = {p: sbuf[a * per:(a + b) * per] for p, (a, b) in self.send_seg.items()}
        recvs = {p: rbuf[a * per:(a + b) * per] for p, (a, b) in self.recv_seg.items()}
        wait = self.comm.exchange(sends, recvs)
        core.halo_update(self.plan_id, x_names, y_names, nk)  # same-GPU neighbours, overlaps the transfer
        wait()
        core.halo_unpack(self.unpack_id, x_names, y_names, nk, rbuf)

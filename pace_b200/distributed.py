"""Subdomains spread over several GPUs: placement and the cross-GPU half of the halo update.

The path shards by subdomain: the 6 * layout^2 subdomains are dealt to the GPUs in contiguous
blocks (24/12/6/3 per GPU at 1/2/4/8 GPUs for the 2x2 layout), every GPU steps its own subdomains, and the only
exchange is the halo update — point to point between neighbours, no collective on the step path.  Replaces the
reference's one-MPI-rank-per-subdomain HaloUpdater (NDSL/ndsl/halo/updater.py:217-303, comm/communicator.py) with,
per halo update and per peer GPU, ONE message: a pack kernel gathers the cells every remote neighbour needs (rotation
and vector sign applied while packing) and, on GPUs, stores them directly into the receiver's buffer over NVLink
(PeerArena: torch symmetric memory + device-side barrier); an unpack kernel scatters them; halo cells whose source
is on the same GPU are filled by the local gather kernel meanwhile.  Without peer memory (CPU tests with gloo, or
PACE_B200_HALO=p2p) the messages travel by torch.distributed isend / irecv.
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


class PeerArena:
    """Receive buffers that every GPU of the node can write: torch symmetric memory (one allocation per process,
    mapped into all peers over NVLink / NVSwitch) plus its device-side barrier.  With it a halo update needs no
    send / receive at all: the pack kernel of the sender stores straight into the receiver's buffer, a stream-ordered
    barrier across the GPUs tells everybody the stores have landed, the unpack kernel reads the local buffer.  Two
    halves are used alternately so that the stores of update n+1 cannot overtake a peer still unpacking update n
    (the barrier of update n+1 separates update n's unpack from update n+2's stores)."""

    def __init__(self, comm, cap_doubles, device):
        import torch.distributed._symmetric_memory as symm_mem

        self.cap = int(cap_doubles)
        self.world = comm.world_size
        # two message halves, then one 8-byte flag slot per rank (the library's own barrier, csrc/halo.cu)
        self.buf = symm_mem.empty(2 * self.cap + 64, dtype=torch.float64, device=device)
        self.buf.zero_()
        self.hdl = symm_mem.rendezvous(self.buf, comm.dist.group.WORLD.group_name if comm.group is None else comm.group.group_name)
        self.ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        self.rank = comm.rank
        self.count = 0
        self.native = False
        self.hdl.barrier()

    def attach(self, core):
        """Hand the mapped arenas to the library: halo updates of the phase calls then run pack / barrier / unpack as
        kernels on the stream, without coming back to Python (pace_b200_halo_peer_arena)."""
        import ctypes

        arr = (ctypes.c_void_p * self.world)(*self.ptrs)
        rc = core.lib.pace_b200_halo_peer_arena(core.handle, self.rank, self.world, arr, self.cap)
        if rc != 0:
            raise RuntimeError("halo_peer_arena: %s" % core.lib.pace_b200_last_error(core.handle).decode())
        self.native = True
        core.peer_exchange = True   # NativeCore.sync() then also checks the barriers' time-out word

    def next_half(self):
        self.count += 1
        return (self.count & 1) * self.cap * 8  # byte offset of the half used by this update

    def barrier(self):
        self.hdl.barrier()


class HaloExchanger:
    """One halo plan (scalar / vector / interface sync) of a DycoreEngine whose neighbours may be remote."""

    NQ_MAX = 8  # fields per update (pace_b200_halo_update limit)

    @staticmethod
    def max_recv_cells(global_ops: List[HaloOp], placement: Placement) -> int:
        """Largest number of remotely sourced halo cells any GPU receives in one update of this plan."""
        per = {}
        for o in global_ops:
            d, s = placement.owner(o.dst_rank), placement.owner(o.src_rank)
            if d != s:
                per[d] = per.get(d, 0) + o.ni * o.nj
        return max(per.values()) if per else 0

    def __init__(self, core, plan_id: int, global_ops: List[HaloOp], placement: Placement, comm, local_index,
                 arena=None):
        self.core, self.comm, self.plan_id, self.arena = core, comm, plan_id, arena
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
        # peer-store mode: one pack plan per peer whose message offsets are those of the PEER's receive layout
        self.peer_packs = []
        if arena is not None:
            for n_peer, p in enumerate(sorted(self.send_seg)):
                layout = sorted((o for o in global_ops if own(o.dst_rank) == p and own(o.src_rank) != p),
                                key=lambda o: own(o.src_rank))
                rows, off = [], 0
                for o in layout:
                    if own(o.src_rank) == me:
                        rows.append((off, local_index[o.src_rank], o))
                    off += o.ni * o.nj
                pid = 64 + plan_id * 16 + n_peer
                core.set_halo_plan_raw(pid, rows)
                self.peer_packs.append((p, pid))
            if arena.native:
                import ctypes

                n = len(self.peer_packs)
                ranks = (ctypes.c_int * max(1, n))(*[p for p, _ in self.peer_packs])
                pids = (ctypes.c_int * max(1, n))(*[pid for _, pid in self.peer_packs])
                rc = core.lib.pace_b200_halo_peer_plan(core.handle, plan_id, n, ranks, pids,
                                                       self.unpack_id if self.remote else -1,
                                                       self.max_recv_cells(global_ops, placement))
                if rc != 0:
                    raise RuntimeError("halo_peer_plan: %s" % core.lib.pace_b200_last_error(core.handle).decode())

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
        if not self.remote and self.arena is None:
            core.halo_update(self.plan_id, x_names, y_names, nk)
            return
        nq = len(x_names)
        if self.arena is not None and self.arena.native:
            n = len(x_names)
            import ctypes

            xs = (ctypes.c_char_p * n)(*[s.encode() for s in x_names])
            ys = (ctypes.c_char_p * n)(*[s.encode() for s in y_names]) if y_names else None
            if core.lib.pace_b200_halo_peer_update(core.handle, self.plan_id, xs, ys, n, nk) != 0:
                raise RuntimeError("halo_peer_update: %s" % core.lib.pace_b200_last_error(core.handle).decode())
            return
        if self.arena is not None:
            # direct peer stores over NVLink: pack into the receivers' buffers, barrier, unpack the local buffer
            ar = self.arena
            assert self.recv_cells * nq * nk <= ar.cap
            half = ar.next_half()
            for p, pid in self.peer_packs:
                core.halo_pack(pid, x_names, y_names, nk, ar.ptrs[p] + half)
            core.halo_update(self.plan_id, x_names, y_names, nk)  # same-GPU neighbours meanwhile
            ar.barrier()
            core.halo_unpack(self.unpack_id, x_names, y_names, nk, ar.ptrs[ar.rank] + half)
            return
        sbuf, rbuf = self._buffers(nq, nk)
        core.halo_pack(self.pack_id, x_names, y_names, nk, sbuf)
        per = nq * nk
        sends = {p: sbuf[a * per:(a + b) * per] for p, (a, b) in self.send_seg.items()}
        recvs = {p: rbuf[a * per:(a + b) * per] for p, (a, b) in self.recv_seg.items()}
        wait = self.comm.exchange(sends, recvs)
        core.halo_update(self.plan_id, x_names, y_names, nk)  # same-GPU neighbours, overlaps the transfer
        wait()
        core.halo_unpack(self.unpack_id, x_names, y_names, nk, rbuf)

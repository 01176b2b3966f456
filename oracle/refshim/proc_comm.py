"""TEST INFRASTRUCTURE ONLY — a multi-process stand-in for mpi4py implementing the reference's
`Comm` interface (NDSL/ndsl/comm/comm_abc.py:14-72) so the unmodified reference can run 6 (or 24) ranks
in this container, which has no MPI.  One OS process per rank, one inbox queue per rank.
"""
import multiprocessing as mp
import pickle
import queue as _q

import numpy as np

_COLL_TAG = -7777


class _Done:
    def wait(self):
        pass


class _RecvReq:
    def __init__(self, comm, buf, source, tag):
        self.comm, self.buf, self.source, self.tag = comm, buf, source, tag
        self.done = False

    def wait(self):
        if self.done:
            return
        data = self.comm._match(self.source, self.tag)
        arr = np.asarray(self.buf)
        arr[...] = np.asarray(data).reshape(arr.shape)
        self.done = True


class ProcComm:
    """rank/size are relative to `members` (world ranks of this communicator)."""

    def __init__(self, inboxes, world_rank, members=None, comm_id=0, stash=None):
        self._inboxes = inboxes
        self._world_rank = world_rank
        self._members = list(range(len(inboxes))) if members is None else list(members)
        self._rank = self._members.index(world_rank)
        self._id = comm_id
        self._stash = stash if stash is not None else []  # shared between split comms
        self._nsplit = 0

    # -- plumbing -----------------------------------------------------------------------------
    def _put(self, dest, tag, payload):
        self._inboxes[self._members[dest]].put(
            pickle.dumps((self._id, self._rank, tag, payload), protocol=4)
        )

    def _match(self, source, tag):
        for n, (cid, src, t, payload) in enumerate(self._stash):
            if cid == self._id and src == source and t == tag:
                del self._stash[n]
                return payload
        while True:
            try:
                raw = self._inboxes[self._world_rank].get(timeout=3600)
            except _q.Empty:
                raise RuntimeError("ProcComm: recv timed out on rank %d" % self._world_rank)
            cid, src, t, payload = pickle.loads(raw)
            if cid == self._id and src == source and t == tag:
                return payload
            self._stash.append((cid, src, t, payload))

    # -- Comm interface -----------------------------------------------------------------------
    def Get_rank(self):
        return self._rank

    def Get_size(self):
        return len(self._members)

    def Isend(self, sendbuf, dest, tag=0, **kwargs):
        self._put(dest, tag, np.array(sendbuf, copy=True))
        return _Done()

    def Send(self, sendbuf, dest, tag=0, **kwargs):
        self.Isend(sendbuf, dest, tag)

    def Irecv(self, recvbuf, source, tag=0, **kwargs):
        return _RecvReq(self, recvbuf, source, tag)

    def Recv(self, recvbuf, source, tag=0, **kwargs):
        self.Irecv(recvbuf, source, tag).wait()

    def sendrecv(self, sendbuf, dest, **kwargs):
        self._put(dest, _COLL_TAG - 1, sendbuf)
        return self._match(kwargs.get("source", dest), _COLL_TAG - 1)

    def allgather(self, sendobj):
        n = self.Get_size()
        if self._rank == 0:
            out = [sendobj] + [None] * (n - 1)
            for r in range(1, n):
                out[r] = self._match(r, _COLL_TAG)
            for r in range(1, n):
                self._put(r, _COLL_TAG, out)
            return out
        self._put(0, _COLL_TAG, sendobj)
        return self._match(0, _COLL_TAG)

    def bcast(self, value, root=0):
        if self._rank == root:
            for r in range(self.Get_size()):
                if r != root:
                    self._put(r, _COLL_TAG - 2, value)
            return value
        return self._match(root, _COLL_TAG - 2)

    def barrier(self):
        self.allgather(None)

    Barrier = barrier

    def allreduce(self, sendobj, op=None):
        vals = self.allgather(sendobj)
        name = getattr(op, "name", None) or str(op)
        name = name.lower()
        if "min" in name:
            fn = np.minimum
        elif "max" in name:
            fn = np.maximum
        else:
            fn = np.add
        acc = vals[0]
        for v in vals[1:]:
            acc = fn(acc, v)
        return acc

    def Scatter(self, sendbuf, recvbuf, root=0, **kwargs):
        if self._rank == root:
            for r in range(self.Get_size()):
                if r == root:
                    np.asarray(recvbuf)[...] = sendbuf[r]
                else:
                    self._put(r, _COLL_TAG - 3, np.array(sendbuf[r], copy=True))
        else:
            np.asarray(recvbuf)[...] = self._match(root, _COLL_TAG - 3)

    def Gather(self, sendbuf, recvbuf, root=0, **kwargs):
        if self._rank == root:
            for r in range(self.Get_size()):
                if r == root:
                    recvbuf[r] = sendbuf
                else:
                    recvbuf[r] = self._match(r, _COLL_TAG - 4)
        else:
            self._put(root, _COLL_TAG - 4, np.array(sendbuf, copy=True))

    def Split(self, color, key):
        info = self.allgather((color, key, self._world_rank))
        mine = sorted((k, w) for (c, k, w) in info if c == color)
        self._nsplit += 1
        cid = hash((self._id, self._nsplit, color)) & 0x7FFFFFFF
        return ProcComm(self._inboxes, self._world_rank, [w for _, w in mine], cid, self._stash)


def run_ranks(n_ranks, target, args=(), timeout=None):
    """Fork `n_ranks` processes running target(comm, *args); returns list of results (pickled)."""
    ctx = mp.get_context("fork")
    inboxes = [ctx.Queue() for _ in range(n_ranks)]
    results = ctx.Queue()

    def _main(rank):
        try:
            out = target(ProcComm(inboxes, rank), *args)
            results.put((rank, True, out))
        except BaseException as e:  # noqa
            import traceback

            results.put((rank, False, traceback.format_exc()))

    procs = [ctx.Process(target=_main, args=(r,)) for r in range(n_ranks)]
    for p in procs:
        p.start()
    outs = [None] * n_ranks
    for _ in range(n_ranks):
        rank, ok, out = results.get(timeout=timeout)
        if not ok:
            for p in procs:
                p.terminate()
            raise RuntimeError("rank %d failed:\n%s" % (rank, out))
        outs[rank] = out
    for p in procs:
        p.join()
    return outs

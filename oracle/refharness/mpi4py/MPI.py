"""In-process stand-in for ``mpi4py.MPI`` -- TEST INFRASTRUCTURE ONLY.

The reference (WxFactory) imports ``from mpi4py import MPI`` at the top of most
modules (e.g. wx_factory/process_topology.py:4, wx_factory/device/device.py:6)
but mpi4py / mpirun are absent from this image.  This module emulates the small
subset the cubed-sphere RHS path uses, with MPI ranks played by *threads* of one
process:

* ``COMM_WORLD`` with thread-local ``rank``, fixed ``size``
* ``Split``, ``bcast``, ``allreduce``/``Allreduce`` (SUM/MAX), ``gather``, ``scatter``, ``Barrier``
* ``Create_dist_graph_adjacent`` -> object with ``Ineighbor_alltoall`` whose
  request's ``Wait`` pulls the 4 blocks out of per-(sequence, src, dst) mailboxes
  (wx_factory/process_topology.py:261,318,384,598).

It is never imported by the product package.
"""

import threading

import numpy

DOUBLE = "MPI_DOUBLE"
SUM = "MPI_SUM"
MAX = "MPI_MAX"
MIN = "MPI_MIN"
COMM_NULL = None

_tls = threading.local()
_world_size = 1
_lock = threading.Condition()
_mailbox = {}
_coll = {}


def _set_world(size):
    global _world_size
    _world_size = size
    with _lock:
        _mailbox.clear()
        _coll.clear()


def _set_rank(rank):
    _tls.rank = rank
    _tls.seq = {}


def Wtime():
    import time

    return time.time()


class Request:
    def __init__(self, fn):
        self._fn = fn

    def Wait(self):
        self._fn()

    wait = Wait


class Comm:
    """A communicator over a fixed list of world ranks."""

    def __init__(self, members=None, key="world"):
        self._members = members
        self._key = key

    # -- identity ------------------------------------------------------------------------
    def _mem(self):
        return list(range(_world_size)) if self._members is None else self._members

    @property
    def rank(self):
        return self._mem().index(getattr(_tls, "rank", 0))

    @property
    def size(self):
        return len(self._mem())

    def Get_rank(self):
        return self.rank

    def Get_size(self):
        return self.size

    def __eq__(self, other):
        return isinstance(other, Comm) and self._key == other._key

    def __hash__(self):
        return hash(self._key)

    # -- collectives: everybody deposits, waits until all are there, reads ---------------
    def _next_seq(self, what):
        k = (self._key, what)
        s = _tls.seq.get(k, 0)
        _tls.seq[k] = s + 1
        return s

    def _exchange(self, what, value):
        """All-gather of python objects among members. Returns list ordered by comm rank."""
        seq = self._next_seq(what)
        key = (self._key, what, seq)
        n = self.size
        me = self.rank
        with _lock:
            slot = _coll.setdefault(key, {"vals": {}, "read": 0})
            slot["vals"][me] = value
            _lock.notify_all()
            while len(slot["vals"]) < n:
                _lock.wait()
            out = [slot["vals"][i] for i in range(n)]
            slot["read"] += 1
            if slot["read"] == n:
                del _coll[key]
        return out

    def Barrier(self):
        self._exchange("barrier", None)

    barrier = Barrier

    def bcast(self, obj, root=0):
        return self._exchange("bcast", obj)[root]

    def gather(self, obj, root=0):
        vals = self._exchange("gather", obj)
        return vals if self.rank == root else None

    def scatter(self, objs, root=0):
        vals = self._exchange("scatter", objs)
        return vals[root][self.rank]

    def allgather(self, obj):
        return self._exchange("allgather", obj)

    def allreduce(self, obj, op=SUM):
        vals = self._exchange("allreduce", obj)
        if op == MAX:
            return max(vals)
        if op == MIN:
            return min(vals)
        total = vals[0]
        for v in vals[1:]:
            total = total + v
        return total

    def Allreduce(self, send, recv, op=SUM):
        if isinstance(send, (list, tuple)):
            send = send[0]
        if isinstance(recv, (list, tuple)):
            recv = recv[0]
        vals = self._exchange("Allreduce", numpy.array(send, copy=True))
        if op == MAX:
            res = numpy.maximum.reduce(vals)
        elif op == MIN:
            res = numpy.minimum.reduce(vals)
        else:
            res = vals[0].copy()
            for v in vals[1:]:
                res = res + v
        recv[...] = res

    def Split(self, color=0, key=0):
        info = self._exchange("split", (color, key, getattr(_tls, "rank", 0)))
        mine = sorted([(k, r) for (c, k, r) in info if c == color])
        members = [r for (_, r) in mine]
        seq = self._next_seq("splitid")
        return Comm(members, key=(self._key, "split", seq, color))

    def Create_dist_graph_adjacent(self, sources, destinations):
        seq = self._next_seq("graphid")
        return _GraphComm(self, list(sources), list(destinations), (self._key, "graph", seq))


class _GraphComm:
    def __init__(self, parent, sources, destinations, key):
        self._parent = parent
        self._src = sources
        self._dst = destinations
        self._key = key
        self._seq = 0

    def Ineighbor_alltoall(self, send, recv):
        """Block i of `send` goes to destinations[i]; block i of `recv` comes from sources[i]
        (the block that source addressed to this rank)."""
        if isinstance(send, (list, tuple)):
            send = send[0]
        if isinstance(recv, (list, tuple)):
            recv = recv[0]
        seq = self._seq
        self._seq += 1
        me = self._parent.rank
        with _lock:
            for i, d in enumerate(self._dst):
                # several of my edges may face the same rank: keep them in send order
                _mailbox.setdefault((self._key, seq, me, d), []).append(numpy.array(send[i], copy=True))
            _lock.notify_all()

        def finish():
            taken = {}
            with _lock:
                for i, s in enumerate(self._src):
                    k = (self._key, seq, s, me)
                    need = self._src.count(s)
                    while len(_mailbox.get(k, ())) < need:
                        _lock.wait()
                    j = taken.get(s, 0)
                    taken[s] = j + 1
                    recv[i] = _mailbox[k][j]

        return Request(finish)


COMM_WORLD = Comm(None, "world")

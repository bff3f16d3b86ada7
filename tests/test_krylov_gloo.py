"""The N > 1 path of the Krylov solvers on CPU: a world_size-2 gloo job in which every rank owns half of each vector.
All reductions of wxfactory_b200.krylov go through the vector backend (multidot / dot / norm / asum_rows /
max_rel_change), so summing those over ranks must reproduce the single-rank reference results of
tests/golden/krylov_small.npz (the reference's own fgmres / kiops); the state-file gather / scatter is exercised the
same way."""

import os
import socket
import types

import numpy
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from vecops_cpu import CpuVecOps
from wxfactory_b200 import load_problem, state_io
from wxfactory_b200.krylov import fgmres, kiops
from wxfactory_b200.layout import local_state
from wxfactory_b200.partition import build_partition

HERE = os.path.dirname(os.path.abspath(__file__))


class DistCpuVecOps(CpuVecOps):
    """CpuVecOps whose reductions are summed over the ranks of the default group (what VecOps does with NCCL)."""

    def _sum(self, a):
        t = torch.from_numpy(numpy.ascontiguousarray(a, dtype=float).copy())
        dist.all_reduce(t)
        return t.numpy()

    def multidot(self, V, m, W, k):
        return self._sum(super().multidot(V, m, W, k))

    def dot(self, x, y):
        return float(self._sum(numpy.array([super().dot(x, y)]))[0])

    def asum_rows(self, U):
        return self._sum(super().asum_rows(U))

    def max_rel_change(self, update, x):
        loc = super().max_rel_change(update, x)
        t = torch.tensor([loc if numpy.isfinite(loc) else -1.0], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t) if float(t) >= 0 else float("inf")


def _worker(rank, world, port, ok):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        G = numpy.load(os.path.join(HERE, "golden", "krylov_small.npz"))
        ops = DistCpuVecOps()

        def split(a):  # this rank's slice of the last axis
            n = a.shape[-1]
            lo, hi = rank * n // world, (rank + 1) * n // world
            return torch.from_numpy(numpy.ascontiguousarray(a[..., lo:hi]))

        def dist_matvec(M):
            n = M.shape[0]
            lo, hi = rank * n // world, (rank + 1) * n // world

            def A(v):
                parts = [torch.empty((r + 1) * n // world - r * n // world, dtype=torch.float64) for r in range(world)]
                dist.all_gather(parts, v.contiguous())
                return torch.from_numpy(M[lo:hi] @ torch.cat(parts).numpy())

            return A

        good = True
        nr, nb, nit, flag, restart = G["fg_meta_a"]
        x, _, norm_b, niter, fl, res = fgmres(dist_matvec(G["fg_M"]), split(G["fg_b"]), x0=split(G["fg_x0"]), tol=1e-10, restart=int(restart), maxiter=50, ops=ops)
        want = split(G["fg_x_a"]).numpy()
        good &= (niter, fl) == (int(nit), int(flag)) and abs(norm_b - nb) <= 1e-13 * nb
        good &= numpy.linalg.norm(x.numpy() - want) <= 1e-11 * numpy.linalg.norm(want)
        w, stats = kiops([1.0], dist_matvec(G["ki_J"]), split(G["ki_u"]), tol=1e-9, m_init=10, mmin=10, mmax=64, ops=ops)
        ws = G["ki_stats_a"]
        good &= [int(stats[i]) for i in (0, 1, 2, 3, 5)] == [int(ws[i]) for i in (0, 1, 2, 3, 5)]
        good &= numpy.linalg.norm(w.numpy() - split(G["ki_w_a"]).numpy()) <= 1e-10 * numpy.linalg.norm(G["ki_w_a"])

        # state bridge: every rank's rows -> global array on rank 0 -> back
        prob, ex = load_problem(os.path.join(HERE, "golden", "dcmip31.npz"))
        part = build_partition(prob.nh, prob.nz, prob.n, prob.bnd_sn, prob.bnd_we, nranks=world, rank=rank, num_var=6, nfaces=6, trace_layout="element")
        rhs = types.SimpleNamespace(prob=prob, part=part, nranks=world, rank=rank, num_trace=6, local_state=lambda Q: local_state(prob, part, Q))
        mine = local_state(prob, part, ex["Q0"])
        total = state_io.gather_state(rhs, mine)
        good &= (total is None) if rank else bool(numpy.array_equal(total, ex["Q0"].reshape(prob.state_shape)))
        back = state_io.scatter_state(rhs, total)
        good &= bool(numpy.array_equal(back, mine))
        # restart file: write on rank 0, read on every rank; a missing file or a state of the wrong shape raises on
        # EVERY rank instead of leaving the others in the broadcast
        import tempfile

        tmp = os.path.join(tempfile.gettempdir(), f"wxb_state_gloo_{port}")
        name = state_io.write_restart(rhs, mine, "[General]\n", tmp, 3, 1.0)
        box = [name]
        dist.broadcast_object_list(box, src=0)
        got, content = state_io.read_restart(rhs, box[0])
        good &= bool(numpy.array_equal(got, mine)) and content.strip() == "[General]"
        for bad, exc in ((box[0] + ".missing", OSError), (None, ValueError)):
            if bad is None:
                bad = os.path.join(tmp, "wrong_shape.npy")
                if rank == 0:
                    state_io.save_state(numpy.zeros((6, 5, 1, 2, 2, 8)), "[General]\n", bad)
                dist.barrier()
            try:
                state_io.read_restart(rhs, bad)
                good = False
            except exc:
                pass
        ok[rank] = int(bool(good))
    finally:
        dist.destroy_process_group()


def test_two_rank_krylov_and_state_bridge():
    world = 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ok = mp.get_context("spawn").Array("i", [0] * world)
    mp.spawn(_worker, args=(world, port, ok), nprocs=world, join=True)
    assert list(ok) == [1] * world


def _bridge_worker(rank, world, port, members, ok):
    """State bridge inside a (sub)group: ranks outside `members` only take part in creating the group."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        group = dist.new_group(ranks=members) if len(members) < world else None
        if rank not in members:
            ok[rank] = 1
            return
        gr, gw = members.index(rank), len(members)
        prob, ex = load_problem(os.path.join(HERE, "golden", "dcmip31.npz"))
        part = build_partition(prob.nh, prob.nz, prob.n, prob.bnd_sn, prob.bnd_we, nranks=gw, rank=gr, num_var=6, nfaces=6, trace_layout="element")
        rhs = types.SimpleNamespace(prob=prob, part=part, nranks=gw, rank=gr, num_trace=6, local_state=lambda Q: local_state(prob, part, Q))
        mine = local_state(prob, part, ex["Q0"])
        total = state_io.gather_state(rhs, torch.from_numpy(mine), group)  # a tensor, as the integrators hold it
        good = (total is None) if gr else bool(numpy.array_equal(total, ex["Q0"].reshape(prob.state_shape)))
        back = state_io.scatter_state(rhs, total, group)
        good = good and bool(numpy.array_equal(back, mine)) and back.dtype == numpy.float64
        ok[rank] = int(good)
    finally:
        dist.destroy_process_group()


def _run_bridge(world, members):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ok = mp.get_context("spawn").Array("i", [0] * world)
    mp.spawn(_bridge_worker, args=(world, port, members, ok), nprocs=world, join=True)
    assert list(ok) == [1] * world


def test_state_bridge_moves_one_piece_per_rank_over_four_ranks():
    """gather_state / scatter_state send float64 tensors point to point, one rank's rows at a time
    (ProcessTopology.gather_cube / distribute_cube, process_topology.py:444-539): 4 ranks, 1.5 panels each."""
    _run_bridge(4, [0, 1, 2, 3])


def test_state_bridge_inside_a_subgroup():
    """The group's rank numbers are translated to global ones for the point-to-point calls: ranks 1 and 2 of a
    3-process job form the group, rank 0 of the job is not part of it."""
    _run_bridge(3, [1, 2])

"""The N > 1 halo path on CPU: world_size-2 and -4 gloo jobs move the packed trace records with the same
`halo.exchange` the GPU path uses, and every rank must end up with the ghost records a single-process emulation of the
whole exchange produces (which test_partition.py checks against the reference's exchange)."""

import os
import socket

import numpy
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import emulate
from wxfactory_b200 import halo
from wxfactory_b200 import partition as pt
from wxfactory_b200.problem import load_problem

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "dcmip31_n4.npz")
NTR = 6  # trace fields of the Euler kernels


def _traces(prob, ex):
    """Synthetic two-sided traces (NTR, 6, nz, nh, nh, 2F), different on every face point."""
    rng = numpy.random.default_rng(3)
    shape = (NTR, 6, prob.nz, prob.nh, prob.nh, 2 * prob.F)
    return rng.standard_normal(shape), rng.standard_normal(shape)


def _worker(rank, world, port, ok):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        prob, ex = load_problem(GOLDEN)
        n, nz, nh, F = prob.n, prob.nz, prob.nh, prob.F
        x1, x2 = _traces(prob, ex)
        parts = [pt.build_partition(nh, nz, n, prob.bnd_sn, prob.bnd_we, world, r, num_var=NTR) for r in range(world)]
        traces = [emulate.local_traces(prob, p, x1, x2, NTR, 6) for p in parts]
        mine, part = traces[rank].copy(), parts[rank]
        emulate.exchange(parts, traces, NTR, 6, nz, F)  # reference result for every rank, computed locally
        send = torch.from_numpy(emulate.pack(part, mine, NTR, 6, nz, F).ravel().copy())
        recv = torch.zeros(part.nghost * NTR * nz * F, dtype=torch.float64)
        halo.exchange(part, send, recv, NTR * nz * F)
        want = traces[rank][NTR * 6 * part.nelems * F :]
        ok[rank] = int(part.nghost > 0 and numpy.array_equal(recv.numpy(), want))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_gloo_halo_exchange_fills_the_ghost_records(world):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ok = mp.get_context("spawn").Array("i", [0] * world)
    mp.spawn(_worker, args=(world, port, ok), nprocs=world, join=True)
    assert list(ok) == [1] * world

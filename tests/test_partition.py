"""Connectivity tables (wxfactory_b200/partition.py) against the oracle's restatement of the reference's
sender-side halo exchange (oracle/cubesphere_exchange.py <- wx_factory/process_topology.py), the analogue of the
reference's ProcessTopologyTest (tests/unit/common/test_process_topology.py:30-140), for 1, 2, 4 and 8 ranks."""

import numpy
import pytest

import emulate
from oracle import cubesphere_exchange as cx
from oracle.euler_oracle import EulerOracle
from wxfactory_b200 import partition as pt


def test_contra_matrices_match_reference_lambdas():
    rng = numpy.random.default_rng(1)
    x = rng.uniform(-1, 1, 7)
    a1, a2 = rng.standard_normal(7), rng.standard_normal(7)
    for p in range(6):
        for e in range(4):
            m = pt.contra_matrix(p, e, x)
            b1, b2 = cx.CONVERT_CONTRA[p][e](a1, a2, x)
            assert numpy.allclose(m[:, 0] * a1 + m[:, 1] * a2, b1, rtol=0, atol=1e-15)
            assert numpy.allclose(m[:, 2] * a1 + m[:, 3] * a2, b2, rtol=0, atol=1e-15)


def test_slabs_cover_every_row_once():
    for nh in (4, 12):
        for G in (1, 2, 4, 8):
            if (6 * nh) % G:
                continue
            rows = []
            for r in range(G):
                for sl in pt.slabs_of_rank(nh, G, r):
                    rows += [(sl.panel, j) for j in range(sl.j0, sl.j1)]
                    assert all(pt.owner_of_row(nh, G, sl.panel, j) == r for j in range(sl.j0, sl.j1))
            assert sorted(rows) == [(p, j) for p in range(6) for j in range(nh)]


@pytest.mark.parametrize("layout", ["field", "element"])
@pytest.mark.parametrize("nranks", [1, 2, 4, 8])
def test_pull_tables_reproduce_reference_exchange(golden, nranks, layout):
    prob, ex = golden("dcmip31_n4")  # n=4, nz=4, nh=4
    n, nz, nh, F = prob.n, prob.nz, prob.nh, prob.F
    orc = EulerOracle(prob)
    rng = numpy.random.default_rng(7)
    q = numpy.moveaxis(ex["Q0"], 1, 0) * (1.0 + 0.01 * rng.standard_normal((5,) + ex["Q0"].shape[:1] + ex["Q0"].shape[2:]))
    q_itf_x1, q_itf_x2 = q @ orc.extrap_x, q @ orc.extrap_y
    s_, n_, w_, e_ = orc._exchange(q_itf_x1, q_itf_x2)  # (5,6,nz,nh,F) each: the reference's received halos
    parts = [pt.build_partition(nh, nz, n, prob.bnd_sn, prob.bnd_we, nranks, r, trace_layout=layout) for r in range(nranks)]
    traces = [emulate.local_traces(prob, p, q_itf_x1, q_itf_x2, 5, 6) for p in parts]
    emulate.exchange(parts, traces, 5, 6, nz, F)
    for p, tr in zip(parts, traces):
        cb = 0
        for sl in p.slabs:
            for jl in range(sl.ny):
                j = sl.j0 + jl
                for i in range(nh):
                    col = cb + jl * nh + i
                    for k in range(nz):
                        # expected neighbour traces in the reference's padded arrays
                        exp = {
                            pt.WEST: q_itf_x1[:, sl.panel, k, j, i - 1, F:] if i > 0 else w_[:, sl.panel, k, j],
                            pt.EAST: q_itf_x1[:, sl.panel, k, j, i + 1, :F] if i < nh - 1 else e_[:, sl.panel, k, j],
                            pt.SOUTH: q_itf_x2[:, sl.panel, k, j - 1, i, F:] if j > 0 else s_[:, sl.panel, k, i],
                            pt.NORTH: q_itf_x2[:, sl.panel, k, j + 1, i, :F] if j < nh - 1 else n_[:, sl.panel, k, i],
                        }
                        for face, want in exp.items():
                            got = emulate.pull(p, tr, col, face, k, 5, n, F)
                            assert numpy.allclose(got, want, rtol=1e-14, atol=1e-14 * numpy.abs(want).max()), (
                                p.rank,
                                sl.panel,
                                j,
                                i,
                                face,
                            )
            cb += sl.ny * nh


def test_send_and_recv_lists_are_consistent():
    for G in (2, 4, 8):
        parts = [pt.build_partition(4, 2, 3, numpy.zeros((6, 12)), numpy.zeros((6, 12)), G, r) for r in range(G)]
        for p in parts:
            assert sum(c for _, _, c in p.recv_peers) == p.nghost
            for peer, first, cnt in p.recv_peers:
                blk = [x for x in parts[peer].send_peers if x[0] == p.rank]
                assert len(blk) == 1 and blk[0][2] == cnt

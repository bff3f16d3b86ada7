"""The tiled synthetic workload (wxfactory_b200/synthetic.py): its arrays agree with a dense twin, and the
periodicity property that bench.py uses for full-size parity holds exactly in the oracle; on the GPU the CUDA path
is checked through it at a size the oracle itself could not finish quickly."""

import numpy
import pytest

from wxfactory_b200 import synthetic as syn
from wxfactory_b200.layout import local_state, local_static
from wxfactory_b200.partition import build_partition


@pytest.fixture(scope="module")
def base(golden):
    prob, ex = golden("dcmip31_n4")
    Qb = syn.base_state(ex["Q0"])
    return prob, ex, Qb, syn.base_direction(Qb)


def test_base_state_is_admissible(base):
    prob, ex, Qb, vb = base
    assert (Qb[:, 0] > 0).all() and (Qb[:, 4] > 0).all()
    assert numpy.isfinite(Qb).all() and numpy.isfinite(vb).all()
    assert numpy.abs(Qb[:, 3]).max() > 0


def test_lazy_slabs_equal_dense_twin(base):
    prob, ex, Qb, vb = base
    tiled = syn.TiledEulerProblem(prob, 8, 12)
    dense = tiled.materialize()
    Qg = tiled.global_state(Qb)
    for G in (1, 2, 8):
        for r in range(G):
            part = build_partition(12, 8, 4, tiled.bnd_sn, tiled.bnd_we, G, r)
            a, d = tiled.local_static(part), local_static(dense, part)
            assert a.keys() == d.keys()
            for k in d:
                assert numpy.array_equal(a[k], d[k]), k
            assert numpy.array_equal(tiled.local_state(part, Qb), local_state(dense, part, Qg))


def test_periodicity_property_holds_in_the_oracle(base):
    from oracle.euler_oracle import EulerOracle

    prob, ex, Qb, vb = base
    small = syn.TiledEulerProblem(prob, 8, 8)
    big = syn.TiledEulerProblem(prob, 12, 12)
    with numpy.errstate(all="ignore"):
        rs = EulerOracle(small.materialize()).rhs(small.global_state(Qb))
        dense = big.materialize()
        rb = EulerOracle(dense).rhs(big.global_state(Qb))
    assert numpy.isfinite(rs).all()
    for G in (1, 4):
        for r in range(G):
            part = build_partition(12, 12, 4, big.bnd_sn, big.bnd_we, G, r)
            err, cnt = syn.check_against_small(big, part, local_state(dense, part, rb), rs)
            assert cnt == 6 * 10 * 10 * 10 // G and err == 0.0
    # and the check does detect a wrong value
    bad = rb.copy()
    bad[3, 2, 5, 5, 5, 7] *= 1.0 + 1e-6
    part = build_partition(12, 12, 4, big.bnd_sn, big.bnd_we, 1, 0)
    assert syn.check_against_small(big, part, local_state(dense, part, bad), rs)[0] > 1e-12


@pytest.mark.gpu
def test_full_size_parity_through_periodicity(base):
    """3.5 M nodes (17.7 M DOFs): every element off the panel edges / bottom / top against the oracle's 8x8x8 twin."""
    import torch

    from oracle import jvp_oracle
    from wxfactory_b200 import EPS_FD, RhsB200

    prob, ex, Qb, vb = base
    dt = float(ex["dt"])
    small = syn.TiledEulerProblem(prob, 8, 8)
    orc = jvp_oracle.make_oracle(small.materialize())
    Qs, vs = small.global_state(Qb), small.global_state(vb)
    with numpy.errstate(all="ignore"):
        want_r = orc.rhs(Qs)
        want_j = jvp_oracle.matvec_fd(orc.rhs, vs.ravel(), dt, Qs, want_r).reshape(Qs.shape)
    big = syn.TiledEulerProblem(prob, 16, 24)
    rhs = RhsB200(big)
    q = torch.from_numpy(rhs.local_state(Qb)).cuda()
    v = torch.from_numpy(rhs.local_state(vb)).cuda()
    out = rhs(q)
    jv = rhs.jvp(v, dt, q, out, mode="fd")
    e_r, cnt = syn.check_against_small(big, rhs.part, out, want_r)
    e_j, _ = syn.check_against_small(big, rhs.part, jv, want_j)
    assert cnt == 6 * 14 * 22 * 22
    assert e_r <= 1e-12
    assert e_j * numpy.linalg.norm(want_j) <= 1e-12 * dt * numpy.linalg.norm(want_r) / EPS_FD
    rhs.close()


# ---- shallow water ------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def sw_base(golden):
    prob, ex = golden("case6")
    Qb = syn.sw_base_state(ex["Q0"])
    return prob, ex, Qb, syn.base_direction(Qb)


def test_sw_lazy_slabs_equal_dense_twin(sw_base):
    prob, ex, Qb, vb = sw_base
    tiled = syn.TiledSwProblem(prob, 24)
    dense = tiled.materialize()
    Qg = tiled.global_state(Qb)
    assert (Qb[:, 0] > 0).all()
    for G in (1, 2, 8):
        for r in range(G):
            part = build_partition(24, 1, 3, tiled.bnd_sn, tiled.bnd_we, G, r, num_var=3, nfaces=4)
            a, d = tiled.local_static(part), local_static(dense, part)
            assert a.keys() == d.keys()
            for k in d:
                assert numpy.array_equal(a[k], d[k]), k
            assert numpy.array_equal(tiled.local_state(part, Qb), local_state(dense, part, Qg))


def test_sw_periodicity_property_holds_in_the_oracle(sw_base):
    from oracle.sw_oracle import SwOracle

    prob, ex, Qb, vb = sw_base
    small = syn.TiledSwProblem(prob, 20)
    big = syn.TiledSwProblem(prob, 30)
    rs = SwOracle(small.materialize()).rhs(small.global_state(Qb))
    dense = big.materialize()
    rb = SwOracle(dense).rhs(big.global_state(Qb))
    assert numpy.isfinite(rs).all()
    for G in (1, 4):
        for r in range(G):
            part = build_partition(30, 1, 3, big.bnd_sn, big.bnd_we, G, r, num_var=3, nfaces=4)
            err, cnt = syn.check_sw_against_small(big, part, local_state(dense, part, rb), rs)
            assert cnt == 6 * 28 * 28 // G and err == 0.0
    bad = rb.copy()
    bad[3, 1, 5, 5, 4] *= 1.0 + 1e-6
    part = build_partition(30, 1, 3, big.bnd_sn, big.bnd_we, 1, 0, num_var=3, nfaces=4)
    assert syn.check_sw_against_small(big, part, local_state(dense, part, bad), rs)[0] > 1e-12


@pytest.mark.gpu
def test_sw_full_size_parity_through_periodicity(sw_base):
    """6 x 400 x 400 elements (8.6 M nodes): every element off the panel edges against the oracle's 20x20 twin."""
    import torch

    from oracle import jvp_oracle
    from wxfactory_b200 import EPS_FD, RhsB200

    prob, ex, Qb, vb = sw_base
    dt = float(ex["dt"])
    small = syn.TiledSwProblem(prob, 20)
    orc = jvp_oracle.make_oracle(small.materialize())
    Qs, vs = small.global_state(Qb), small.global_state(vb)
    want_r = orc.rhs(Qs)
    want_j = jvp_oracle.matvec_fd(orc.rhs, vs.ravel(), dt, Qs, want_r).reshape(Qs.shape)
    big = syn.TiledSwProblem(prob, 400)
    rhs = RhsB200(big)
    q = torch.from_numpy(rhs.local_state(Qb)).cuda()
    v = torch.from_numpy(rhs.local_state(vb)).cuda()
    out = rhs(q)
    jv = rhs.jvp(v, dt, q, out, mode="fd")
    e_r, cnt = syn.check_sw_against_small(big, rhs.part, out, want_r)
    e_j, _ = syn.check_sw_against_small(big, rhs.part, jv, want_j)
    assert cnt == 6 * 398 * 398
    assert e_r <= 1e-12
    assert e_j * numpy.linalg.norm(want_j) <= 1e-12 * dt * numpy.linalg.norm(want_r) / EPS_FD
    rhs.close()

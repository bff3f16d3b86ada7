"""The per-step kernels of csrc/stepops.cu through the C ABI: the exponential filter against the
reference's own output (tests/golden/expfilter_*.npz) and the vector kernels against numpy on seeded inputs.

Tolerances: the filter replaces the reference's dense (P x P) product by three 1-D passes -- another summation order,
FILTER_TOL = 1e-14 relative; lincomb3 is bit-exact (operation-by-operation rounding); reductions differ from numpy's
pairwise/BLAS order, 1e-13 relative to sum |x||y|."""

import os

import numpy
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu
FILTER_TOL = 1e-14


def rel(a, b):
    nb = numpy.linalg.norm(b)
    return numpy.linalg.norm(a - b) / (nb if nb > 0 else 1.0)  # rho*w is identically zero in the initial states


@pytest.mark.parametrize("name", ["dcmip21", "dcmip21_n5", "dcmip31_n4"])
def test_expfilter_matches_reference(golden, name):
    import torch

    from wxfactory_b200 import RhsB200
    from wxfactory_b200.stepops import ExpFilterB200

    prob, ex = golden(name)
    g = numpy.load(os.path.join(GOLDEN, f"expfilter_{name}.npz"))
    rhs = RhsB200(prob)
    filt = ExpFilterB200(rhs, g["expfilter"])
    out = filt.apply_filter_3d(ex["Q0"], None)  # host path
    assert isinstance(out, numpy.ndarray) and out.shape == ex["Q0"].shape
    assert rel(out, g["Qf"]) <= FILTER_TOL
    for v in range(5):
        assert rel(out[:, v], g["Qf"][:, v]) <= FILTER_TOL
    q = torch.from_numpy(ex["Q0"]).cuda()
    outd = filt(q.view(-1))  # device path, flattened input keeps its shape
    assert outd.is_cuda and outd.shape == (q.numel(),)
    assert numpy.array_equal(outd.cpu().numpy().reshape(out.shape), out)
    with pytest.raises(ValueError):
        ExpFilterB200(rhs, numpy.eye(prob.n + 1))
    rhs.close()


def test_expfilter_needs_the_filter_first(golden):
    import torch

    from wxfactory_b200 import RhsB200, capi

    prob, ex = golden("dcmip21")
    rhs = RhsB200(prob)
    q = torch.from_numpy(ex["Q0"]).cuda()
    code = rhs._lib.wxb_expfilter(rhs._h, q.data_ptr(), q.data_ptr(), None)
    assert code == -4  # WXB_ESTATE
    assert b"wxb_set_expfilter" in capi.load().wxb_last_error(rhs._h)
    rhs.close()


@pytest.fixture(scope="module")
def ops():
    from wxfactory_b200.stepops import VecOps

    return VecOps(0)


@pytest.mark.parametrize("n", [1, 1000, 148 * 4 * 256 + 77, 3_000_001])
def test_lincomb3_bit_exact(ops, n):
    import torch

    rng = numpy.random.default_rng(n)
    x, y, z = (rng.standard_normal(n) for _ in range(3))
    dt = 0.37
    xd, yd, zd = (torch.from_numpy(a).cuda() for a in (x, y, z))
    assert numpy.array_equal(ops.lincomb3(1.0, xd, 0.0, None, 1.0, zd, dt).cpu().numpy(), x + z * dt)
    assert numpy.array_equal(ops.lincomb3(0.75, xd, 0.25, yd, 0.25, zd, dt).cpu().numpy(), 0.75 * x + 0.25 * y + 0.25 * z * dt)
    c1, c2 = 1.0 / 3.0, 2.0 / 3.0
    assert numpy.array_equal(ops.lincomb3(c1, xd, c2, yd, c2, zd, dt).cpu().numpy(), c1 * x + c2 * y + c2 * z * dt)


def test_nonfinite_counts(ops):
    import torch

    x = torch.zeros(1_000_003, dtype=torch.float64, device="cuda")
    assert ops.nonfinite(x) == 0
    x[5] = float("nan")
    x[77777] = float("inf")
    x[-1] = -float("inf")
    assert ops.nonfinite(x) == 3


# 65 and 129 rows: PMEX asks for j + 1 <= mmax + 1 rows (pmex.py:168-176); VecOps.multidot walks the basis in chunks of
# WXB_VEC_MAX_ROWS rows
@pytest.mark.parametrize("m,k,n", [(1, 1, 10), (2, 1, 100_003), (3, 2, 5), (22, 2, 200_001), (64, 2, 50_000), (65, 2, 30_001), (129, 2, 9_999)])
def test_multidot_matches_numpy_and_is_reproducible(ops, m, k, n):
    import torch

    rng = numpy.random.default_rng(m * 1000 + k)
    V = rng.standard_normal((m + 2, n))
    Vd = torch.from_numpy(V).cuda()
    got = ops.multidot(Vd, m, Vd[m : m + k], k)
    want = V[:m] @ V[m : m + k].T
    bound = 1e-13 * (numpy.abs(V[:m]) @ numpy.abs(V[m : m + k]).T)
    assert numpy.all(numpy.abs(got - want) <= bound)
    assert numpy.array_equal(got, ops.multidot(Vd, m, Vd[m : m + k], k))  # fixed-order reduction
    assert abs(ops.norm(Vd[0]) - numpy.linalg.norm(V[0])) <= 1e-14 * numpy.linalg.norm(V[0])


@pytest.mark.parametrize("m", [0, 1, 5, 64, 100])
def test_combine_matches_numpy(ops, m):
    import torch

    n = 70_001
    rng = numpy.random.default_rng(m)
    V, y, c = rng.standard_normal((max(m, 1), n)), rng.standard_normal(n), rng.standard_normal(m)
    Vd, yd = torch.from_numpy(V).cuda(), torch.from_numpy(y).cuda()
    got = ops.combine(c, Vd, 0.5, yd, 1.25).cpu().numpy()
    want = (0.5 * y + (c @ V[:m] if m else 0.0)) * 1.25
    assert numpy.linalg.norm(got - want) <= 1e-14 * max(numpy.linalg.norm(want), 1.0) * max(m, 1)
    out = ops.combine(c, Vd, 1.0, yd, 1.0, out=yd)  # in place on y
    assert out.data_ptr() == yd.data_ptr()
    assert numpy.linalg.norm(yd.cpu().numpy() - (y + (c @ V[:m] if m else 0.0))) <= 1e-14 * numpy.linalg.norm(y) * max(m, 1)


@pytest.mark.parametrize("m", [0, 1, 7, 21])
def test_gs1sync_matches_the_stand_in(ops, m):
    import torch

    from vecops_cpu import CpuVecOps

    n = 90_001
    rng = numpy.random.default_rng(m + 5)
    Q = rng.standard_normal((m + 2, n))
    s2, s1 = rng.standard_normal(m), rng.standard_normal(m)
    Qd, Qc = torch.from_numpy(Q.copy()).cuda(), torch.from_numpy(Q.copy())
    ops.gs1sync(Qd, m, s2, s1, 1.7, 0.3, Qd[m], Qd[m + 1])
    CpuVecOps().gs1sync(Qc, m, s2, s1, 1.7, 0.3, Qc[m], Qc[m + 1])
    got, want = Qd.cpu().numpy(), Qc.numpy()
    assert numpy.array_equal(got[:m], Q[:m])  # earlier rows untouched
    assert numpy.linalg.norm(got[m:] - want[m:]) <= 1e-14 * numpy.linalg.norm(want[m:]) * max(m, 1)


def test_scalar_reductions(ops):
    import torch

    rng = numpy.random.default_rng(11)
    U = rng.standard_normal((3, 200_003))
    Ud = torch.from_numpy(U).cuda()
    got = ops.asum_rows(Ud)
    want = numpy.abs(U).sum(axis=1)
    assert numpy.all(numpy.abs(got - want) <= 1e-13 * want)
    x, u = U[0].copy(), U[1] * 1e-3
    x[::7] = 0.0
    got = ops.max_rel_change(torch.from_numpy(u).cuda(), torch.from_numpy(x).cuda())
    nz = x != 0
    assert got == float(numpy.max(numpy.abs(u[nz] / x[nz])))  # a maximum of correctly rounded quotients: exact
    assert ops.max_rel_change(Ud[1], torch.zeros_like(Ud[0])) == float("inf")


def test_vector_entry_points_reject_bad_arguments():
    from wxfactory_b200 import capi

    lib = capi.load()
    assert lib.wxb_vec_lincomb3(10, 1.0, None, 0.0, None, 1.0, None, 1.0, None, None) == -1
    assert lib.wxb_vec_multidot(10, 65, None, 10, 1, None, 10, None, None, None) == -1
    assert lib.wxb_vec_multidot(10, 2, None, 10, 3, None, 10, None, None, None) == -1
    assert lib.wxb_vec_combine(10, 65, None, None, 10, 0.0, None, 1.0, None, None) == -1
    assert lib.wxb_vec_reduce(10, 2, None, None, None, None, None) == -1

"""Host logic of the device-resident Krylov solvers (wxfactory_b200/krylov.py) against the reference's own fgmres /
kiops results on seeded small systems (tests/golden/krylov_small.npz, made by oracle/refharness/make_step_golden.py),
with the CPU stand-in for the vector kernels (tests/vecops_cpu.py)."""

import os

import numpy
import torch

from vecops_cpu import CpuVecOps
from wxfactory_b200.krylov import fgmres, kiops, pmex

G = numpy.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "krylov_small.npz"))


def _t(a):
    return torch.from_numpy(numpy.ascontiguousarray(a, dtype=float))


def test_fgmres_reproduces_reference_iterations_and_solution():
    M = G["fg_M"]
    A = lambda v: _t(M @ v.numpy())  # noqa: E731
    dM = numpy.diag(M).copy()
    for tag in ("a", "b", "h", "p"):  # h: Hegedus rescaling of the initial guess (fgmres.py:166-172), p: Jacobi preconditioner
        nr, nb, nit, flag, restart = G[f"fg_meta_{tag}"]
        x, norm_r, norm_b, niter, fl, res = fgmres(A, _t(G["fg_b"]), x0=_t(G["fg_x0"]), tol=1e-10, restart=int(restart), maxiter=50, ops=CpuVecOps(),
                                                   hegedus=(tag == "h"), preconditioner=(lambda v: _t(v.numpy() / dM)) if tag == "p" else None)
        assert (niter, fl) == (int(nit), int(flag))
        assert abs(norm_b - nb) <= 1e-13 * nb
        want = G[f"fg_res_{tag}"]
        got = numpy.array([r[0] for r in res])
        assert got.shape == want.shape
        # the history spans 10 decades; compare each entry relative to itself, loosely at the converged end
        assert numpy.all(numpy.abs(got - want) <= 1e-6 * want + 1e-13)
        assert numpy.linalg.norm(x.numpy() - G[f"fg_x_{tag}"]) <= 1e-12 * numpy.linalg.norm(G[f"fg_x_{tag}"])


def test_fgmres_error_behaviour():
    import pytest

    with pytest.raises(ValueError):
        fgmres(lambda v: v, _t(numpy.ones(5)), restart=20, ops=CpuVecOps())  # b shorter than the restart length
    x, nr, nb, nit, flag, res = fgmres(lambda v: v, _t(numpy.zeros(50)), restart=5, ops=CpuVecOps())  # b = 0: immediate return
    assert (nr, nb, nit, flag) == (0.0, 0.0, 0, 0) and float(x.abs().max()) == 0.0


def test_kiops_reproduces_reference_substeps_and_result():
    J = G["ki_J"]
    A = lambda v: _t(J @ v.numpy())  # noqa: E731
    for tag, taus, kw in (("a", [1.0], dict(m_init=10, mmin=10, mmax=64)), ("b", [0.25, 0.5, 1.0], dict(m_init=4, mmin=4, mmax=24, task1=True)),
                          ("c", [-0.05], dict(m_init=10, mmin=10, mmax=64))):  # c: backward in time
        w, stats = kiops(taus, A, _t(G["ki_u"]), tol=1e-9, ops=CpuVecOps(), **kw)
        want = G[f"ki_stats_{tag}"]
        assert [int(stats[i]) for i in (0, 1, 2, 3, 5)] == [int(want[i]) for i in (0, 1, 2, 3, 5)]
        assert abs(stats[4] - want[4]) <= 1e-3 * want[4]
        ref = G[f"ki_w_{tag}"]
        assert numpy.linalg.norm(w.numpy() - ref) <= 1e-11 * numpy.linalg.norm(ref)


def test_pmex_reproduces_reference_substeps_and_result():
    J = G["ki_J"]
    A = lambda v: _t(J @ v.numpy())  # noqa: E731
    for tag, taus, kw in (("a", [1.0], dict(m_init=10, mmin=10, mmax=64)), ("b", [0.25, 0.5, 1.0], dict(m_init=4, mmin=4, mmax=24, task1=True)),
                          ("c", [-0.05], dict(m_init=10, mmin=10, mmax=64))):
        w, stats = pmex(taus, A, _t(G["ki_u"]), tol=1e-9, ops=CpuVecOps(), **kw)
        want = G[f"pm_stats_{tag}"]
        assert [int(stats[i]) for i in (0, 1, 2, 3, 5, 6)] == [int(want[i]) for i in (0, 1, 2, 3, 5, 6)]
        assert abs(stats[4] - want[4]) <= 1e-3 * want[4]
        ref = G[f"pm_w_{tag}"]
        assert numpy.linalg.norm(w.numpy() - ref) <= 1e-11 * numpy.linalg.norm(ref)


def test_single_row_input_raises_like_the_reference():
    import pytest

    for solver in (kiops, pmex):
        with pytest.raises(ValueError):  # the reference's zero-padding of a single row fails in numpy (kiops.py:78-81)
            solver([1.0], lambda v: v, _t(G["ki_u"][:1]), ops=CpuVecOps())


def test_pmex_reaches_mmax_rows():
    """PMEX's one reduction per Krylov step covers j + 1 rows with j up to mmax, i.e. one more row than the
    basis size limit; the vector backend must take it (the CUDA backend walks the basis in chunks of WXB_VEC_MAX_ROWS
    rows, tests/test_gpu_stepops.py).  A stiff operator at a tight tolerance drives the adaptive Krylov size to mmax."""
    rng = numpy.random.default_rng(5)
    n = 400
    J = -numpy.diag(numpy.logspace(0, 4, n)) + 0.01 * rng.standard_normal((n, n))
    u = numpy.vstack([rng.standard_normal(n), rng.standard_normal(n)])
    seen = []

    class Recording(CpuVecOps):
        def multidot(self, V, m, W, k):
            seen.append(m)
            return super().multidot(V, m, W, k)

    mmax = 16
    w, stats = pmex([1.0], lambda v: _t(J @ v.numpy()), _t(u), tol=1e-7, m_init=8, mmin=8, mmax=mmax, ops=Recording())
    assert max(seen) == mmax + 1
    import scipy.linalg

    # phi_0(J) u0 + phi_1(J) u1 through the augmented matrix exponential
    A = numpy.zeros((n + 1, n + 1))
    A[:n, :n] = J
    A[:n, n] = u[1]
    want = scipy.linalg.expm(A)[:n] @ numpy.concatenate([u[0], [1.0]])
    assert numpy.linalg.norm(w.numpy()[0] - want) <= 1e-5 * numpy.linalg.norm(want)


def test_pmex_rejects_an_unusable_error_estimate():
    """the reference's pmex takes math.log / math.ceil of its error estimate and raises when that is NaN or
    infinite (pmex.py:283-300); the port raises a ValueError instead of carrying the NaN into the Krylov size."""
    import pytest

    n = 50
    bad = lambda v: _t(numpy.full(n, numpy.nan))  # noqa: E731
    u = numpy.vstack([numpy.ones(n), numpy.ones(n)])
    with numpy.errstate(all="ignore"), pytest.raises(ValueError):
        pmex([1.0], bad, _t(u), tol=1e-7, m_init=4, mmin=4, mmax=8, ops=CpuVecOps())

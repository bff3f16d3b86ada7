"""Golden vectors for the per-step operators and the callers' solvers -- TEST INFRASTRUCTURE ONLY.

    python oracle/refharness/make_step_golden.py [filter] [solvers] [integrators]

Everything is produced by the UNMODIFIED reference run here (harness.py):

* ``tests/golden/expfilter_<case>.npz``   ``DFROperators.expfilter`` (n x n) and
  ``ops.apply_filter_3d(Q0, metric)`` (wx_factory/geometry/operators.py:257-261) per rank, for dcmip21 (n = 3, the
  config that sets expfilter_apply = 1) and its n = 5 twin, and dcmip31 at n = 4 with the same filter settings;
* ``tests/golden/krylov_small.npz``       the reference's ``fgmres`` (wx_factory/solvers/fgmres.py:98-281) on a seeded
  dense system and ``kiops`` (wx_factory/solvers/kiops.py:10-346) on a seeded matrix with three phi-vectors:
  inputs, solutions, iteration counts, residual history / stats;
* ``tests/golden/steps_<integrator>_<case>.npz``  final state after N = 10 steps of the reference's own ``Ros2`` +
  fgmres, ``Epi(2)`` + kiops, ``Epi(3)`` + kiops and ``Epi(3)`` + pmex (config/case6.ini) around its own RHS on 6 thread-ranks (the drift report of
  SURVEY.md section 8d), with the per-step solver iteration counts.
"""

import os
import sys
import types

import numpy

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(os.path.dirname(_HERE))
sys.path.insert(0, _HERE)
sys.path.insert(0, _ROOT)

import harness  # noqa: E402
from make_golden import CONFIGS  # noqa: E402

GOLDEN = os.path.join(_ROOT, "tests", "golden")

FILTER_CASES = {
    "dcmip21": ("dcmip21", None),
    "dcmip21_n5": ("dcmip21_n5", None),
    "dcmip31_n4": ("dcmip31_n4", {"expfilter_apply": 1, "expfilter_order": 4, "expfilter_cutoff": 0.5, "expfilter_strength": 0.1}),
}

# name: (integrator, config, steps, dt fraction, config overrides)
STEP_CASES = {
    "ros2_dcmip21": ("ros2", "dcmip21", 10, 0.2, {"gmres_restart": 20, "tolerance": 1e-7}),
    # jacobian_method = fd: the finite-difference JVP of the north star (the schema default is the complex step)
    "epi2_dcmip31": ("epi2", "dcmip31", 10, 0.2, {"tolerance": 1e-7, "exponential_solver": "kiops", "jacobian_method": "fd"}),
    "epi3_dcmip21": ("epi3", "dcmip21", 10, 0.2, {"tolerance": 1e-7, "exponential_solver": "kiops", "jacobian_method": "fd"}),
    # config/case6.ini as shipped: Epi(3) + PMEX on the shallow-water Rossby-Haurwitz wave (BASELINE configs[0])
    # a higher-order multistep member (two previous steps, phi_1..phi_3): exercises the history deque and the coefficient rows
    "epi4_dcmip21": ("epi4", "dcmip21", 6, 0.2, {"tolerance": 1e-7, "exponential_solver": "kiops", "jacobian_method": "fd"}),
    "epi3_case6": ("epi3", "case6", 10, 1.0, {"tolerance": 1e-7, "exponential_solver": "pmex", "jacobian_method": "fd"}),
}


def make_filter():
    for name, (cfg_name, extra) in FILTER_CASES.items():
        ini, overrides, patch = CONFIGS[cfg_name]
        ov = dict(overrides or {})
        ov.update(extra or {})
        st = harness.build(harness.read_ini(ini), overrides=ov, ic_patch=patch)
        assert st[0].ops.expfilter_apply
        Qf = numpy.stack(harness.run_ranks(6, lambda r: st[r].ops.apply_filter_3d(st[r].Q, st[r].metric)))
        path = os.path.join(GOLDEN, f"expfilter_{name}.npz")
        numpy.savez_compressed(path, expfilter=numpy.asarray(st[0].ops.expfilter, dtype=float), Qf=Qf)
        Q0 = numpy.stack([s.Q for s in st])
        print(f"expfilter {name}: |Qf-Q0|/|Q0| = {numpy.linalg.norm(Qf-Q0)/numpy.linalg.norm(Q0):.3e}, {os.path.getsize(path)/1e6:.2f} MB")


def make_solvers():
    harness.bootstrap()

    def work(_r):
        from device import CpuDevice
        from mpi4py import MPI
        from solvers.fgmres import fgmres
        from solvers.kiops import kiops
        from solvers.pmex import pmex

        dev = CpuDevice(comm=MPI.COMM_WORLD)
        rng = numpy.random.default_rng(20261105)
        n = 300
        M = numpy.eye(n) * 4.0 + rng.standard_normal((n, n)) / numpy.sqrt(n)
        b = rng.standard_normal(n)
        x0 = rng.standard_normal(n)
        out = {"fg_M": M, "fg_b": b, "fg_x0": x0}
        dM = numpy.diag(M).copy()
        for tag, restart in (("a", 20), ("b", 6), ("h", 10), ("p", 8)):  # h: Hegedus rescaling, p: Jacobi-preconditioned
            x, nr, nb, nit, flag, res = fgmres(lambda v: M @ v, b, x0=x0, tol=1e-10, restart=restart, maxiter=50, device=dev, hegedus=(tag == "h"),
                                               preconditioner=(lambda v: v / dM) if tag == "p" else None)
            out[f"fg_x_{tag}"] = x
            out[f"fg_meta_{tag}"] = numpy.array([nr, nb, nit, flag, restart], dtype=float)
            out[f"fg_res_{tag}"] = numpy.array([r[0] for r in res])
        nk = 200
        J = -numpy.diag(numpy.linspace(0.1, 30.0, nk)) + 0.5 * rng.standard_normal((nk, nk)) / numpy.sqrt(nk)
        u = rng.standard_normal((3, nk))
        out["ki_J"], out["ki_u"] = J, u
        for tag, taus, kw in (("a", [1.0], dict(m_init=10, mmin=10, mmax=64)), ("b", [0.25, 0.5, 1.0], dict(m_init=4, mmin=4, mmax=24, task1=True)),
                              ("c", [-0.05], dict(m_init=10, mmin=10, mmax=64))):  # c: backward in time (sgn = -1), a single phi_0 vector below
            w, stats = kiops(taus, lambda v: J @ v, u, tol=1e-9, device=dev, **kw)
            out[f"ki_w_{tag}"] = w
            out[f"ki_stats_{tag}"] = numpy.array([float(s) for s in stats])
            w, stats = pmex(taus, lambda v: J @ v, u, tol=1e-9, device=dev, **kw)
            out[f"pm_w_{tag}"] = w
            out[f"pm_stats_{tag}"] = numpy.array([float(s) for s in stats])
        # the reference's own unit-test inputs (tests/unit/solvers/test_kiops.py, test_pmex.py, test_fgmres.py:
        # random.Random(5646459), 64 x 64 entries uniform in [-1000, 1000], identity matvec for the phi solvers)
        import random

        rnd = random.Random(5646459)
        U = numpy.array([[rnd.uniform(-1000.0, 1000.0) for _ in range(64)] for _ in range(64)])
        out["ut_U"] = U
        w, stats = kiops([1.0], lambda v: v, U, 1e-7, device=dev)
        out["ut_kiops_w"], out["ut_kiops_stats"] = w, numpy.array([float(s) for s in stats])
        w, stats = pmex([1.0], lambda v: v, U, 1e-7, device=dev)
        out["ut_pmex_w"], out["ut_pmex_stats"] = w, numpy.array([float(s) for s in stats])
        rnd = random.Random(5646459)
        bv = numpy.array([rnd.uniform(-1000.0, 1000.0) for _ in range(64)])
        Am = numpy.array([[rnd.uniform(-1000.0, 1000.0) for _ in range(64)] for _ in range(64)])
        x, nr, nb, nit, flag, res = fgmres(lambda v: Am @ v, bv, tol=1e-7, device=dev)
        out["ut_fg_b"], out["ut_fg_A"], out["ut_fg_x"] = bv, Am, x
        out["ut_fg_meta"] = numpy.array([nr, nb, nit, flag], dtype=float)
        return out

    out = harness.run_ranks(1, work)[0]
    path = os.path.join(GOLDEN, "krylov_small.npz")
    numpy.savez_compressed(path, **out)
    print("krylov_small:", {k: (v.tolist() if v.size < 8 else v.shape) for k, v in out.items() if "meta" in k or "stats" in k})


def make_integrators():
    harness.bootstrap()
    sys.modules.setdefault("netCDF4", types.ModuleType("netCDF4"))  # imported by the output manager only
    from integrators.epi import Epi
    from integrators.ros2 import Ros2

    only = os.environ.get("WXB_STEP_CASES")
    for name, (kind, cfg_name, nsteps, frac, extra) in STEP_CASES.items():
        if only and name not in only.split(","):
            continue
        ini, overrides, patch = CONFIGS[cfg_name]
        ov = dict(overrides or {})
        ov.update(extra)
        st = harness.build(harness.read_ini(ini), overrides=ov, ic_patch=patch)
        dt = float(st[0].config.dt) * frac

        def per_rank(r):
            cfg = st[r].config
            if kind == "ros2":
                integ = Ros2(cfg, st[r].rhs.full, device=st[r].device)
            else:
                integ = Epi(cfg, int(kind[3]), st[r].rhs.full, device=st[r].device)
            Q = st[r].Q
            its = []
            for _ in range(nsteps):
                integ.__prestep__(Q, dt)
                Q = integ.__step__(Q, dt)
                its.append(integ.solver_info.total_num_it if integ.solver_info is not None else -1)  # None: multistep start-up step
            return Q, its

        res = harness.run_ranks(6, per_rank)
        QN = numpy.stack([q for q, _ in res])
        path = os.path.join(GOLDEN, f"steps_{name}.npz")
        numpy.savez_compressed(path, QN=QN, nsteps=numpy.int64(nsteps), dt=numpy.float64(dt), iterations=numpy.array(res[0][1]),
                               tolerance=numpy.float64(extra["tolerance"]), gmres_restart=numpy.int64(extra.get("gmres_restart", 0)),
                               exponential_solver=numpy.str_(extra.get("exponential_solver", "")))
        Q0 = numpy.stack([s.Q for s in st])
        print(f"{name}: {nsteps} steps of dt={dt}, iterations {res[0][1]}, |QN-Q0|/|Q0| = {numpy.linalg.norm(QN-Q0)/numpy.linalg.norm(Q0):.3e}, "
              f"{os.path.getsize(path)/1e6:.2f} MB")


if __name__ == "__main__":
    numpy.seterr(all="ignore")
    what = sys.argv[1:] or ["filter", "solvers", "integrators"]
    if "filter" in what:
        make_filter()
    if "solvers" in what:
        make_solvers()
    if "integrators" in what:
        make_integrators()

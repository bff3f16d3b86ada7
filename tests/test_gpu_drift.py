"""Drift report: N steps of SSP-RK3 with the B200 right-hand side against the trajectory the
UNMODIFIED reference produced with its own Tvdrk3 integrator and RHS (oracle/refharness/make_drift.py ->
tests/golden/tvdrk3_*.npz).  The three stage formulas are wx_factory/integrators/tvdrk3.py:13-17; everything stays
on the device (30 RHS evaluations, no host round trip).  The oracle is put through the same steps on the CPU in the
non-gpu test, which pins the golden trajectories themselves."""

import os

import numpy
import pytest

from conftest import GOLDEN

CASES = ["case6", "dcmip21", "dcmip31_n4"]
DRIFT_TOL = 1e-12  # relative L2 difference of the whole state after the last step


def ssprk3(rhs, Q, dt, nsteps):
    for _ in range(nsteps):
        Q1 = Q + rhs(Q) * dt
        Q2 = 0.75 * Q + 0.25 * Q1 + 0.25 * rhs(Q1) * dt
        Q = 1.0 / 3.0 * Q + 2.0 / 3.0 * Q2 + 2.0 / 3.0 * rhs(Q2) * dt
    return Q


def _traj(name):
    with numpy.load(os.path.join(GOLDEN, f"tvdrk3_{name}.npz")) as z:
        return z["QN"], int(z["nsteps"]), float(z["dt"])


@pytest.mark.parametrize("name", ["case6", "dcmip21"])
def test_oracle_reproduces_the_reference_trajectory(golden, name):
    from oracle import jvp_oracle

    prob, ex = golden(name)
    QN, nsteps, dt = _traj(name)
    orc = jvp_oracle.make_oracle(prob)
    with numpy.errstate(all="ignore"):
        got = ssprk3(orc.rhs, ex["Q0"], dt, nsteps)
    assert numpy.array_equal(got, QN)


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_state_drift_after_n_steps(golden, name):
    import torch

    from wxfactory_b200 import RhsB200

    prob, ex = golden(name)
    QN, nsteps, dt = _traj(name)
    rhs = RhsB200(prob)
    Q = ssprk3(rhs, torch.from_numpy(ex["Q0"]).cuda(), dt, nsteps)
    got = Q.cpu().numpy()
    drift = numpy.linalg.norm(got - QN) / numpy.linalg.norm(QN)
    per_var = [numpy.linalg.norm(got[:, v] - QN[:, v]) / max(numpy.linalg.norm(QN[:, v]), 1e-300) for v in range(prob.num_var)]
    print(f"drift {name}: {drift:.3e} after {nsteps} steps; per variable " + " ".join(f"{x:.1e}" for x in per_var))
    assert drift <= DRIFT_TOL
    rhs.close()

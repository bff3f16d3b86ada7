"""ORACLE (test infrastructure, CPU, numpy) -- matrix-free Jacobian-vector products.

Restates wx_factory/solvers/matvec.py:
  matvec_fun(..., method='fd')  :57-68   J v ~= dt * (RHS(Q + eps v) - RHS(Q)) / eps,
                                         eps = sqrt(finfo(float32).eps), unscaled
  matvec_rat                    :76-88   v - 0.5 * (FD J v)        (Ros2 system matrix I - dt/2 J)
The complex-step variant (:59-61) is excluded by design (north star: FD JVP).
Checker only; pinned against the reference's own matvec outputs in tests/golden.
"""

import math

import numpy

EPS_FD = math.sqrt(numpy.finfo(numpy.float32).eps)


def matvec_fd(rhs_fn, vec, dt, Q, rhs0):
    Qvec = Q + EPS_FD * vec.reshape(Q.shape)
    jac = dt * (rhs_fn(Qvec) - rhs0) / EPS_FD
    return jac.flatten()


def matvec_rat(rhs_fn, vec, dt, Q, rhs0):
    Qvec = Q + EPS_FD * vec.reshape(Q.shape)
    jac = dt * (rhs_fn(Qvec) - rhs0) / EPS_FD
    return vec - 0.5 * jac.flatten()


def make_oracle(prob):
    """EulerOracle or SwOracle for the given problem."""
    if prob.equations == "euler":
        from .euler_oracle import EulerOracle

        return EulerOracle(prob)
    from .sw_oracle import SwOracle

    return SwOracle(prob)

"""ORACLE (test infrastructure, CPU, numpy) -- per-step exponential filter of the 3-D state.

Restatement of ``DFROperators.apply_filter_3d`` (wx_factory/geometry/operators.py:257-261) with the dense operator the
reference builds (operators.py:110-119):

    filter_x = kron(I_{n^2}, F).T ; filter_y = kron(I_n, kron(F, I_n)).T ; filter_z = kron(F, I_{n^2}).T
    expfilter_new = (filter_x @ filter_y) @ filter_z
    Q -> ((sqrtG * Q) @ expfilter_new) * inv_sqrtG

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import it.  Parity status: PINNED against the
reference run here (tests/golden/expfilter_*.npz made by oracle/refharness/make_step_golden.py;
tests/test_oracle_vs_golden.py)."""

import numpy


def dense_filter(F):
    n = F.shape[0]
    I2, I3 = numpy.eye(n), numpy.eye(n * n)
    fx = numpy.kron(I3, F).T
    fy = numpy.kron(I2, numpy.kron(F, I2)).T
    fz = numpy.kron(F, I3).T
    return (fx @ fy) @ fz


def apply_filter_3d(prob, F, Q):
    """Q: (6, 5, nz, nh, nh, P) (any shape of that size) -> filtered state in the same shape."""
    shape = Q.shape
    q = numpy.asarray(Q, dtype=float).reshape(prob.state_shape)
    sg = prob.sqrtg.reshape((6, 1) + prob.state_shape[2:])
    return (((sg * q) @ dense_filter(numpy.asarray(F, dtype=float))) * (1.0 / sg)).reshape(shape)

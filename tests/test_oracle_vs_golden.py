"""The numpy oracle against the golden vectors produced by the reference itself
(oracle/refharness/make_golden.py).  The oracle follows the reference operation by operation,
so agreement is expected to be bit-exact on the machine that made the fixtures; 1e-14 relative
leaves room for a different BLAS summation order elsewhere."""

import numpy
import pytest

from conftest import EXTRA_ORACLE_PINS, GOLDEN_NAMES
from oracle import jvp_oracle

TOL = 1e-14


def rel(a, b):
    return numpy.linalg.norm(a - b) / numpy.linalg.norm(b)


@pytest.mark.parametrize("name", GOLDEN_NAMES + EXTRA_ORACLE_PINS)
def test_rhs_matches_reference(golden, name):
    prob, ex = golden(name)
    with numpy.errstate(all="ignore"):
        out = jvp_oracle.make_oracle(prob).rhs(ex["Q0"])
    assert out.shape == ex["rhs0"].shape
    assert rel(out, ex["rhs0"]) <= TOL


@pytest.mark.parametrize("name", GOLDEN_NAMES + EXTRA_ORACLE_PINS)
def test_jvp_matches_reference(golden, name):
    prob, ex = golden(name)
    orc = jvp_oracle.make_oracle(prob)
    dt = float(ex["dt"])
    with numpy.errstate(all="ignore"):
        fd = jvp_oracle.matvec_fd(orc.rhs, ex["v"].ravel(), dt, ex["Q0"], ex["rhs0"])
        rat = jvp_oracle.matvec_rat(orc.rhs, ex["v"].ravel(), dt, ex["Q0"], ex["rhs0"])
    # FD amplifies RHS rounding by 1/eps: state the tolerance relative to dt*|rhs|/eps
    scale = dt * numpy.linalg.norm(ex["rhs0"]) / jvp_oracle.EPS_FD
    assert numpy.linalg.norm(fd - ex["jv_fd"].ravel()) <= TOL * scale
    assert numpy.linalg.norm(rat - ex["jv_rat"].ravel()) <= TOL * max(scale, numpy.linalg.norm(ex["v"]))


def test_flat_input_shape_is_preserved(golden):
    prob, ex = golden("dcmip21")
    orc = jvp_oracle.make_oracle(prob)
    out = orc.rhs(ex["Q0"].ravel())
    assert out.shape == (ex["Q0"].size,)
    assert rel(out, ex["rhs0"].ravel()) <= TOL


@pytest.mark.parametrize("name", ["dcmip21", "dcmip21_n5", "dcmip31_n4"])
def test_filter_oracle_matches_reference(golden, name):
    import os

    from conftest import GOLDEN
    from oracle import filter_oracle

    prob, ex = golden(name)
    g = numpy.load(os.path.join(GOLDEN, f"expfilter_{name}.npz"))
    out = filter_oracle.apply_filter_3d(prob, g["expfilter"], ex["Q0"])
    assert rel(out, g["Qf"]) <= TOL

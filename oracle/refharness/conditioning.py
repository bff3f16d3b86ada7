"""Rounding sensitivity of the reference RHS on every golden state -- TEST INFRASTRUCTURE ONLY.

    python oracle/refharness/conditioning.py        # writes tests/golden/conditioning.json

The numpy oracle reproduces the reference bit for bit (test_oracle_vs_golden.py), so evaluating the SAME
formulas in 80-bit extended precision (numpy.longdouble) measures how far the reference's own float64 result is
from the exactly-rounded one: ``ref_rounding = |rhs0_ref - rhs0_ld| / |rhs0_ld|``.  States at rest in hydrostatic
balance (DCMIP 3-1 at high order) make the RHS a small difference of large terms; there the reference itself carries
~1e-10 of rounding noise, and a different (sum-factorised) summation order cannot agree with it more closely than
that.  tests/test_gpu_parity.py uses ``max(1e-12, ref_rounding)`` as the per-case bound.
"""

import json
import os
import sys

import numpy

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, _ROOT)

from oracle import jvp_oracle  # noqa: E402
from wxfactory_b200.problem import load_problem  # noqa: E402

NAMES = ["case6", "case5", "galewsky", "dcmip31", "dcmip21", "dcmip31_n4", "dcmip31_n5", "dcmip31_n6", "dcmip21_n5"]


def _to_ld(o):
    for k, v in list(vars(o).items()):
        if isinstance(v, numpy.ndarray) and v.dtype == numpy.float64:
            setattr(o, k, v.astype(numpy.longdouble))
        elif isinstance(v, list) and v and isinstance(v[0], numpy.ndarray) and v[0].dtype == numpy.float64:
            setattr(o, k, [x.astype(numpy.longdouble) for x in v])


def main():
    numpy.seterr(all="ignore")
    out = {"numpy": numpy.__version__, "longdouble_bits": int(numpy.finfo(numpy.longdouble).nmant) + 1, "cases": {}}
    for name in NAMES:
        prob, ex = load_problem(os.path.join(_ROOT, "tests", "golden", name + ".npz"))
        o = jvp_oracle.make_oracle(prob)
        r64 = o.rhs(ex["Q0"])
        assert numpy.array_equal(r64, ex["rhs0"]), name
        _to_ld(o)
        rld = o.rhs(ex["Q0"].astype(numpy.longdouble))
        assert rld.dtype == numpy.longdouble, (name, rld.dtype)
        nrm = float(numpy.linalg.norm(rld.astype(float)))
        allv = float(numpy.linalg.norm((ex["rhs0"] - rld).astype(float))) / nrm
        per = []
        for v in range(prob.num_var):
            d = float(numpy.linalg.norm((ex["rhs0"][:, v] - rld[:, v]).astype(float)))
            per.append(d / max(float(numpy.linalg.norm(rld[:, v].astype(float))), 1e-300))
        out["cases"][name] = {"ref_rounding": allv, "ref_rounding_per_var": per}
        print(name, f"{allv:.3e}", " ".join(f"{x:.2e}" for x in per))
    with open(os.path.join(_ROOT, "tests", "golden", "conditioning.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()

"""The table-driven exp / log of wxfactory_b200/csrc/fastmath.cuh, built for the host with g++ and compared with libm in
long double on 4 M random arguments over the ranges the kernels use (x in e^-12 .. e^12, and around 1 for the log; exp
arguments in -24 .. 24).  Bounds: log within 0.55 ulp (its k*ln2 + log c sum is exact: the interface fluxes amplify the
rounding of exactly these logs), exp within 1.3 ulp; non-positive / non-finite arguments give NaN (so a state
that has blown up still trips the callers' non-finite check)."""

import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))


def test_fast_exp_log_accuracy(tmp_path):
    exe = str(tmp_path / "fastmath_check")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-I", os.path.join(HERE, "..", "wxfactory_b200", "csrc"),
                           os.path.join(HERE, "fastmath_check.cpp"), "-o", exe])
    out = subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split()
    vals = dict(zip(out[0::2], out[1::2]))
    assert float(vals["log_ulp"]) <= 0.55
    assert float(vals["log_abs"]) <= 1.2e-16
    assert float(vals["exp_ulp"]) <= 1.3
    assert vals["special"] == "1"

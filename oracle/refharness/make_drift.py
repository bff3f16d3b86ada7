"""Golden trajectories for the drift report -- TEST INFRASTRUCTURE ONLY.

    python oracle/refharness/make_drift.py        # writes tests/golden/tvdrk3_<case>.npz

Runs the UNMODIFIED reference: its own ``Tvdrk3.__step__`` (wx_factory/integrators/tvdrk3.py:12-19) around its own
``RhsBundle.full`` on 6 thread-ranks, N steps from the initial state with a stable fraction of the config's dt, and stores the final
state.  tests/test_gpu_drift.py integrates the same steps with the B200 RHS (drift report).
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

# name: (steps, fraction of the config's dt: the configs' steps are sized for the exponential integrators, far beyond
# the explicit stability limit of SSP-RK3)
CASES = {"case6": (10, 0.05), "dcmip21": (10, 0.01), "dcmip31_n4": (10, 0.01)}


def main():
    numpy.seterr(all="ignore")
    harness.bootstrap()
    sys.modules.setdefault("netCDF4", types.ModuleType("netCDF4"))  # imported by the output manager only
    from integrators.tvdrk3 import Tvdrk3

    for name, (nsteps, frac) in CASES.items():
        ini, overrides, patch = CONFIGS[name]
        st = harness.build(harness.read_ini(ini), overrides=overrides, ic_patch=patch)
        dt = float(st[0].config.dt) * frac

        def per_rank(r):
            integ = Tvdrk3(st[r].config, st[r].rhs.full, device=st[r].device)
            Q = st[r].Q
            for _ in range(nsteps):
                Q = integ.__step__(Q, dt)
            return Q

        QN = numpy.stack(harness.run_ranks(6, per_rank))
        path = os.path.join(_ROOT, "tests", "golden", f"tvdrk3_{name}.npz")
        numpy.savez_compressed(path, QN=QN, nsteps=numpy.int64(nsteps), dt=numpy.float64(dt))
        Q0 = numpy.stack([s.Q for s in st])
        print(f"{name}: {nsteps} steps of dt={dt}, |QN-Q0|/|Q0| = {numpy.linalg.norm(QN-Q0)/numpy.linalg.norm(Q0):.3e}, "
              f"{os.path.getsize(path)/1e6:.2f} MB")


if __name__ == "__main__":
    main()

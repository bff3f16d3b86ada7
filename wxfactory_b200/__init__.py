"""wxfactory_b200 -- B200-native DFR right-hand side and FD Jacobian-vector product for WxFactory's cubed sphere.

Only what the hot path needs: the problem container (static inputs), the partition / connectivity tables,
the ctypes binding of libwxb200.so (hand-written sm_100a CUDA behind a C ABI, include/wxb200.h) and the
host-side mirror of the reference's RHS / matvec interface; around it the per-step operators
(stepops: exponential filter, stage algebra, Krylov kernels), the device-resident solvers and integrators that call
the path (krylov: fgmres, kiops; integrators: Tvdrk3, Ros2, Epi) and the restart-file bridge (state_io).
"""

from .device import B200Device  # noqa: F401
from .problem import EulerProblem, SwProblem, load_problem, save_problem  # noqa: F401
from .rhs import EPS_FD, MatvecOpB200, MatvecOpFd, MatvecOpRat, RhsB200, RhsBundleB200, matvec_fun, matvec_rat  # noqa: F401
from .stepops import ExpFilterB200, VecOps  # noqa: F401
from . import integrators, krylov, state_io  # noqa: F401

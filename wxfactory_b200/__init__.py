"""wxfactory_b200 -- B200-native DFR right-hand side and FD Jacobian-vector product for WxFactory's cubed sphere.

Only what the hot path needs: the problem container (static inputs), the partition / connectivity tables,
the ctypes binding of libwxb200.so (hand-written sm_100a CUDA behind a C ABI, include/wxb200.h) and the
host-side mirror of the reference's RHS / matvec interface.
"""

from .device import B200Device  # noqa: F401
from .problem import EulerProblem, SwProblem, load_problem, save_problem  # noqa: F401
from .rhs import EPS_FD, MatvecOpB200, MatvecOpFd, MatvecOpRat, RhsB200, RhsBundleB200, matvec_fun, matvec_rat  # noqa: F401

"""TEST INFRASTRUCTURE ONLY: ``RhsB200`` with its C calls served by the numpy oracle.

The CPU tests that drive the reference's own, unmodified integrators and solvers over the B200 Python surface need the
host-side logic of ``wxfactory_b200.rhs`` (shape handling, fresh result arrays, ``matvec_fun`` dispatch on ``__self__``,
``MatvecOp`` binding, launch accounting) without a GPU.  ``OracleRhsB200`` keeps every Python line of ``RhsB200`` and
replaces only what sits below the C ABI: a stand-in library object whose ``wxb_*_host`` entry points read / write the
same raw host pointers and evaluate ``oracle/`` instead of launching kernels.  Never imported by the package."""

import ctypes

import numpy

from oracle import jvp_oracle
from wxfactory_b200 import rhs as rhs_mod
from wxfactory_b200.partition import build_partition


def _arr(ptr, n):
    return numpy.ctypeslib.as_array((ctypes.c_double * n).from_address(int(ptr)))


class _OracleLib:
    """The subset of libwxb200.so the host path of RhsB200 calls (include/wxb200.h), answered by the oracle."""

    def __init__(self, prob):
        self.orc = jvp_oracle.make_oracle(prob)
        self.shape = prob.state_shape
        self.n = int(numpy.prod(self.shape))
        self.lq = self.lr = None
        self.calls = []

    def _rhs(self, q):
        with numpy.errstate(all="ignore"):
            return self.orc.rhs(q.reshape(self.shape)).reshape(-1)

    def _jvp(self, q, v, r0, dt, eps, mode):
        with numpy.errstate(all="ignore"):
            jac = dt * (self._rhs(q + eps * v) - r0) / eps  # wx_factory/solvers/matvec.py:57-66
        return jac if mode == 0 else v - 0.5 * jac

    def wxb_rhs_host(self, h, q, out):
        self.calls.append("rhs_host")
        _arr(out, self.n)[:] = self._rhs(_arr(q, self.n))
        return 0

    def wxb_jvp_host(self, h, q, v, r0, dt, eps, mode, out):
        self.calls.append("jvp_host")
        _arr(out, self.n)[:] = self._jvp(_arr(q, self.n), _arr(v, self.n), _arr(r0, self.n), dt, eps, mode)
        return 0

    def wxb_linearize_host(self, h, q, out):
        self.calls.append("linearize_host")
        self.lq = _arr(q, self.n).copy()
        self.lr = self._rhs(self.lq)
        _arr(out, self.n)[:] = self.lr
        return 0

    def wxb_bind_host(self, h, q, r0):
        self.calls.append("bind_host")
        self.lq, self.lr = _arr(q, self.n).copy(), _arr(r0, self.n).copy()
        return 0

    def wxb_jvp_bound_host(self, h, v, dt, eps, mode, out):
        self.calls.append("jvp_bound_host")
        _arr(out, self.n)[:] = self._jvp(self.lq, _arr(v, self.n), self.lr, dt, eps, mode)
        return 0

    def wxb_set_expfilter(self, h, filt):
        n = int(round(numpy.prod(self.shape[-1:]) ** (1.0 / 3.0)))
        self.expfilter = _arr(filt, n * n).reshape(n, n).copy()  # kept; the filter itself runs on device tensors only
        return 0

    def wxb_handle_fused(self, h):
        return 1

    def wxb_timings(self, h, ms):
        return -4

    def wxb_destroy(self, h):
        return 0

    def wxb_last_error(self, h):
        return b""


class OracleRhsB200(rhs_mod.RhsB200):
    """Same Python surface, no device: see the module docstring."""

    def __init__(self, problem, nranks=1, rank=0, device=0, process_group=None):
        if nranks != 1:
            raise ValueError("the oracle-backed stand-in serves single-rank handles")
        self.prob, self.nranks, self.rank, self.device_index, self.group = problem, 1, 0, device, process_group
        euler = problem.equations == "euler"
        self.num_var = problem.num_var
        self.sweep = True
        self._fused = (True, True)
        self.path_name = "sweep" if euler else "tile"
        self.num_trace = 6 if euler else 3
        self.part = build_partition(problem.nh, problem.nz if euler else 1, problem.n, problem.bnd_sn, problem.bnd_we, num_var=self.num_trace,
                                    nfaces=6 if euler else 4, trace_layout="element" if euler else "field")
        self._lib = _OracleLib(problem)
        self._h = object()
        self.size = self._lib.n
        self.timestamps, self.timings = [], []
        self._have_timing = False
        self.launches = 0
        self._torch = None

    def _host_result(self):
        return numpy.empty(self.size)

    def close(self):
        self._h = None

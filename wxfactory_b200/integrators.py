"""Device-resident time integrators around ``RhsB200``.

Same class names, constructor arguments and ``step`` / ``__step__`` / ``__prestep__`` protocol as the reference
(wx_factory/integrators/integrator.py:17-110, tvdrk3.py, ros2.py, epi.py), so a driver written for the reference's
integrators runs unchanged; the state, every stage vector and the Krylov bases stay on the GPU and each full-state
combination is one launch of csrc/stepops.cu (``VecOps``).  ``param`` is the reference's ``Configuration`` or any object
with the attributes read here (``tolerance``, ``gmres_restart``, ``exponential_solver``, ``verbose_solver``).

The state ``Q`` is a CUDA float64 tensor of any shape of the right size (what ``RhsB200.__call__`` accepts); the new
state comes back in the same shape.
"""

import math
from collections import deque
from time import time

from .krylov import fgmres, kiops, pmex
from .rhs import MatvecOpFd, MatvecOpRat, RhsB200
from .stepops import VecOps


class SolverInfo:
    """wx_factory/solvers/solver_info.py: what an integrator reports about its last step."""

    def __init__(self, flag=0, total_solve_time=0.0, total_num_it=0, iterations=None):
        self.flag, self.total_solve_time, self.total_num_it = flag, total_solve_time, total_num_it
        self.iterations = iterations if iterations is not None else []


class Integrator:
    """integrator.py:17-110 without the output manager and preconditioners (outside the hot path)."""

    def __init__(self, param, rhs, ops=None, filter=None, **kwargs):  # noqa: A002
        owner = getattr(rhs, "__self__", rhs)
        if isinstance(owner, RhsB200):
            self.rhs_obj = owner
            self.ops = ops if ops is not None else VecOps(owner.device_index, owner.group, owner.nranks)
        else:  # a foreign RHS callable: the vector backend must come with it (the CPU tests check the host logic so)
            if ops is None:
                raise TypeError("the B200 integrators need an RhsB200 (or its .full) as right-hand side")
            self.rhs_obj, self.ops = rhs, ops
        self.filter = filter  # optional ExpFilterB200 applied after every step (simulation.py:148)
        self.verbose_solver = getattr(param, "verbose_solver", 0)
        self.solver_info = None
        self.last_solver_info = None
        self.sim_time = -1.0
        self.failure_flag = 0
        self.num_completed_steps = 0
        self.latest_time = 0.0

    def __prestep__(self, Q, dt):
        pass

    def __step__(self, Q, dt):
        raise NotImplementedError

    def step(self, Q, dt):
        """Advance the system forward in time (integrator.py:69-110); also applies the per-step filter and the
        NaN check the reference's driver does right after (simulation.py:148,399-408)."""
        t0 = time()
        self.__prestep__(Q, dt)
        result = self.__step__(Q, dt)
        if self.filter is not None:
            result = self.filter.apply_filters(result, None, None, dt)
        self.latest_time = time() - t0
        self.sim_time += dt
        self.num_completed_steps += 1
        # integrator.py:107-108 resets solver_info after handing it to the output manager; here there is no output manager:
        # the step's report stays readable as last_solver_info and solver_info is reset like the reference's
        self.last_solver_info, self.solver_info = self.solver_info, None
        return result

    def check_finite(self, Q):
        """simulation.py:399-408: raise if the state holds a NaN (or Inf)."""
        bad = self.ops.nonfinite(Q.reshape(-1))
        if bad:
            raise ValueError(f"NaN detected in the state ({bad} non-finite values)")


class Tvdrk3(Integrator):
    """SSP-RK3 (tvdrk3.py:7-20).  Each stage is one RHS evaluation + one fused combination whose roundings are
    those of the reference's numpy expressions.

    ``graph=True`` (single-rank ``RhsB200`` only): the 9 launches of a step (3 x traces + main, 3 combinations) are
    captured once into a CUDA graph and replayed -- for small grids, where a step is launch-bound.  The graph is
    re-captured when ``dt`` or the state size changes.  While it is in use the handle's per-evaluation timing ring holds
    events recorded inside the graph: ``RhsB200.timing_history`` is not available for those evaluations."""

    def __init__(self, param, rhs, graph=False, **kwargs):
        super().__init__(param, rhs, **kwargs)
        self.rhs = rhs
        self.graph = bool(graph)
        self._g = None
        if self.graph and not (isinstance(self.rhs_obj, RhsB200) and self.rhs_obj.nranks == 1):
            raise ValueError("graph=True needs a single-rank RhsB200")

    def _stages(self, q, dt, f):
        o = self.ops
        q1 = o.lincomb3(1.0, q, 0.0, None, 1.0, f(q), dt)                           # Q + rhs(Q)*dt
        q2 = o.lincomb3(0.75, q, 0.25, q1, 0.25, f(q1), dt)                         # 0.75 Q + 0.25 Q1 + 0.25 rhs(Q1) dt
        return o.lincomb3(1.0 / 3.0, q, 2.0 / 3.0, q2, 2.0 / 3.0, f(q2), dt, out=q1)  # 1/3 Q + 2/3 Q2 + 2/3 rhs(Q2) dt

    def _capture(self, q, dt):
        r = self.rhs_obj
        torch = r._tm()
        f = lambda x: r._eval_device(x, None, None, 0.0, 0.0, -1)  # noqa: E731  (no timing queries inside a capture)
        gq = q.clone()
        side = torch.cuda.Stream(device=q.device)
        side.wait_stream(torch.cuda.current_stream(q.device))
        with torch.cuda.stream(side):  # one eager pass on a side stream: allocations and lazy set-up happen here
            self._stages(gq, dt, f)
        torch.cuda.current_stream(q.device).wait_stream(side)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            out = self._stages(gq, dt, f)
        r._have_timing = False  # the ring's latest events live in the graph
        self._g = (g, gq, out, float(dt), q.numel())

    def __step__(self, Q, dt):
        q = Q.reshape(-1)
        self.solver_info = SolverInfo(total_num_it=1)
        if self.graph:
            if self._g is None or self._g[3] != float(dt) or self._g[4] != q.numel():
                self._capture(q, dt)
            g, gq, out, _, _ = self._g
            gq.copy_(q)
            g.replay()
            return out.clone().view(Q.shape)
        return self._stages(q, dt, self.rhs_obj).view(Q.shape)


class Ros2(Integrator):
    """Rosenbrock-Euler with the matrix-free system  (I - dt/2 J) x = b  solved by FGMRES (ros2.py:13-89)."""

    def __init__(self, param, rhs_handle, **kwargs):
        super().__init__(param, rhs_handle, **kwargs)
        self.rhs_handle = rhs_handle
        self.tol = param.tolerance
        self.gmres_restart = param.gmres_restart
        self.linear_solver = getattr(param, "linear_solver", "fgmres")
        if self.linear_solver != "fgmres":
            raise ValueError("only linear_solver = fgmres is available on the B200 path")
        self.preconditioner = None

    def __prestep__(self, Q, dt):
        rhs = self.rhs_obj(Q)
        self.Q_flat = Q.reshape(-1)
        self.A = MatvecOpRat(dt, Q, rhs, self.rhs_obj)
        # b = A(Q) + rhs*dt (ros2.py:30)
        self.b = self.ops.lincomb3(1.0, self.A(self.Q_flat), 0.0, None, 1.0, rhs.reshape(-1), dt)

    def __step__(self, Q, dt):
        maxiter = 20000 // self.gmres_restart
        t0 = time()
        Qnew, norm_r, norm_b, num_iter, flag, residuals = fgmres(
            self.A, self.b, x0=self.Q_flat, tol=self.tol, restart=self.gmres_restart, maxiter=maxiter,
            preconditioner=self.preconditioner, verbose=self.verbose_solver, ops=self.ops,
        )
        self.solver_info = SolverInfo(flag, time() - t0, num_iter, residuals)
        self.failure_flag = flag
        return Qnew.view(Q.shape)


_EPI_COEFS = {  # epi.py:44-61
    2: [[]],
    3: [[2 / 3]],
    4: [[-3 / 10, 3 / 40], [32 / 5, -11 / 10]],
    5: [[-4 / 5, 2 / 5, -4 / 45], [12, -9 / 2, 8 / 9], [3, 0, -1 / 3]],
    6: [[-49 / 60, 351 / 560, -359 / 1260, 367 / 6720], [92 / 7, -99 / 14, 176 / 63, -1 / 2], [485 / 21, -151 / 14, 23 / 9, -31 / 168]],
}


class Epi(Integrator):
    """Multistep exponential propagation iterative integrators of order 2..6 with KIOPS or PMEX (epi.py:30-363)."""

    def __init__(self, param, order, rhs, init_method=None, init_substeps=1, **kwargs):
        super().__init__(param, rhs, **kwargs)
        self.rhs = rhs
        # the reference's schema default is the complex-step product (epi.py:103-104 passes param.jacobian_method on); this
        # path provides the finite-difference one only: a Configuration handed over unchanged must not silently get another
        # Jacobian approximation
        method = getattr(param, "jacobian_method", "fd")
        if method != "fd":
            raise TypeError(f"jacobian_method = {method!r}: the B200 path implements the finite-difference JVP only (set jacobian_method = fd)")
        self.tol = param.tolerance
        self.krylov_size = 1
        self.exponential_solver = getattr(param, "exponential_solver", "kiops")
        if self.exponential_solver not in ("kiops", "pmex"):
            raise ValueError(f"Unrecognized exponential solver {self.exponential_solver} (the B200 path provides kiops and pmex)")
        if order not in _EPI_COEFS:
            raise ValueError(f"Unsupported order {order} for EPI method")
        self.A = _EPI_COEFS[order]
        k = len(self.A)
        self.n_prev = len(self.A[0])
        if order == 2:
            k -= 1
        self.max_phi = k + 1
        self.previous_Q, self.previous_rhs = deque(), deque()
        self.dt = 0.0
        if init_method or self.n_prev == 0:
            self.init_method = init_method
        else:
            self.init_method = Epi(param, 2, rhs, **{k: v for k, v in kwargs.items() if k != "filter"})  # the filter runs once, in the outer step
        self.init_substeps = init_substeps

    def __step__(self, Q, dt):
        o, f = self.ops, self.rhs_obj
        if self.dt and abs(self.dt - dt) > 1e-10:  # dt changed: restart the multistep history
            self.previous_Q, self.previous_rhs = deque(), deque()
        self.dt = dt
        if len(self.previous_Q) < self.n_prev:
            self.previous_Q.appendleft(Q)
            self.previous_rhs.appendleft(f(Q))
            dt /= self.init_substeps
            for _ in range(self.init_substeps):
                Q = self.init_method.step(Q, dt)
            return Q
        rhs = f(Q)
        matvec = MatvecOpFd(dt, Q, rhs, f)
        n = rhs.numel()
        vec = o.zeros(self.max_phi + 1, n)
        vec[1].copy_(rhs.reshape(-1))
        if self.n_prev:
            unit = MatvecOpFd(1.0, Q, rhs, f)
            for i in range(self.n_prev):
                # R(y_{n-i}) = (rhs_{n-i} - rhs) - J (Q_{n-i} - Q)   (epi.py:108-112)
                dq = o.combine((-1.0,), Q.reshape(1, -1), 1.0, self.previous_Q[i].reshape(-1))
                r = o.combine((-1.0, -1.0), _rows(o, rhs.reshape(-1), unit(dq)), 1.0, self.previous_rhs[i].reshape(-1))
                for k, row in enumerate(self.A, start=2):
                    o.combine((row[i],), r.view(1, -1), 1.0, vec[k], 1.0, out=vec[k])
        if self.exponential_solver == "pmex":  # epi.py:314-315: default m_init / mmin, no Krylov-size memory
            phiv, stats = pmex([1.0], matvec, vec, tol=self.tol, mmax=64, task1=False, ops=o)
        else:  # epi.py:325-338
            phiv, stats = kiops([1.0], matvec, vec, tol=self.tol, m_init=self.krylov_size, mmin=16, mmax=64, task1=False, ops=o)
            self.krylov_size = math.floor(0.7 * stats[5] + 0.3 * self.krylov_size)
        self.kiops_stats = stats
        self.solver_info = SolverInfo(total_num_it=stats[2])
        if self.n_prev > 0:
            self.previous_Q.pop()
            self.previous_Q.appendleft(Q)
            self.previous_rhs.pop()
            self.previous_rhs.appendleft(rhs)
        # Q + phiv*dt (epi.py:360)
        return o.lincomb3(1.0, Q.reshape(-1), 0.0, None, 1.0, phiv.reshape(-1), dt).view(Q.shape)


def _rows(ops, a, b):
    """Two vectors as the rows of one (2, n) array (for a two-term combination)."""
    m = ops.zeros(2, a.numel())
    m[0].copy_(a)
    m[1].copy_(b)
    return m

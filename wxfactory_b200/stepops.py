"""Per-step operators around the RHS, backed by csrc/stepops.cu through the C ABI.

* ``ExpFilterB200``: ``DFROperators.apply_filter_3d`` / ``apply_filters`` (wx_factory/geometry/operators.py:235-261).
* ``VecOps``: the vector algebra the integrators and Krylov solvers do on full-state arrays, on CUDA tensors
  (wx_factory/integrators/tvdrk3.py:13-17, ros2.py:30; wx_factory/solvers/fgmres.py:16-72, kiops.py:176-211,
  global_operations.py:15-34).  One object per (device, process group); every method enqueues on torch's current
  stream; only ``multidot`` / ``norm`` / ``nonfinite`` read a result back (the algorithms' own synchronisation points).

No CPU fallback: the methods raise without a CUDA device.  (tests/vecops_cpu.py is a CPU stand-in used to check
the solvers' host logic on CPU; it is test infrastructure and is never imported from here.)
"""

import ctypes
import math

import numpy

from . import capi


def _sp(torch, device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class VecOps:
    """Device vector algebra through libwxb200.so.  ``group``: torch.distributed process group over which dot
    products are summed (the reference's ``comm.allreduce``); None = single rank."""

    def __init__(self, device=0, group=None, world_size=1):
        import torch

        if not torch.cuda.is_available():
            raise RuntimeError("wxfactory_b200 needs a CUDA device: there is no CPU fallback")
        self.torch = torch
        self.lib = capi.load()
        self.device = torch.device("cuda", int(device))
        self.group, self.world_size = group, int(world_size)
        self._scratch = torch.empty(int(self.lib.wxb_vec_multidot_scratch(capi.WXB_VEC_MAX_ROWS, 2)), dtype=torch.float64, device=self.device)
        self._small = torch.empty(2 * capi.WXB_VEC_MAX_ROWS, dtype=torch.float64, device=self.device)
        self._count = torch.zeros(1, dtype=torch.int64, device=self.device)
        self.launches = 0

    max_rows = capi.WXB_VEC_MAX_ROWS

    # ---- allocation -------------------------------------------------------------------------------
    def zeros(self, *shape):
        return self.torch.zeros(*shape, dtype=self.torch.float64, device=self.device)

    def copy(self, x):
        return x.clone()

    def empty_like(self, x):
        return self.torch.empty_like(x)

    def _chk(self, code, what):
        if code != 0:
            raise (ValueError if code == -1 else capi.WxbError)(f"{what} failed ({code})")

    # ---- elementwise ------------------------------------------------------------------------------
    def lincomb3(self, a, x, b, y, c, z, dt, out=None):
        """((a x) + (b y)) + ((c z) dt), rounded operation by operation like the numpy expression; y may be None."""
        out = self.torch.empty_like(x) if out is None else out
        self._chk(
            self.lib.wxb_vec_lincomb3(x.numel(), float(a), x.data_ptr(), float(b), y.data_ptr() if y is not None else None, float(c),
                                      z.data_ptr(), float(dt), out.data_ptr(), _sp(self.torch, self.device)),
            "wxb_vec_lincomb3",
        )
        self.launches += 1
        return out

    def nonfinite(self, x):
        """Number of NaN/Inf entries (simulation.py:399-408); synchronises."""
        self._chk(self.lib.wxb_vec_nonfinite(x.numel(), x.data_ptr(), self._count.data_ptr(), _sp(self.torch, self.device)), "wxb_vec_nonfinite")
        self.launches += 1
        if self.world_size > 1:
            import torch.distributed as dist

            dist.all_reduce(self._count, group=self.group)
        return int(self._count.item())

    # ---- reductions -------------------------------------------------------------------------------
    def multidot(self, V, m, W, k):
        """(m, k) numpy array of V[i] . W[c] summed over ranks.  V: 2-D tensor (rows, N) whose first m rows are used,
        W: 2-D tensor (>= k rows, N) or 1-D (k = 1)."""
        n = V.shape[-1]
        ldw = W.stride(0) if W.dim() == 2 else n
        ldv = V.stride(0) if V.dim() == 2 else n
        if m * k > self._small.numel():  # PMEX reaches mmax + 1 rows (wx_factory/solvers/pmex.py:150-230)
            self._small = self.torch.empty(2 * m * k, dtype=self.torch.float64, device=self.device)
        out = self._small[: m * k]
        # the kernel takes up to WXB_VEC_MAX_ROWS rows per call: longer bases go in chunks of rows, each chunk's
        # (rows x k) block landing at its place in `out`
        for r0 in range(0, m, self.max_rows):
            mr = min(self.max_rows, m - r0)
            self._chk(
                self.lib.wxb_vec_multidot(n, mr, V.data_ptr() + 8 * r0 * ldv, ldv, k, W.data_ptr(), ldw, out[r0 * k :].data_ptr(),
                                          self._scratch.data_ptr(), _sp(self.torch, self.device)),
                "wxb_vec_multidot",
            )
            self.launches += (mr + 7) // 8 + 1  # MD_ROWS rows per pass + the final reduction (csrc/stepops.cu)
        if self.world_size > 1:
            import torch.distributed as dist

            dist.all_reduce(out, group=self.group)
        return out.cpu().numpy().reshape(m, k).copy()

    def dot(self, x, y):
        return float(self.multidot(x.view(1, -1), 1, y.view(1, -1), 1)[0, 0])

    def norm(self, x):
        """global_norm (wx_factory/solvers/global_operations.py:15-27)."""
        return float(numpy.sqrt(self.dot(x, x)))

    # ---- basis updates ----------------------------------------------------------------------------
    def combine(self, coef, V, b0=0.0, y=None, s=1.0, out=None):
        """out = (b0 y + sum_i coef[i] V[i]) * s."""
        coef = numpy.ascontiguousarray(coef, dtype=numpy.float64).reshape(-1)
        m = coef.size
        n = V.shape[-1] if m else y.numel()
        if out is None:
            out = self.torch.empty(n, dtype=self.torch.float64, device=self.device)
        if m > self.max_rows:  # long combinations (KIOPS with m up to 128) accumulate in chunks of rows
            k = self.max_rows
            self.combine(coef[:k], V[:k], b0, y, 1.0, out=out)
            for i in range(k, m, k):
                last = i + k >= m
                self.combine(coef[i : i + k], V[i : i + k], 1.0, out, s if last else 1.0, out=out)
            return out
        self._chk(
            self.lib.wxb_vec_combine(n, m, coef.ctypes.data if m else None, V.data_ptr() if m else None, V.stride(0) if m and V.dim() == 2 else n,
                                     float(b0), y.data_ptr() if y is not None else None, float(s), out.data_ptr(), _sp(self.torch, self.device)),
            "wxb_vec_combine",
        )
        self.launches += 1
        return out

    def _reduce(self, kind, x, y, slot):
        self._chk(
            self.lib.wxb_vec_reduce(x.numel(), kind, x.data_ptr(), y.data_ptr() if y is not None else None, self._small[slot : slot + 1].data_ptr(),
                                    self._scratch.data_ptr(), _sp(self.torch, self.device)),
            "wxb_vec_reduce",
        )
        self.launches += 2

    def asum_rows(self, U):
        """1-norms of the rows of U summed over ranks (kiops.py:111-116; once per call, on the few phi-function vectors)."""
        k = U.shape[0]
        for i in range(k):
            self._reduce(0, U[i], None, i)
        a = self._small[:k]
        if self.world_size > 1:
            import torch.distributed as dist

            dist.all_reduce(a, group=self.group)
        return a.cpu().numpy().copy()

    def max_rel_change(self, update, x):
        """max |update / x| over x != 0 (stagnation test of fgmres.py:257-262, once per restart cycle); the maximum over
        ranks, so that every rank takes the same decision (the reference takes the local one)."""
        self._reduce(1, x, update, 0)
        t = self._small[:1]
        if self.world_size > 1:
            import torch.distributed as dist

            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        loc = float(t.item())
        return loc if loc >= 0.0 else math.inf

    def scale(self, x, s, out=None):
        return self.combine((), x, 1.0, x, s, out=x if out is None else out)

    def gs1sync(self, Q, m, s2, s1, norm, r, q2, q1):
        """Rows j-2 (q2) and j-1 (q1) of fgmres.py:60-69 in one pass over Q[:m]."""
        s2 = numpy.ascontiguousarray(s2, dtype=numpy.float64)
        s1 = numpy.ascontiguousarray(s1, dtype=numpy.float64)
        self._chk(
            self.lib.wxb_vec_gs1sync(q2.numel(), m, s2.ctypes.data if m else None, s1.ctypes.data if m else None, Q.data_ptr() if m else None,
                                     Q.stride(0), float(norm), float(r), q2.data_ptr(), q1.data_ptr(), _sp(self.torch, self.device)),
            "wxb_vec_gs1sync",
        )
        self.launches += 1


class ExpFilterB200:
    """The per-step exponential filter of a 3-D Euler run on the device of ``rhs`` (an ``RhsB200``).

    ``expfilter``: the (n, n) nodal filter ``DFROperators.expfilter`` (operators.py:110-112).  Mirrors
    ``DFROperators.apply_filter_3d(Q, metric)`` (``metric`` is ignored: sqrtG lives on the handle) and
    ``apply_filters(Q, geom, metric, dt)`` (operators.py:235-261).  numpy in -> numpy out, CUDA tensor in -> CUDA
    tensor out, any shape of the right size."""

    def __init__(self, rhs, expfilter):
        self.rhs = rhs
        f = numpy.ascontiguousarray(expfilter, dtype=numpy.float64)
        n = rhs.prob.n
        if f.shape != (n, n):
            raise ValueError(f"expfilter must be ({n}, {n}), got {f.shape}")
        if rhs.prob.equations != "euler":
            raise TypeError("The 3D filter can only be applied on a CubedSphere3D geometry")  # operators.py:104
        self.expfilter = f
        rhs._check(rhs._lib.wxb_set_expfilter(rhs._h, f.ctypes.data), "wxb_set_expfilter")
        self.expfilter_apply = True
        self.launches = 0

    def apply_filter_3d(self, Q, metric=None):
        r = self.rhs
        if isinstance(Q, numpy.ndarray):
            torch = r._tm()
            a = r._as_host(Q, "Q")
            qd = torch.from_numpy(a.reshape(-1)).to(f"cuda:{r.device_index}", non_blocking=True)
            return r._to_host(self._apply(qd)).reshape(Q.shape)
        qd, shape = r._as_device(Q, "Q")
        return self._apply(qd).view(shape)

    def _apply(self, qd):
        r = self.rhs
        torch = r._tm()
        out = torch.empty_like(qd)
        r._check(r._lib.wxb_expfilter(r._h, qd.data_ptr(), out.data_ptr(), _sp(torch, qd.device)), "wxb_expfilter")
        self.launches += 1
        return out

    def apply_filters(self, Q, geom=None, metric=None, dt=None):
        return self.apply_filter_3d(Q, metric) if self.expfilter_apply else Q

    __call__ = apply_filter_3d

"""CPU stand-in for wxfactory_b200.stepops.VecOps -- TEST INFRASTRUCTURE ONLY (never imported by the package).

Same methods on torch CPU tensors with plain numpy arithmetic, so that the host logic of wxfactory_b200.krylov
(FGMRES, KIOPS) can be checked against the reference's golden results without a GPU.  The CUDA kernels themselves
are checked against this stand-in and against numpy in tests/test_gpu_stepops.py."""

import math

import numpy
import torch


class CpuVecOps:
    max_rows = 64
    launches = 0

    def zeros(self, *shape):
        return torch.zeros(*shape, dtype=torch.float64)

    def copy(self, x):
        return x.clone()

    def lincomb3(self, a, x, b, y, c, z, dt, out=None):
        s = x.numpy() if a == 1.0 else a * x.numpy()
        if y is not None:
            s = s + b * y.numpy()
        r = torch.from_numpy(s + (c * z.numpy()) * dt)
        if out is not None:
            out.copy_(r)
            return out
        return r

    def nonfinite(self, x):
        return int((~numpy.isfinite(x.numpy())).sum())

    def multidot(self, V, m, W, k):
        Wm = W.numpy().reshape(-1, V.shape[-1])[:k]
        return (V.numpy().reshape(-1, V.shape[-1])[:m] @ Wm.T).reshape(m, k)

    def dot(self, x, y):
        return float(x.numpy() @ y.numpy())

    def norm(self, x):
        return math.sqrt(self.dot(x, x))

    def combine(self, coef, V, b0=0.0, y=None, s=1.0, out=None):
        coef = numpy.asarray(coef, dtype=float).reshape(-1)
        m = coef.size
        acc = coef @ V.numpy().reshape(-1, V.shape[-1])[:m] if m else 0.0
        if y is not None:
            acc = b0 * y.numpy() + acc
        r = torch.from_numpy(numpy.asarray(acc * s, dtype=float))
        if out is not None:
            out.copy_(r)
            return out
        return r

    def scale(self, x, s, out=None):
        return self.combine((), None, 1.0, x, s, out=x if out is None else out)

    def asum_rows(self, U):
        return numpy.abs(U.numpy()).sum(axis=1)

    def max_rel_change(self, update, x):
        u, xx = update.numpy(), x.numpy()
        nz = xx != 0
        return float(numpy.max(numpy.abs(u[nz] / xx[nz]))) if nz.any() else math.inf

    def gs1sync(self, Q, m, s2, s1, norm, r, q2, q1):
        Qm = Q.numpy()[:m]
        a2 = q2.numpy() - (Qm.T @ numpy.asarray(s2) if m else 0.0)
        a1 = q1.numpy() - (Qm.T @ numpy.asarray(s1) if m else 0.0)
        a2 = a2 / norm
        q2.copy_(torch.from_numpy(a2))
        q1.copy_(torch.from_numpy((a1 - a2 * r) / norm))

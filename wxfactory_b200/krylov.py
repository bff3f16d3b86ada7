"""Device-resident Krylov solvers for the callers of the RHS / JVP.

``fgmres``, ``kiops`` and ``pmex`` keep the reference's signatures, return values and scalar control flow
(wx_factory/solvers/fgmres.py:98-281, wx_factory/solvers/kiops.py:10-346, wx_factory/solvers/pmex.py:8-374) so that ``Ros2`` and ``Epi`` can call
them unchanged; what differs is where the work happens: the Krylov bases live on the GPU and every full-length
operation is one launch of csrc/stepops.cu through ``VecOps`` (multi-dot with a fixed-order reduction, fused
Gram-Schmidt row updates, linear combinations).  Only the small projected matrices (R, T, K, H, Givens rotations,
``expm`` of H) are on the host, fed by the one device->host read per orthogonalisation the algorithms need anyway
(the reference's ``comm.allreduce``).

``ops`` is the vector backend (``stepops.VecOps`` on a CUDA device; the CPU tests pass a numpy stand-in to check the
host logic).  Vectors are 1-D arrays of the backend (CUDA float64 tensors).
"""

import math
from time import time

import numpy
import scipy.linalg

__all__ = ["fgmres", "kiops", "pmex"]


def _numpy_inf_nan(fn):
    """Run fn with numpy's divide / invalid warnings off: the adaptivity formulas of KIOPS / PMEX rely on inf and nan
    propagating like in the reference (e.g. log(0) when an error estimate vanishes) and must not change the caller's
    error state."""
    import functools

    @functools.wraps(fn)
    def wrapped(*args, **kwargs):
        with numpy.errstate(divide="ignore", invalid="ignore"):
            return fn(*args, **kwargs)

    return wrapped


# ------------------------------------------------------------------------------------------------------
# FGMRES with the one-synchronisation Gram-Schmidt
# ------------------------------------------------------------------------------------------------------
def _ortho_1sync(ops, V, R, T, K, j):
    """fgmres.py:16-72: make row j-1 of V orthogonal to rows < j-1 and normalise row j-2, with ONE reduction
    (V[:j] @ V[j-2:j].T) and one fused update pass.  R, T, K: host matrices carried between calls."""
    if j < 2:
        return -1.0
    g = ops.multidot(V, j, V[j - 2 : j], 2)  # (j, 2), summed over ranks
    s = g[: j - 2, 0].copy()
    R[: j - 1, j - 1] = g[: j - 1, 1]
    norm = math.sqrt(g[j - 2, 0] - s @ s)
    R[j - 2, j - 2] = norm
    R[j - 2, j - 1] -= s @ R[: j - 2, j - 1]
    R[j - 2, j - 1] /= norm
    T[: j - 2, j - 2] = s / norm
    if j > 2:
        Lp = numpy.tril(T[: j - 2, : j - 2].T, -1) + numpy.eye(j - 2)
        r3 = numpy.linalg.solve(Lp, s)
        R[: j - 2, j - 2] = K[: j - 2, j - 3] + r3
        K[: j - 1, j - 2] = (R[: j - 1, j - 1] - R[: j - 1, 1 : j - 1] @ r3) / norm
    else:
        K[: j - 1, j - 2] = R[: j - 1, j - 1] / norm
    # rows j-2 and j-1 in one pass over V[:j-2] (fgmres.py:60-69)
    ops.gs1sync(V, j - 2, s, R[: j - 2, j - 1].copy(), norm, R[j - 2, j - 1], V[j - 2], V[j - 1])
    return norm


def _rotg(a, b):
    """fgmres.py:75-95."""
    if abs(b) == 0:
        return 1.0, 0.0, a
    if abs(a) == 0:
        return 0.0, 1.0, b
    scl = min(abs(a), abs(b))
    sigma = math.copysign(1.0, a) if abs(a) > abs(b) else math.copysign(1.0, b)
    r = sigma * (scl * math.sqrt((a / scl) ** 2 + (b / scl) ** 2))
    return a / r, b / r, r


def fgmres(A, b, x0=None, tol=1e-5, restart=20, maxiter=None, preconditioner=None, hegedus=False, verbose=0, prefix="",
           device=None, ops=None):
    """Solve A x = b (fgmres.py:98-281).  Returns (x, norm_r, norm_b, niter, flag, residuals) like the reference.

    ``A`` / ``preconditioner``: callables on flat device vectors (e.g. ``MatvecOpRat`` of an ``RhsB200``)."""
    if ops is None:
        raise ValueError("fgmres needs a vector backend (stepops.VecOps)")
    n = int(b.shape[0])
    if n <= restart:
        raise ValueError("The b vector should be longer than the number of restart")
    if restart + 2 > ops.max_rows:
        raise ValueError(f"restart must be at most {ops.max_rows - 2}")
    t_start = time()
    niter = 0
    if maxiter is None:
        maxiter = n * 10
    x = ops.zeros(n) if x0 is None else ops.copy(x0)
    norm_b = ops.norm(b)
    if norm_b == 0.0:
        return ops.zeros(n), 0.0, 0.0, 0, 0, [(0.0, time() - t_start, 0.0)]
    tol_relative = tol * norm_b
    Ax0 = A(x)
    residuals = []
    if hegedus:
        d = ops.dot(Ax0, Ax0)
        if d != 0.0:
            x = ops.combine((), None, 1.0, x0, ops.dot(b, Ax0) / d)
            Ax0 = A(x)
    r = ops.combine((-1.0,), Ax0.view(1, -1), 1.0, b)  # b - A x
    norm_r = ops.norm(r)
    residuals.append((norm_r / norm_b, time() - t_start, 0.0))

    V = ops.zeros(restart + 2, n)  # Krylov basis, row-major (fgmres.py:189)
    Z = ops.zeros(restart + 1, n)  # preconditioned vectors
    for _outer in range(maxiter):
        H = numpy.zeros((restart + 2, restart + 2))
        R = numpy.zeros((restart + 2, restart + 2))
        T = numpy.zeros((restart + 2, restart + 2))
        K = numpy.zeros((restart + 2, restart + 2))
        rot = []
        V.zero_()
        ops.combine((), None, 1.0, r, 1.0 / norm_r, out=V[0])
        Z[0].copy_(V[0] if preconditioner is None else preconditioner(V[0]))
        V[1].copy_(A(Z[0]))
        v_norm = _ortho_1sync(ops, V, R, T, K, 2)
        g = numpy.zeros(restart + 2)
        g[0] = norm_r
        inner = 0
        for inner in range(restart):
            niter += 1
            Z[inner + 1].copy_(V[inner + 1] if preconditioner is None else preconditioner(V[inner + 1]))
            # lagged normalisation (fgmres.py:206): A(Z / v_norm) * v_norm
            t = A(ops.combine((), None, 1.0, Z[inner + 1], 1.0 / v_norm))
            ops.combine((), None, 1.0, t, v_norm, out=V[inner + 2])
            v_norm = _ortho_1sync(ops, V, R, T, K, inner + 3)
            H[inner, : inner + 2] = R[: inner + 2, inner + 1]
            ops.scale(Z[inner + 1], 1.0 / v_norm)
            for k in range(inner):  # previous Givens rotations (fgmres.py:213-214,283-303)
                H[inner, k : k + 2] = rot[k] @ H[inner, k : k + 2]
            if inner != n - 1 and H[inner, inner + 1] != 0:
                c, s, _ = _rotg(H[inner, inner], H[inner, inner + 1])
                Qb = numpy.array([[c, s], [-s, c]])
                rot.append(Qb)
                g[inner : inner + 2] = Qb @ g[inner : inner + 2]
                H[inner, inner] = Qb[0, :] @ H[inner, inner : inner + 2]
                H[inner, inner + 1] = 0.0
            if inner < restart - 1:
                norm_r = abs(g[inner + 1])
                residuals.append((norm_r / norm_b, time() - t_start, 0.0))
                if verbose > 1:
                    print(f"{prefix}norm_r / b = {residuals[-1][0]:.3e}", flush=True)
                if norm_r < tol_relative:
                    break
        y = scipy.linalg.solve_triangular(H[0 : inner + 1, 0 : inner + 1].T, g[0 : inner + 1])
        update = ops.combine(y, Z[: inner + 1])
        x = ops.combine((1.0,), update.view(1, -1), 1.0, x)
        r = ops.combine((-1.0,), A(x).view(1, -1), 1.0, b)
        norm_r = ops.norm(r)
        residuals.append((norm_r / norm_b, time() - t_start, 0.0))
        if verbose > 0:
            print(f"{prefix}res: {norm_r/norm_b:.2e} (iter {niter})", flush=True)
        if ops.max_rel_change(update, x) < 1e-12:  # stagnation (fgmres.py:257-262)
            return x, norm_r, norm_b, niter, -1, residuals
        if norm_r < tol_relative:
            return x, norm_r, norm_b, niter, 0, residuals
    return x, norm_r, norm_b, niter, (0 if norm_r < tol_relative else -1), residuals


# ------------------------------------------------------------------------------------------------------
# KIOPS
# ------------------------------------------------------------------------------------------------------
@_numpy_inf_nan
def kiops(tau_out, A, u, tol=1e-7, m_init=10, mmin=10, mmax=128, iop=2, task1=False, device=None, ops=None):
    """w[i] = phi_0(tau_i A) u[0] + phi_1(tau_i A) u[1] + ... (kiops.py:10-346), adaptive Krylov size and sub-step,
    incomplete orthogonalisation of length ``iop``.  ``u``: (p+1, n) device array.  Returns (w, stats) like the
    reference: stats = (substeps, rejected, Krylov steps, matrix exponentials, error estimate, last Krylov size).

    The augmented basis vectors are (n + p) long in the reference; here their first n entries are rows of a device
    matrix and the p-entry tails stay on the host (they take part in every product with their own tiny sums)."""
    if ops is None:
        raise ValueError("kiops needs a vector backend (stepops.VecOps)")
    tau_out = numpy.atleast_1d(numpy.asarray(tau_out, dtype=float))
    ppo, n = int(u.shape[0]), int(u.shape[1])
    p = ppo - 1
    if p == 0:
        # the reference pads a single row with `row_stack((u, zeros(len(u))))` (kiops.py:78-81, pmex.py:53-56), which
        # raises for any n > 1: same error here (the integrators always pass phi_0 and phi_1 rows)
        raise ValueError("all the input array dimensions except for the concatenation axis must match exactly: u needs at least two rows")
    m = max(mmin, min(m_init, mmax))
    Vn = ops.zeros(mmax + 1, n)           # first n entries of every basis vector
    Va = numpy.zeros((mmax + 1, p))       # the augmented tails
    H = numpy.zeros((mmax + 1, mmax + 1))
    step = krystep = ireject = reject = exps = 0
    sgn = float(numpy.sign(tau_out[-1]))
    tau_now, tau_end = 0.0, float(abs(tau_out[-1]))
    happy = False
    j = 0
    conv = 0.0
    nsteps = len(tau_out)
    w = ops.zeros(nsteps, n)
    w[0].copy_(u[0])
    normU = float(numpy.max(ops.asum_rows(u[1:]))) if u.shape[0] > 1 else 0.0  # kiops.py:111-116
    if ppo > 1 and normU > 0:
        ex = math.ceil(math.log2(normU))
        nu, mu = 2.0 ** (-ex), 2.0**ex
    else:
        nu = mu = 1.0
    u_flip = ops.zeros(p, n)  # nu * flipud(u[1:])
    for k in range(p):
        ops.combine((), None, 1.0, u[p - k], nu, out=u_flip[k])
    tau = tau_end
    gamma, gamma_mmax = (0.2, 0.1) if tau_end > 1 else (0.9, 0.6)
    delta = 1.4
    oldm, oldtau, omega = -1, math.nan, math.nan
    orderold = kestold = True
    order, kest = 1.0, 2.0
    l = 0
    beta = nrm = 0.0

    def dots(rows, jj):  # V[rows] . V[jj] over the full augmented length, summed over ranks
        cnt = rows.stop - rows.start
        return ops.multidot(Vn[rows], cnt, Vn[jj].view(1, -1), 1)[:, 0] + Va[rows] @ Va[jj]

    while tau_now < tau_end:
        if j == 0:  # starting vector of this sub-step (kiops.py:151-168)
            Vn[0].copy_(w[l])
            for k in range(p - 1):
                i = p - k + 1
                Va[0, k] = (tau_now**i) / math.factorial(i) * mu
            Va[0, p - 1] = mu
            beta = math.sqrt(dots(slice(0, 1), 0)[0])
            ops.scale(Vn[0], 1.0 / beta)
            Va[0] /= beta
        while j < m:  # incomplete orthogonalisation (kiops.py:171-212)
            j += 1
            t = A(Vn[j - 1])
            ops.combine(Va[j - 1], u_flip, 1.0, t, 1.0, out=Vn[j])  # augmented matrix-vector product
            Va[j, : p - 1] = Va[j - 1, 1:p]
            Va[j, p - 1] = 0.0
            rows = slice(max(0, j - iop), j)
            h = dots(rows, j)
            H[rows, j - 1] = h
            ops.combine(-h, Vn[rows], 1.0, Vn[j], 1.0, out=Vn[j])
            Va[j] -= Va[rows].T @ h
            nrm = math.sqrt(dots(slice(j, j + 1), j)[0])
            if nrm < tol:
                happy = True
                break
            H[j, j - 1] = nrm
            ops.scale(Vn[j], 1.0 / nrm)
            Va[j] /= nrm
            krystep += 1
        H[0, j] = 1.0  # phi_1 for the error estimate (kiops.py:215-226)
        nrm = float(H[j, j - 1])
        H[j, j - 1] = 0.0
        F = scipy.linalg.expm(sgn * tau * H[: j + 1, : j + 1])
        exps += 1
        H[j, j - 1] = nrm
        if happy:
            omega, err = 0.0, 0.0
            tau_new = min(tau_end - (tau_now + tau), tau)
            m_new = m
            happy = False
        else:
            err = abs(beta * nrm * F[j - 1, j])
            oldomega = omega
            omega = numpy.float64(tau_end * err / (tau * tol))  # numpy scalar: inf/nan instead of exceptions, like the reference
            # order and Krylov-size estimates (kiops.py:244-262)
            if m == oldm and tau != oldtau and ireject >= 1:
                order = max(1.0, float(numpy.log(omega / oldomega) / numpy.log(tau / oldtau)))
                orderold = False
            elif orderold or ireject == 0:
                orderold = True
                order = j / 4
            else:
                orderold = True
            if m != oldm and tau == oldtau and ireject >= 1:
                kest = max(1.1, (omega / oldomega) ** (1 / (oldm - m)))
                kestold = False
            elif kestold or ireject == 0:
                kestold = True
                kest = 2
            else:
                kestold = True
            remaining = tau_end - tau_now if omega > delta else tau_end - (tau_now + tau)
            same_tau = min(remaining, tau)
            tau_opt = tau * (gamma / omega) ** (1 / order)
            tau_opt = min(remaining, max(tau / 5, min(5 * tau, tau_opt)))
            m_opt = numpy.ceil(j + numpy.log(omega / gamma) / numpy.log(kest))
            m_opt = max(mmin, min(mmax, max(math.floor(3 / 4 * m), min(m_opt, math.ceil(4 / 3 * m)))))
            if j == mmax:
                if omega > delta:
                    m_new = j
                    tau_new = tau * (gamma_mmax / omega) ** (1 / order)
                    tau_new = min(tau_end - tau_now, max(tau / 5, tau_new))
                else:
                    tau_new, m_new = tau_opt, m
            else:
                m_new, tau_new = m_opt, same_tau
        if omega <= delta:  # accepted (kiops.py:297-329)
            reject += ireject
            step += 1
            blown = sum(1 for k in range(l, nsteps) if abs(tau_out[k]) < abs(tau_now + tau))
            if blown:
                w[l + blown].copy_(w[l])
                for k in range(blown):
                    F2 = scipy.linalg.expm(sgn * (tau_out[l + k] - tau_now) * H[:j, :j])
                    ops.combine(beta * F2[:j, 0], Vn[:j], out=w[l + k])
                l += blown
            ops.combine(beta * F[:j, 0], Vn[:j], out=w[l])
            tau_now += tau
            j = 0
            ireject = 0
            conv += err
        else:
            ireject += 1
            H[0, j] = 0.0
        oldtau, tau = tau, tau_new
        oldm, m = m, int(m_new)
    if task1:
        for k in range(nsteps):
            ops.scale(w[k], 1.0 / tau_out[k])
    return w, (step, reject, krystep, exps, conv, m)


# ------------------------------------------------------------------------------------------------------
# PMEX: KIOPS with a low-synchronisation (2-step Gauss-Seidel) orthogonalisation and a norm estimate
# ------------------------------------------------------------------------------------------------------
@_numpy_inf_nan
def pmex(tau_out, A, u, tol=1e-7, delta=1.2, m_init=10, mmin=10, mmax=128, reuse_info=True, task1=False, device=None, ops=None):
    """pmex.py:8-374: same phi-function problem and sub-stepping as ``kiops``; every Krylov step needs ONE reduction
    (V[:j+1] @ V[j-1:j+1].T, pmex.py:168-176) from which the projection coefficients and, when it is safe, the norm of
    the new vector (in extended precision, pmex.py:198-221) follow.  Returns (w, stats) with the reference's 7 stats:
    (substeps, rejected, Krylov steps, matrix exponentials, error estimate, last Krylov size, norms that needed their
    own reduction)."""
    if ops is None:
        raise ValueError("pmex needs a vector backend (stepops.VecOps)")
    tau_out = numpy.atleast_1d(numpy.asarray(tau_out, dtype=float))
    ppo, n = int(u.shape[0]), int(u.shape[1])
    p = ppo - 1
    if p == 0:
        # the reference pads a single row with `row_stack((u, zeros(len(u))))` (kiops.py:78-81, pmex.py:53-56), which
        # raises for any n > 1: same error here (the integrators always pass phi_0 and phi_1 rows)
        raise ValueError("all the input array dimensions except for the concatenation axis must match exactly: u needs at least two rows")
    step = krystep = ireject = reject = exps = reg_comm_nrm = 0
    sgn = float(numpy.sign(tau_out[-1]))
    tau_now, tau_end = 0.0, float(abs(tau_out[-1]))
    happy = False
    j = 0
    conv = 0.0
    nsteps = len(tau_out)
    m = max(mmin, min(m_init, mmax))
    Vn = ops.zeros(mmax + 1, n)
    Va = numpy.zeros((mmax + 1, p))
    H = numpy.zeros((mmax + 1, mmax + 1))
    Minv, M, N = numpy.eye(mmax), numpy.eye(mmax), numpy.zeros((mmax, mmax))
    w = ops.zeros(nsteps, n)
    w[0].copy_(u[0])
    normU = float(numpy.max(ops.asum_rows(u[1:]))) if u.shape[0] > 1 else 0.0
    if ppo > 1 and normU > 0:
        ex = math.ceil(math.log2(normU))
        nu, mu = 2.0 ** (-ex), 2.0**ex
    else:
        nu = mu = 1.0
    u_flip = ops.zeros(p, n)
    for k in range(p):
        ops.combine((), None, 1.0, u[p - k], nu, out=u_flip[k])
    tau = tau_end
    gamma, gamma_mmax = (0.2, 0.1) if tau_end > 1 else (0.9, 0.6)
    old_m, old_tau, ohm = -1, math.nan, math.nan
    kestold = True
    kest = 2.0
    l = 0
    beta = nrm = 0.0
    while tau_now < tau_end:
        if j == 0:
            H[:, :] = 0.0
            Vn[0].copy_(w[l])
            for k in range(p - 1):
                i = p - k + 1
                Va[0, k] = (tau_now**i) / math.factorial(i) * mu
            Va[0, p - 1] = mu
            beta = math.sqrt(ops.dot(Vn[0], Vn[0]) + Va[0] @ Va[0])
            ops.scale(Vn[0], 1.0 / beta)
            Va[0] /= beta
        while j < m:
            j += 1
            t = A(Vn[j - 1])
            ops.combine(Va[j - 1], u_flip, 1.0, t, 1.0, out=Vn[j])
            Va[j, : p - 1] = Va[j - 1, 1:p]
            Va[j, p - 1] = 0.0
            # the one reduction of the step (pmex.py:168-176)
            g = ops.multidot(Vn, j + 1, Vn[j - 1 : j + 1], 2) + Va[: j + 1] @ Va[j - 1 : j + 1].T
            if j > 1:
                M[j - 1, : j - 1] = g[: j - 1, 0]
                N[: j - 1, j - 1] = -g[: j - 1, 0]
                Minv[j - 1, : j - 1] = -g[: j - 1, 0].T @ Minv[: j - 1, : j - 1]
            rhs = (numpy.eye(j) + N[:j, :j] @ Minv[:j, :j]) @ g[:j, 1]
            # same call as the reference (pmex.py:190-192), including its default `lower=False`
            sol = scipy.linalg.solve_triangular(M[:j, :j], rhs, unit_diagonal=True, check_finite=False)
            ops.combine(-sol, Vn[:j], 1.0, Vn[j], 1.0, out=Vn[j])
            Va[j] -= sol @ Va[:j]
            sum_sqrd = numpy.sum(numpy.array(g[:j, 1], numpy.longdouble) ** 2)  # extended precision (pmex.py:198-200)
            if g[-1, 1] < sum_sqrd:
                curr_nrm = math.sqrt(ops.dot(Vn[j], Vn[j]) + Va[j] @ Va[j])
                reg_comm_nrm += 1
            else:
                curr_nrm = float(numpy.sqrt(g[-1, 1] - sum_sqrd))
            if curr_nrm < tol:
                happy = True
                break
            ops.scale(Vn[j], 1.0 / curr_nrm)
            Va[j] /= curr_nrm
            H[j, j - 1] = curr_nrm
            H[:j, j - 1] = sol
            krystep += 1
        H[0, j] = 1.0
        nrm = float(H[j, j - 1])
        H[j, j - 1] = 0.0
        F_half = scipy.linalg.expm(sgn * 0.5 * tau * H[: j + 1, : j + 1])
        F = F_half @ F_half
        exps += 1
        H[j, j - 1] = nrm
        if happy:
            ohm, err = 0.0, 0.0
            tau_new = min(tau_end - (tau_now + tau), tau)
            m_new = m
            happy = False
        else:
            err_half = abs(beta * nrm * F_half[j - 1, j])
            err = abs(beta * nrm * F[j - 1, j])
            old_ohm = ohm
            ohm = numpy.float64(tau_end * err / (tau * tol))
            if not numpy.isfinite(ohm) or ohm <= 0.0:
                # pmex.py:283-300 uses math.log / math.ceil here, which raise on a vanishing, infinite or NaN error estimate;
                # numpy would carry the NaN into the Krylov size instead
                raise ValueError(f"pmex: error estimate {float(err)!r} gives no usable step-size control (omega = {float(ohm)!r})")
            order = float(numpy.log(numpy.float64(err) / err_half) / math.log(2))
            if m != old_m and tau == old_tau and ireject >= 1:
                kest = max(1.1, (ohm / old_ohm) ** (1 / (old_m - m)))
                kestold = False
            elif kestold or ireject == 0:
                kest = 2
                kestold = True
            else:
                kestold = True
            remaining = tau_end - tau_now if ohm > delta else tau_end - (tau_now + tau)
            same_tau = min(remaining, tau)
            tau_opt = tau * (gamma / ohm) ** (1 / order)
            tau_opt = min(remaining, max(tau / 5, min(5 * tau, tau_opt)))
            m_opt = numpy.ceil(j + numpy.log(ohm / gamma) / numpy.log(kest))
            m_opt = max(mmin, min(mmax, max(math.floor(3 / 4 * m), min(m_opt, math.ceil(4 / 3 * m)))))
            if j == mmax:
                if ohm > delta:
                    m_new = j
                    tau_new = tau * (gamma_mmax / ohm) ** (1 / order)
                    tau_new = min(tau_end - tau_now, max(tau / 5, tau_new))
                else:
                    tau_new, m_new = tau_opt, m
            else:
                m_new = m if same_tau < tau else m_opt  # tau was cut to land on the end: keep m (pmex.py:301-305)
                tau_new = same_tau
        if ohm <= delta:
            reject += ireject
            step += 1
            blown = sum(1 for k in range(l, nsteps) if abs(tau_out[k]) < abs(tau_now + tau))
            if blown:
                w[l + blown].copy_(w[l])
                for k in range(blown):
                    F2 = scipy.linalg.expm(sgn * (tau_out[l + k] - tau_now) * H[:j, :j])
                    ops.combine(beta * F2[:j, 0], Vn[:j], out=w[l + k])
                l += blown
            ops.combine(beta * F[:j, 0], Vn[:j], out=w[l])
            tau_now += tau
            j = 0
            ireject = 0
            conv += err
        else:
            ireject += 1
            H[0, j] = 0.0
        old_tau, tau = tau, tau_new
        old_m, m = m, int(m_new)
    if task1:
        for k in range(nsteps):
            ops.scale(w[k], 1.0 / tau_out[k])
    return w, (step, reject, krystep, exps, conv, m, reg_comm_nrm)

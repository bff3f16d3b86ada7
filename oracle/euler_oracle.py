"""ORACLE (test infrastructure, CPU, numpy) -- 3-D Euler DFR right-hand side on the cubed sphere.

A restatement of the reference's algorithm for this path, phase by phase, with the same
dense Kronecker operator GEMMs and numpy ufuncs the reference uses, all 6 panels in one
process (the 6 MPI ranks of the reference's minimum layout):

    RHS.__call__                      wx_factory/rhs/rhs.py:75-122
    solution_extrapolation            wx_factory/rhs/rhs_dfr.py:50-71
    start/end_communication           wx_factory/rhs/rhs_dfr.py:141-201  (-> cubesphere_exchange.py)
    pointwise_fluxes                  wx_factory/pde/pde_euler_cubesphere.py:72-124
    flux_divergence_partial           wx_factory/rhs/rhs_dfr.py:25-34,89-104
    riemann_fluxes                    wx_factory/rhs/rhs_dfr.py:203-302,
                                      wx_factory/pde/pde_euler_cubesphere.py:126-201,
                                      wx_factory/pde/fluxes.py:150-222,326-403,507-582
    flux_divergence                   wx_factory/rhs/rhs_dfr.py:106-139
    forcing_terms                     wx_factory/pde/pde_euler_cubesphere.py:203-290,
                                      wx_factory/init/dcmip.py:676-755
    operators (kron expansion)        wx_factory/geometry/operators.py:161-183,79-80

It is the checker, never the product: only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import it.  Parity status: PINNED against the
reference itself run in this container (tests/golden/*.npz made by oracle/refharness/make_golden.py;
test_oracle_vs_golden.py), since the reference holds no golden vector for this path.
"""

import numpy

from . import cubesphere_exchange as cx

# wx_factory/common/definitions.py:5-12,22-26
p0 = 100000.0
Rd = 287.05
cpd = 1005.46
cvd = cpd - Rd
gravity = 9.80616
heat_capacity_ratio = cpd / cvd
idx_rho, idx_rho_u1, idx_rho_u2, idx_rho_w, idx_rho_theta = 0, 1, 2, 3, 4


class EulerOracle:
    def __init__(self, prob):
        self.p = prob
        n, P, F = prob.n, prob.P, prob.F
        self.n, self.P, self.F = n, P, F
        self.nz, self.nh = prob.nz, prob.nh
        I2, I3 = numpy.identity(n), numpy.identity(n * n)
        kron, vstack = numpy.kron, numpy.vstack
        # operators.py:161-165
        self.extrap_x = vstack((kron(I3, prob.ext_w), kron(I3, prob.ext_e))).T
        self.extrap_y = vstack((kron(I2, kron(prob.ext_w, I2)), kron(I2, kron(prob.ext_e, I2)))).T
        self.extrap_z = vstack((kron(prob.ext_w, I3), kron(prob.ext_e, I3))).T
        # operators.py:167-169
        self.derivative_x = kron(I3, prob.diff).T
        self.derivative_y = kron(I2, kron(prob.diff, I2)).T
        self.derivative_z = kron(prob.diff, I3).T
        # operators.py:171-183
        cw, ce = prob.corr[:, 0], prob.corr[:, 1]
        self.correction_WE = vstack((kron(I3, cw), kron(I3, ce)))
        self.correction_SN = vstack((kron(I2, kron(cw, I2)), kron(I2, kron(ce, I2))))
        self.correction_DU = vstack((kron(cw, I3), kron(ce, I3)))
        # operators.py:79-80
        self.highfilter_k = kron(prob.hfilt.T, I3)

        nz, nh = self.nz, self.nh
        sh = (6, nz, nh, nh, P)
        self.sqrtG = prob.sqrtg.reshape(sh)
        self.inv_sqrtG = 1.0 / self.sqrtG  # metric3d.py:1157
        self.h = prob.h.reshape((6, 6) + sh[1:])
        self.chris = prob.chris.reshape((6, 27) + sh[1:])
        self.inv_dzdeta = prob.inv_dzdeta.reshape(sh)
        self.damp = None if prob.damp is None else prob.damp.reshape((6, 4) + sh[1:])
        # two-sided, halo-padded interface metric as the reference holds it
        # (cubed_sphere_3d.py:191-194,795-897): outer ghost faces are 0
        self.fm_i = self._two_sided(prob.fm_i, -2)  # (6,4,nz,nh,nh+2,2F)
        self.fm_j = self._two_sided(prob.fm_j, -3)
        self.fm_k = self._two_sided(prob.fm_k, -4)
        # edge coordinates in the (n_h, n, n) halo_side_shape (cubed_sphere_3d.py:197-204)
        self.bsn = [numpy.tile(prob.bnd_sn[p].reshape(nh, n), n).reshape(nh, n, n) for p in range(6)]
        self.bwe = [numpy.tile(prob.bnd_we[p].reshape(nh, n), n).reshape(nh, n, n) for p in range(6)]

    def _two_sided(self, fm, axis):
        F = self.F
        a = numpy.moveaxis(fm, axis, -2)  # (..., ne+1, F)
        ne = a.shape[-2] - 1
        out = numpy.zeros(a.shape[:-2] + (ne + 2, 2 * F))
        out[..., 1:, :F] = a
        out[..., :-1, F:] = a
        return numpy.ascontiguousarray(numpy.moveaxis(out, -2, axis))

    # ------------------------------------------------------------------------------------------
    def _exchange(self, q_itf_x1, q_itf_x2):
        """rhs_dfr.py:141-201: returns q_itf_s, q_itf_n (5,6,nz,nx,F) and q_itf_w, q_itf_e (5,6,nz,ny,F)."""
        F, n, nz, nh = self.F, self.n, self.nz, self.nh
        base = (nz, nh, n, n)  # get_base_shape(..., halo_side_shape), process_topology.py:542-561
        flip = (-3, -1)

        def edges_of(v):
            return [
                [
                    q_itf_x2[v][p][:, 0, :, :F].reshape(base),
                    q_itf_x2[v][p][:, -1, :, F:].reshape(base),
                    q_itf_x1[v][p][:, :, 0, :F].reshape(base),
                    q_itf_x1[v][p][:, :, -1, F:].reshape(base),
                ]
                for p in range(6)
            ]

        out = [numpy.empty((5, 6, nz, nh, F)) for _ in range(4)]  # s, n, w, e
        for v in (idx_rho, idx_rho_theta):
            recv = cx.exchange_scalars(edges_of(v), flip)
            for p in range(6):
                for e in range(4):
                    out[e][v, p] = recv[p][e].reshape(nz, nh, F)
        e1, e2, e3 = edges_of(idx_rho_u1), edges_of(idx_rho_u2), edges_of(idx_rho_w)
        vec = [[(e1[p][e], e2[p][e], e3[p][e]) for e in range(4)] for p in range(6)]
        coords = [[self.bsn[p], self.bsn[p], self.bwe[p], self.bwe[p]] for p in range(6)]
        recv = cx.exchange_vectors(vec, coords, flip)
        for p in range(6):
            for e in range(4):
                for c, v in enumerate((idx_rho_u1, idx_rho_u2, idx_rho_w)):
                    out[e][v, p] = recv[p][e][c].reshape(nz, nh, F)
        return out

    # ------------------------------------------------------------------------------------------
    def _rusanov(self, q, u, pres, fm, d, axis):
        """fluxes.py:150-222 (d=2), :326-403 (d=0), :507-582 (d=1), written once: `axis` is the
        (negative) position of the direction's padded element axis in the field arrays."""
        F = self.F
        mv = lambda a: numpy.moveaxis(a, axis, -2)  # noqa: E731  -> (..., ne+2, 2F)
        q, u, pres, fm = mv(q), mv(u), mv(pres), mv(fm)
        hi = numpy.s_[..., :-1, F:]  # "east / north(elem) / down": max-side face of element e
        lo = numpy.s_[..., 1:, :F]  # "west / south(elem) / up":   min-side face of element e+1
        sqrtG, h_d0, h_d1, h_d2 = fm[:, 0], fm[:, 1], fm[:, 2], fm[:, 3]
        h_dd = fm[:, 1 + d]
        u_l, u_r = u[hi], u[lo]
        if self.p.advection_only:
            eig_l, eig_r = numpy.abs(u_l), numpy.abs(u_r)
        else:
            eig_l = numpy.abs(u_l) + numpy.sqrt(h_dd[hi] * heat_capacity_ratio * pres[hi] / q[idx_rho][hi])
            eig_r = numpy.abs(u_r) + numpy.sqrt(h_dd[lo] * heat_capacity_ratio * pres[lo] / q[idx_rho][lo])
        eig = numpy.maximum(eig_l, eig_r)

        flux_l = sqrtG[hi] * u_l * q[hi]
        flux_r = sqrtG[lo] * u_r * q[lo]
        wadv_l = flux_l[idx_rho_w].copy()
        wadv_r = flux_r[idx_rho_w].copy()
        flux_l[idx_rho_u1] += sqrtG[hi] * h_d0[hi] * pres[hi]
        flux_l[idx_rho_u2] += sqrtG[hi] * h_d1[hi] * pres[hi]
        flux_l[idx_rho_w] += sqrtG[hi] * h_d2[hi] * pres[hi]
        flux_r[idx_rho_u1] += sqrtG[lo] * h_d0[lo] * pres[lo]
        flux_r[idx_rho_u2] += sqrtG[lo] * h_d1[lo] * pres[lo]
        flux_r[idx_rho_w] += sqrtG[lo] * h_d2[lo] * pres[lo]
        wpres_l = sqrtG[hi] * h_d2[hi] * pres[hi]
        wpres_r = sqrtG[lo] * h_d2[lo] * pres[lo]

        flux = numpy.zeros_like(q)
        wadv = numpy.zeros_like(pres)
        wpres = numpy.zeros_like(pres)
        flux[hi] = 0.5 * (flux_l + flux_r - eig * sqrtG[hi] * (q[lo] - q[hi]))
        flux[lo] = flux[hi]
        wadv[hi] = 0.5 * (wadv_l + wadv_r - eig * sqrtG[hi] * (q[idx_rho_w][lo] - q[idx_rho_w][hi]))
        wadv[lo] = wadv[hi]
        wpres[hi] = 0.5 * (wpres_l + wpres_r) / pres[hi]
        wpres[lo] = 0.5 * (wpres_l + wpres_r) / pres[lo]
        # interior (non-ghost) elements only, back in the original axis order (rhs_dfr.py:289-302)
        back = lambda a: numpy.moveaxis(a[..., 1:-1, :], -2, axis)  # noqa: E731
        return back(flux), back(wadv), back(wpres)

    # ------------------------------------------------------------------------------------------
    def rhs(self, Q):
        """Q: (6, 5, nz, ny, nx, P) (or any shape of that size) -> dQ/dt in the same shape."""
        given_shape = Q.shape
        F, P, nz, nh = self.F, self.P, self.nz, self.nh
        q = numpy.moveaxis(numpy.asarray(Q, dtype=float).reshape(self.p.state_shape), 1, 0)  # (5,6,nz,ny,nx,P)
        sqrtG, h = self.sqrtG, self.h
        H = lambda k: h[:, k]  # noqa: E731   H_PAIRS: 0=11 1=12 2=13 3=22 4=23 5=33

        # 1. extrapolation (rhs_dfr.py:50-71)
        q_itf_x1 = q @ self.extrap_x
        q_itf_x2 = q @ self.extrap_y
        q_itf_x3 = q @ self.extrap_z
        log_rho = numpy.log(q[idx_rho])
        log_rt = numpy.log(q[idx_rho_theta])
        q_itf_x1[idx_rho] = numpy.exp(log_rho @ self.extrap_x)
        q_itf_x1[idx_rho_theta] = numpy.exp(log_rt @ self.extrap_x)
        q_itf_x2[idx_rho] = numpy.exp(log_rho @ self.extrap_y)
        q_itf_x2[idx_rho_theta] = numpy.exp(log_rt @ self.extrap_y)
        q_itf_x3[idx_rho] = numpy.exp(log_rho @ self.extrap_z)
        q_itf_x3[idx_rho_theta] = numpy.exp(log_rt @ self.extrap_z)

        # 2./5. halo (rhs_dfr.py:141-201)
        q_itf_s, q_itf_n, q_itf_w, q_itf_e = self._exchange(q_itf_x1, q_itf_x2)

        # 3. pointwise fluxes (pde_euler_cubesphere.py:72-124)
        rho = q[idx_rho]
        u1 = q[idx_rho_u1] / rho
        u2 = q[idx_rho_u2] / rho
        w = q[idx_rho_w] / rho
        f_x1 = sqrtG * u1 * q
        f_x2 = sqrtG * u2 * q
        f_x3 = sqrtG * w * q
        wadv_x1 = sqrtG * u1 * q[idx_rho_w]
        wadv_x2 = sqrtG * u2 * q[idx_rho_w]
        wadv_x3 = sqrtG * w * q[idx_rho_w]
        pressure = p0 * numpy.exp((cpd / cvd) * numpy.log((Rd / p0) * q[idx_rho_theta]))
        f_x1[idx_rho_u1] += sqrtG * H(0) * pressure
        f_x1[idx_rho_u2] += sqrtG * H(1) * pressure
        f_x1[idx_rho_w] += sqrtG * H(2) * pressure
        wpres_x1 = sqrtG * H(2)
        f_x2[idx_rho_u1] += sqrtG * H(1) * pressure
        f_x2[idx_rho_u2] += sqrtG * H(3) * pressure
        f_x2[idx_rho_w] += sqrtG * H(4) * pressure
        wpres_x2 = sqrtG * H(4)
        f_x3[idx_rho_u1] += sqrtG * H(2) * pressure
        f_x3[idx_rho_u2] += sqrtG * H(4) * pressure
        f_x3[idx_rho_w] += sqrtG * H(5) * pressure
        wpres_x3 = sqrtG * H(5)
        logp = numpy.log(pressure)

        # 4. interior derivative (rhs_dfr.py:25-34,89-104)
        Dx, Dy, Dz = self.derivative_x, self.derivative_y, self.derivative_z
        df1, df2, df3 = f_x1 @ Dx, f_x2 @ Dy, f_x3 @ Dz
        w1_adv, w1_pa, w1_pb = wadv_x1 @ Dx, wpres_x1 @ Dx, logp @ Dx
        w2_adv, w2_pa, w2_pb = wadv_x2 @ Dy, wpres_x2 @ Dy, logp @ Dy
        w3_adv, w3_pa, w3_pb = wadv_x3 @ Dz, wpres_x3 @ Dz, logp @ Dz

        # 6. Riemann (rhs_dfr.py:203-302)
        full_x1 = numpy.ones((5, 6, nz, nh, nh + 2, 2 * F))
        full_x2 = numpy.ones((5, 6, nz, nh + 2, nh, 2 * F))
        full_x3 = numpy.ones((5, 6, nz + 2, nh, nh, 2 * F))
        full_x1[..., 1:-1, :] = q_itf_x1
        full_x2[..., 1:-1, :, :] = q_itf_x2
        full_x3[..., 1:-1, :, :, :] = q_itf_x3
        full_x1[..., 0, F:] = q_itf_w
        full_x1[..., -1, :F] = q_itf_e
        full_x2[..., 0, :, F:] = q_itf_s
        full_x2[..., -1, :, :F] = q_itf_n
        full_x3[..., 0, :, :, F:] = full_x3[..., 1, :, :, :F]
        full_x3[..., 0, :, :, :F] = full_x3[..., 0, :, :, F:]
        full_x3[..., -1, :, :, :F] = full_x3[..., -2, :, :, F:]
        full_x3[..., -1, :, :, F:] = full_x3[..., -1, :, :, :F]

        # pde_euler_cubesphere.py:145-160
        u1_itf = full_x1[idx_rho_u1] / full_x1[idx_rho]
        u2_itf = full_x2[idx_rho_u2] / full_x2[idx_rho]
        w_itf = full_x3[idx_rho_w] / full_x3[idx_rho]
        w_itf[..., 0, :, :, :F] = 0.0
        w_itf[..., 0, :, :, F:] = -w_itf[..., 1, :, :, :F]
        w_itf[..., -1, :, :, F:] = 0.0
        w_itf[..., -1, :, :, :F] = -w_itf[..., -2, :, :, F:]
        p_x1 = p0 * numpy.exp((cpd / cvd) * numpy.log(full_x1[idx_rho_theta] * (Rd / p0)))
        p_x2 = p0 * numpy.exp((cpd / cvd) * numpy.log(full_x2[idx_rho_theta] * (Rd / p0)))
        p_x3 = p0 * numpy.exp((cpd / cvd) * numpy.log(full_x3[idx_rho_theta] * (Rd / p0)))

        f_itf_x1, wadv_itf_x1, wpres_itf_x1 = self._rusanov(full_x1, u1_itf, p_x1, self.fm_i, 0, -2)
        f_itf_x2, wadv_itf_x2, wpres_itf_x2 = self._rusanov(full_x2, u2_itf, p_x2, self.fm_j, 1, -3)
        f_itf_x3, wadv_itf_x3, wpres_itf_x3 = self._rusanov(full_x3, w_itf, p_x3, self.fm_k, 2, -4)
        p_itf_x1 = p_x1[..., 1:-1, :]
        p_itf_x2 = p_x2[..., 1:-1, :, :]
        p_itf_x3 = p_x3[..., 1:-1, :, :, :]

        # 7. correction + assembly (rhs_dfr.py:106-139)
        WE, SN, DU = self.correction_WE, self.correction_SN, self.correction_DU
        df1 += f_itf_x1 @ WE
        df2 += f_itf_x2 @ SN
        df3 += f_itf_x3 @ DU
        logp_i, logp_j, logp_k = numpy.log(p_itf_x1), numpy.log(p_itf_x2), numpy.log(p_itf_x3)

        w1_adv += wadv_itf_x1 @ WE
        w1_pa += wpres_itf_x1 @ WE
        w1_pa *= pressure
        w1_pb += logp_i @ WE
        w1_pb *= pressure * wpres_x1
        w_df1 = w1_adv + w1_pa + w1_pb
        w2_adv += wadv_itf_x2 @ SN
        w2_pa += wpres_itf_x2 @ SN
        w2_pa *= pressure
        w2_pb += logp_j @ SN
        w2_pb *= pressure * wpres_x2
        w_df2 = w2_adv + w2_pa + w2_pb
        w3_adv += wadv_itf_x3 @ DU
        w3_pa += wpres_itf_x3 @ DU
        w3_pa *= pressure
        w3_pb += logp_k @ DU
        w3_pb *= pressure * wpres_x3
        w_df3 = w3_adv + w3_pa + w3_pb

        rhs = -self.inv_sqrtG * (df1 + df2 + df3)
        rhs[idx_rho_w] = -self.inv_sqrtG * (w_df1 + w_df2 + w_df3)

        # 8. forcing (pde_euler_cubesphere.py:203-290)
        forcing = numpy.zeros_like(rhs)
        c = self.chris
        for i, row in enumerate((idx_rho_u1, idx_rho_u2, idx_rho_w)):
            c01, c02, c03, c11, c12, c13, c22, c23, c33 = (c[:, 9 * i + k] for k in range(9))
            forcing[row] = (
                2.0 * rho * (c01 * u1 + c02 * u2 + c03 * w)
                + c11 * (rho * u1 * u1 + H(0) * pressure)
                + 2.0 * c12 * (rho * u1 * u2 + H(1) * pressure)
                + 2.0 * c13 * (rho * u1 * w + H(2) * pressure)
                + c22 * (rho * u2 * u2 + H(3) * pressure)
                + 2.0 * c23 * (rho * u2 * w + H(4) * pressure)
                + c33 * (rho * w * w + H(5) * pressure)
            )
        forcing[idx_rho_w] += self.inv_dzdeta * gravity * self.inv_sqrtG * ((sqrtG * rho) @ self.highfilter_k)
        if self.damp is not None:  # dcmip.py:676-755
            dw = rho / self.p.damp_tau0 * self.damp[:, 0]
            forcing[idx_rho_u1] += dw * (u1 - self.damp[:, 1])
            forcing[idx_rho_u2] += dw * (u2 - self.damp[:, 2])
            forcing[idx_rho_w] += dw * (w - self.damp[:, 3])
        rhs -= forcing
        if self.p.advection_only:  # rhs_dfr.py:307-313
            rhs[...] = 0.0
        return numpy.ascontiguousarray(numpy.moveaxis(rhs, 0, 1)).reshape(given_shape)

    __call__ = rhs

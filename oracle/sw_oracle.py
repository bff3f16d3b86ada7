"""ORACLE (test infrastructure, CPU, numpy) -- shallow-water DFR right-hand side on the cubed sphere.

Restates ``RhsShallowWater.__compute_rhs__`` (wx_factory/rhs/rhs_sw.py:58-240) for all 6 panels
in one process, with the reference's dense Kronecker operators (wx_factory/geometry/operators.py:191-208)
and the halo of wx_factory/process_topology.py (-> cubesphere_exchange.py).

Checker only (see euler_oracle.py header).  Parity status: PINNED against the reference itself
run in this container (tests/golden/case6*.npz, galewsky*.npz).
"""

import numpy

from . import cubesphere_exchange as cx

gravity = 9.80616  # wx_factory/common/definitions.py:12
idx_h, idx_hu1, idx_hu2 = 0, 1, 2  # definitions.py:15-19


class SwOracle:
    def __init__(self, prob):
        self.p = prob
        n, nh = prob.n, prob.nh
        self.n, self.nh = n, nh
        ident = numpy.identity(n)
        kron, vstack = numpy.kron, numpy.vstack
        # operators.py:191-208 (2-D branch)
        self.extrap_x = vstack((kron(ident, prob.ext_w), kron(ident, prob.ext_e))).T
        self.extrap_y = vstack((kron(prob.ext_w, ident), kron(prob.ext_e, ident))).T
        self.derivative_x = kron(ident, prob.diff).T
        self.derivative_y = kron(prob.diff, ident).T
        cw, ce = prob.corr[:, 0], prob.corr[:, 1]
        self.correction_SN = vstack((kron(cw, ident), kron(ce, ident)))
        self.correction_WE = vstack((kron(ident, cw), kron(ident, ce)))
        sh = (6, nh, nh, n * n)
        self.sqrtG = prob.sqrtg.reshape(sh)
        self.inv_sqrtG = 1.0 / self.sqrtG
        self.H11, self.H12, self.H22 = (prob.h[:, k].reshape(sh) for k in range(3))
        self.chr = [prob.chris[:, k].reshape(sh) for k in range(8)]
        self.fm_i = self._two_sided(prob.fm_i, -2)  # (6,3,nh,nh+2,2n)
        self.fm_j = self._two_sided(prob.fm_j, -3)  # (6,3,nh+2,nh,2n)

    def _two_sided(self, fm, axis):
        n = self.n
        a = numpy.moveaxis(fm, axis, -2)
        ne = a.shape[-2] - 1
        out = numpy.zeros(a.shape[:-2] + (ne + 2, 2 * n))
        out[..., 1:, :n] = a
        out[..., :-1, n:] = a
        return numpy.ascontiguousarray(numpy.moveaxis(out, -2, axis))

    def rhs(self, Qin):
        given_shape = Qin.shape
        n, nh = self.n, self.nh
        p = self.p
        Q = numpy.moveaxis(numpy.asarray(Qin, dtype=float).reshape(p.state_shape), 1, 0)  # (3,6,ny,nx,P)
        sqrtG = self.sqrtG

        Qu = Q.copy()
        if p.hsurf is not None:
            Qu[idx_h] += p.hsurf.reshape(sqrtG.shape)

        var_itf_i = numpy.zeros((3, 6, nh, nh + 2, 2 * n))
        var_itf_i[:, :, :, 1:-1, :] = Qu @ self.extrap_x
        var_itf_j = numpy.zeros((3, 6, nh + 2, nh, 2 * n))
        var_itf_j[:, :, 1:-1, :, :] = Qu @ self.extrap_y

        Qu[idx_hu1] /= Q[idx_h]
        Qu[idx_hu2] /= Q[idx_h]
        u1, u2 = Qu[idx_hu1], Qu[idx_hu2]

        # halo (rhs_sw.py:103-117): edge data flattened to one line of nh*n points
        def edges_of(v):
            return [
                [
                    var_itf_j[v, q, 1, :, :n].reshape(nh * n),
                    var_itf_j[v, q, -2, :, n:].reshape(nh * n),
                    var_itf_i[v, q, :, 1, :n].reshape(nh * n),
                    var_itf_i[v, q, :, -2, n:].reshape(nh * n),
                ]
                for q in range(6)
            ]

        e1, e2 = edges_of(idx_hu1), edges_of(idx_hu2)
        vec = [[(e1[q][e], e2[q][e]) for e in range(4)] for q in range(6)]
        coords = [[p.bnd_sn[q], p.bnd_sn[q], p.bnd_we[q], p.bnd_we[q]] for q in range(6)]
        rv = cx.exchange_vectors(vec, coords, -1)
        rs = cx.exchange_scalars(edges_of(idx_h), -1)
        for q in range(6):
            for c, v in enumerate((idx_hu1, idx_hu2)):
                var_itf_j[v, q, 0, :, n:] = rv[q][cx.SOUTH][c].reshape(nh, n)
                var_itf_j[v, q, -1, :, :n] = rv[q][cx.NORTH][c].reshape(nh, n)
                var_itf_i[v, q, :, 0, n:] = rv[q][cx.WEST][c].reshape(nh, n)
                var_itf_i[v, q, :, -1, :n] = rv[q][cx.EAST][c].reshape(nh, n)
            var_itf_j[idx_h, q, 0, :, n:] = rs[q][cx.SOUTH].reshape(nh, n)
            var_itf_j[idx_h, q, -1, :, :n] = rs[q][cx.NORTH].reshape(nh, n)
            var_itf_i[idx_h, q, :, 0, n:] = rs[q][cx.WEST].reshape(nh, n)
            var_itf_i[idx_h, q, :, -1, :n] = rs[q][cx.EAST].reshape(nh, n)

        # pointwise fluxes (rhs_sw.py:120-131)
        flux_x1 = numpy.empty_like(Q)
        flux_x2 = numpy.empty_like(Q)
        flux_x1[idx_h] = sqrtG * Q[idx_hu1]
        flux_x2[idx_h] = sqrtG * Q[idx_hu2]
        hsq = Q[idx_h] ** 2
        flux_x1[idx_hu1] = sqrtG * (Q[idx_hu1] * u1 + 0.5 * gravity * self.H11 * hsq)
        flux_x2[idx_hu1] = sqrtG * (Q[idx_hu1] * u2 + 0.5 * gravity * self.H12 * hsq)
        flux_x1[idx_hu2] = sqrtG * (Q[idx_hu2] * u1 + 0.5 * gravity * self.H12 * hsq)
        flux_x2[idx_hu2] = sqrtG * (Q[idx_hu2] * u2 + 0.5 * gravity * self.H22 * hsq)
        df1 = flux_x1 @ self.derivative_x
        df2 = flux_x2 @ self.derivative_y

        if p.hsurf is not None:
            var_itf_i[idx_h] -= p.hsurf_itf_i
            var_itf_j[idx_h] -= p.hsurf_itf_j

        west = numpy.s_[..., 1:, :n]
        east = numpy.s_[..., :-1, n:]
        south = numpy.s_[..., 1:, :, :n]
        north = numpy.s_[..., :-1, :, n:]

        # AUSM along x1 (rhs_sw.py:170-187)
        sg, h11, h21 = self.fm_i[:, 0], self.fm_i[:, 1], self.fm_i[:, 2]
        with numpy.errstate(invalid="ignore", divide="ignore"):
            a = numpy.sqrt(gravity * var_itf_i[idx_h] * h11)
            tmp = var_itf_i[idx_h] * a
            m = numpy.where(tmp != 0.0, var_itf_i[idx_hu1] / tmp, 0.0)
        big_M = 0.25 * ((m[east] + 1.0) ** 2 - (m[west] - 1.0) ** 2)
        flux_x1_itf = numpy.zeros_like(var_itf_i)
        flux_x1_itf[east] = sg[east] * (
            numpy.maximum(0.0, big_M) * a[east] * var_itf_i[east] + numpy.minimum(0.0, big_M) * a[west] * var_itf_i[west]
        )
        p11 = sg * (0.5 * gravity) * h11 * var_itf_i[idx_h] ** 2
        p21 = sg * (0.5 * gravity) * h21 * var_itf_i[idx_h] ** 2
        flux_x1_itf[idx_hu1][east] += 0.5 * ((1.0 + m[east]) * p11[east] + (1.0 - m[west]) * p11[west])
        flux_x1_itf[idx_hu2][east] += 0.5 * ((1.0 + m[east]) * p21[east] + (1.0 - m[west]) * p21[west])
        flux_x1_itf[west] = flux_x1_itf[east]

        # AUSM along x2 (rhs_sw.py:189-207)
        sg, h12, h22 = self.fm_j[:, 0], self.fm_j[:, 1], self.fm_j[:, 2]
        with numpy.errstate(invalid="ignore", divide="ignore"):
            a = numpy.sqrt(gravity * var_itf_j[idx_h] * h22)
            m = var_itf_j[idx_hu2] / (var_itf_j[idx_h] * a)
        m[numpy.where(numpy.isnan(m))] = 0.0
        big_M = 0.25 * ((m[north] + 1.0) ** 2 - (m[south] - 1.0) ** 2)
        flux_x2_itf = numpy.zeros_like(var_itf_j)
        flux_x2_itf[north] = sg[north] * (
            numpy.maximum(0.0, big_M) * a[north] * var_itf_j[north] + numpy.minimum(0.0, big_M) * a[south] * var_itf_j[south]
        )
        p12 = sg * (0.5 * gravity) * h12 * var_itf_j[idx_h] ** 2
        p22 = sg * (0.5 * gravity) * h22 * var_itf_j[idx_h] ** 2
        flux_x2_itf[idx_hu1][north] += 0.5 * ((1.0 + m[north]) * p12[north] + (1.0 - m[south]) * p12[south])
        flux_x2_itf[idx_hu2][north] += 0.5 * ((1.0 + m[north]) * p22[north] + (1.0 - m[south]) * p22[south])
        flux_x2_itf[south] = flux_x2_itf[north]

        df1 += flux_x1_itf[:, :, :, 1:-1, :] @ self.correction_WE
        df2 += flux_x2_itf[:, :, 1:-1, :, :] @ self.correction_SN

        if p.hsurf is None:
            dz1 = dz2 = 0.0
        else:
            dz1, dz2 = p.dzdx1.reshape(sqrtG.shape), p.dzdx2.reshape(sqrtG.shape)

        # forcing (rhs_sw.py:221-235)
        c1_01, c1_02, c1_11, c1_12, c2_01, c2_02, c2_12, c2_22 = self.chr
        forcing = numpy.zeros_like(Q)
        forcing[idx_hu1] = (
            2.0 * (c1_01 * Q[idx_hu1] + c1_02 * Q[idx_hu2])
            + c1_11 * Q[idx_hu1] * u1
            + 2.0 * c1_12 * Q[idx_hu1] * u2
            + gravity * Q[idx_h] * (self.H11 * dz1 + self.H12 * dz2)
        )
        forcing[idx_hu2] = (
            2.0 * (c2_01 * Q[idx_hu1] + c2_02 * Q[idx_hu2])
            + 2.0 * c2_12 * Q[idx_hu1] * u2
            + c2_22 * Q[idx_hu2] * u2
            + gravity * Q[idx_h] * (self.H12 * dz1 + self.H22 * dz2)
        )
        rhs = self.inv_sqrtG * (-df1 - df2) - forcing
        return numpy.ascontiguousarray(numpy.moveaxis(rhs, 0, 1)).reshape(given_shape)

    __call__ = rhs

"""Build an ``EulerProblem`` / ``SwProblem`` from the reference's own host objects.

This is the setup-time half of the drop-in boundary: the reference constructs
``geometry``, ``operators`` and ``metric`` (wx_factory/simulation/simulation.py:96-103) and hands
them to the RHS constructor (wx_factory/rhs/rhs.py:14-25, wx_factory/rhs/rhs_sw.py:18-36); we read
the arrays the RHS uses from those objects and keep them in compact form (see problem.py).

``ranks`` is a sequence of 6 objects (one per panel, in rank order) with attributes
``geom, ops, metric, config`` and optionally ``topo`` -- duck-typed, nothing from the reference
is imported here.
"""

import math

import numpy

from .problem import H_PAIRS, EulerProblem, SwProblem

_gravity = 9.80616  # wx_factory/common/definitions.py:12


def _single_valued(two_sided, axis, F, what):
    """two_sided: (..., ne+2, ..., 2F) padded along `axis` (negative, counting the last axis as -1).
    Returns the (.., ne+1, .., F) unique-face array after checking both sides agree."""
    a = numpy.moveaxis(two_sided, axis, -2)  # (..., ne+2, 2F)
    lo = a[..., 1:, :F]  # min-side face of elements 1..ne+1
    hi = a[..., :-1, F:]  # max-side face of elements 0..ne
    if not numpy.array_equal(lo, hi):
        raise ValueError(f"interface metric {what} is not single-valued across faces")
    return numpy.ascontiguousarray(numpy.moveaxis(lo, -2, axis))


def _schar_damping_static(geom, metric, shear):
    """State-independent pieces of dcmip_schar_damping (wx_factory/init/dcmip.py:676-755):
    profile = sin^2(pi/2 (z-zh)/(ztop-zh)) for z > zh else 0; (u1,u2,u3)_ref contravariant."""
    T0, Ueq, Zh = 300.0, 20.0, 20000.0
    Cs = 2.5e-4 if shear else 0.0
    lat = geom.polar[1, ...]
    z = geom.polar[2, ...]
    prof = numpy.sin(numpy.pi / 2 * (z - Zh) / (geom.ztop - Zh)) ** 2
    prof[z <= Zh] = 0.0
    Tref = T0 * (1 - Cs * Ueq**2 / _gravity * numpy.sin(lat) ** 2)
    uref = Ueq * numpy.cos(lat) * (2 * T0 / Tref * Cs * z + Tref / T0) ** 0.5
    u1r, u2r, u3r = geom.wind2contra(uref, 0.0, 0.0, metric)
    shape = prof.shape
    return numpy.stack([prof, numpy.broadcast_to(u1r, shape), numpy.broadcast_to(u2r, shape), numpy.broadcast_to(u3r, shape)])


def euler_problem_from_reference(ranks) -> EulerProblem:
    if len(ranks) != 6:
        raise ValueError("need the 6 per-panel reference objects (6-rank layout)")
    g0, o0, c0 = ranks[0].geom, ranks[0].ops, ranks[0].config
    n = int(c0.num_solpts)
    nz = int(g0.num_elements_x3)
    nh = int(g0.num_elements_x1)
    P, F = n**3, n**2
    E = nz * nh * nh
    case = int(c0.case_number)

    sqrtg, h, chris, idz, fmi, fmj, fmk, bsn, bwe, damp = ([] for _ in range(10))
    for r in ranks:
        m, g = r.metric, r.geom
        sqrtg.append(numpy.asarray(m.sqrtG_new).reshape(E, P))
        hc = numpy.asarray(m.h_contra_new)
        for a, b in H_PAIRS:
            if not numpy.array_equal(hc[a, b], hc[b, a]):
                raise ValueError("h_contra is not symmetric")
        h.append(numpy.stack([hc[a, b].reshape(E, P) for a, b in H_PAIRS]))
        chris.append(numpy.asarray(m.christoffel).reshape(27, E, P))
        idz.append(numpy.asarray(m.inv_dzdeta_new).reshape(E, P))
        for d, (lst, sq, hh, ax) in enumerate(
            (
                (fmi, m.sqrtG_itf_i_new, m.h_contra_itf_i_new, -2),
                (fmj, m.sqrtG_itf_j_new, m.h_contra_itf_j_new, -3),
                (fmk, m.sqrtG_itf_k_new, m.h_contra_itf_k_new, -4),
            )
        ):
            comps = [_single_valued(numpy.asarray(sq), ax, F, f"sqrtG dir {d}")]
            for b in range(3):
                comps.append(_single_valued(numpy.asarray(hh[d, b]), ax, F, f"h[{d},{b}]"))
            lst.append(numpy.stack(comps))
        bsn.append(numpy.asarray(g.boundary_sn).reshape(nh * n))
        bwe.append(numpy.asarray(g.boundary_we).reshape(nh * n))
        if case in (21, 22):
            damp.append(_schar_damping_static(g, m, shear=(case == 22)).reshape(4, E, P))

    return EulerProblem(
        n=n,
        nz=nz,
        nh=nh,
        advection_only=bool(case < 13),
        case_number=case,
        ext_w=numpy.asarray(o0.extrap_west, dtype=float).copy(),
        ext_e=numpy.asarray(o0.extrap_east, dtype=float).copy(),
        diff=numpy.asarray(o0.diff_solpt, dtype=float).copy(),
        corr=numpy.asarray(o0.correction, dtype=float).copy(),
        hfilt=numpy.asarray(o0.highfilter, dtype=float).copy(),
        sqrtg=numpy.stack(sqrtg),
        h=numpy.stack(h),
        chris=numpy.stack(chris),
        inv_dzdeta=numpy.stack(idz),
        fm_i=numpy.stack(fmi),
        fm_j=numpy.stack(fmj),
        fm_k=numpy.stack(fmk),
        bnd_sn=numpy.stack(bsn),
        bnd_we=numpy.stack(bwe),
        damp=numpy.stack(damp) if damp else None,
    )


def sw_problem_from_reference(ranks) -> SwProblem:
    if len(ranks) != 6:
        raise ValueError("need the 6 per-panel reference objects (6-rank layout)")
    g0, o0, c0 = ranks[0].geom, ranks[0].ops, ranks[0].config
    n = int(c0.num_solpts)
    nh = int(g0.num_elements_x1)
    E, P = nh * nh, n * n
    sqrtg, h, chris, fmi, fmj, bsn, bwe = ([] for _ in range(7))
    topo = {k: [] for k in ("hsurf", "hsurf_itf_i", "hsurf_itf_j", "dzdx1", "dzdx2")}
    for r in ranks:
        m, g = r.metric, r.geom
        if not numpy.array_equal(m.H_contra_12, m.H_contra_21):
            raise ValueError("H_contra is not symmetric")
        sqrtg.append(numpy.asarray(m.sqrtG).reshape(E, P))
        h.append(numpy.stack([numpy.asarray(x).reshape(E, P) for x in (m.H_contra_11, m.H_contra_12, m.H_contra_22)]))
        chris.append(
            numpy.stack(
                [
                    numpy.asarray(x).reshape(E, P)
                    for x in (
                        m.christoffel_1_01,
                        m.christoffel_1_02,
                        m.christoffel_1_11,
                        m.christoffel_1_12,
                        m.christoffel_2_01,
                        m.christoffel_2_02,
                        m.christoffel_2_12,
                        m.christoffel_2_22,
                    )
                ]
            )
        )
        fmi.append(
            numpy.stack(
                [_single_valued(numpy.asarray(x), -2, n, "2d itf_i") for x in (m.sqrtG_itf_i, m.H_contra_11_itf_i, m.H_contra_21_itf_i)]
            )
        )
        fmj.append(
            numpy.stack(
                [_single_valued(numpy.asarray(x), -3, n, "2d itf_j") for x in (m.sqrtG_itf_j, m.H_contra_12_itf_j, m.H_contra_22_itf_j)]
            )
        )
        bsn.append(numpy.asarray(g.boundary_sn).reshape(nh * n))
        bwe.append(numpy.asarray(g.boundary_we).reshape(nh * n))
        t = getattr(r, "topo", None)
        if t is not None:
            topo["hsurf"].append(numpy.asarray(t.hsurf).reshape(E, P))
            topo["hsurf_itf_i"].append(numpy.asarray(t.hsurf_itf_i))
            topo["hsurf_itf_j"].append(numpy.asarray(t.hsurf_itf_j))
            topo["dzdx1"].append(numpy.asarray(t.dzdx1).reshape(E, P))
            topo["dzdx2"].append(numpy.asarray(t.dzdx2).reshape(E, P))
    kw = {k: numpy.stack(v) for k, v in topo.items() if len(v) == 6}
    return SwProblem(
        n=n,
        nh=nh,
        case_number=int(c0.case_number),
        ext_w=numpy.asarray(o0.extrap_west, dtype=float).copy(),
        ext_e=numpy.asarray(o0.extrap_east, dtype=float).copy(),
        diff=numpy.asarray(o0.diff_solpt, dtype=float).copy(),
        corr=numpy.asarray(o0.correction, dtype=float).copy(),
        sqrtg=numpy.stack(sqrtg),
        h=numpy.stack(h),
        chris=numpy.stack(chris),
        fm_i=numpy.stack(fmi),
        fm_j=numpy.stack(fmj),
        bnd_sn=numpy.stack(bsn),
        bnd_we=numpy.stack(bwe),
        **kw,
    )


def problem_from_reference(ranks):
    eq = str(ranks[0].config.equations).lower()
    return sw_problem_from_reference(ranks) if eq == "shallow_water" else euler_problem_from_reference(ranks)

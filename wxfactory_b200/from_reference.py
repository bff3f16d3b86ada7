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


def _euler_panel(geom, ops, metric, config):
    """The arrays of ONE panel (= one reference rank of the 6-rank layout) the Euler right-hand side reads."""
    n = int(config.num_solpts)
    nz = int(geom.num_elements_x3)
    nh = int(geom.num_elements_x1)
    P, F = n**3, n**2
    E = nz * nh * nh
    case = int(config.case_number)
    m, g = metric, geom
    out = {"equations": "euler", "n": n, "nz": nz, "nh": nh, "case_number": case}
    out["sqrtg"] = numpy.asarray(m.sqrtG_new).reshape(E, P)
    hc = numpy.asarray(m.h_contra_new)
    for a, b in H_PAIRS:
        if not numpy.array_equal(hc[a, b], hc[b, a]):
            raise ValueError("h_contra is not symmetric")
    out["h"] = numpy.stack([hc[a, b].reshape(E, P) for a, b in H_PAIRS])
    out["chris"] = numpy.asarray(m.christoffel).reshape(27, E, P)
    out["inv_dzdeta"] = numpy.asarray(m.inv_dzdeta_new).reshape(E, P)
    for d, (key, sq, hh, ax) in enumerate(
        (
            ("fm_i", m.sqrtG_itf_i_new, m.h_contra_itf_i_new, -2),
            ("fm_j", m.sqrtG_itf_j_new, m.h_contra_itf_j_new, -3),
            ("fm_k", m.sqrtG_itf_k_new, m.h_contra_itf_k_new, -4),
        )
    ):
        comps = [_single_valued(numpy.asarray(sq), ax, F, f"sqrtG dir {d}")]
        for b in range(3):
            comps.append(_single_valued(numpy.asarray(hh[d, b]), ax, F, f"h[{d},{b}]"))
        out[key] = numpy.stack(comps)
    out["bnd_sn"] = numpy.asarray(g.boundary_sn).reshape(nh * n)
    out["bnd_we"] = numpy.asarray(g.boundary_we).reshape(nh * n)
    if case in (21, 22):
        out["damp"] = _schar_damping_static(g, m, shear=(case == 22)).reshape(4, E, P)
    for k, attr in (("ext_w", "extrap_west"), ("ext_e", "extrap_east"), ("diff", "diff_solpt"), ("corr", "correction"), ("hfilt", "highfilter")):
        out[k] = numpy.asarray(getattr(ops, attr), dtype=float).copy()
    return out


def _sw_panel(geom, ops, metric, config, topo):
    """The arrays of one panel the shallow-water right-hand side reads."""
    n = int(config.num_solpts)
    nh = int(geom.num_elements_x1)
    E, P = nh * nh, n * n
    m, g = metric, geom
    if not numpy.array_equal(m.H_contra_12, m.H_contra_21):
        raise ValueError("H_contra is not symmetric")
    out = {"equations": "shallow_water", "n": n, "nh": nh, "case_number": int(config.case_number)}
    out["sqrtg"] = numpy.asarray(m.sqrtG).reshape(E, P)
    out["h"] = numpy.stack([numpy.asarray(x).reshape(E, P) for x in (m.H_contra_11, m.H_contra_12, m.H_contra_22)])
    out["chris"] = numpy.stack(
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
    out["fm_i"] = numpy.stack([_single_valued(numpy.asarray(x), -2, n, "2d itf_i") for x in (m.sqrtG_itf_i, m.H_contra_11_itf_i, m.H_contra_21_itf_i)])
    out["fm_j"] = numpy.stack([_single_valued(numpy.asarray(x), -3, n, "2d itf_j") for x in (m.sqrtG_itf_j, m.H_contra_12_itf_j, m.H_contra_22_itf_j)])
    out["bnd_sn"] = numpy.asarray(g.boundary_sn).reshape(nh * n)
    out["bnd_we"] = numpy.asarray(g.boundary_we).reshape(nh * n)
    if topo is not None:
        out["hsurf"] = numpy.asarray(topo.hsurf).reshape(E, P)
        out["hsurf_itf_i"] = numpy.asarray(topo.hsurf_itf_i)
        out["hsurf_itf_j"] = numpy.asarray(topo.hsurf_itf_j)
        out["dzdx1"] = numpy.asarray(topo.dzdx1).reshape(E, P)
        out["dzdx2"] = numpy.asarray(topo.dzdx2).reshape(E, P)
    for k, attr in (("ext_w", "extrap_west"), ("ext_e", "extrap_east"), ("diff", "diff_solpt"), ("corr", "correction")):
        out[k] = numpy.asarray(getattr(ops, attr), dtype=float).copy()
    return out


def panel_arrays(geom, operators, metric, config, topo=None):
    """What ONE reference rank contributes to the problem: a dict of plain numpy arrays (picklable, so that an MPI
    gather can move it), read from the objects ``Simulation.__init__`` builds on every rank
    (wx_factory/simulation/simulation.py:96-103).  Anything that needs the rank's own objects (the state-independent
    part of the DCMIP 2-x damping calls ``geom.wind2contra``) is evaluated here, on the rank."""
    eq = str(config.equations).lower()
    if eq == "shallow_water":
        return _sw_panel(geom, operators, metric, config, topo)
    if eq == "euler":
        return _euler_panel(geom, operators, metric, config)
    raise ValueError(f"Unrecognized combination of equations ({eq}) and geometry ({type(geom)})")


def problem_from_panels(panels):
    """Six ``panel_arrays`` dicts (rank order = panel order) -> EulerProblem / SwProblem."""
    if len(panels) != 6:
        raise ValueError("need the arrays of the 6 panels (6-rank layout)")
    p0 = panels[0]
    st = lambda k: numpy.stack([p[k] for p in panels])  # noqa: E731
    if p0["equations"] == "euler":
        return EulerProblem(
            n=p0["n"], nz=p0["nz"], nh=p0["nh"], advection_only=bool(p0["case_number"] < 13), case_number=p0["case_number"],
            ext_w=p0["ext_w"], ext_e=p0["ext_e"], diff=p0["diff"], corr=p0["corr"], hfilt=p0["hfilt"],
            sqrtg=st("sqrtg"), h=st("h"), chris=st("chris"), inv_dzdeta=st("inv_dzdeta"),
            fm_i=st("fm_i"), fm_j=st("fm_j"), fm_k=st("fm_k"), bnd_sn=st("bnd_sn"), bnd_we=st("bnd_we"),
            damp=st("damp") if "damp" in p0 else None,
        )
    kw = {k: st(k) for k in ("hsurf", "hsurf_itf_i", "hsurf_itf_j", "dzdx1", "dzdx2") if all(k in p for p in panels)}
    return SwProblem(
        n=p0["n"], nh=p0["nh"], case_number=p0["case_number"],
        ext_w=p0["ext_w"], ext_e=p0["ext_e"], diff=p0["diff"], corr=p0["corr"],
        sqrtg=st("sqrtg"), h=st("h"), chris=st("chris"), fm_i=st("fm_i"), fm_j=st("fm_j"), bnd_sn=st("bnd_sn"), bnd_we=st("bnd_we"),
        **kw,
    )


def gather_to_one_process(geom, operators, metric, config, topo=None, comm=None, root=0):
    """Called by EVERY rank of the reference's 6-rank job with its own objects: the per-panel arrays travel to `root`
    through ``comm.gather`` (mpi4py API; default ``MPI.COMM_WORLD``), which gets the assembled problem; the other ranks
    get None.  The reference's minimum is one rank per panel (wx_factory/process_topology.py:69-82); tiles smaller than
    a panel are not supported here."""
    if comm is None:
        from mpi4py import MPI

        comm = MPI.COMM_WORLD
    if comm.Get_size() != 6:
        raise ValueError(f"Wrong number of PEs ({comm.Get_size()}): this bridge gathers the 6-rank layout (one rank per panel)")
    panels = comm.gather(panel_arrays(geom, operators, metric, config, topo), root=root)
    return problem_from_panels(panels) if comm.Get_rank() == root else None


def problem_from_reference(ranks):
    """``ranks``: the 6 per-panel objects with attributes geom, ops, metric, config[, topo], all in this process."""
    return problem_from_panels([panel_arrays(r.geom, r.ops, r.metric, r.config, getattr(r, "topo", None)) for r in ranks])

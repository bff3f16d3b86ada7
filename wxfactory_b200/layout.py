"""Local (per-rank) views of a 6-panel problem: static arrays for wxb_create and state scatter/gather.

A rank's state is the concatenation of its slabs, each in the reference's per-rank layout
``(num_var, nz, ny_slab, nx, P)`` (wx_factory/geometry/cubed_sphere_3d.py:186-190); with one rank and six
whole-panel slabs this is simply ``numpy.stack`` of the six reference rank states.
"""

import numpy

from .partition import Partition


def _is_euler(prob):
    return prob.equations == "euler"


def local_state(prob, part: Partition, Qglobal):
    """(6, nvar, [nz,] nh, nh, P) -> flat local state of `part`."""
    Q = numpy.asarray(Qglobal).reshape(prob.state_shape)
    pieces = []
    for sl in part.slabs:
        if _is_euler(prob):
            pieces.append(Q[sl.panel][:, :, sl.j0 : sl.j1].ravel())
        else:
            pieces.append(Q[sl.panel][:, sl.j0 : sl.j1].ravel())
    return numpy.concatenate(pieces)


def scatter_back(prob, part: Partition, local, Qglobal):
    """Write a flat local state into the matching rows of a global (6, ...) array."""
    Q = Qglobal.reshape(prob.state_shape)
    off = 0
    for sl in part.slabs:
        if _is_euler(prob):
            shp = (prob.num_var, prob.nz, sl.ny, prob.nh, prob.P)
            cnt = int(numpy.prod(shp))
            Q[sl.panel][:, :, sl.j0 : sl.j1] = local[off : off + cnt].reshape(shp)
        else:
            shp = (prob.num_var, sl.ny, prob.nh, prob.P)
            cnt = int(numpy.prod(shp))
            Q[sl.panel][:, sl.j0 : sl.j1] = local[off : off + cnt].reshape(shp)
        off += cnt
    return Qglobal


def _node_field(prob, part, arr, lead):
    """arr: (6, *lead, E, P) panel-stacked per-node field -> (*lead, E_local*P) in local element order."""
    nz = prob.nz if _is_euler(prob) else 1
    nh, P = prob.nh, prob.P
    a = arr.reshape((6,) + lead + (nz, nh, nh, P))
    parts = [a[sl.panel][..., :, sl.j0 : sl.j1, :, :].reshape(lead + (-1,)) for sl in part.slabs]
    return numpy.ascontiguousarray(numpy.concatenate(parts, axis=-1))


def local_static(prob, part: Partition):
    """dict of contiguous float64/int32 arrays named like wxb_static_arrays' fields."""
    euler = _is_euler(prob)
    out = {
        "ext_w": prob.ext_w,
        "ext_e": prob.ext_e,
        "diff": prob.diff,
        "corr": prob.corr,
        "sqrtg": _node_field(prob, part, prob.sqrtg, ()),
    }
    if euler:
        out["hfilt"] = prob.hfilt
        out["h"] = _node_field(prob, part, prob.h, (6,))
        out["chris"] = _node_field(prob, part, prob.chris, (27,))
        out["inv_dzdeta"] = _node_field(prob, part, prob.inv_dzdeta, ())
        if prob.damp is not None:
            out["damp"] = _node_field(prob, part, prob.damp, (4,))
        fi, fj, fk = [], [], []
        for sl in part.slabs:
            fi.append(prob.fm_i[sl.panel][:, :, sl.j0 : sl.j1].ravel())  # (4,nz,ny,nx+1,F)
            fj.append(prob.fm_j[sl.panel][:, :, sl.j0 : sl.j1 + 1].ravel())  # (4,nz,ny+1,nx,F)
            fk.append(prob.fm_k[sl.panel][:, :, sl.j0 : sl.j1].ravel())  # (4,nz+1,ny,nx,F)
        out["fm_i"], out["fm_j"], out["fm_k"] = (numpy.concatenate(x) for x in (fi, fj, fk))
    else:
        out["h"] = _node_field(prob, part, prob.h, (3,))
        out["chris"] = _node_field(prob, part, prob.chris, (8,))
        fi, fj = [], []
        for sl in part.slabs:
            fi.append(prob.fm_i[sl.panel][:, sl.j0 : sl.j1].ravel())  # (3,ny,nx+1,n)
            fj.append(prob.fm_j[sl.panel][:, sl.j0 : sl.j1 + 1].ravel())  # (3,ny+1,nx,n)
        out["fm_i"], out["fm_j"] = numpy.concatenate(fi), numpy.concatenate(fj)
        if prob.hsurf is not None:
            out["topo"] = numpy.stack(
                [_node_field(prob, part, x, ()) for x in (prob.hsurf, prob.dzdx1, prob.dzdx2)]
            )
            n = prob.n
            ti, tj = [], []
            for sl in part.slabs:
                a = prob.hsurf_itf_i[sl.panel][sl.j0 : sl.j1]  # (ny, nx+2, 2n) two-sided
                ti.append(numpy.stack([a[:, :-1, n:], a[:, 1:, :n]]).ravel())  # L side, R side per face
                b = prob.hsurf_itf_j[sl.panel][sl.j0 : sl.j1 + 2]  # padded rows j0..j1+1
                tj.append(numpy.stack([b[:-1, :, n:], b[1:, :, :n]]).ravel())
            out["topo_itf"] = numpy.concatenate(ti + tj)
    out["nbr"] = numpy.ascontiguousarray(part.nbr, dtype=numpy.int32)
    out["xform"] = numpy.ascontiguousarray(part.xform, dtype=numpy.float64)
    out["send"] = numpy.ascontiguousarray(part.send, dtype=numpy.int32)
    return {k: (v if v.dtype == numpy.int32 else numpy.ascontiguousarray(v, dtype=numpy.float64)) for k, v in out.items()}

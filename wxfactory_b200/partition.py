"""Partition of the cubed sphere over 1/2/4/8 GPUs and the static face-connectivity tables.

The reference decomposes into 6*N^2 square tiles, one per MPI rank (wx_factory/process_topology.py:69-82),
which excludes 1, 2, 4 and 8 ranks.  Here the 6*nh rows of element columns (panel-major, row j, all i) are cut
into G contiguous groups; a rank owns 1..6 *slabs* ``(panel, j0:j1)`` whose arrays keep the reference's
per-rank layout ``(num_var, nz, ny_slab, nx, n^3)``.  The vertical is never split (neither does the reference).

Connectivity is resolved on the host, once, into a *pull* table: for every local element column and each of
its 4 horizontal faces, where the facing neighbour's trace lives (local trace array or a ghost record received
from another rank), whether the along-edge node order is reversed, and which 2x2 matrix maps the neighbour's
(u1, u2) components into the local panel basis.  The panel tables restate the reference's:

    all_neighbors    wx_factory/process_topology.py:105-113
    flips            wx_factory/process_topology.py:128-135   (sender reverses along the edge)
    convert_contras  wx_factory/process_topology.py:138-175   (sender converts to the receiver's basis)

so that GPU code has no panel logic at all.
"""

from dataclasses import dataclass, field

import numpy

WEST, EAST, SOUTH, NORTH = 0, 1, 2, 3  # kernel face order
# reference edge order is S, N, W, E (process_topology.py:12-15)
_REF_EDGE = {SOUTH: 0, NORTH: 1, WEST: 2, EAST: 3}
_FROM_REF = {0: SOUTH, 1: NORTH, 2: WEST, 3: EAST}

_NEIGHBORS = [  # [panel][S, N, W, E]
    [5, 4, 3, 1],
    [5, 4, 0, 2],
    [5, 4, 1, 3],
    [5, 4, 2, 0],
    [0, 2, 3, 1],
    [2, 0, 3, 1],
]
_FLIPS = [  # [panel][S, N, W, E]
    [False, False, False, False],
    [True, False, False, False],
    [True, True, False, False],
    [False, True, False, False],
    [False, True, True, False],
    [True, False, False, True],
]
# convert_contras[panel][S,N,W,E] as 2x2 matrices with entries in {0, +-1, +-c}, c = 2x/(1+x^2):
# each entry is (constant, coefficient of c)
_I, _O = (1.0, 0.0), (0.0, 0.0)
_M = (-1.0, 0.0)
_C, _NC = (0.0, 1.0), (0.0, -1.0)
_CONTRA = [
    [((_I, _C), (_O, _I)), ((_I, _NC), (_O, _I)), ((_I, _O), (_C, _I)), ((_I, _O), (_NC, _I))],
    [((_O, _I), (_M, _NC)), ((_O, _M), (_I, _NC)), ((_I, _O), (_C, _I)), ((_I, _O), (_NC, _I))],
    [((_M, _NC), (_O, _M)), ((_M, _C), (_O, _M)), ((_I, _O), (_C, _I)), ((_I, _O), (_NC, _I))],
    [((_O, _M), (_I, _C)), ((_O, _I), (_M, _C)), ((_I, _O), (_C, _I)), ((_I, _O), (_NC, _I))],
    [((_I, _C), (_O, _I)), ((_M, _C), (_O, _M)), ((_NC, _M), (_I, _O)), ((_NC, _I), (_M, _O))],
    [((_M, _NC), (_O, _M)), ((_I, _NC), (_O, _I)), ((_C, _I), (_M, _O)), ((_C, _M), (_I, _O))],
]


def contra_matrix(panel, ref_edge, coord):
    """2x2 matrices (…,4) = (m11, m12, m21, m22) of convert_contras[panel][ref_edge] at `coord`."""
    coord = numpy.asarray(coord, dtype=float)
    c = 2.0 * coord / (1.0 + coord**2)
    rows = _CONTRA[panel][ref_edge]
    out = numpy.empty(coord.shape + (4,))
    for r in range(2):
        for s in range(2):
            k0, k1 = rows[r][s]
            out[..., 2 * r + s] = k0 + k1 * c
    return out


@dataclass
class Slab:
    panel: int
    j0: int
    j1: int

    @property
    def ny(self):
        return self.j1 - self.j0


def slabs_of_rank(nh, nranks, rank):
    """Contiguous rows [rank*R, (rank+1)*R) of the 6*nh panel-major rows, split at panel boundaries."""
    total = 6 * nh
    if total % nranks != 0:
        raise ValueError(f"6*nh = {total} rows cannot be split evenly over {nranks} ranks")
    rows = total // nranks
    lo, hi = rank * rows, (rank + 1) * rows
    out = []
    while lo < hi:
        p = lo // nh
        end = min(hi, (p + 1) * nh)
        out.append(Slab(p, lo - p * nh, end - p * nh))
        lo = end
    return out


def owner_of_row(nh, nranks, panel, j):
    return (panel * nh + j) // ((6 * nh) // nranks)


@dataclass
class Partition:
    """Everything rank `rank` of `nranks` needs to know about who its neighbours are."""

    nh: int
    nz: int
    n: int
    nranks: int
    rank: int
    slabs: list
    ncols: int = 0
    nelems: int = 0
    nbr: numpy.ndarray = None  # (4, ncols, 4) int32: base, kstride, vstride, flags
    xform: numpy.ndarray = None  # (nxform, n, 4)
    # ghost records, grouped by peer: recv_peers[i] = (peer rank, first record, count)
    recv_peers: list = field(default_factory=list)
    send_peers: list = field(default_factory=list)
    send: numpy.ndarray = None  # (nsend,) int32: local col*4 + face, in send-buffer order
    nghost: int = 0
    trace_layout: str = "field"

    @property
    def nsend(self):
        return 0 if self.send is None else int(self.send.size)


def _facing(panel, face, m, nh):
    """The column and face on the other side of panel edge (`panel`, `face`) at along-edge position m.
    Returns (q, jq, iq, face_q, flip, ref_edge_q)."""
    e = _REF_EDGE[face]
    q = _NEIGHBORS[panel][e]
    eq = _NEIGHBORS[q].index(panel)  # the neighbour's edge that faces us (each panel's 4 neighbours are distinct)
    flip = _FLIPS[q][eq]  # the SENDER reverses its data if its own flag is set (process_topology.py:313,379)
    mq = nh - 1 - m if flip else m
    fq = _FROM_REF[eq]
    if fq == SOUTH:
        jq, iq = 0, mq
    elif fq == NORTH:
        jq, iq = nh - 1, mq
    elif fq == WEST:
        jq, iq = mq, 0
    else:
        jq, iq = mq, nh - 1
    return q, jq, iq, fq, flip, eq


def _ghost_requests(nh, nranks, rank):
    """Ordered list of (owner, panel, j, i, face_of_owner_column) this rank must receive."""
    req = []
    seen = set()
    for sl in slabs_of_rank(nh, nranks, rank):
        p = sl.panel
        for j in range(sl.j0, sl.j1):
            for i in range(nh):
                for face in (WEST, EAST, SOUTH, NORTH):
                    tgt = _neighbour_column(p, j, i, face, nh)
                    q, jq, iq, fq = tgt[:4]
                    own = owner_of_row(nh, nranks, q, jq)
                    key = (q, jq, iq, fq)
                    if own != rank and key not in seen:
                        seen.add(key)
                        req.append((own,) + key)
    req.sort(key=lambda t: t[0])  # stable: grouped by owner, otherwise in discovery order
    return req


def _neighbour_column(p, j, i, face, nh):
    """(q, jq, iq, face_q, flip, ref_edge_q or None) of the column facing face `face` of column (p, j, i)."""
    if face == WEST:
        if i > 0:
            return p, j, i - 1, EAST, False, None
        return _facing(p, WEST, j, nh)
    if face == EAST:
        if i < nh - 1:
            return p, j, i + 1, WEST, False, None
        return _facing(p, EAST, j, nh)
    if face == SOUTH:
        if j > 0:
            return p, j - 1, i, NORTH, False, None
        return _facing(p, SOUTH, i, nh)
    if j < nh - 1:
        return p, j + 1, i, SOUTH, False, None
    return _facing(p, NORTH, i, nh)


def build_partition(nh, nz, n, bnd_sn, bnd_we, nranks=1, rank=0, num_var=5, nfaces=6, trace_layout="field"):
    """bnd_sn / bnd_we: (6, nh*n) edge coordinates of every panel (problem.bnd_sn / bnd_we).
    For shallow water pass nz=1, num_var=3, nfaces=4.

    trace_layout: how the device trace array is ordered, which fixes (base, kstride, vstride) of the pull table
    (all in rows of F doubles):
      "field"   [field][face][E] + ghost records [rec][field][k]       (shallow-water kernels)
      "element" [E][face][field] + ghost records [rec][k][field]       (Euler kernels: an element's 36 rows and a
                face's 6 fields are contiguous, the distance between fields is always one row)"""
    if trace_layout not in ("field", "element"):
        raise ValueError(f"unknown trace layout {trace_layout!r}")
    elem_major = trace_layout == "element"
    slabs = slabs_of_rank(nh, nranks, rank)
    part = Partition(nh=nh, nz=nz, n=n, nranks=nranks, rank=rank, slabs=slabs)
    part.trace_layout = trace_layout
    nx = nh
    # local numbering
    cbase, ebase, loc = [], [], {}
    c = e = 0
    for s, sl in enumerate(slabs):
        cbase.append(c)
        ebase.append(e)
        loc[sl.panel, sl.j0] = s
        c += sl.ny * nx
        e += nz * sl.ny * nx
    part.ncols, part.nelems = c, e
    E = e

    def local_slab(q, jq):
        for s, sl in enumerate(slabs):
            if sl.panel == q and sl.j0 <= jq < sl.j1:
                return s
        return -1

    ghosts = _ghost_requests(nh, nranks, rank)
    ghost_id = {g[1:]: k for k, g in enumerate(ghosts)}
    part.nghost = len(ghosts)
    start = 0
    while start < len(ghosts):
        own = ghosts[start][0]
        end = start
        while end < len(ghosts) and ghosts[end][0] == own:
            end += 1
        part.recv_peers.append((own, start, end - start))
        start = end

    ghost_base = num_var * nfaces * E  # in units of F doubles
    nbr = numpy.zeros((4, part.ncols, 4), dtype=numpy.int32)
    xrows = []
    for s, sl in enumerate(slabs):
        p = sl.panel
        for jl in range(sl.ny):
            j = sl.j0 + jl
            for i in range(nx):
                col = cbase[s] + jl * nx + i
                for face in (WEST, EAST, SOUTH, NORTH):
                    q, jq, iq, fq, flip, eq = _neighbour_column(p, j, i, face, nh)
                    flags = 1 if flip else 0
                    if eq is not None:
                        # sender = panel q, its edge eq, its own coordinate at the sending node
                        coords = (bnd_sn if eq in (0, 1) else bnd_we)[q]
                        mq = iq if eq in (0, 1) else jq
                        a_recv = numpy.arange(n)
                        a_send = (n - 1 - a_recv) if flip else a_recv
                        xrows.append(contra_matrix(q, eq, coords[mq * n + a_send]))
                        flags |= len(xrows) << 8
                    sq = local_slab(q, jq)
                    if sq >= 0:
                        slq = slabs[sq]
                        e0 = ebase[sq] + (jq - slq.j0) * nx + iq  # the neighbour column's element at level 0
                        if elem_major:
                            base = (e0 * nfaces + fq) * num_var
                            kstride, vstride = slq.ny * nx * nfaces * num_var, 1
                        else:
                            base = fq * E + e0
                            kstride, vstride = slq.ny * nx, nfaces * E
                    else:
                        g = ghost_id[q, jq, iq, fq]
                        base = ghost_base + g * num_var * nz
                        kstride, vstride = (num_var, 1) if elem_major else (1, nz)
                    nbr[face, col] = (base, kstride, vstride, flags)
    part.nbr = nbr
    part.xform = numpy.stack(xrows) if xrows else numpy.zeros((0, n, 4))

    # what the others need from us, in THEIR ghost order
    send = []
    for peer in range(nranks):
        if peer == rank:
            continue
        mine = [g for g in _ghost_requests(nh, nranks, peer) if g[0] == rank]
        if not mine:
            continue
        part.send_peers.append((peer, len(send), len(mine)))
        for _, q, jq, iq, fq in mine:
            s = local_slab(q, jq)
            send.append((cbase[s] + (jq - slabs[s].j0) * nx + iq) * 4 + fq)
    part.send = numpy.asarray(send, dtype=numpy.int32)
    return part

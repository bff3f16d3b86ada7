"""numpy emulation of the device-side *pull* of neighbour traces through the partition tables
(what face_point() in csrc/euler3d.cu does), used to check the tables on CPU."""

import numpy

from wxfactory_b200.partition import build_partition


def _local_view(part, tr, nvar, nfaces, F):
    """View of the local part of a trace array indexed [var, face, element, F] whatever the layout."""
    E = part.nelems
    if part.trace_layout == "element":
        return tr[: nvar * nfaces * E * F].reshape(E, nfaces, nvar, F).transpose(2, 1, 0, 3)
    return tr[: nvar * nfaces * E * F].reshape(nvar, nfaces, E, F)


def local_traces(prob, part, q_itf_x1, q_itf_x2, nvar, nfaces):
    """Trace array of rank `part` in its layout ("field": [var][face][E][F], "element": [E][face][var][F]; + ghost
    tail) from panel-stacked two-sided traces q_itf_x1/x2: (nvar, 6, nz, nh, nh, 2F)."""
    F = q_itf_x1.shape[-1] // 2
    nz = q_itf_x1.shape[2]
    E = part.nelems
    tr = numpy.zeros((nvar * nfaces * E + part.nghost * nvar * nz) * F)
    loc = _local_view(part, tr, nvar, nfaces, F)
    e0 = 0
    for sl in part.slabs:
        cnt = nz * sl.ny * prob.nh
        a1 = q_itf_x1[:, sl.panel, :, sl.j0 : sl.j1].reshape(nvar, cnt, 2 * F)
        a2 = q_itf_x2[:, sl.panel, :, sl.j0 : sl.j1].reshape(nvar, cnt, 2 * F)
        loc[:, 0, e0 : e0 + cnt] = a1[..., :F]
        loc[:, 1, e0 : e0 + cnt] = a1[..., F:]
        loc[:, 2, e0 : e0 + cnt] = a2[..., :F]
        loc[:, 3, e0 : e0 + cnt] = a2[..., F:]
        e0 += cnt
    return tr


def pack(part, tr, nvar, nfaces, nz, F):
    """Send buffer of `part` (what pack_kernel does): [rec][var][k][F], or [rec][k][var][F] for the element layout."""
    loc = _local_view(part, tr, nvar, nfaces, F)
    cb, eb = [], []
    c = e = 0
    for sl in part.slabs:
        cb.append(c)
        eb.append(e)
        c += sl.ny * part.nh
        e += nz * sl.ny * part.nh
    elem = part.trace_layout == "element"
    out = numpy.zeros((part.nsend, nz, nvar, F) if elem else (part.nsend, nvar, nz, F))
    for r, cf in enumerate(part.send):
        col, f = int(cf) >> 2, int(cf) & 3
        s = max(i for i in range(len(cb)) if col >= cb[i])
        cl = col - cb[s]
        for k in range(nz):
            row = loc[:, f, eb[s] + k * part.slabs[s].ny * part.nh + cl]
            if elem:
                out[r, k] = row
            else:
                out[r, :, k] = row
    return out


def exchange(parts, traces, nvar, nfaces, nz, F):
    """Move packed records into the peers' ghost tails (a record keeps its internal order)."""
    sends = [pack(p, t, nvar, nfaces, nz, F) for p, t in zip(parts, traces)]
    for p, t in zip(parts, traces):
        ghost = t[nvar * nfaces * p.nelems * F :].reshape(p.nghost, nvar * nz * F)
        for peer, first, cnt in p.recv_peers:
            src = parts[peer]
            blk = [x for x in src.send_peers if x[0] == p.rank]
            assert len(blk) == 1 and blk[0][2] == cnt
            ghost[first : first + cnt] = sends[peer][blk[0][1] : blk[0][1] + cnt].reshape(cnt, -1)


def pull(part, tr, col, face, k, nvar, n, F, euler=True):
    """Neighbour trace (nvar, F) facing `face` of local column `col` at level k."""
    base, ks, vs, flags = (int(x) for x in part.nbr[face, col])
    out = numpy.empty((nvar, F))
    for v in range(nvar):
        o = (base + v * vs + k * ks) * F
        out[v] = tr[o : o + F]
    if flags & 1:
        out = out.reshape(nvar, F // n, n)[:, :, ::-1].reshape(nvar, F)
    xid = flags >> 8
    if xid:
        m = part.xform[xid - 1]  # (n, 4) in receiver node order
        a = out.reshape(nvar, F // n, n)
        b1 = m[:, 0] * a[1] + m[:, 1] * a[2]
        b2 = m[:, 2] * a[1] + m[:, 3] * a[2]
        a[1], a[2] = b1, b2
        out = a.reshape(nvar, F)
    return out

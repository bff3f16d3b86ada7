"""numpy emulation of the device-side *pull* of neighbour traces through the partition tables
(what face_point() in csrc/euler3d.cu does), used to check the tables on CPU."""

import numpy

from wxfactory_b200.partition import build_partition


def local_traces(prob, part, q_itf_x1, q_itf_x2, nvar, nfaces):
    """Trace array [var][face][E][F] (+ ghost tail) of rank `part` from panel-stacked two-sided traces
    q_itf_x1/x2: (nvar, 6, nz, nh, nh, 2F)."""
    F = q_itf_x1.shape[-1] // 2
    nz = q_itf_x1.shape[2]
    E = part.nelems
    tr = numpy.zeros((nvar * nfaces * E + part.nghost * nvar * nz) * F)
    loc = tr[: nvar * nfaces * E * F].reshape(nvar, nfaces, E, F)
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
    """Send buffer [rec][var][k][F] of `part` (what pack_kernel does)."""
    E = part.nelems
    loc = tr[: nvar * nfaces * E * F].reshape(nvar, nfaces, E, F)
    cb, eb = [], []
    c = e = 0
    for sl in part.slabs:
        cb.append(c)
        eb.append(e)
        c += sl.ny * part.nh
        e += nz * sl.ny * part.nh
    out = numpy.zeros((part.nsend, nvar, nz, F))
    for r, cf in enumerate(part.send):
        col, f = int(cf) >> 2, int(cf) & 3
        s = max(i for i in range(len(cb)) if col >= cb[i])
        cl = col - cb[s]
        for k in range(nz):
            out[r, :, k] = loc[:, f, eb[s] + k * part.slabs[s].ny * part.nh + cl]
    return out


def exchange(parts, traces, nvar, nfaces, nz, F):
    """Move packed records into the peers' ghost tails."""
    sends = [pack(p, t, nvar, nfaces, nz, F) for p, t in zip(parts, traces)]
    for p, t in zip(parts, traces):
        ghost = t[nvar * nfaces * p.nelems * F :].reshape(p.nghost, nvar, nz, F)
        for peer, first, cnt in p.recv_peers:
            src = parts[peer]
            blk = [x for x in src.send_peers if x[0] == p.rank]
            assert len(blk) == 1 and blk[0][2] == cnt
            ghost[first : first + cnt] = sends[peer][blk[0][1] : blk[0][1] + cnt]


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

"""Synthetic high-resolution cubed-sphere Euler problems (BASELINE.json configs[4]).

The reference's metric set-up does not scale to 10^7..10^8 nodes (numeric Christoffel solve), and it
does not exist on the GPU box.  The throughput workload therefore takes a small problem BUILT BY THE REFERENCE
(tests/golden/dcmip31_n4.npz: DCMIP 3-1 physics, n = 4, 4x4x4 elements per panel) and *tiles* its per-element
metric, face metric and edge coordinates along the three element axes: element (k, j, i) of the big grid uses base
element (k mod bz, j mod bh, i mod bh); face l of an axis with ne elements uses base face (l mod b), the last
one base face b.  Cost-wise the values are irrelevant; what matters is that every array has the real size and the
real access pattern, and that the numbers are physically admissible (positive density/pressure/metric).

Tiling gives a size-independent parity property for free: with a tiled state, the RHS of every element that does
not touch a panel edge, the bottom or the top depends only on its residues (k mod bz, j mod bh, i mod bh) and its
panel, so a full-size device result can be checked element by element against the numpy oracle run on a SMALL tiled
problem (2bz x 2bh x 2bh per panel) -- see ``check_against_small``.

Only the slabs a rank owns are ever materialised (``local_static`` / ``local_state``), so 1/2/4/8-rank runs build
1/N of the arrays each.
"""

import numpy

from .layout import local_static as _dense_local_static
from .problem import EulerProblem


def _face_index(ne, b):
    """base face index of faces 0..ne of an axis with ne elements tiled from b elements."""
    idx = numpy.arange(ne + 1) % b
    idx[ne] = b
    return idx


class TiledEulerProblem:
    """Duck-types the attributes ``RhsB200`` reads from an ``EulerProblem``; arrays are produced per partition."""

    equations = "euler"
    num_var = 5
    hsurf = None

    def __init__(self, base: EulerProblem, nz: int, nh: int):
        if base.equations != "euler":
            raise ValueError("the synthetic sweep is defined for the 3-D Euler equations")
        self.base = base
        self.n, self.nz, self.nh = base.n, int(nz), int(nh)
        self.bz, self.bh = base.nz, base.nh
        self.P, self.F = base.P, base.F
        self.advection_only = base.advection_only
        self.case_number = base.case_number
        self.damp = base.damp
        self.damp_tau0 = base.damp_tau0
        for k in ("ext_w", "ext_e", "diff", "corr", "hfilt"):
            setattr(self, k, getattr(base, k))
        n = self.n
        em = numpy.arange(self.nh) % self.bh
        node = (em[:, None] * n + numpy.arange(n)[None, :]).ravel()
        self.bnd_sn = numpy.ascontiguousarray(base.bnd_sn[:, node])
        self.bnd_we = numpy.ascontiguousarray(base.bnd_we[:, node])

    # ---- sizes ---------------------------------------------------------------------------------------
    @property
    def elems_per_panel(self):
        return self.nz * self.nh * self.nh

    @property
    def panel_shape(self):
        return (5, self.nz, self.nh, self.nh, self.P)

    @property
    def state_shape(self):
        return (6,) + self.panel_shape

    @property
    def num_nodes(self):
        return 6 * self.elems_per_panel * self.P

    @property
    def num_dofs(self):
        return 5 * self.num_nodes

    # ---- tiling helpers ------------------------------------------------------------------------------
    def _kji(self, sl):
        kk = numpy.arange(self.nz) % self.bz
        jj = numpy.arange(sl.j0, sl.j1) % self.bh
        ii = numpy.arange(self.nh) % self.bh
        return kk, jj, ii

    def _node_field(self, part, arr, lead):
        """arr (6, *lead, bz*bh*bh, P) -> (*lead, E_local*P)."""
        a = arr.reshape((6,) + lead + (self.bz, self.bh, self.bh, self.P))
        counts = [self.nz * sl.ny * self.nh * self.P for sl in part.slabs]
        out = numpy.empty(lead + (sum(counts),))  # filled slab by slab: the peak is the result + one slab (24 GB at s = 1)
        off = 0
        for sl, cnt in zip(part.slabs, counts):
            kk, jj, ii = self._kji(sl)
            blk = a[sl.panel][..., kk[:, None, None], jj[None, :, None], ii[None, None, :], :]
            out[..., off : off + cnt] = blk.reshape(lead + (-1,))
            off += cnt
            del blk
        return out

    def local_static(self, part):
        """Same dict as layout.local_static for the slabs of `part`."""
        b = self.base
        out = {
            "ext_w": b.ext_w,
            "ext_e": b.ext_e,
            "diff": b.diff,
            "corr": b.corr,
            "hfilt": b.hfilt,
            "sqrtg": self._node_field(part, b.sqrtg, ()),
            "h": self._node_field(part, b.h, (6,)),
            "chris": self._node_field(part, b.chris, (27,)),
            "inv_dzdeta": self._node_field(part, b.inv_dzdeta, ()),
        }
        if b.damp is not None:
            out["damp"] = self._node_field(part, b.damp, (4,))
        fi, fj, fk = [], [], []
        kf = _face_index(self.nz, self.bz)
        jf_all = _face_index(self.nh, self.bh)
        if_ = _face_index(self.nh, self.bh)
        for sl in part.slabs:
            kk, jj, ii = self._kji(sl)
            jf = jf_all[sl.j0 : sl.j1 + 1]
            fi.append(b.fm_i[sl.panel][:, kk[:, None, None], jj[None, :, None], if_[None, None, :], :].ravel())
            fj.append(b.fm_j[sl.panel][:, kk[:, None, None], jf[None, :, None], ii[None, None, :], :].ravel())
            fk.append(b.fm_k[sl.panel][:, kf[:, None, None], jj[None, :, None], ii[None, None, :], :].ravel())
        out["fm_i"], out["fm_j"], out["fm_k"] = (numpy.concatenate(x) for x in (fi, fj, fk))
        out["nbr"] = numpy.ascontiguousarray(part.nbr, dtype=numpy.int32)
        out["xform"] = numpy.ascontiguousarray(part.xform, dtype=numpy.float64)
        out["send"] = numpy.ascontiguousarray(part.send, dtype=numpy.int32)
        return {k: (v if v.dtype == numpy.int32 else numpy.ascontiguousarray(v, dtype=numpy.float64)) for k, v in out.items()}

    def local_state(self, part, base_Q):
        """Tile a base state (6, 5, bz, bh, bh, P) into this rank's flat local state."""
        q = numpy.asarray(base_Q).reshape((6, 5, self.bz, self.bh, self.bh, self.P))
        pieces = []
        for sl in part.slabs:
            kk, jj, ii = self._kji(sl)
            pieces.append(q[sl.panel][:, kk[:, None, None], jj[None, :, None], ii[None, None, :], :].ravel())
        return numpy.concatenate(pieces)

    # ---- small dense twin for the oracle ---------------------------------------------------------------
    def materialize(self) -> EulerProblem:
        """The whole tiled problem as a dense EulerProblem (use at oracle-sized grids only)."""
        b = self.base

        class _All:  # a one-rank partition covering the six panels
            slabs = None

        from .partition import slabs_of_rank

        allp = _All()
        allp.slabs = slabs_of_rank(self.nh, 1, 0)
        E, P = self.elems_per_panel, self.P

        def per_panel(arr, lead):
            a = self._node_field(allp, arr, lead).reshape(lead + (6, E, P))
            return numpy.ascontiguousarray(numpy.moveaxis(a, len(lead), 0))

        kk = numpy.arange(self.nz) % self.bz
        hh = numpy.arange(self.nh) % self.bh
        kf, hf = _face_index(self.nz, self.bz), _face_index(self.nh, self.bh)
        ix = numpy.ix_
        fm_i = numpy.stack([b.fm_i[p][ix(range(4), kk, hh, hf)] for p in range(6)])
        fm_j = numpy.stack([b.fm_j[p][ix(range(4), kk, hf, hh)] for p in range(6)])
        fm_k = numpy.stack([b.fm_k[p][ix(range(4), kf, hh, hh)] for p in range(6)])
        return EulerProblem(
            n=self.n,
            nz=self.nz,
            nh=self.nh,
            advection_only=b.advection_only,
            case_number=b.case_number,
            ext_w=b.ext_w,
            ext_e=b.ext_e,
            diff=b.diff,
            corr=b.corr,
            hfilt=b.hfilt,
            sqrtg=per_panel(b.sqrtg, ()),
            h=per_panel(b.h, (6,)),
            chris=per_panel(b.chris, (27,)),
            inv_dzdeta=per_panel(b.inv_dzdeta, ()),
            fm_i=fm_i,
            fm_j=fm_j,
            fm_k=fm_k,
            bnd_sn=self.bnd_sn,
            bnd_we=self.bnd_we,
            damp=None if b.damp is None else per_panel(b.damp, (4,)),
            damp_tau0=b.damp_tau0,
        )

    def global_state(self, base_Q):
        from .partition import slabs_of_rank

        class _All:
            slabs = slabs_of_rank(self.nh, 1, 0)

        return self.local_state(_All, base_Q).reshape(self.state_shape)


def base_state(base_Q0, seed=20261104, amplitude=1e-3):
    """A non-trivial admissible state on the base block: the reference's initial state times (1 + amplitude*N(0,1)),
    with small random momenta added so that no term of the RHS vanishes identically."""
    rng = numpy.random.default_rng(seed)
    Q = numpy.array(base_Q0, dtype=numpy.float64, copy=True)
    Q *= 1.0 + amplitude * rng.standard_normal(Q.shape)
    rho = Q[:, 0]
    Q[:, 1] += rho * 1e-7 * rng.standard_normal(rho.shape)  # contravariant u1, u2 are O(1e-6) rad/s-scaled
    Q[:, 2] += rho * 1e-7 * rng.standard_normal(rho.shape)
    Q[:, 3] += rho * 1e-1 * rng.standard_normal(rho.shape)
    return Q


def base_direction(base_Q, seed=20261105):
    """JVP direction: standard normal x 1e-3 x per-variable RMS of the state."""
    rng = numpy.random.default_rng(seed)
    v = rng.standard_normal(base_Q.shape)
    for var in range(base_Q.shape[1]):
        rms = numpy.sqrt(numpy.mean(base_Q[:, var] ** 2))
        v[:, var] *= 1e-3 * (rms if rms > 0 else 1.0)
    return v


def residue_map(ne, b):
    """For an axis of ne elements tiled from b: (positions that are not on the boundary, the position in a 2b-element
    small twin that has the same residue and is not on its boundary either)."""
    pos = numpy.arange(1, ne - 1)
    r = pos % b
    small = numpy.where(r == 0, b, r)
    return pos, small


def check_against_small(tiled: TiledEulerProblem, part, local_out, small_out):
    """Largest per-variable relative deviation (L2 over all compared elements, worst of the 5 variables) between a full-size local result
    and the small twin's result of the same residue class, over every element that is not on a panel edge, the
    bottom or the top.  `local_out` flat local array (numpy or torch, any device); `small_out` the oracle's
    (6, 5, 2bz, 2bh, 2bh, P) result.  Returns (max relative error, number of elements compared)."""
    import torch

    out = local_out if isinstance(local_out, torch.Tensor) else torch.from_numpy(numpy.ascontiguousarray(local_out))
    dev = out.device
    small = torch.from_numpy(numpy.ascontiguousarray(small_out)).to(dev)
    nz, nh, P = tiled.nz, tiled.nh, tiled.P
    kpos, ksm = residue_map(nz, tiled.bz)
    ipos, ism = residue_map(nh, tiled.bh)
    num = torch.zeros(5, dtype=torch.float64, device=dev)
    den = torch.zeros(5, dtype=torch.float64, device=dev)
    off = 0
    count = 0
    t = lambda a: torch.from_numpy(numpy.ascontiguousarray(a)).to(dev)  # noqa: E731
    for sl in part.slabs:
        cnt = 5 * nz * sl.ny * nh * P
        blk = out[off : off + cnt].view(5, nz, sl.ny, nh, P)
        off += cnt
        jglob = numpy.arange(sl.j0, sl.j1)
        keep = (jglob >= 1) & (jglob <= nh - 2)
        if not keep.any():
            continue
        jl = numpy.nonzero(keep)[0]
        r = jglob[keep] % tiled.bh
        jsm = numpy.where(r == 0, tiled.bh, r)
        got = blk[:, t(kpos)][:, :, t(jl)][:, :, :, t(ipos)]
        want = small[sl.panel][:, t(ksm)][:, :, t(jsm)][:, :, :, t(ism)]
        num += ((got - want) ** 2).sum(dim=(1, 2, 3, 4))
        den += (want**2).sum(dim=(1, 2, 3, 4))
        count += len(kpos) * len(jl) * len(ipos)
    if count == 0:
        return 0.0, 0
    per_var = torch.sqrt(num / torch.clamp(den, min=1e-300))
    return float(per_var.max()), count


# ---------------------------------------------------------------------------------------------------------------
# Shallow water (BASELINE.json configs[0]/[1]: case6 / galewsky "at high element count")
class TiledSwProblem:
    """Tiled shallow-water problem: element (j, i) of the big panel uses base element (j mod bh, i mod bh) of a
    problem built by the reference (tests/golden/case6.npz or galewsky.npz, no topography).  Duck-types what
    ``RhsB200`` reads from a ``SwProblem``."""

    equations = "shallow_water"
    num_var = 3
    hsurf = None

    def __init__(self, base, nh: int):
        if base.equations == "euler" or base.hsurf is not None:
            raise ValueError("the shallow-water tiling takes a shallow-water base problem without topography")
        self.base = base
        self.n, self.nh, self.bh = base.n, int(nh), base.nh
        self.P, self.F = base.P, base.F
        self.case_number = base.case_number
        for k in ("ext_w", "ext_e", "diff", "corr"):
            setattr(self, k, getattr(base, k))
        n = self.n
        em = numpy.arange(self.nh) % self.bh
        node = (em[:, None] * n + numpy.arange(n)[None, :]).ravel()
        self.bnd_sn = numpy.ascontiguousarray(base.bnd_sn[:, node])
        self.bnd_we = numpy.ascontiguousarray(base.bnd_we[:, node])

    @property
    def elems_per_panel(self):
        return self.nh * self.nh

    @property
    def panel_shape(self):
        return (3, self.nh, self.nh, self.P)

    @property
    def state_shape(self):
        return (6,) + self.panel_shape

    @property
    def num_nodes(self):
        return 6 * self.elems_per_panel * self.P

    @property
    def num_dofs(self):
        return 3 * self.num_nodes

    def _ji(self, sl):
        return numpy.arange(sl.j0, sl.j1) % self.bh, numpy.arange(self.nh) % self.bh

    def _node_field(self, part, arr, lead):
        a = arr.reshape((6,) + lead + (self.bh, self.bh, self.P))
        out = []
        for sl in part.slabs:
            jj, ii = self._ji(sl)
            out.append(a[sl.panel][..., jj[:, None], ii[None, :], :].reshape(lead + (-1,)))
        return numpy.ascontiguousarray(numpy.concatenate(out, axis=-1))

    def local_static(self, part):
        b = self.base
        out = {
            "ext_w": b.ext_w,
            "ext_e": b.ext_e,
            "diff": b.diff,
            "corr": b.corr,
            "sqrtg": self._node_field(part, b.sqrtg, ()),
            "h": self._node_field(part, b.h, (3,)),
            "chris": self._node_field(part, b.chris, (8,)),
        }
        hf = _face_index(self.nh, self.bh)
        fi, fj = [], []
        for sl in part.slabs:
            jj, ii = self._ji(sl)
            jf = hf[sl.j0 : sl.j1 + 1]
            fi.append(b.fm_i[sl.panel][:, jj[:, None], hf[None, :], :].ravel())
            fj.append(b.fm_j[sl.panel][:, jf[:, None], ii[None, :], :].ravel())
        out["fm_i"], out["fm_j"] = numpy.concatenate(fi), numpy.concatenate(fj)
        out["nbr"] = numpy.ascontiguousarray(part.nbr, dtype=numpy.int32)
        out["xform"] = numpy.ascontiguousarray(part.xform, dtype=numpy.float64)
        out["send"] = numpy.ascontiguousarray(part.send, dtype=numpy.int32)
        return {k: (v if v.dtype == numpy.int32 else numpy.ascontiguousarray(v, dtype=numpy.float64)) for k, v in out.items()}

    def local_state(self, part, base_Q):
        q = numpy.asarray(base_Q).reshape((6, 3, self.bh, self.bh, self.P))
        pieces = []
        for sl in part.slabs:
            jj, ii = self._ji(sl)
            pieces.append(q[sl.panel][:, jj[:, None], ii[None, :], :].ravel())
        return numpy.concatenate(pieces)

    def _all(self):
        from .partition import slabs_of_rank

        class _All:
            slabs = slabs_of_rank(self.nh, 1, 0)

        return _All

    def materialize(self):
        """The whole tiled problem as a dense SwProblem (oracle-sized grids only)."""
        from .problem import SwProblem

        b, allp = self.base, self._all()
        E, P = self.elems_per_panel, self.P

        def per_panel(arr, lead):
            a = self._node_field(allp, arr, lead).reshape(lead + (6, E, P))
            return numpy.ascontiguousarray(numpy.moveaxis(a, len(lead), 0))

        hh = numpy.arange(self.nh) % self.bh
        hf = _face_index(self.nh, self.bh)
        ix = numpy.ix_
        return SwProblem(
            n=self.n,
            nh=self.nh,
            case_number=b.case_number,
            ext_w=b.ext_w,
            ext_e=b.ext_e,
            diff=b.diff,
            corr=b.corr,
            sqrtg=per_panel(b.sqrtg, ()),
            h=per_panel(b.h, (3,)),
            chris=per_panel(b.chris, (8,)),
            fm_i=numpy.stack([b.fm_i[p][ix(range(3), hh, hf)] for p in range(6)]),
            fm_j=numpy.stack([b.fm_j[p][ix(range(3), hf, hh)] for p in range(6)]),
            bnd_sn=self.bnd_sn,
            bnd_we=self.bnd_we,
        )

    def global_state(self, base_Q):
        return self.local_state(self._all(), base_Q).reshape(self.state_shape)


def sw_base_state(base_Q0, seed=20261106, amplitude=1e-3):
    """Admissible perturbed shallow-water state on the base block (h > 0 kept by the multiplicative noise)."""
    rng = numpy.random.default_rng(seed)
    Q = numpy.array(base_Q0, dtype=numpy.float64, copy=True)
    Q *= 1.0 + amplitude * rng.standard_normal(Q.shape)
    h = Q[:, 0]
    Q[:, 1] += h * 1e-7 * rng.standard_normal(h.shape)
    Q[:, 2] += h * 1e-7 * rng.standard_normal(h.shape)
    return Q


def check_sw_against_small(tiled: TiledSwProblem, part, local_out, small_out):
    """Shallow-water counterpart of ``check_against_small``: elements not on a panel edge are compared with the
    2bh x 2bh twin's element of the same residue class.  Returns (worst per-variable relative L2 error, count)."""
    import torch

    out = local_out if isinstance(local_out, torch.Tensor) else torch.from_numpy(numpy.ascontiguousarray(local_out))
    dev = out.device
    small = torch.from_numpy(numpy.ascontiguousarray(small_out)).to(dev)
    nh, P = tiled.nh, tiled.P
    ipos, ism = residue_map(nh, tiled.bh)
    num = torch.zeros(3, dtype=torch.float64, device=dev)
    den = torch.zeros(3, dtype=torch.float64, device=dev)
    off = count = 0
    t = lambda a: torch.from_numpy(numpy.ascontiguousarray(a)).to(dev)  # noqa: E731
    for sl in part.slabs:
        cnt = 3 * sl.ny * nh * P
        blk = out[off : off + cnt].view(3, sl.ny, nh, P)
        off += cnt
        jglob = numpy.arange(sl.j0, sl.j1)
        keep = (jglob >= 1) & (jglob <= nh - 2)
        if not keep.any():
            continue
        jl = numpy.nonzero(keep)[0]
        r = jglob[keep] % tiled.bh
        jsm = numpy.where(r == 0, tiled.bh, r)
        got = blk[:, t(jl)][:, :, t(ipos)]
        want = small[sl.panel][:, t(jsm)][:, :, t(ism)]
        num += ((got - want) ** 2).sum(dim=(1, 2, 3))
        den += (want**2).sum(dim=(1, 2, 3))
        count += len(jl) * len(ipos)
    if count == 0:
        return 0.0, 0
    return float(torch.sqrt(num / torch.clamp(den, min=1e-300)).max()), count

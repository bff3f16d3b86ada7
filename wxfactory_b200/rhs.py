"""Host-side mirror of the reference's RHS objects for the cubed sphere, backed by libwxb200.so.

Drop-in surface (same names, argument meaning and error behaviour as the reference):

* ``RhsB200.__call__(q) -> dq/dt`` / ``.full(q)``: ``RHS.__call__`` (wx_factory/rhs/rhs.py:75-125) and
  ``RhsShallowWater.__call__`` (wx_factory/rhs/rhs_sw.py:38-56).  ``q`` may have any shape of the right size
  (Krylov solvers pass flattened vectors, wx_factory/solvers/matvec.py:65); the result comes back in the
  caller's shape as a fresh array the caller owns.
* ``.timings / .retrieve_last_times() / .clear_timings() / .print_times()``: wx_factory/rhs/rhs.py:68-73,190-213.
* ``matvec_fun(vec, dt, Q, rhs, rhs_handle, method)`` and ``matvec_rat`` / ``MatvecOpRat`` with the reference's
  positional signatures (wx_factory/solvers/matvec.py:36-88); when ``rhs_handle`` is a ``RhsB200`` the fused
  device JVP is used, otherwise the generic formula.
* ``RhsBundleB200``: ``RhsBundle`` (wx_factory/rhs/rhs_selector.py:18-84) built from the reference's own
  geometry / operators / metric objects.

Arrays: numpy ``float64`` in -> numpy out (host path: H2D, kernels, D2H inside the call: what an unmodified
reference integrator sees); anything exporting ``__dlpack__`` on the CUDA device (e.g. a torch tensor) in ->
torch CUDA tensor out (device-resident path).  complex128 input (the reference's complex-step JVP) raises
``TypeError``: this path implements the finite-difference JVP only (BASELINE.json north_star).
"""

import ctypes
import math
import os

import numpy

from . import capi
from .layout import local_state, local_static
from .partition import build_partition

EPS_FD = math.sqrt(numpy.finfo(numpy.float32).eps)  # wx_factory/solvers/matvec.py:64


class _DevPtr:
    """Expose a raw device pointer through __cuda_array_interface__ so torch can view it (no copy)."""

    def __init__(self, ptr, count):
        self.__cuda_array_interface__ = {
            "shape": (int(count),),
            "typestr": "<f8",
            "data": (int(ptr), False),
            "version": 3,
            "strides": None,
        }


class RhsB200:
    """dQ/dt = RHS(Q) on the cubed sphere for one rank of a 1/2/4/8-way partition."""

    def __init__(self, problem, nranks=1, rank=0, device=0, process_group=None, defer_comm=False):
        self.prob = problem
        self.nranks, self.rank = int(nranks), int(rank)
        self.device_index = int(device)
        self.group = process_group
        euler = problem.equations == "euler"
        self.num_var = problem.num_var
        # Euler: the fused column-sweep kernel (csrc/euler_sweep.cu, default) or the traces + main pair of
        # csrc/euler3d.cu (WXB_EULER_PATH=split); read by wxb_create
        self._euler = euler
        # trace fields per face point: the Euler kernels carry 6 (csrc/euler3d.cu, WXB_TRACE_FIELDS)
        self.num_trace = 6 if euler else 3
        nz = problem.nz if euler else 1
        self.part = build_partition(
            problem.nh,
            nz,
            problem.n,
            problem.bnd_sn,
            problem.bnd_we,
            nranks=self.nranks,
            rank=self.rank,
            num_var=self.num_trace,
            nfaces=6 if euler else 4,
            trace_layout="element" if euler else "field",  # what csrc/euler3d.cu / csrc/sw2d.cu index
        )
        self._lib = capi.load()
        cfg = capi.WxbConfig()
        cfg.equations = capi.EQ_EULER3D if euler else capi.EQ_SW
        cfg.n, cfg.nz, cfg.nx = problem.n, nz, problem.nh
        cfg.nslabs = len(self.part.slabs)
        if cfg.nslabs > capi.WXB_MAX_SLABS:
            raise ValueError(f"too many slabs ({cfg.nslabs})")
        for s, sl in enumerate(self.part.slabs):
            cfg.slab_ny[s] = sl.ny
        cfg.nghost, cfg.nsend = self.part.nghost, self.part.nsend
        cfg.nxform = int(self.part.xform.shape[0])
        cfg.advection_only = int(bool(getattr(problem, "advection_only", False)))
        cfg.has_damping = int(euler and problem.damp is not None)
        cfg.has_topo = int((not euler) and problem.hsurf is not None)
        cfg.damp_tau0 = float(getattr(problem, "damp_tau0", 25.0))
        make_static = getattr(problem, "local_static", None)  # synthetic.TiledEulerProblem builds its slabs lazily
        self._static = make_static(self.part) if make_static else local_static(problem, self.part)  # keep alive during create
        arrs = capi.WxbStaticArrays()
        for name, typ in capi._ARRAY_FIELDS:
            a = self._static.get(name)
            setattr(arrs, name, capi.ptr(a, typ) if a is not None and a.size else typ())
        h = ctypes.c_void_p()
        code = self._lib.wxb_create(ctypes.byref(cfg), ctypes.byref(arrs), self.device_index, ctypes.byref(h))
        capi.check(self._lib, None, code, "wxb_create")
        self._h = h
        self._static = None
        # one kernel path per handle, chosen at create (csrc/wxb_api.cu: build): fused (Euler "sweep", shallow water "tile")
        # or the traces + main pair ("split")
        fused = bool(self._lib.wxb_handle_fused(self._h))
        self._fused = (fused, fused)  # (RHS, JVP)
        self.sweep = fused
        self.path_name = ("sweep" if euler else "tile") if fused else "split"
        self.size = int(self._lib.wxb_state_size(self._h))
        self.timestamps = []
        self.timings = []
        self._have_timing = False
        self.launches = 0  # kernels launched by this object (bench.py's gpu_launches)
        self._torch = None
        self._send_t = self._recv_t = None
        self._comm_stream = None
        # multi-rank handles: the halo exchange runs inside the library (wxb_comm_init / wxb_eval_dist: NCCL send / recv on
        # the handle's own stream).  WXB_HALO=torch keeps the torch.distributed point-to-point exchange of halo.py.
        self._lib_comm = False
        if not defer_comm:
            self.init_comm()

    def init_comm(self):
        """Collective over the ranks of the handle's group (the constructor calls it unless ``defer_comm``: a caller that
        staggers the construction of its ranks -- the set-up holds the static arrays on the host -- joins afterwards)."""
        if self.nranks > 1 and not self._lib_comm and os.environ.get("WXB_HALO", "lib") != "torch":
            self._init_lib_comm()

    # ---- lifetime ---------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            self._lib.wxb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    # ---- helpers ----------------------------------------------------------------------------------
    @property
    def shape(self):
        return (self.size,)

    def local_state(self, Qglobal):
        """Extract this rank's flat state from a (6, num_var, ...) global array (for a tiled synthetic problem:
        from the base block's state)."""
        tiled = getattr(self.prob, "local_state", None)
        if tiled is not None:
            return tiled(self.part, Qglobal)
        return local_state(self.prob, self.part, Qglobal)

    def _check(self, code, what):
        capi.check(self._lib, self._h, code, what)

    def _kernels_per_eval(self, jvp=False):
        adv = self.prob.equations == "euler" and self.prob.advection_only
        if adv:
            return 0
        multi = (1 if self.part.nsend else 0) + (1 if self.part.nghost else 0)
        if self._fused[1 if jvp else 0]:  # one fused launch (two on a multi-rank handle: inner, outer tiles) + the edge traces
            return 1 + multi
        return 2 + multi  # traces (+ pack) and main (inner, outer)

    def _tm(self):
        if self._torch is None:
            import torch

            if not torch.cuda.is_available():
                raise RuntimeError("wxfactory_b200 needs a CUDA device: there is no CPU fallback")
            self._torch = torch
        return self._torch

    def _as_device(self, x, name):
        """Return (flat contiguous float64 CUDA tensor, original shape) of a device array-like."""
        torch = self._tm()
        if not isinstance(x, torch.Tensor):
            x = torch.from_dlpack(x)
        if x.is_complex():
            raise TypeError("complex state: the B200 path implements the finite-difference JVP only (jacobian_method = fd)")
        if x.dtype != torch.float64:
            raise TypeError(f"{name} must be float64, got {x.dtype}")
        if not x.is_cuda or x.device.index != self.device_index:
            raise ValueError(f"{name} must live on cuda:{self.device_index}")
        if x.numel() != self.size:
            raise ValueError(f"{name} has {x.numel()} values, this RHS expects {self.size}")
        return x.contiguous().view(-1), tuple(x.shape)

    def _as_host(self, x, name):
        if numpy.iscomplexobj(x):
            raise TypeError("complex state: the B200 path implements the finite-difference JVP only (jacobian_method = fd)")
        a = numpy.ascontiguousarray(x, dtype=numpy.float64)
        if a.size != self.size:
            raise ValueError(f"{name} has {a.size} values, this RHS expects {self.size}")
        return a

    def _host_result(self):
        """A fresh numpy array the caller owns, backed by page-locked memory from torch's caching host allocator
        (so the D2H copy runs at full PCIe speed and a later H2D of the same array -- e.g. as the JVP's cached
        rhs -- does too)."""
        torch = self._tm()
        return torch.empty(self.size, dtype=torch.float64, pin_memory=True).numpy()

    def _to_host(self, t):
        """Device result -> numpy (through a reusable pinned buffer; synchronises)."""
        torch = self._tm()
        out = torch.empty(t.numel(), dtype=torch.float64, pin_memory=True)  # torch's caching host allocator: a page-locked
        out.copy_(t.view(-1), non_blocking=True)                           # block is re-used once its last owner let go
        torch.cuda.current_stream(t.device).synchronize()
        return out.numpy()

    def timing_history(self, max_evals=256):
        """Device durations of the most recent evaluations, oldest first: array (k, 3) in ms
        [traces(+pack), main (incl. the halo wait on a multi-rank handle), total] and array (k,) of modes
        (-1 RHS, 0 FD-JVP, 1 rational JVP).  Synchronises."""
        ms = (ctypes.c_float * (3 * max_evals))()
        md = (ctypes.c_int * max_evals)()
        k = self._lib.wxb_timing_history(self._h, max_evals, ctypes.byref(ms), ctypes.byref(md))
        if k < 0:
            self._check(k, "wxb_timing_history")
        return numpy.array(ms[: 3 * k], dtype=numpy.float64).reshape(k, 3), numpy.array(md[:k], dtype=numpy.int64)

    # ---- halo exchange between ranks ---------------------------------------------------------------------
    def _init_lib_comm(self):
        """Join the library's own NCCL communicator: rank 0 makes the unique id, torch.distributed ships it."""
        import torch.distributed as dist

        torch = self._tm()
        if not (dist.is_available() and dist.is_initialized()) or dist.get_backend(self.group) != "nccl":
            return  # CPU / gloo jobs (tests) keep the torch exchange
        idbuf = (ctypes.c_char * capi.WXB_NCCL_ID_BYTES)()
        if self.rank == 0:
            self._check(self._lib.wxb_comm_unique_id(ctypes.byref(idbuf)), "wxb_comm_unique_id")
        box = [bytes(idbuf.raw)]
        dist.broadcast_object_list(box, src=dist.get_global_rank(self.group, 0) if self.group is not None else 0, group=self.group)
        idbuf.raw = box[0]

        def lists(peers):
            arr = numpy.asarray(peers, dtype=numpy.int32).reshape(-1, 3)
            return len(peers), [numpy.ascontiguousarray(arr[:, c]) for c in range(3)]

        ns, (sp, sf, sc) = lists(self.part.send_peers)
        nr, (rp, rf, rc) = lists(self.part.recv_peers)
        torch.cuda.set_device(self.device_index)
        self._check(
            self._lib.wxb_comm_init(self._h, ctypes.byref(idbuf), self.rank, self.nranks, ns, sp.ctypes.data, sf.ctypes.data, sc.ctypes.data,
                                    nr, rp.ctypes.data, rf.ctypes.data, rc.ctypes.data),
            "wxb_comm_init",
        )
        self._lib_comm = True

    def _exchange(self, stream):
        torch = self._tm()
        if self._send_t is None:
            ns, nr = int(self._lib.wxb_send_size(self._h)), int(self._lib.wxb_recv_size(self._h))
            self._send_t = torch.as_tensor(_DevPtr(self._lib.wxb_send_buffer(self._h), ns), device=f"cuda:{self.device_index}")
            self._recv_t = torch.as_tensor(_DevPtr(self._lib.wxb_recv_buffer(self._h), nr), device=f"cuda:{self.device_index}")
            euler = self.prob.equations == "euler"
            self._rec = self.num_trace * (self.prob.nz if euler else 1) * self.prob.F
        from .halo import exchange

        with torch.cuda.stream(stream):
            exchange(self.part, self._send_t, self._recv_t, self._rec, self.group)

    def _host_entry_points(self):
        """The library's host entry points (chunked upload / sweeps / download, csrc/wxb_api.cu: pipelined_eval) serve a
        handle without remote neighbours, and a multi-rank handle once the library owns the exchange (init_comm): the
        call is then collective over the ranks like every evaluation."""
        return (self.part.nghost == 0 and self.part.nsend == 0) or bool(self._lib_comm)

    # ---- evaluation -------------------------------------------------------------------------------
    def _eval_device(self, q, v, rhs0, dt, eps, mode):
        torch = self._tm()
        out = torch.empty(self.size, dtype=torch.float64, device=q.device)
        stream = torch.cuda.current_stream(q.device)
        sp = ctypes.c_void_p(stream.cuda_stream)
        vp = ctypes.c_void_p(v.data_ptr()) if v is not None else None
        rp = ctypes.c_void_p(rhs0.data_ptr()) if rhs0 is not None else None
        if self.part.nghost == 0 and self.part.nsend == 0:
            if mode < 0:
                self._check(self._lib.wxb_rhs(self._h, q.data_ptr(), out.data_ptr(), sp), "wxb_rhs")
            else:
                self._check(self._lib.wxb_jvp(self._h, q.data_ptr(), vp, rp, dt, eps, mode, out.data_ptr(), sp), "wxb_jvp")
        elif self._lib_comm:
            self._check(self._lib.wxb_eval_dist(self._h, q.data_ptr(), vp, rp, dt, eps, mode, out.data_ptr(), sp), "wxb_eval_dist")
        else:
            if self._comm_stream is None:
                self._comm_stream = torch.cuda.Stream(device=q.device)
            self._check(self._lib.wxb_traces(self._h, q.data_ptr(), vp, eps, sp), "wxb_traces")
            cs = self._comm_stream
            cs.wait_stream(stream)
            self._exchange(cs)  # overlaps with the inner elements below
            self._check(self._lib.wxb_finish(self._h, q.data_ptr(), vp, rp, dt, eps, mode, 1, out.data_ptr(), sp), "wxb_finish")
            stream.wait_stream(cs)
            self._check(self._lib.wxb_finish(self._h, q.data_ptr(), vp, rp, dt, eps, mode, 2, out.data_ptr(), sp), "wxb_finish")
        self.launches += self._kernels_per_eval(mode >= 0)
        self._have_timing = True
        return out

    def __call__(self, q):
        if len(self.timestamps) > 0:
            self.retrieve_last_times()
        if isinstance(q, numpy.ndarray):
            a = self._as_host(q, "q")
            if not self._host_entry_points():  # torch-side exchange: H2D, evaluate, D2H
                torch = self._tm()
                qd = torch.from_numpy(a.reshape(-1)).to(f"cuda:{self.device_index}", non_blocking=True)
                res = self._eval_device(qd, None, None, 0.0, 0.0, -1)
                self.timestamps = [True]
                return self._to_host(res).reshape(q.shape)
            out = self._host_result()
            self._check(self._lib.wxb_rhs_host(self._h, a.ctypes.data, out.ctypes.data), "wxb_rhs_host")
            self.launches += self._kernels_per_eval()
            self._have_timing = True
            self.timestamps = [True]
            return out.reshape(q.shape)
        qd, shape = self._as_device(q, "q")
        out = self._eval_device(qd, None, None, 0.0, 0.0, -1)
        self.timestamps = [True]
        return out.view(shape)

    def full(self, q):
        return self.__call__(q)

    def jvp(self, vec, dt, Q, rhs0, mode="fd", eps=EPS_FD):
        """Fused FD Jacobian-vector product: dt*(RHS(Q+eps*vec)-rhs0)/eps ('fd') or vec - 0.5*that ('rat').
        Returns a flattened array like the reference's matvec_fun / matvec_rat."""
        m = {"fd": capi.JVP_FD, "rat": capi.JVP_RAT}.get(mode)
        if m is None:
            raise ValueError(f"unknown JVP mode {mode!r} (complex-step is not available on this path)")
        if isinstance(Q, numpy.ndarray):
            q, v, r = self._as_host(Q, "Q"), self._as_host(vec, "vec"), self._as_host(rhs0, "rhs")
            if not self._host_entry_points():
                torch = self._tm()
                dev = f"cuda:{self.device_index}"
                qd, vd, rd = (torch.from_numpy(x.reshape(-1)).to(dev, non_blocking=True) for x in (q, v, r))
                return self._to_host(self._eval_device(qd, vd, rd, float(dt), float(eps), m))
            out = self._host_result()
            self._check(
                self._lib.wxb_jvp_host(self._h, q.ctypes.data, v.ctypes.data, r.ctypes.data, float(dt), float(eps), m, out.ctypes.data),
                "wxb_jvp_host",
            )
            self.launches += self._kernels_per_eval(True)
            return out.reshape(-1)
        q, _ = self._as_device(Q, "Q")
        v, _ = self._as_device(vec, "vec")
        r, _ = self._as_device(rhs0, "rhs")
        return self._eval_device(q, v, r, float(dt), float(eps), m)

    # ---- linearisation point kept on the device (MatvecOp, wx_factory/solvers/matvec.py:7-34,71-73) --
    def linearize(self, Q, dt, mode="fd", eps=EPS_FD):
        """``rhs = self(Q)`` plus the operator ``vec -> matvec_fun(vec, dt, Q, rhs, self)`` (``mode="rat"``:
        ``matvec_rat``) in one call: returns a ``MatvecOpB200`` whose ``.rhs`` is RHS(Q) in Q's shape.  With a numpy
        ``Q`` on a single-rank handle, Q and RHS(Q) stay on the device, so every later application moves only
        ``vec`` up and the product down (the integrators build one operator per time step and apply it once per
        Krylov iteration: wx_factory/integrators/ros2.py:29)."""
        if mode not in ("fd", "rat"):
            raise ValueError(f"unknown JVP mode {mode!r} (complex-step is not available on this path)")
        if isinstance(Q, numpy.ndarray) and self._host_entry_points():
            a = self._as_host(Q, "Q")
            out = self._host_result()
            self._check(self._lib.wxb_linearize_host(self._h, a.ctypes.data, out.ctypes.data), "wxb_linearize_host")
            self.launches += self._kernels_per_eval()
            self._have_timing = True
            self.timestamps = [True]
            self._lin_token = getattr(self, "_lin_token", 0) + 1
            return MatvecOpB200(self, dt, Q, out.reshape(Q.shape), mode, eps, token=self._lin_token)
        if isinstance(Q, numpy.ndarray):
            # multi-rank handle, host arrays: Q goes up once, RHS(Q) comes down once and BOTH stay on the device with the
            # operator, so that every application moves only vec up and the product down (2 uploads per step, not 4)
            torch = self._tm()
            qd = torch.from_numpy(self._as_host(Q, "Q").reshape(-1)).to(f"cuda:{self.device_index}", non_blocking=True)
            rd = self._eval_device(qd, None, None, 0.0, 0.0, -1)
            self.timestamps = [True]
            op = MatvecOpB200(self, dt, Q, self._to_host(rd).reshape(Q.shape), mode, eps, token=None)
            op._dev = (qd, rd)
            return op
        return MatvecOpB200(self, dt, Q, self(Q), mode, eps, token=None)

    def _bind(self, Q, rhs0):
        q, r = self._as_host(Q, "Q"), self._as_host(rhs0, "rhs")
        self._check(self._lib.wxb_bind_host(self._h, q.ctypes.data, r.ctypes.data), "wxb_bind_host")
        self._lin_token = getattr(self, "_lin_token", 0) + 1
        return self._lin_token

    def _jvp_bound(self, vec, dt, mode, eps):
        v = self._as_host(vec, "vec")
        out = self._host_result()
        m = capi.JVP_FD if mode == "fd" else capi.JVP_RAT
        self._check(self._lib.wxb_jvp_bound_host(self._h, v.ctypes.data, float(dt), float(eps), m, out.ctypes.data), "wxb_jvp_bound_host")
        self.launches += self._kernels_per_eval(True)
        return out

    # ---- timing vocabulary of the reference (rhs.py:68-73,190-213) ---------------------------------
    def clear_timings(self):
        self.timestamps = []
        self.timings = []

    def retrieve_last_times(self):
        if not self._have_timing:
            return
        ms = (ctypes.c_float * 3)()
        if self._lib.wxb_timings(self._h, ctypes.byref(ms)) == 0:
            self.timings.append([ms[0] * 1e-3, ms[1] * 1e-3, ms[2] * 1e-3])

    def print_times(self):
        if not self.timings:
            return
        t = numpy.array(self.timings) * 1000.0
        print(
            f"RHS times (device):\n  Traces (extrapolation+pack): {t[:, 0].mean():7.3f} ms\n"
            f"  Main (flux, Riemann, lift, forcing): {t[:, 1].mean():7.3f} ms\n  Total: {t[:, 2].mean():7.3f} ms"
        )


def matvec_fun(vec, dt, Q, rhs, rhs_handle, method="fd"):
    """wx_factory/solvers/matvec.py:36-68 with the same positional signature (FD only)."""
    if method == "complex":
        raise TypeError("complex-step JVP is not available on the B200 path: set jacobian_method = fd")
    owner = getattr(rhs_handle, "__self__", rhs_handle)
    if isinstance(owner, (RhsB200, _PerRankRhs)):
        return owner.jvp(vec, dt, Q, rhs, mode="fd")
    Qvec = Q + EPS_FD * vec.reshape(Q.shape)
    jac = dt * (rhs_handle(Qvec) - rhs) / EPS_FD
    return jac.flatten()


def matvec_rat(vec, dt, Q, rhs, rhs_handle):
    """wx_factory/solvers/matvec.py:76-88."""
    owner = getattr(rhs_handle, "__self__", rhs_handle)
    if isinstance(owner, (RhsB200, _PerRankRhs)):
        return owner.jvp(vec, dt, Q, rhs, mode="rat")
    Qvec = Q + EPS_FD * vec.reshape(Q.shape)
    jac = dt * (rhs_handle(Qvec) - rhs) / EPS_FD
    return vec - 0.5 * jac.flatten()


class MatvecOpB200:
    """``MatvecOp`` (wx_factory/solvers/matvec.py:7-27: ``dtype``, ``shape``, ``size``, ``__call__(vec)``) for a
    ``RhsB200``.  With host arrays (single-rank handle, or a multi-rank one whose exchange the library owns) the
    linearisation point (Q, rhs) is uploaded once -- at
    ``RhsB200.linearize`` or at the first application -- and kept on the device; the operator is then a snapshot of
    (Q, rhs) at that moment (the reference's closure would see later in-place changes of Q; its integrators build
    a new operator every step instead).  Device arrays and multi-rank handles go through ``RhsB200.jvp``."""

    def __init__(self, owner, dt, Q, rhs_vec, mode="fd", eps=EPS_FD, token=None):
        self.owner, self.dt, self.Q, self.rhs, self.mode, self.eps = owner, dt, Q, rhs_vec, mode, eps
        self.dtype, self.shape = Q.dtype, tuple(Q.shape)
        self.size = int(math.prod(self.shape))
        self._token = token
        self._host = isinstance(Q, numpy.ndarray) and isinstance(rhs_vec, numpy.ndarray) and owner._host_entry_points()

    def matvec(self, vec):
        return self(vec)

    _dev = None  # (Q, RHS(Q)) kept on the device by RhsB200.linearize on a multi-rank handle

    def __call__(self, vec):
        o = self.owner
        if self._dev is not None and isinstance(vec, numpy.ndarray):
            torch = o._tm()
            vd = torch.from_numpy(o._as_host(vec, "vec").reshape(-1)).to(self._dev[0].device, non_blocking=True)
            m = capi.JVP_FD if self.mode == "fd" else capi.JVP_RAT
            return o._to_host(o._eval_device(self._dev[0], vd, self._dev[1], float(self.dt), float(self.eps), m))
        if not (self._host and isinstance(vec, numpy.ndarray)):
            return o.jvp(vec, self.dt, self.Q, self.rhs, mode=self.mode, eps=self.eps)
        if self._token is None or self._token != getattr(o, "_lin_token", 0):
            self._token = o._bind(self.Q, self.rhs)  # another operator took the handle's slot in between
        return o._jvp_bound(vec, self.dt, self.mode, self.eps)


class MatvecOpRat(MatvecOpB200):
    """wx_factory/solvers/matvec.py:71-73 (same constructor); falls back to the generic formula for a foreign RHS."""

    def __init__(self, dt, Q, rhs_vec, rhs_handle):
        owner = getattr(rhs_handle, "__self__", rhs_handle)
        self.rhs_handle = rhs_handle
        if isinstance(owner, RhsB200):
            super().__init__(owner, dt, Q, rhs_vec, mode="rat")
        else:
            self.owner, self.dt, self.Q, self.rhs = None, dt, Q, rhs_vec
            self.dtype, self.shape = Q.dtype, tuple(Q.shape)
            self.size = int(math.prod(self.shape))

    @property
    def rhs_vec(self):
        return self.rhs

    def __call__(self, vec):
        if self.owner is None:
            return matvec_rat(vec, self.dt, self.Q, self.rhs, self.rhs_handle)
        return super().__call__(vec)


class MatvecOpFd(MatvecOpRat):
    """The finite-difference operator ``vec -> matvec_fun(vec, dt, Q, rhs, rhs_handle, "fd")`` the EPI / KIOPS
    integrators wrap in a lambda (wx_factory/integrators/epi.py), as an object with the same constructor as
    ``MatvecOpRat``."""

    def __init__(self, dt, Q, rhs_vec, rhs_handle):
        super().__init__(dt, Q, rhs_vec, rhs_handle)
        self.mode = "fd"

    def __call__(self, vec):
        if self.owner is None:
            return matvec_fun(vec, self.dt, self.Q, self.rhs, self.rhs_handle, "fd")
        return MatvecOpB200.__call__(self, vec)


class _PerRankRhs:
    """``RhsBundle.full`` of ONE rank of the reference's 6-rank MPI job when the whole sphere lives on one B200 owned by
    rank 0: ``rhs(q_local)`` gathers the six local states on rank 0 (``comm.gather``), evaluates there through the C
    ABI and scatters dQ/dt back, so every rank gets its own panel's result with ``RHS.__call__``'s contract (any shape
    in, same shape out, fresh array; wx_factory/rhs/rhs.py:75-122).  Collective: every rank must call it the same number
    of times, as with the reference (wx_factory/rhs/rhs_euler.py:180-181)."""

    def __init__(self, comm, owner, local_size, root=0):
        self.comm, self.owner, self.root, self.local_size = comm, owner, root, int(local_size)
        self.rank = comm.Get_rank()
        self.timings, self.timestamps = [], []

    def _collective(self, arrays, fn):
        """Gather the tuple of local arrays of every rank, apply fn(list of stacked arrays) on the root, scatter."""
        shape = arrays[0].shape
        for x in arrays:
            if numpy.iscomplexobj(x):
                raise TypeError("complex state: the B200 path implements the finite-difference JVP only (jacobian_method = fd)")
            if x.size != self.local_size:
                raise ValueError(f"array has {x.size} values, this rank's RHS expects {self.local_size}")
        parts = self.comm.gather(tuple(numpy.asarray(x, dtype=numpy.float64).reshape(-1) for x in arrays), root=self.root)
        out = None
        if self.rank == self.root:
            stacked = [numpy.stack([p[i] for p in parts]) for i in range(len(arrays))]
            res = numpy.asarray(fn(*stacked)).reshape(len(parts), -1)
            out = [res[r].copy() for r in range(len(parts))]
        return self.comm.scatter(out, root=self.root).reshape(shape)

    def __call__(self, q):
        return self._collective((q,), lambda Q: self.owner(Q))

    def full(self, q):
        return self(q)

    def jvp(self, vec, dt, Q, rhs0, mode="fd", eps=EPS_FD):
        vec = numpy.asarray(vec)
        out = self._collective((vec.reshape(numpy.shape(Q)), Q, rhs0), lambda V, QQ, R: self.owner.jvp(V, dt, QQ, R, mode=mode, eps=eps))
        return out.reshape(-1)

    def retrieve_last_times(self):
        if self.rank == self.root:
            self.owner.retrieve_last_times()
            self.timings = self.owner.timings

    def clear_timings(self):
        self.timings, self.timestamps = [], []
        if self.rank == self.root:
            self.owner.clear_timings()

    def print_times(self):
        if self.rank == self.root:
            self.owner.print_times()


class RhsBundleB200:
    """wx_factory/rhs/rhs_selector.py:18-84: ``.full`` is the RHS callable handed to every integrator
    (wx_factory/simulation/simulation.py:336-395).  Two ways to build it:

    * ``RhsBundleB200(geom, operators, metric, topo, ptopo, param, fields_shape)`` -- the reference's own argument list,
      called by EVERY rank of its 6-rank MPI job where ``RhsBundle(...)`` stands (simulation.py:111).  The static
      arrays are gathered on rank 0 (``from_reference.gather_to_one_process``, ``ptopo.comm``), which owns the device
      handle; ``.full(q_local)`` is collective (gather states, evaluate, scatter).
    * ``RhsBundleB200(ranks, device=0)`` -- all six per-panel objects (geom, ops, metric, config[, topo]) in one
      process; ``.full`` takes the stacked state ``(6, nvar, ...)`` or anything of that size.
    """

    rhs_class = None  # the class that owns the device handle (RhsB200); tests substitute an oracle-backed stand-in

    def __init__(self, geom, operators=None, metric=None, topo=None, ptopo=None, param=None, fields_shape=None, debug=False, device=0, comm=None):
        from .from_reference import gather_to_one_process, problem_from_reference

        make = self.rhs_class or RhsB200

        ops0 = None
        if operators is None and param is None:  # all six panels' objects in this process
            ranks = geom
            eq = str(ranks[0].config.equations).lower()
            if eq not in ("euler", "shallow_water"):
                raise ValueError(f"Unrecognized combination of equations ({eq}) and geometry")
            self.problem = problem_from_reference(ranks)
            self.owner = make(self.problem, device=device)
            self.full = self.owner
            self.shape = self.problem.state_shape
            ops0 = getattr(ranks[0], "ops", None)
        else:  # one rank of the reference's 6-rank job
            eq = str(param.equations).lower()
            if eq not in ("euler", "shallow_water"):
                raise ValueError(f"Unrecognized combination of equations ({eq}) and geometry ({type(geom)})")
            if comm is None:
                comm = ptopo.comm
            self.shape = tuple(fields_shape)
            self.problem = gather_to_one_process(geom, operators, metric, param, topo, comm=comm)
            self.owner = make(self.problem, device=device) if self.problem is not None else None
            self.full = _PerRankRhs(comm, self.owner, int(numpy.prod(self.shape)))
            ops0 = operators
        # the per-step exponential filter of the same operators object, when the configuration applies one
        # (DFROperators.expfilter_apply, wx_factory/geometry/operators.py:101-119)
        self.filter = None
        if eq == "euler" and getattr(ops0, "expfilter_apply", False) and self.owner is not None:
            from .stepops import ExpFilterB200

            self.filter = ExpFilterB200(self.owner, numpy.asarray(ops0.expfilter, dtype=float))

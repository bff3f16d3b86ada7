"""Run the UNMODIFIED reference (WxFactory, /root/reference) in this container -- TEST INFRASTRUCTURE ONLY.

Used to (1) validate the numpy restatement in ``oracle/`` and (2) generate the golden
fixtures under ``tests/golden/`` (see ``make_golden.py``).  Nothing here travels to the
GPU box at run time: ``/root/reference`` does not exist there.

What it does, mirroring ``Simulation.__init__`` (wx_factory/simulation/simulation.py:52-131)
without the output manager (netCDF4 absent) and without the integrator:

    device -> ProcessTopology -> CubedSphere2D/3D -> DFROperators -> init_state_vars -> RhsBundle

on 6*N^2 thread-ranks (``mpi4py`` stand-in in this directory), with a no-op stub for
``compiler.compile_kernels`` (its import-time ``get_processor_name()`` raises on this
box: wx_factory/compiler/compile_kernels.py:51-72; the cubed-sphere PDE never touches
the compiled module: wx_factory/pde/pde_euler_cubesphere.py:65-66).
"""

import os
import sys
import threading
import types

REF_ROOT = os.environ.get("WX_REFERENCE_ROOT", "/root/reference")
_HERE = os.path.dirname(os.path.abspath(__file__))


def available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "wx_factory"))


_bootstrapped = False


def bootstrap():
    """Put the stand-ins and the reference on sys.path (idempotent)."""
    global _bootstrapped
    if _bootstrapped:
        return
    if not available():
        raise RuntimeError(f"reference not found at {REF_ROOT}")
    sys.path.insert(0, _HERE)  # our mpi4py stand-in wins
    sys.path.insert(1, os.path.join(REF_ROOT, "wx_factory"))
    stub = types.ModuleType("compiler.compile_kernels")
    stub.compile = lambda *a, **k: None
    stub.load_module = lambda *a, **k: None
    pkg = types.ModuleType("compiler")
    pkg.compile_kernels = stub
    pkg.__path__ = []
    sys.modules["compiler"] = pkg
    sys.modules["compiler.compile_kernels"] = stub
    _bootstrapped = True


class RankState:
    """Everything one reference rank owns."""

    def __init__(self):
        self.config = None
        self.device = None
        self.ptopo = None
        self.geom = None
        self.ops = None
        self.metric = None
        self.topo = None
        self.Q = None
        self.rhs = None


def run_ranks(nranks, fn):
    """Run fn(rank) on `nranks` threads acting as MPI ranks; returns list of results. Re-raises."""
    bootstrap()
    from mpi4py import MPI

    MPI._set_world(nranks)
    out = [None] * nranks
    err = [None] * nranks

    def work(r):
        MPI._set_rank(r)
        try:
            out[r] = fn(r)
        except BaseException as e:  # noqa: BLE001
            import traceback

            err[r] = traceback.format_exc()

    threads = [threading.Thread(target=work, args=(r,), daemon=True) for r in range(nranks)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for r, e in enumerate(err):
        if e is not None:
            raise RuntimeError(f"rank {r} failed:\n{e}")
    return out


def make_config(ini_text: str, overrides: dict | None = None):
    bootstrap()
    cwd = os.getcwd()
    os.chdir(REF_ROOT)
    try:
        from common import Configuration, ConfigurationSchema, default_schema_path, readfile

        schema = ConfigurationSchema(readfile(default_schema_path))
        config = Configuration(ini_text, schema)
    finally:
        os.chdir(cwd)
    for k, v in (overrides or {}).items():
        setattr(config, k, v)
        if k == "num_elements_horizontal":
            config.num_elements_horizontal_total = v
    return config


def read_ini(name: str) -> str:
    with open(os.path.join(REF_ROOT, "config", name)) as f:
        return f.read()


def build(ini_text: str, nranks: int = 6, overrides: dict | None = None, ic_patch=None):
    """Build the reference objects for every rank. Returns list[RankState]."""
    import copy
    import math

    base_config = make_config(ini_text, overrides)

    def per_rank(r):
        from mpi4py import MPI
        from device import CpuDevice
        from geometry import CubedSphere2D, CubedSphere3D, DFROperators
        from init.init_state_vars import init_state_vars
        from process_topology import ProcessTopology
        from rhs.rhs_selector import RhsBundle

        st = RankState()
        cfg = copy.deepcopy(base_config)
        lines = int(math.isqrt(nranks // 6))
        cfg.num_elements_horizontal = cfg.num_elements_horizontal_total // lines  # simulation.py:209-223
        st.config = cfg
        st.device = CpuDevice(comm=MPI.COMM_WORLD)
        st.ptopo = ProcessTopology(st.device, comm=MPI.COMM_WORLD)
        if cfg.equations == "shallow_water":
            st.geom = CubedSphere2D(
                cfg.num_elements_horizontal, cfg.num_solpts, cfg.lambda0, cfg.phi0, cfg.alpha0, st.ptopo, cfg, st.device
            )
        else:
            st.geom = CubedSphere3D(
                cfg.num_elements_horizontal,
                cfg.num_elements_vertical,
                cfg.num_solpts,
                cfg.lambda0,
                cfg.phi0,
                cfg.alpha0,
                cfg.ztop,
                st.ptopo,
                cfg,
                st.device,
            )
        st.ops = DFROperators(st.geom, cfg, st.device)
        if ic_patch is not None:
            ic_patch()
        st.Q, st.topo, st.metric = init_state_vars(st.geom, st.ops, cfg)
        st.rhs = RhsBundle(st.geom, st.ops, st.metric, st.topo, st.ptopo, cfg, st.Q.shape)
        return st

    return run_ranks(nranks, per_rank)


def eval_rhs(states, Qs):
    """Evaluate the reference RHS collectively: Qs[r] is rank r's state. Returns list of arrays."""
    return run_ranks(len(states), lambda r: states[r].rhs.full(Qs[r]))


def eval_matvec_fd(states, vs, dt, Qs, rhs0s):
    """Reference FD Jacobian-vector product (wx_factory/solvers/matvec.py:36-68, method='fd')."""
    bootstrap()
    from solvers.matvec import matvec_fun

    return run_ranks(
        len(states), lambda r: matvec_fun(vs[r], dt, Qs[r], rhs0s[r], states[r].rhs.full, method="fd")
    )


def eval_matvec_rat(states, vs, dt, Qs, rhs0s):
    """Reference Ros2 system matrix product (wx_factory/solvers/matvec.py:76-88)."""
    bootstrap()
    from solvers.matvec import matvec_rat

    return run_ranks(len(states), lambda r: matvec_rat(vs[r], dt, Qs[r], rhs0s[r], states[r].rhs.full))

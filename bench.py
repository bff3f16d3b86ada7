#!/usr/bin/env python
"""Throughput benchmark of the B200 cubed-sphere DG right-hand side + fused FD Jacobian-vector product.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

Metric (BASELINE.json): FP64 DOF-updates/s.  One *step* = one RHS evaluation dQ/dt = RHS(Q) followed by one fused
finite-difference Jacobian-vector product J(Q) v around it (what an EPI/KIOPS or Rosenbrock/FGMRES integrator does per
Krylov vector), i.e. 2 x (5 x nodes) DOF-updates.  Workload = BASELINE.json configs[4], the synthetic
high-resolution Euler state at the largest single-GPU size of SURVEY.md section 8(d), S2 s = 1: n = 4, 24 x 96 x 96
elements per panel (85 M nodes, 425 M DOFs, ~30 GB resident) at one GPU and nh ~ 96*sqrt(N) at N so that the elements per
GPU stay fixed (weak scaling; s = 2 at N = 8): metric, operators and state of a reference-built DCMIP 3-1 problem tiled
along the element axes (wxfactory_b200/synthetic.py).  The four .ini configs of BASELINE.json hold 16 k .. 146 k DOFs --
microseconds of HBM time, launch-bound -- and are parity-test cases, not bench lines.

Protocol: W warm-up steps, then --repeats (5) timed loops of exactly K steps each, every loop bracketed by
barrier + synchronize and timed with CUDA events, max over ranks; `value` / `ms_per_step` are the MEDIAN loop.  A further
back-to-back loop of >= 2 s gives the `sustained` object (clocks sampled separately).  `sweep` holds the smaller S2 sizes
(s = 1/4, 1/2) at N = 1; `strong` (N > 1) the fixed s = 1 problem cut N ways in the same run.

JSON line keys (one line, rank 0): the base contract's, plus
  roofline      algorithmic HBM bytes of one RHS evaluation (B_rhs(n) = 360 + 96/n bytes per node) over the
                evaluation's device time (the sweep kernel, or traces + main on the split path), measured with the C
                ABI's own CUDA events on the launching stream for every RHS launch of the timed loops (median = burst
                figure; `sustained` = mean over the last evaluations of the >= 2 s loop);
                roofline_jvp is the same for the JVP launches (B_jvp(n) = 440 + 96/n).
  e2e           the same step through the public API with HOST numpy arrays (pinned): H2D of every input and D2H
                of every result inside the timed region.
  cpu_baseline  the numpy oracle (bit-identical restatement of the reference's numpy RHS) on a bounded sample of the
                same workload (same tiling at 8 x 8 x 8 elements per panel), one single-threaded replica per core.
  parity        full-size check of the device result against the oracle through the tiling periodicity property.

`--impl reference` times only the CPU arm (rank 0; other ranks exit) and prints the same line with "impl": "reference".
"""

import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BASE_FIXTURE = os.path.join(ROOT, "tests", "golden", "dcmip31_n4.npz")
METRIC = "DG RHS+JVP DOF-updates/s FP64"
UNIT = "DOF-updates/s"
SMALL_NZ, SMALL_NH = 8, 8  # the oracle-sized twin of the workload (2 x the base block per axis)
L2_BYTES = 126 * 2**20


def nh_for(ngpus, nh1):
    """Elements per panel side that keep 6*nh^2/N (elements per GPU) ~fixed and 6*nh divisible by N."""
    nh = nh1 * math.sqrt(ngpus)
    step = 4
    cands = [int(round(nh / step)) * step + d for d in (-4, 0, 4)]
    cands = [c for c in cands if c > 0 and (6 * c) % ngpus == 0]
    return min(cands, key=lambda c: abs(c - nh))


def b_rhs(n):
    return 360.0 + 96.0 / n  # bytes per node per RHS evaluation


def b_jvp(n):
    return 440.0 + 96.0 / n  # + v and the cached base RHS


# DRAM bytes per node and evaluation actually moved by the Euler kernels (dram__bytes_read.sum + dram__bytes_write.sum,
# n = 4) come from an `ncu --set full` capture summarised under profiles/ and recorded in profiles/ncu_traffic.json
# together with a hash of the kernel sources (scripts/ncu_traffic.py).  A capture made from other sources is stale: the
# line then says "traffic": null.
KERNEL_SOURCES = ["euler_sweep.cu", "euler3d.cu", "fastmath.cuh", "wxb_internal.cuh"]


def kernel_source_hash():
    import hashlib

    h = hashlib.sha256()
    for f in KERNEL_SOURCES:
        with open(os.path.join(ROOT, "wxfactory_b200", "csrc", f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def ncu_traffic(path_name):
    """(bytes per node {rhs, jvp}, source note) for the Euler path in use, or (None, why)."""
    fn = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(fn):
        return None, "no profiles/ncu_traffic.json"
    with open(fn) as f:
        rec = json.load(f)
    ent = rec.get(path_name)
    if not ent:
        return None, f"no capture of the {path_name} path"
    if ent.get("source_hash") != kernel_source_hash():
        return None, f"stale: {ent.get('report')} was captured from other kernel sources"
    return {"rhs": ent["rhs_bytes"] / ent["nodes"], "jvp": ent["jvp_bytes"] / ent["nodes"]}, ent.get("report")


# ------------------------------------------------------------------------------------------------------------
# CPU arm: the numpy oracle on the small twin, one single-threaded replica per core (spawned before CUDA exists)
_W = {}


def _cpu_init():
    import numpy

    from oracle import jvp_oracle
    from wxfactory_b200 import synthetic as syn
    from wxfactory_b200.problem import load_problem

    numpy.seterr(all="ignore")
    base, ex = load_problem(BASE_FIXTURE)
    Qb = syn.base_state(ex["Q0"])
    vb = syn.base_direction(Qb)
    small = syn.TiledEulerProblem(base, SMALL_NZ, SMALL_NH)
    _W["orc"] = jvp_oracle.make_oracle(small.materialize())
    _W["Q"] = small.global_state(Qb)
    _W["v"] = small.global_state(vb).ravel()
    _W["dt"] = float(ex["dt"])
    _W["jvp"] = jvp_oracle.matvec_fd


def _cpu_step(_):
    t = time.perf_counter()
    r0 = _W["orc"].rhs(_W["Q"])
    _W["jvp"](_W["orc"].rhs, _W["v"], _W["dt"], _W["Q"], r0)
    return time.perf_counter() - t


def cpu_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_cpu_arm(steps, warmup, max_workers=64):
    """Returns dict(value, ms_per_step, cores, sample, dofs)."""
    import multiprocessing as mp

    workers = max(1, min(cpu_cores(), max_workers))
    saved = {k: os.environ.get(k) for k in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS")}
    for k in saved:
        os.environ[k] = "1"
    try:
        ctx = mp.get_context("spawn")
        with ctx.Pool(workers, initializer=_cpu_init) as pool:
            for _ in range(max(1, warmup)):
                pool.map(_cpu_step, range(workers), chunksize=1)
            t0 = time.perf_counter()
            for _ in range(steps):
                pool.map(_cpu_step, range(workers), chunksize=1)
            dt = time.perf_counter() - t0
    finally:
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    dofs = 5 * 6 * SMALL_NZ * SMALL_NH * SMALL_NH * 64
    return {
        "value": workers * 2 * dofs * steps / dt,
        "ms_per_step": 1e3 * dt / steps,
        "cores": workers,
        "dofs": dofs,
        "sample": (
            f"{workers} single-threaded replicas of the numpy oracle, each {steps} steps (RHS + FD-JVP) on the same "
            f"tiled workload at {SMALL_NZ}x{SMALL_NH}x{SMALL_NH} elements/panel (n=4, {dofs} DOFs)"
        ),
    }


# ------------------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clock / throttle-reason samples while a timed region runs (B200_PROFILING.md recipe)."""

    FIELDS = (
        "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    )

    def __init__(self, gpus):
        self.gpus = set(gpus)
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE,
                stderr=subprocess.DEVNULL,
                text=True,
            )
        except OSError:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._pump, daemon=True)
        self.thread.start()

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def summary(self, windows):
        """windows: list of (t0, t1) perf_counter intervals that were timed."""
        sm, smax, reasons, power = [], 0.0, set(), 0.0
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, line in self.rows:
            if not any(a <= t <= b for a, b in windows):
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                if int(f[0]) not in self.gpus:
                    continue
                sm.append(float(f[1]))
                smax = max(smax, float(f[2]))
                power = max(power, float(f[3]))
            except ValueError:
                continue
            for nm, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {
            "sm_mhz": statistics.median(sm) if sm else None,
            "sm_max_mhz": smax or None,
            "reasons": sorted(reasons),
            "samples": len(sm),
            "power_w_max": power or None,
        }


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------------------
def _median(xs):
    return float(statistics.median(xs))


class Workload:
    """One tiled Euler problem on this rank: handle, device-resident and pinned host copies of the state / direction."""

    def __init__(self, base, ex, nz, nh, world, rank, local_rank, host=True):
        import torch

        from wxfactory_b200 import RhsB200
        from wxfactory_b200 import synthetic as syn

        self.nz, self.nh = nz, nh
        self.tiled = syn.TiledEulerProblem(base, nz, nh)
        t0 = time.perf_counter()
        # a rank holds ~34 GB of host arrays while its handle is created (s = 1): at most 4 ranks do so at a time
        self.rhs = None
        for g in range(0, world, 4):
            if g <= rank < g + 4:
                self.rhs = RhsB200(self.tiled, nranks=world, rank=rank, device=local_rank, defer_comm=True)
            if world > 4:
                import torch.distributed as dist

                dist.barrier()
        self.rhs.init_comm()
        self.size = self.rhs.size
        self.nodes_local = self.size // 5
        Qb = syn.base_state(ex["Q0"])
        vb = syn.base_direction(Qb)
        self.Qb, self.vb = Qb, vb
        dev = torch.device("cuda", local_rank)
        if host:  # host copies live in pinned memory: they are what the reference's integrators would hold
            pin = lambda: torch.empty(self.size, dtype=torch.float64, pin_memory=True)  # noqa: E731
            self.q_h, self.v_h = pin(), pin()
            self.q_h.numpy()[:] = self.rhs.local_state(Qb)
            self.v_h.numpy()[:] = self.rhs.local_state(vb)
            self.q_d, self.v_d = self.q_h.to(dev), self.v_h.to(dev)
        else:
            self.q_d = torch.from_numpy(self.rhs.local_state(Qb)).to(dev)
            self.v_d = torch.from_numpy(self.rhs.local_state(vb)).to(dev)
        self.dt = float(ex["dt"])
        self.setup_s = time.perf_counter() - t0

    def step_device(self):
        out = self.rhs(self.q_d)
        return self.rhs.jvp(self.v_d, self.dt, self.q_d, out, mode="fd")

    def close(self):
        self.rhs.close()
        self.q_d = self.v_d = self.q_h = self.v_h = None


def run_b200(args):
    import numpy
    import torch
    import torch.distributed as dist

    from wxfactory_b200 import EPS_FD
    from wxfactory_b200 import synthetic as syn
    from wxfactory_b200.problem import load_problem

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the B200 path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def timed_loop(fn, steps):
        """Exactly `steps` calls between two CUDA events, barrier + synchronize on both sides; ms, max over ranks."""
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        w0 = time.perf_counter()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        return max_over_ranks(e0.elapsed_time(e1)), (w0 - 0.05, time.perf_counter() + 0.05)

    # ---- workload -----------------------------------------------------------------------------------------
    base, ex = load_problem(BASE_FIXTURE)
    n = base.n
    nz = args.nz
    nh = args.nh if args.nh > 0 else nh_for(world, args.nh1)
    wl = Workload(base, ex, nz, nh, world, rank, local_rank)
    rhs, tiled, size, nodes_local, dt = wl.rhs, wl.tiled, wl.size, wl.nodes_local, wl.dt
    q_d, v_d, q_h, v_h = wl.q_d, wl.v_d, wl.q_h, wl.v_h
    r0_h = torch.empty(size, dtype=torch.float64, pin_memory=True)
    total_nodes = tiled.num_nodes
    total_dofs = tiled.num_dofs
    state_bytes = size * 8
    path_name = rhs.path_name

    # ---- full-size parity through the tiling periodicity property (untimed) ---------------------------------
    small_ref = None

    def parity_of(w):
        """rhs / jvp of workload w against the oracle on the 8x8x8 twin, over every comparable element."""
        nonlocal small_ref
        from oracle import jvp_oracle

        if small_ref is None:
            small = syn.TiledEulerProblem(base, SMALL_NZ, SMALL_NH)
            orc = jvp_oracle.make_oracle(small.materialize())
            Qs, vs = small.global_state(w.Qb), small.global_state(w.vb)
            with numpy.errstate(all="ignore"):
                want_r = orc.rhs(Qs)
                want_j = jvp_oracle.matvec_fd(orc.rhs, vs.ravel(), w.dt, Qs, want_r).reshape(Qs.shape)
            small_ref = (want_r, want_j)
        want_r, want_j = small_ref
        r = w.rhs(w.q_d)
        j = w.rhs.jvp(w.v_d, w.dt, w.q_d, r, mode="fd")
        torch.cuda.synchronize()
        e_r, cnt = syn.check_against_small(w.tiled, w.rhs.part, r, want_r)
        e_j, _ = syn.check_against_small(w.tiled, w.rhs.part, j, want_j)
        # the FD formula amplifies RHS rounding by dt/eps: state the JVP error relative to dt*|rhs|/eps
        jscale = w.dt * float(numpy.linalg.norm(want_r)) / EPS_FD / max(float(numpy.linalg.norm(want_j)), 1e-300)
        e_r, e_j, cnt = max_over_ranks(e_r), max_over_ranks(e_j / jscale), int(max_over_ranks(float(cnt)))
        return {
            "rhs_rel_err": e_r,
            "jvp_err_over_dt_rhs_eps": e_j,
            "elements_checked_per_rank": cnt,
            "tolerance": 1e-12,
            "ok": bool(e_r <= 1e-12 and e_j <= 1e-12),
            "ranks": world,
            "how": "every element off the panel edges / bottom / top vs the numpy oracle on the 8x8x8 twin (tiling periodicity)",
        }

    parity = None if args.skip_check else parity_of(wl)
    rhs0_d = rhs(q_d)
    r0_h.copy_(rhs0_d)
    torch.cuda.synchronize()

    host_call_s = [0.0, 0.0]  # wall time inside the two public calls of the end-to-end step (both return synchronised)

    def step_host():
        t0 = time.perf_counter()
        op = rhs.linearize(q_h.numpy(), dt, mode="fd")  # RHS(Q) comes back; (Q, RHS(Q)) stay as the operator's point
        t1 = time.perf_counter()
        jv = op(v_h.numpy())
        host_call_s[0] += t1 - t0
        host_call_s[1] += time.perf_counter() - t1
        return op.rhs, jv

    sampler = ClockSampler(range(world)) if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.3)

    # ---- device-resident timed loops: --repeats loops of exactly --steps steps, the median loop is the headline ------
    for _ in range(args.warmup):
        wl.step_device()
    launches0 = rhs.launches
    loop_ms, loop_windows, t_rhs, t_jvp = [], [], [], []
    for _ in range(args.repeats):
        ms, win = timed_loop(wl.step_device, args.steps)
        loop_ms.append(ms)
        loop_windows.append(win)
        hist, modes = rhs.timing_history(min(256, 2 * args.steps))
        t_rhs.append(hist[modes < 0])
        t_jvp.append(hist[modes >= 0])
    launches = (rhs.launches - launches0) // args.repeats
    ms_dev = _median(loop_ms)
    t_rhs, t_jvp = numpy.concatenate(t_rhs), numpy.concatenate(t_jvp)

    # ---- sustained: one back-to-back loop of >= --sustain-s seconds ---------------------------------------------------
    sustained = None
    sus_win = None
    if args.sustain_s > 0:
        sus_steps = max(args.steps, int(math.ceil(args.sustain_s * 1e3 / (ms_dev / args.steps))))
        ms_sus, sus_win = timed_loop(wl.step_device, sus_steps)
        hist, modes = rhs.timing_history(256)
        sustained = {"steps": sus_steps, "seconds": ms_sus * 1e-3, "ms_per_step": ms_sus / sus_steps,
                     "value": 2.0 * total_dofs * sus_steps / (ms_sus * 1e-3), "unit": UNIT,
                     "_rhs_ms": float(hist[modes < 0][:, 2].mean()), "_jvp_ms": float(hist[modes >= 0][:, 2].mean())}

    # ---- end to end through the public API with host arrays ------------------------------------------------
    e2e_steps = max(3, min(args.steps, args.e2e_steps))
    for _ in range(2):
        step_host()
    host_call_s[0] = host_call_s[1] = 0.0
    ms_e2e, e2e_win = timed_loop(step_host, e2e_steps)
    if sampler:
        time.sleep(0.2)
        sampler.stop()

    peak, peak_src = measured_peaks()
    traffic, traffic_src = ncu_traffic(path_name) if n == 4 else (None, "captured at n = 4 only")
    kernels = {k: (["euler_sweep<4>"] if rhs._fused[i] else ["euler_traces<4>", "euler_main<4>"]) for i, k in enumerate(("rhs", "jvp"))}
    line = None
    if rank == 0:
        def roof(tt, bytes_per_node, what, key, sus_ms):
            if len(tt) == 0:
                return None
            ev_ms = _median(tt[:, 2])
            ach = bytes_per_node * nodes_local / (ev_ms * 1e-3) / 1e9
            out = {
                "bound": "hbm",
                "achieved": ach,
                "peak": peak,
                "unit": "GB/s",
                "frac": ach / peak,
                "traffic": traffic[key] * nodes_local if traffic else None,
                "traffic_source": traffic_src,
                "kernels": kernels[key],
                "launch_ms": {"traces": _median(tt[:, 0]), "main": _median(tt[:, 1]), "evaluation": ev_ms},
                "algorithmic_bytes_per_node": bytes_per_node,
                "nodes_per_launch": nodes_local,
                "launches_timed": int(len(tt)),
                "peak_source": peak_src,
                "what": what + " (burst: median over the launches of the timed loops)",
            }
            if sus_ms is not None:
                a2 = bytes_per_node * nodes_local / (sus_ms * 1e-3) / 1e9
                out["sustained"] = {"achieved": a2, "frac": a2 / peak, "evaluation_ms": sus_ms,
                                    "what": f"mean over the last evaluations of the {sustained['seconds']:.1f} s back-to-back loop"}
            return out

        clocks = sampler.summary(loop_windows + [e2e_win]) if sampler else None
        if sustained is not None:
            rhs_sus, jvp_sus = sustained.pop("_rhs_ms"), sustained.pop("_jvp_ms")
            sustained["clocks"] = sampler.summary([sus_win]) if sampler else None
        else:
            rhs_sus = jvp_sus = None
        line = {
            "metric": METRIC,
            "value": 2.0 * total_dofs * args.steps / (ms_dev * 1e-3),
            "unit": UNIT,
            "n_gpus": world,
            "steps": args.steps,
            "warmup": args.warmup,
            "ms_per_step": ms_dev / args.steps,
            "higher_is_better": True,
            "scaling": "weak",
            "vs_baseline": None,
            "dtype": "f64",
            "data": "synthetic (reference-built DCMIP 3-1 n=4 metric/state tiled along the element axes)",
            "config": {
                "workload": f"configs[4] synthetic cubed-sphere Euler: n=4, {nz}x{nh}x{nh} elements/panel x 6 panels",
                "num_solpts": n,
                "elements_per_panel": [nz, nh, nh],
                "nodes": total_nodes,
                "dofs": total_dofs,
                "step": "1 RHS + 1 fused FD-JVP = 2 x dofs DOF-updates",
                "kernel_path": path_name,
                "partition": f"{world} rank(s), contiguous rows of element columns; face-trace halo over NCCL p2p",
                "l2": f"state {state_bytes / 2**20:.0f} MiB + metric {nodes_local * 35 * 8 / 2**20:.0f} MiB per GPU > L2 ({L2_BYTES >> 20} MiB): no flush needed",
                "protocol": f"{args.repeats} loops of {args.steps} steps, median loop reported; all loops ms: " + ", ".join(f"{m:.3f}" for m in loop_ms),
                "parity_ok": None if parity is None else parity["ok"],
            },
            "e2e": {
                "value": 2.0 * total_dofs * e2e_steps / (ms_e2e * 1e-3),
                "unit": UNIT,
                "h2d_bytes_per_step": 2 * state_bytes,
                "d2h_bytes_per_step": 2 * state_bytes,
                "steps": e2e_steps,
                "ms_per_step": ms_e2e / e2e_steps,
                "host_call_ms": {"linearize": host_call_s[0] * 1e3 / e2e_steps, "apply": host_call_s[1] * 1e3 / e2e_steps,
                                 "what": "rank 0's wall time inside the two calls of a step (a multi-rank call ends when the slowest neighbour has sent its records)"},
                "how": (
                    "op = RhsB200.linearize(Q numpy, dt) [Q up, RHS(Q) down, both kept on the device as the operator's "
                    "linearisation point, as MatvecOp captures them]; op(v numpy) [v up, J v down] -> wxb_linearize_host / "
                    "wxb_jvp_bound_host, state moved in chunks of levels on 3 streams; pinned host arrays in, numpy out"
                    if world == 1
                    else "the same two calls on every rank (collective): level chunks up, the records other ranks need formed and exchanged "
                    "chunk by chunk over NCCL inside the library, sweeps and downloads behind them (wxb_api.cu: pipelined_eval)"
                    if world <= 2 and os.environ.get("WXB_HOST_DIST", "pipe") == "pipe" or os.environ.get("WXB_HOST_DIST") == "pipe"
                    else "the same two calls on every rank (collective): whole arrays up, wxb_eval_dist (exchange overlapped with the inner "
                    "tiles), whole result down -- with more than two GPUs copying at once the host side of the box is the limit and "
                    "overlapping the two directions buys nothing (DESIGN section 6; WXB_HOST_DIST=pipe forces the level pipeline)"
                ),
            },
            "gpu_launches": int(launches),
            "roofline": roof(t_rhs, b_rhs(n), "one RHS evaluation", "rhs", rhs_sus),
            "roofline_jvp": roof(t_jvp, b_jvp(n), "one fused FD-JVP evaluation (cached base RHS)", "jvp", jvp_sus),
            "sustained": sustained,
            "clocks": clocks,
            "parity": parity,
            "setup_s": wl.setup_s,
        }
    if world == 1 and line is not None and not args.skip_stepops:
        line["stepops"] = measure_stepops(rhs, q_d, rhs0_d, peak)
    del rhs0_d, q_d, v_d, q_h, v_h, r0_h
    wl.close()
    torch.cuda.empty_cache()

    def quick(nzq, nhq, reps=3):
        """A second problem size in the same process: parity, then `reps` loops of --steps steps (median)."""
        w = Workload(base, ex, nzq, nhq, world, rank, local_rank, host=False)
        par = None if args.skip_check else parity_of(w)
        for _ in range(args.warmup):
            w.step_device()
        ms = []
        tr, tj = [], []
        for _ in range(reps):
            m, _ = timed_loop(w.step_device, args.steps)
            ms.append(m)
            hist, modes = w.rhs.timing_history(min(256, 2 * args.steps))
            tr.append(hist[modes < 0][:, 2])
            tj.append(hist[modes >= 0][:, 2])
        m = _median(ms)
        er, ej = _median(numpy.concatenate(tr)), _median(numpy.concatenate(tj))
        out = {
            "elements_per_panel": [nzq, nhq, nhq], "nodes": w.tiled.num_nodes, "dofs": w.tiled.num_dofs,
            "value": 2.0 * w.tiled.num_dofs * args.steps / (m * 1e-3), "unit": UNIT, "ms_per_step": m / args.steps,
            "rhs_ms": er, "jvp_ms": ej, "kernel_path": w.rhs.path_name,
            "rhs_frac": b_rhs(n) * w.nodes_local / (er * 1e-3) / 1e9 / peak,
            "jvp_frac": b_jvp(n) * w.nodes_local / (ej * 1e-3) / 1e9 / peak,
            "parity": par,
        }
        w.close()
        del w
        torch.cuda.empty_cache()
        return out

    if world == 1 and not args.skip_sweep:
        sw = [quick(max(1, nz // 4), max(4, nh // 4)), quick(max(1, nz // 2), max(4, nh // 2))]
        if line is not None:
            line["sweep"] = {"what": "SURVEY 8(d) S2 at s = 1/4 and s = 1/2 of the headline size, same protocol (3 loops, median)", "sizes": sw}
    if world > 1 and not args.skip_strong and (6 * args.nh1) % world == 0:
        # strong scaling: the fixed N = 1 problem (S2 s = 1) cut `world` ways, same run
        st = quick(args.nz, args.nh1)
        if line is not None:
            st["what"] = (f"strong scaling: the fixed {args.nz}x{args.nh1}x{args.nh1} problem of the N = 1 line split over {world} ranks; "
                          "efficiency = value / (N x the N = 1 line's value)")
            line["strong"] = st
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return line


def measure_stepops(rhs, q_d, r_d, peak, reps=10):
    """Side measurement: the per-step kernels of csrc/stepops.cu at the bench size, each timed
    with CUDA events over `reps` launches after 3 warm-up launches, as algorithmic bytes / time against the measured HBM
    peak, with the numpy expression the reference evaluates timed beside it on one host core on a 1/16 sample.
    Inputs (0.42 GB state, 8.9 GB Krylov basis) are larger than L2."""
    import numpy
    import torch

    from wxfactory_b200.stepops import ExpFilterB200, VecOps

    ops = VecOps(rhs.device_index)
    n = q_d.numel()
    reps = int(os.environ.get("WXB_STEPOPS_REPS", reps))  # 1 under ncu: every kernel captured once
    out = {"state_values": n, "reps": reps}

    def timed(fn):
        for _ in range(3 if reps > 1 else 0):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    def cpu_timed(fn):
        fn()
        t0 = time.perf_counter()
        fn()
        return (time.perf_counter() - t0) * 1e3

    def entry(ms, nbytes, cpu_ms, cpu_frac, note):
        ach = nbytes / (ms * 1e-3) / 1e9
        return {"ms": ms, "bytes": nbytes, "achieved_gbs": ach, "frac_of_hbm_peak": ach / peak,
                "cpu_numpy_ms_scaled": cpu_ms / cpu_frac, "cpu_sample": f"1/{int(1 / cpu_frac)} of the vectors, 1 core", "what": note}

    ns = n // 16
    xs, ys, zs = (numpy.random.default_rng(i).standard_normal(ns) for i in range(3))
    # exponential filter: read 5 variables + sqrtG, write 5 (88 B per node)
    P = rhs.prob.n ** 3
    F1 = numpy.eye(rhs.prob.n) * 0.9 + 0.1 / rhs.prob.n
    filt = ExpFilterB200(rhs, F1)
    ms = timed(lambda: filt(q_d))
    dense = numpy.kron(numpy.kron(F1, F1), F1)
    qs = xs[: (ns // (5 * P)) * 5 * P].reshape(5, -1, P)
    sgs = numpy.abs(ys[: qs.shape[1] * P]).reshape(1, -1, P) + 1.0
    out["expfilter"] = entry(ms, (n + n + n // 5) * 8, cpu_timed(lambda: ((sgs * qs) @ dense) * (1.0 / sgs)), qs.size / n,
                             "wxb_expfilter: ((sqrtG Q) @ expfilter_new) * inv_sqrtG as three 1-D passes")
    # SSP-RK3 stage combination: 3 reads + 1 write
    tmp = torch.empty_like(q_d)
    q1_d = q_d * 1.01  # a third, distinct input: the kernel reads 3 arrays and writes 1
    ms = timed(lambda: ops.lincomb3(0.75, q_d, 0.25, q1_d, 0.25, r_d, 0.5, out=tmp))
    out["lincomb3"] = entry(ms, 4 * n * 8, cpu_timed(lambda: 0.75 * xs + 0.25 * ys + 0.25 * zs * 0.5), ns / n, "wxb_vec_lincomb3: 0.75 Q + 0.25 Q1 + 0.25 rhs dt")
    # Krylov basis of restart 20 (22 rows): 1-sync Gram-Schmidt product and fused row update, solution update
    m = 22
    V = torch.empty(m, n, dtype=torch.float64, device=q_d.device)
    for i in range(m):
        V[i].copy_(q_d if i % 2 else r_d)
        V[i].mul_(1.0 + 0.01 * i)
    Vs = numpy.random.default_rng(9).standard_normal((m, ns // 4))
    ms = timed(lambda: ops.multidot(V, m, V[m - 2 :], 2))
    out["multidot_22x2"] = entry(ms, m * n * 8, cpu_timed(lambda: Vs @ Vs[m - 2 :].T), Vs.shape[1] / n,
                                 "wxb_vec_multidot: Q[:22] @ Q[20:22].T (fgmres.py:38) incl. the 44-value read-back")
    c = numpy.linspace(0.1, 0.2, m - 2)

    def gs_cpu():
        a = Vs[m - 2] - Vs[: m - 2].T @ c
        b = Vs[m - 1] - Vs[: m - 2].T @ c
        a /= 1.7
        return (b - a * 0.3) / 1.7

    ms = timed(lambda: ops.gs1sync(V, m - 2, c, c, 1.0, 0.0, V[m - 2], V[m - 1]))
    out["gs1sync_20"] = entry(ms, (m + 2) * n * 8, cpu_timed(gs_cpu), Vs.shape[1] / n, "wxb_vec_gs1sync: rows 20 and 21 against rows 0..19 (fgmres.py:60-69)")
    ms = timed(lambda: ops.combine(c, V, 1.0, q_d, 1.0, out=tmp))
    out["combine_20"] = entry(ms, (m - 2 + 2) * n * 8, cpu_timed(lambda: Vs[m - 1] + c @ Vs[: m - 2]), Vs.shape[1] / n, "wxb_vec_combine: x + Z[:20].T y (fgmres.py:246-249)")
    out["launches"] = ops.launches + filt.launches
    return out


def run_shallow_water(args):
    """Side measurement (rank 0, one GPU): BASELINE.json configs[1] -- galewsky (n = 3) "at high element count",
    the reference's 30x30 panels tiled to --sw-nh.  Same protocol as the headline workload: device time of every
    evaluation from the C ABI's events, roofline against B_sw(n) = 144 + 48/n (RHS) / 192 + 48/n (JVP, + v and the
    cached base RHS) bytes per node, parity through the tiling periodicity against the numpy oracle."""
    import numpy
    import torch

    from oracle import jvp_oracle
    from wxfactory_b200 import EPS_FD, RhsB200, synthetic as syn
    from wxfactory_b200.problem import load_problem

    base, ex = load_problem(os.path.join(ROOT, "tests", "golden", "galewsky.npz"))
    Qb = syn.sw_base_state(ex["Q0"])
    vb = syn.base_direction(Qb)
    dt = float(ex["dt"])
    tiled = syn.TiledSwProblem(base, args.sw_nh)
    rhs = RhsB200(tiled)
    q = torch.from_numpy(rhs.local_state(Qb)).cuda()
    v = torch.from_numpy(rhs.local_state(vb)).cuda()
    r0 = rhs(q)
    jv = rhs.jvp(v, dt, q, r0)
    small = syn.TiledSwProblem(base, 2 * base.nh)
    orc = jvp_oracle.make_oracle(small.materialize())
    Qs, vs = small.global_state(Qb), small.global_state(vb)
    wr = orc.rhs(Qs)
    wj = jvp_oracle.matvec_fd(orc.rhs, vs.ravel(), dt, Qs, wr).reshape(Qs.shape)
    er, cnt = syn.check_sw_against_small(tiled, rhs.part, r0, wr)
    ej, _ = syn.check_sw_against_small(tiled, rhs.part, jv, wj)
    ej *= float(numpy.linalg.norm(wj)) / (dt * float(numpy.linalg.norm(wr)) / EPS_FD)
    steps = max(5, min(args.steps, 20))
    for _ in range(3):
        rhs.jvp(v, dt, q, rhs(q))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        rhs.jvp(v, dt, q, rhs(q))
    e1.record()
    torch.cuda.synchronize()
    hist, modes = rhs.timing_history(2 * steps)
    nodes, n = tiled.num_nodes, tiled.n
    peak, peak_src = measured_peaks()

    def roof(tt, bpn, what):
        sec = float(tt[:, 2].mean()) * 1e-3
        ach = bpn * nodes / sec / 1e9
        return {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                "kernels": [f"sw_tile<{n}>"] if rhs._fused[0] else [f"sw_traces<{n}>", f"sw_main<{n}>"],
                "launch_ms": {"traces": float(tt[:, 0].mean()), "main": float(tt[:, 1].mean()), "evaluation": float(tt[:, 2].mean())},
                "algorithmic_bytes_per_node": bpn, "nodes_per_launch": nodes, "launches_timed": int(len(tt)), "what": what}

    out = {
        "workload": f"configs[1] galewsky shallow water: n={n}, {args.sw_nh}x{args.sw_nh} elements/panel x 6 panels (reference's 30x30 panel tiled)",
        "nodes": nodes,
        "dofs": 3 * nodes,
        "value": 2.0 * 3 * nodes * steps / (e0.elapsed_time(e1) * 1e-3),
        "unit": UNIT,
        "ms_per_step": e0.elapsed_time(e1) / steps,
        "roofline": roof(hist[modes < 0], 144 + 48 / n, "one shallow-water RHS evaluation"),
        "roofline_jvp": roof(hist[modes >= 0], 192 + 48 / n, "one fused FD-JVP evaluation (cached base RHS)"),
        "parity": {"rhs_rel_err": er, "jvp_err_over_dt_rhs_eps": ej, "elements_checked": cnt, "tolerance": 1e-12,
                   "ok": bool(er <= 1e-12 and ej <= 1e-12)},
    }
    rhs.close()
    return out


def reference_line(args):
    """`--impl reference`: the CPU arm alone, on the arm's own config/metric/unit."""
    nh = args.nh if args.nh > 0 else nh_for(args.gpus, args.nh1)
    cpu = run_cpu_arm(args.steps, args.warmup)
    return {
        "impl": "reference",
        "metric": METRIC,
        "value": cpu["value"],
        "unit": UNIT,
        "n_gpus": args.gpus,
        "steps": args.steps,
        "warmup": args.warmup,
        "ms_per_step": cpu["ms_per_step"],
        "higher_is_better": True,
        "scaling": "weak",
        "vs_baseline": None,
        "dtype": "f64",
        "data": "synthetic (reference-built DCMIP 3-1 n=4 metric/state tiled along the element axes)",
        "config": {
            "workload": f"configs[4] synthetic cubed-sphere Euler: n=4, {args.nz}x{nh}x{nh} elements/panel x 6 panels",
            "step": "1 RHS + 1 FD-JVP = 2 x dofs DOF-updates",
            "note": "each step is a bounded sample of the workload (see cpu_baseline.sample); throughput is per DOF-update",
        },
        "cpu_baseline": {"value": cpu["value"], "unit": UNIT, "cores": cpu["cores"], "kind": "port", "sample": cpu["sample"]},
        "e2e": {"value": cpu["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def main():
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["b200", "reference"], default="b200")
    ap.add_argument("--nz", type=int, default=24, help="elements per column (SURVEY 8d S2 s = 1: 24)")
    ap.add_argument("--nh1", type=int, default=96, help="elements per panel side at one GPU (S2 s = 1: 96; scaled by sqrt(N))")
    ap.add_argument("--repeats", type=int, default=5, help="timed loops of --steps steps; the median loop is reported")
    ap.add_argument("--sustain-s", type=float, default=2.0, help="length of the back-to-back `sustained` loop in seconds (0: skip)")
    ap.add_argument("--skip-sweep", action="store_true", help="skip the smaller S2 sizes (s = 1/4, 1/2) at N = 1")
    ap.add_argument("--skip-strong", action="store_true", help="skip the strong-scaling measurement at N > 1")
    ap.add_argument("--nh", type=int, default=0, help="override the elements per panel side")
    ap.add_argument("--e2e-steps", type=int, default=10)
    ap.add_argument("--cpu-steps", type=int, default=10)
    ap.add_argument("--skip-check", action="store_true")
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--skip-sw", action="store_true", help="skip the shallow-water side measurement (configs[1])")
    ap.add_argument("--sw-nh", type=int, default=720, help="shallow-water elements per panel side")
    ap.add_argument("--skip-stepops", action="store_true", help="skip the side measurement of the per-step kernels (filter, stage algebra, Krylov)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3  # timing rule: at least 3 warm-up steps

    rank = int(os.environ.get("RANK", "0"))
    if args.impl == "reference":
        if rank != 0:
            return 0
        print(json.dumps(reference_line(args)), flush=True)
        return 0

    if args.gpus > 1 and "WORLD_SIZE" not in os.environ:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 1000), os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd)

    line = run_b200(args)
    if rank == 0:
        if args.gpus == 1 and not args.skip_sw:
            line["shallow_water"] = run_shallow_water(args)
        if args.gpus == 1 and not args.skip_cpu:
            cpu = run_cpu_arm(args.cpu_steps, 2)
            line["cpu_baseline"] = {"value": cpu["value"], "unit": UNIT, "cores": cpu["cores"], "kind": "port", "sample": cpu["sample"]}
        print(json.dumps(line), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())

"""Shallow-water throughput at high element count (BASELINE.json configs[1]: galewsky, n = 3, tiled from the
reference's 30x30 panels) -- run on the GPU box.  Device time, DOF-updates/s and HBM-roofline fraction of one RHS and
one fused FD-JVP evaluation, with the tiled-periodicity parity check.  One JSON object per line.
    python scripts/sw_bench.py [base=galewsky] [nh,nh,...]"""
import json, os, sys
import numpy, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import jvp_oracle
from wxfactory_b200 import EPS_FD, RhsB200, load_problem
from wxfactory_b200 import synthetic as syn

PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0


def run(base_name, nh, steps=20):
    base, ex = load_problem(os.path.join(ROOT, "tests", "golden", base_name + ".npz"))
    Qb = syn.sw_base_state(ex["Q0"]); vb = syn.base_direction(Qb); dt = float(ex["dt"])
    tiled = syn.TiledSwProblem(base, nh)
    rhs = RhsB200(tiled)
    q = torch.from_numpy(rhs.local_state(Qb)).cuda(); v = torch.from_numpy(rhs.local_state(vb)).cuda()
    r0 = rhs(q); jv = rhs.jvp(v, dt, q, r0)
    small = syn.TiledSwProblem(base, 2 * base.nh)
    orc = jvp_oracle.make_oracle(small.materialize())
    Qs, vs = small.global_state(Qb), small.global_state(vb)
    wr = orc.rhs(Qs); wj = jvp_oracle.matvec_fd(orc.rhs, vs.ravel(), dt, Qs, wr).reshape(Qs.shape)
    er, cnt = syn.check_sw_against_small(tiled, rhs.part, r0, wr)
    ej, _ = syn.check_sw_against_small(tiled, rhs.part, jv, wj)
    ej *= float(numpy.linalg.norm(wj)) / (dt * float(numpy.linalg.norm(wr)) / EPS_FD)
    for _ in range(3):
        rhs.jvp(v, dt, q, rhs(q))
    torch.cuda.synchronize()
    for _ in range(steps):
        rhs.jvp(v, dt, q, rhs(q))
    hist, modes = rhs.timing_history(2 * steps)
    nodes, n = tiled.num_nodes, tiled.n
    R, J = hist[modes < 0], hist[modes >= 0]
    t_r, t_j = R[:, 2].mean() * 1e-3, J[:, 2].mean() * 1e-3
    out = {"equations": "shallow_water", "base": base_name, "n": n, "nh": nh, "nodes": nodes,
           "rhs_ms": t_r * 1e3, "jvp_ms": t_j * 1e3,
           "rhs_kernels_ms": [float(R[:, 0].mean()), float(R[:, 1].mean())],
           "jvp_kernels_ms": [float(J[:, 0].mean()), float(J[:, 1].mean())],
           "rhs_gdof_s": 3 * nodes / t_r / 1e9, "jvp_gdof_s": 3 * nodes / t_j / 1e9,
           "rhs_frac_hbm": (144 + 48 / n) * nodes / t_r / 1e9 / PEAK,
           "jvp_frac_hbm": (192 + 48 / n) * nodes / t_j / 1e9 / PEAK,
           "parity_rhs": er, "parity_jvp": ej, "elements_checked": cnt}
    print(json.dumps(out), flush=True)
    rhs.close()


if __name__ == "__main__":
    base = sys.argv[1] if len(sys.argv) > 1 else "galewsky"
    sizes = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [240, 960]
    for nh in sizes:
        run(base, nh)

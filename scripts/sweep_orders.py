"""Order sweep n = 2..6 (run on the GPU box): device time, DOF-updates/s and HBM-roofline fraction of one RHS and one
fused FD-JVP evaluation at (a) the reference authors' rhs_benchmark shapes (tests/rhs_benchmark/run.sh:67-71 in the
reference: 60^3 nodes per panel, 1.296 M nodes) and (b) ~10 M nodes, with the tiled-periodicity parity check at every
point.  Writes one JSON object per line to stdout."""
import json, os, sys, time
import numpy, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import jvp_oracle
from wxfactory_b200 import EPS_FD, RhsB200, load_problem
from wxfactory_b200 import synthetic as syn

PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
BASES = {2: "dcmip31", 3: "dcmip21", 4: "dcmip31_n4", 5: "dcmip31_n5", 6: "dcmip31_n6"}
S1 = {2: (30, 30), 3: (20, 20), 4: (15, 15), 5: (12, 12), 6: (10, 10)}  # (nh, nz)


def big_shape(n):
    # ~10.6 M nodes: 6 * nz * nh^2 * n^3
    nz = {2: 24, 3: 16, 4: 12, 5: 10, 6: 8}[n]
    nh = int(round((10.6e6 / (6 * nz * n**3)) ** 0.5 / 2)) * 2
    return nh, nz


def run(n, nh, nz, label, steps=20):
    base, ex = load_problem(os.path.join(ROOT, "tests", "golden", BASES[n] + ".npz"))
    Qb = syn.base_state(ex["Q0"]); vb = syn.base_direction(Qb); dt = float(ex["dt"])
    tiled = syn.TiledEulerProblem(base, nz, nh)
    rhs = RhsB200(tiled)
    q = torch.from_numpy(rhs.local_state(Qb)).cuda(); v = torch.from_numpy(rhs.local_state(vb)).cuda()
    r0 = rhs(q); jv = rhs.jvp(v, dt, q, r0)
    small = syn.TiledEulerProblem(base, 2 * base.nz, 2 * base.nh)
    orc = jvp_oracle.make_oracle(small.materialize())
    Qs, vs = small.global_state(Qb), small.global_state(vb)
    with numpy.errstate(all="ignore"):
        wr = orc.rhs(Qs); wj = jvp_oracle.matvec_fd(orc.rhs, vs.ravel(), dt, Qs, wr).reshape(Qs.shape)
    er, cnt = syn.check_against_small(tiled, rhs.part, r0, wr)
    ej, _ = syn.check_against_small(tiled, rhs.part, jv, wj)
    ej *= float(numpy.linalg.norm(wj)) / (dt * float(numpy.linalg.norm(wr)) / EPS_FD)
    for _ in range(3):
        rhs.jvp(v, dt, q, rhs(q))
    torch.cuda.synchronize()
    for _ in range(steps):
        rhs.jvp(v, dt, q, rhs(q))
    hist, modes = rhs.timing_history(2 * steps)
    nodes = tiled.num_nodes
    t_r, t_j = hist[modes < 0][:, 2].mean() * 1e-3, hist[modes >= 0][:, 2].mean() * 1e-3
    out = {"n": n, "shape": label, "elements_per_panel": [nz, nh, nh], "nodes": nodes, "kernel_path": rhs.path_name,
           "rhs_ms": t_r * 1e3, "jvp_ms": t_j * 1e3,
           "rhs_gdof_s": 5 * nodes / t_r / 1e9, "jvp_gdof_s": 5 * nodes / t_j / 1e9,
           "rhs_frac_hbm": (360 + 96 / n) * nodes / t_r / 1e9 / PEAK, "jvp_frac_hbm": (440 + 96 / n) * nodes / t_j / 1e9 / PEAK,
           "parity_rhs": er, "parity_jvp": ej, "elements_checked": cnt}
    print(json.dumps(out), flush=True)
    rhs.close()


if __name__ == "__main__":
    orders = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [2, 3, 4, 5, 6]
    quick = len(sys.argv) > 2 and sys.argv[2] == "quick"
    for n in orders:
        if not quick:
            nh, nz = S1[n]
            run(n, nh, nz, "rhs_benchmark")
        nh, nz = big_shape(n)
        run(n, nh, nz, "10M nodes", steps=10 if quick else 20)

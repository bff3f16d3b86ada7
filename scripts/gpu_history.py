"""Per-evaluation device times over a long run (does the main kernel slow down as the run goes on?)."""
import json, os, sys, time
import numpy, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wxfactory_b200 import RhsB200, load_problem
from wxfactory_b200 import synthetic as syn
base, ex = load_problem(os.path.join(ROOT, "tests", "golden", "dcmip31_n4.npz"))
Qb = syn.base_state(ex["Q0"]); vb = syn.base_direction(Qb); dt = float(ex["dt"])
tiled = syn.TiledEulerProblem(base, 12, 48)
rhs = RhsB200(tiled)
q = torch.from_numpy(rhs.local_state(Qb)).cuda(); v = torch.from_numpy(rhs.local_state(vb)).cuda()
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 120
for rep in range(2):
    for _ in range(steps):
        rhs.jvp(v, dt, q, rhs(q))
    hist, modes = rhs.timing_history(256)
    R = hist[modes < 0][:, 1]; J = hist[modes >= 0][:, 1]
    f = lambda a: " ".join("%.3f" % x for x in a)
    print("rep", rep, "RHS main every 10th:", f(R[::10]))
    print("rep", rep, "JVP main every 10th:", f(J[::10]))
    os.system("nvidia-smi --query-gpu=clocks.sm,clocks.mem,power.draw,power.limit,temperature.gpu,clocks_throttle_reasons.active --format=csv,noheader")
    time.sleep(2)

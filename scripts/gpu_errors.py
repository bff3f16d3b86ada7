"""Print the parity margins of the CUDA path on every golden case (run on the GPU box)."""
import json, os, sys
import numpy
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wxfactory_b200 import EPS_FD, RhsB200, load_problem, matvec_fun

cond = json.load(open(os.path.join(ROOT, "tests/golden/conditioning.json")))["cases"]
rel = lambda a, b: float(numpy.linalg.norm(a - b) / max(numpy.linalg.norm(b), 1e-300))
for name in ["case6", "case5", "galewsky", "dcmip31", "dcmip21", "dcmip31_n4", "dcmip31_n5", "dcmip31_n6", "dcmip21_n5"]:
    prob, ex = load_problem(os.path.join(ROOT, "tests/golden", name + ".npz"))
    rhs = RhsB200(prob)
    out = rhs(ex["Q0"])
    dt = float(ex["dt"])
    fd = matvec_fun(ex["v"].ravel(), dt, ex["Q0"], ex["rhs0"], rhs.full, "fd")
    scale = dt * numpy.linalg.norm(ex["rhs0"]) / EPS_FD
    print(f"{name:11s} rhs {rel(out, ex['rhs0']):.2e} (ref rounding {cond[name]['ref_rounding']:.2e})  per var "
          + " ".join(f"{rel(out[:, v], ex['rhs0'][:, v]):.1e}/{cond[name]['ref_rounding_per_var'][v]:.1e}" for v in range(prob.num_var))
          + f"  jvp {numpy.linalg.norm(fd - ex['jv_fd'].ravel()) / scale:.2e}")
    rhs.close()

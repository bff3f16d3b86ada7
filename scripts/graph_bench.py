"""Launch-bound regime: ms per SSP-RK3 step, eager vs one CUDA-graph replay, on the reference's own small grids."""
import json
import os
import sys
import time
import types

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from wxfactory_b200 import RhsB200, integrators, load_problem  # noqa: E402

out = {}
for case in ("dcmip21", "dcmip31", "dcmip31_n4"):
    prob, ex = load_problem(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", case + ".npz"))
    rhs = RhsB200(prob)
    Q0 = torch.from_numpy(ex["Q0"]).cuda()
    res = {"dofs": int(Q0.numel())}
    for tag, kw in (("eager", {}), ("graph", {"graph": True})):
        integ = integrators.Tvdrk3(types.SimpleNamespace(), rhs.full, **kw)
        Q = Q0
        for _ in range(20):
            Q = integ.step(Q, 1e-3)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(200):
            Q = integ.step(Q, 1e-3)
        torch.cuda.synchronize()
        res[tag + "_ms_per_step"] = (time.perf_counter() - t0) * 1e3 / 200
    out[case] = res
    rhs.close()
print(json.dumps(out))

"""Time the Krylov vector kernels alone at the bench's vector length (no RHS handle):
    python scripts/vec_bench.py [n]          (WXB_VEC_NOVEC2=1: scalar multi-dot loads)"""
import json
import os
import sys

import numpy
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from wxfactory_b200.stepops import VecOps  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 53084160
ops = VecOps(0)
m = 22
V = torch.randn(m, n, dtype=torch.float64, device="cuda")


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


out = {"n": n, "novec2": os.environ.get("WXB_VEC_NOVEC2", "0")}
for rows, k in ((22, 2), (8, 2), (2, 1), (1, 1)):
    ms = timed(lambda: ops.multidot(V, rows, V[m - 2 :], k))
    out[f"multidot_{rows}x{k}"] = {"ms": ms, "gbs": rows * n * 8 / ms / 1e6}
g = ops.multidot(V, m, V[m - 2 :], 2)
Vh = V[:, : 1 << 20].cpu().numpy()
out["check"] = float(numpy.abs(ops.multidot(V[:, : 1 << 20], m, V[m - 2 :, : 1 << 20], 2) - Vh @ Vh[m - 2 :].T).max())
print(json.dumps(out))

"""Host <-> device copy bandwidth with all ranks active at once (explains the end-to-end numbers at N > 1):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 scripts/pcie_probe.py"""
import os
import time

import torch
import torch.distributed as dist

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("gloo")
n = 1 << 28  # 2 GiB of doubles
h_in = torch.empty(n, dtype=torch.float64, pin_memory=True).fill_(1.0)
h_out = torch.empty(n, dtype=torch.float64, pin_memory=True).fill_(0.0)
d_a, d_b = torch.empty(n, dtype=torch.float64, device="cuda"), torch.ones(n, dtype=torch.float64, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(kind, active):
    if world > 1:
        dist.barrier()
    if not active:
        return 0.0
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        if kind in ("h2d", "both"):
            with torch.cuda.stream(s1):
                d_a.copy_(h_in, non_blocking=True)
        if kind in ("d2h", "both"):
            with torch.cuda.stream(s2):
                h_out.copy_(d_b, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return 3 * n * 8 / dt / 1e9  # GB/s per direction


res = {}
for kind in ("h2d", "d2h", "both"):
    run(kind, True)  # warm-up
    res[kind + "_alone"] = run(kind, rank == 0)
    res[kind + "_all"] = run(kind, True)
aff = sorted(os.sched_getaffinity(0))
line = f"rank {rank}: " + ", ".join(f"{k} {v:.1f}" for k, v in res.items()) + f" GB/s per direction; cpus {aff[0]}-{aff[-1]} ({len(aff)})"
if world > 1:
    lines = [None] * world
    dist.all_gather_object(lines, line)
    if rank == 0:
        print("\n".join(lines))
    dist.destroy_process_group()
else:
    print(line)

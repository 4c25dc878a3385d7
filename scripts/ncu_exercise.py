#!/usr/bin/env python
"""One launch of each representative kernel (for `ncu --set full` and for the launch list): the three single-GPU variants,
the cluster variant, the PDL geometry, and — through the loopback transport on one GPU — the one-shot, two-shot, tree and
push collective kernels plus the fused sharded scalar reduce. Sizes are kept small: ncu replays every kernel ~40 times."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_mpi_reductions_b200 as b2r  # noqa: E402
from cuda_mpi_reductions_b200.parallel import Communicator  # noqa: E402

torch.cuda.set_device(0)
n = 1 << 24
x = b2r.fill_device(n, torch.float32, seed=1, mode="signed")
xi = b2r.fill_device(n, torch.int32, seed=2, mode="full")
xd = b2r.fill_device(n, torch.float64, seed=3, mode="signed")
xb = b2r.fill_device(n, torch.bfloat16, seed=4, mode="signed")
ws = b2r.Workspace()
for rep in range(2):  # first pass warms up (module load), second pass is what the profiler should look at (-s skips)
    b2r.reduce(xi, "sum", workspace=ws, cfg={"variant": "ldg128"})
    b2r.reduce(x, "min", workspace=ws, cfg={"variant": "ldg256", "cluster": 2})
    b2r.reduce(xd, "sum", workspace=ws, cfg={"variant": "tma"})
    b2r.reduce(xb, "sum", workspace=ws, cfg={"variant": "tma", "pdl": 2})
    b2r.reduce(xb, "max", workspace=ws, cfg={"variant": "ldg128"})
    b2r.dot(x, x, workspace=ws)
    torch.cuda.synchronize()
comms = Communicator.loopback(4, device=0, heap_bytes=64 << 20)
m = 1 << 22
xs = [b2r.fill_device(m, torch.float32, seed=10 + r, mode="signed") for r in range(4)]
for rep in range(2):
    for algo in ("oneshot", "twoshot", "tree", "push"):
        Communicator.group_allreduce(comms, xs, "sum", algo=algo)
    Communicator.group_reduce(comms, xs, "sum", root=1, algo="twoshot")
    Communicator.group_reduce_scalar(comms, xs, "sum", root=0, all=True)
    Communicator.group_check(comms)
for c in comms:
    c.close()
print("ncu_exercise OK")

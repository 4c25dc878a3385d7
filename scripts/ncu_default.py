#!/usr/bin/env python
"""What bench.py launches, small enough for ncu: default launches (safe PDL + dynamic tail + one-atomic int32 finish ->
reduce_ldg_dyn_kernel), early-streaming launches (-> reduce_tma_kernel), and the loopback collectives AUTO now uses (LL, push)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_mpi_reductions_b200 as b2r  # noqa: E402
from cuda_mpi_reductions_b200.parallel import Communicator  # noqa: E402

torch.cuda.set_device(0)
n = 1 << 24
xs = [b2r.fill_device(n, torch.int32, seed=s, mode="sdk") for s in range(4)]
xd = b2r.fill_device(n // 2, torch.float64, seed=9, mode="signed")
ws = b2r.Workspace()
out = torch.zeros((), dtype=torch.int32, device="cuda")
early = b2r.launch_config(pdl=b2r.PDL_INPUT_READY)
for i in range(8):
    b2r.reduce(xs[i % 4], "sum", out=out, workspace=ws)            # default launch
for i in range(4):
    b2r.reduce(xs[i % 4], "sum", out=out, workspace=ws, cfg=early)  # early streaming
xl = b2r.fill_device(1 << 25, torch.int32, seed=7, mode="sdk")
for i in range(3):
    b2r.reduce(xl, "sum", out=out, workspace=ws)                    # 128 MiB: the dynamically scheduled tail engages (>= one tile per CTA)
    b2r.reduce(xd, "sum", workspace=ws)
torch.cuda.synchronize()
comms = Communicator.loopback(4, device=0, heap_bytes=32 << 20)
small = [b2r.fill_device(4096, torch.float32, seed=10 + r, mode="signed") for r in range(4)]
big = [b2r.fill_device(1 << 22, torch.float32, seed=20 + r, mode="signed") for r in range(4)]
for rep in range(2):
    Communicator.group_allreduce(comms, small, "sum")               # AUTO -> LL
    Communicator.group_reduce(comms, big, "sum", root=1, algo="push")
    Communicator.group_reduce_scalar(comms, big, "sum", root=0)
    Communicator.group_check(comms)
for c in comms:
    c.close()
print("ncu_default OK")

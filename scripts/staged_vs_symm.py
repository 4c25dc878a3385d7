#!/usr/bin/env python
"""What does a user with ordinary torch tensors pay? Communicator.allreduce / reduce on torch.empty buffers (staged through the
symmetric heap: one device-to-device copy in, one out, 64 MiB chunks) against the same call on symm_empty buffers (zero copy)
and against NCCL, device-timed, max over ranks. Run under torchrun:
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 scripts/staged_vs_symm.py"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cuda_mpi_reductions_b200.parallel import Communicator  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
comm = Communicator.from_torch_distributed(heap_bytes=(2 << 30) + (64 << 20), timeout_ms=30000)


def timed(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    comm.barrier()
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / iters], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


if rank == 0:
    print(f"# {world} GPUs; us per call (max over ranks); fp32 SUM")
    print(f"# {'coll':9s} {'bytes':>12s} {'symmetric':>10s} {'staged':>10s} {'NCCL':>10s}  staged/symmetric  staged vs NCCL")
for nbytes in (64 << 10, 4 << 20, 64 << 20, 1 << 30):
    n = nbytes // 4
    sx, sy = comm.symm_empty(n, torch.float32), comm.symm_empty(n, torch.float32)
    px, py = torch.empty(n, device=dev), torch.empty(n, device=dev)
    sx.fill_(rank + 1.0)
    px.fill_(rank + 1.0)
    nz = torch.empty(n, device=dev)
    iters = 5 if nbytes >= (1 << 30) else 20
    for coll in ("allreduce", "reduce"):
        if coll == "allreduce":
            t_s = timed(lambda: comm.allreduce(sx, "sum", out=sy), iters)
            t_p = timed(lambda: comm.allreduce(px, "sum", out=py), iters)
            t_n = timed(lambda: dist.all_reduce(nz), iters)
            ok = bool((py == world * (world + 1) / 2).all())
        else:
            t_s = timed(lambda: comm.reduce(sx, "sum", root=0, out=sy), iters)
            t_p = timed(lambda: comm.reduce(px, "sum", root=0, out=py), iters)
            t_n = timed(lambda: dist.reduce(nz, dst=0), iters)
            ok = rank != 0 or bool((py == world * (world + 1) / 2).all())
        comm.check()
        if rank == 0:
            print(f"  {coll:9s} {nbytes:12d} {t_s * 1e3:10.1f} {t_p * 1e3:10.1f} {t_n * 1e3:10.1f}  {t_p / t_s:8.2f}x slower     {t_n / t_p:5.2f}x  {'ok' if ok else 'WRONG'}")
    comm.symm_free_all()
    del px, py, nz
comm.close()
dist.destroy_process_group()

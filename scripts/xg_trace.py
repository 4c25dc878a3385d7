#!/usr/bin/env python
"""Where does a step of the headline loop spend its time? Per-phase timeline of the LAST 8 launches of a long back-to-back run,
from the in-kernel %globaltimer stamps (every CTA stamps start / streamed / combined / exit; the library keeps a ring of 8
launches). Run alone for one GPU, or under torchrun for the fused scalar-sharded reduce across GPUs:

    python scripts/xg_trace.py [--steps 200] [--pdl 2]
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 scripts/xg_trace.py

Per launch and rank it prints (microseconds, relative to the first CTA start of the oldest launch in the ring):
  start     first / last CTA start                 (launch ramp)
  streamed  first / last CTA done streaming        (tail skew)
  combined  first / last CTA past griddepcontrol.wait + block combine  (= previous launch complete)
  exit      last CTA exit                          (ticket; on the root: waits for the peers' packets, combines, acks)
and the period between consecutive launches' last exits — the number the benchmark sees.
%globaltimer is per GPU: columns are comparable within a rank, not across ranks."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_mpi_reductions_b200 as b2r  # noqa: E402
from cuda_mpi_reductions_b200 import _lib  # noqa: E402
import ctypes as C  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=1 << 24)
ap.add_argument("--steps", type=int, default=200)
ap.add_argument("--pdl", type=int, default=2)
ap.add_argument("--graph", action="store_true", help="capture the timed steps in one CUDA graph and replay it")
ap.add_argument("--notrace", action="store_true")
a = ap.parse_args()
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
copies = 8
big = torch.empty(copies * a.n, dtype=torch.int32, device=dev)
inputs = [big[c * a.n:(c + 1) * a.n] for c in range(copies)]
for c, t in enumerate(inputs):
    b2r.fill_device(a.n, torch.int32, seed=1000 * rank + c, mode="sdk", out=t)
out = torch.zeros((), dtype=torch.int32, device=dev)
cfg = b2r.launch_config(pdl=a.pdl)
comm = None
if world > 1:
    import torch.distributed as dist
    from cuda_mpi_reductions_b200.parallel import Communicator

    dist.init_process_group("nccl", device_id=dev)
    comm = Communicator.from_torch_distributed(timeout_ms=30000)
    step, _ = comm.prepare_reduce_scalar_sharded(inputs, "sum", root=0, out=out, cfg=cfg)
    h = C.c_void_p()
    _lib.check(_lib.LIB.b2r_comm_workspace(comm._h, C.byref(h)), "b2r_comm_workspace")
    ws_handle = h
    _lib.check(_lib.LIB.b2r_workspace_trace_enable(ws_handle, 1), "trace_enable")
else:
    ws = b2r.Workspace(dev)
    ws.trace(True)
    ws_handle = ws

    def step(i):
        b2r.reduce(inputs[i % copies], 0, out=out, workspace=ws, cfg=cfg)

side = torch.cuda.Stream(dev)
torch.cuda.synchronize()
torch.cuda.set_stream(side)
if comm is not None:  # the prepared call bound the stream that was current when it was made: make it again on this one
    step, _ = comm.prepare_reduce_scalar_sharded(inputs, "sum", root=0, out=out, cfg=cfg)
if a.notrace:
    _lib.check(_lib.LIB.b2r_workspace_trace_enable(ws_handle.handle if hasattr(ws_handle, "handle") else ws_handle, 0), "trace off")
for i in range(20):
    step(i)
torch.cuda.synchronize()
graph = None
if a.graph:
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=side):
        for i in range(a.steps):
            step(i)
    graph.replay()
    torch.cuda.synchronize()
if comm is not None:
    dist.barrier()
    comm.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
if graph is not None:
    graph.replay()
else:
    for i in range(a.steps):
        step(i)
e1.record()
torch.cuda.synchronize()
us_per_step = e0.elapsed_time(e1) * 1e3 / a.steps
if a.notrace:
    print(f"rank {rank}/{world}: {us_per_step:.2f} us per step over {a.steps} steps (pdl={a.pdl}, graph={bool(graph)}, no trace)", flush=True)
    if comm is not None:
        dist.barrier()
        comm.close()
        dist.destroy_process_group()
    sys.exit(0)
rows = []
for back in range(7, -1, -1):
    t = b2r.trace_last(ws_handle, back=back).double()
    rows.append(t)
t0 = rows[0][:, 0].min()
lines = [f"rank {rank}/{world}: {us_per_step:.2f} us per step over {a.steps} steps (pdl={a.pdl}); last 8 launches, us since the first CTA start of the oldest:",
         "  launch  start(first..last)   streamed(first..last)   combined(first..last)   exit(last)   period"]
prev_exit = None
for k, t in enumerate(rows):
    u = (t - t0) / 1e3
    ex = u[:, 3].max().item()
    per = f"{ex - prev_exit:7.2f}" if prev_exit is not None else "      -"
    lines.append(f"  {k:6d}  {u[:,0].min():7.2f}..{u[:,0].max():7.2f}   {u[:,1].min():8.2f}..{u[:,1].max():7.2f}    {u[:,2].min():8.2f}..{u[:,2].max():7.2f}   {ex:9.2f}   {per}")
    prev_exit = ex
text = "\n".join(lines)
if comm is not None:
    gathered = [None] * world
    dist.all_gather_object(gathered, text)
    if rank == 0:
        print("\n".join(gathered))
    comm.close()
    dist.destroy_process_group()
else:
    print(text)

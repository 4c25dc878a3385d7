#!/usr/bin/env python3
"""One process per GPU under torchrun: the reference's `mpirun reduce` use case (mpi/reduce.c) with b200reduce.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
        examples/allreduce_torchrun.py [--mib 64] [--dtype float32]

torch.distributed (NCCL) is only the plumbing here: it carries the bootstrap id, and serves as the cross-check. The data
path is the library's own kernels over NVLink: a symmetric-heap buffer (zero copy), an in-place allreduce, a reduce to
rank 0, and the fused "every rank reduces its shard, the scalars are combined over NVLink" call."""
import argparse
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_mpi_reductions_b200 as b2r  # noqa: E402
from cuda_mpi_reductions_b200.parallel import Communicator  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mib", type=int, default=64, help="message size per rank")
    ap.add_argument("--dtype", default="float32", choices=["int32", "float32", "float64", "bfloat16"])
    ap.add_argument("--iters", type=int, default=20)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    dtype = getattr(torch, args.dtype)
    count = (args.mib << 20) // torch.empty((), dtype=dtype).element_size()

    comm = Communicator.from_torch_distributed(heap_bytes=2 * (args.mib << 20) + (16 << 20))
    send = comm.symm_empty(count, dtype)  # lives in the symmetric heap: peers read/write it directly
    recv = comm.symm_empty(count, dtype)
    b2r.fill_device(count, dtype, seed=1 + rank, mode="signed", out=send)

    # allreduce, checked against NCCL
    comm.allreduce(send, "sum", out=recv)
    comm.check()
    ref = send.clone()
    dist.all_reduce(ref)
    err = (recv.double() - ref.double()).abs().max().item()
    algo = comm.select_algo(count * send.element_size(), dtype, "sum", allreduce=True)

    # timing: device events, max over ranks
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        comm.allreduce(send, "sum", out=recv)
    comm.barrier()
    e0.record()
    for _ in range(args.iters):
        comm.allreduce(send, "sum", out=recv)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / args.iters], device=dev, dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)

    # reduce to rank 0 and the fused scalar reduction over every rank's shard
    top = comm.reduce(send, "max", root=0)
    total = comm.reduce_scalar_sharded(send, "sum", root=0)
    comm.check()
    if rank == 0:
        nbytes = count * send.element_size()
        print(f"{world} GPUs, {args.mib} MiB {args.dtype} per rank, algorithm '{algo}', multicast={comm.has_multicast}")
        print(f"  allreduce: {ms.item() * 1e3:.1f} us  algbw {nbytes / ms.item() * 1e-6:.1f} GB/s  "
              f"busbw {nbytes / ms.item() * 1e-6 * 2 * (world - 1) / world:.1f} GB/s  max |diff| vs NCCL {err:.3g}")
        print(f"  reduce(max) -> rank 0: first element {top.reshape(-1)[0].item():.6g};  sharded scalar sum: {total.item():.6g}")
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

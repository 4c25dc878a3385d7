"""Multi-GPU sections of bench.py (N > 1), measured in the same process right after the headline region.

collectives : the metric BASELINE.json names for several GPUs — element-wise reduce-to-root and allreduce (MPI_Reduce
              semantics, mpi/reduce.c:76,90) — as bus bandwidth of this library's fused kernels AND of NCCL
              (torch.distributed.reduce / all_reduce) on the same sizes, both device-timed, max over ranks, both checked
              against a closed-form oracle.
config4     : BASELINE.json config 4 — a 1 GiB fp64 array sharded over the GPUs reduced to one scalar at root: the fused
              kernel (b2r_reduce_scalar_sharded) against what a library user writes (torch.sum per GPU + ncclReduce of the
              scalars); restatement of the vendored simpleMultiGPU (simpleMultiGPU.cpp:212-293).
Every timed loop: >= 3 warm-up calls, a device-side barrier, CUDA events on the launching stream, MAX over ranks.
"""
from __future__ import annotations

import torch

_OPS = {"sum": "SUM", "min": "MIN", "max": "MAX"}


def _max_over_ranks(dist, ms: float, dev) -> float:
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def _timed(fn, iters: int, warmup: int, pre) -> float:
    """ms per call on this rank: `pre()` (a device-side barrier) is enqueued right before the first event."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    pre()
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def _pattern(n: int, rank: int, dtype, dev) -> torch.Tensor:
    # small integers: exact in every dtype, so the expected result is a closed form and fp32 sums have one right answer
    return (torch.arange(n, device=dev, dtype=torch.int32) % 251 + rank).to(dtype)


def _expected(n: int, world: int, dtype, dev) -> torch.Tensor:
    return (world * (torch.arange(n, device=dev, dtype=torch.int32) % 251) + world * (world - 1) // 2).to(dtype)


def collectives_section(comm, dist, dev, rank: int, world: int, sizes=(64 << 20, 1 << 30), iters_small: int = 20, iters_big: int = 5) -> dict:
    """comm: a Communicator created with symmetric_only=True and a heap that holds two buffers of max(sizes)."""
    rows = []
    launches = 0
    big = max(sizes)
    send = comm.symm_empty(big // 4, torch.float32)  # reused (viewed) for every dtype / size
    recv = comm.symm_empty(big // 4, torch.float32)
    nccl_send = torch.empty(big // 4, dtype=torch.float32, device=dev)
    nccl_recv = torch.empty(big // 4, dtype=torch.float32, device=dev)
    cases = [("reduce", torch.float32, s) for s in sizes] + [("allreduce", torch.float32, s) for s in sizes]
    cases += [("reduce", torch.int32, sizes[0]), ("allreduce", torch.int32, sizes[0])]

    def measure(coll, dtype, nbytes):
        n = nbytes // 4
        x, y = send[:n].view(dtype), recv[:n].view(dtype)
        nx, ny = nccl_send[:n].view(dtype), nccl_recv[:n].view(dtype)
        x.copy_(_pattern(n, rank, dtype, dev))
        nx.copy_(x)
        want = _expected(n, world, dtype, dev)
        iters = iters_small if nbytes <= (256 << 20) else iters_big

        if coll == "reduce":
            def ours():
                comm.reduce(x, "sum", root=0, out=y)

            def theirs():
                dist.reduce(ny, dst=0, op=dist.ReduceOp.SUM)
        else:
            def ours():
                comm.allreduce(x, "sum", out=y)

            def theirs():
                dist.all_reduce(ny, op=dist.ReduceOp.SUM)

        # ---- correctness of both arms (NCCL reduces in place: feed it a fresh copy)
        y.zero_()
        ours()
        comm.check()
        ok_b2r = bool(torch.equal(y, want)) if (coll == "allreduce" or rank == 0) else True
        ny.copy_(nx)
        theirs()
        torch.cuda.synchronize()
        ok_nccl = bool(torch.equal(ny, want)) if (coll == "allreduce" or rank == 0) else True
        flags = torch.tensor([int(ok_b2r), int(ok_nccl)], device=dev)
        dist.all_reduce(flags, op=dist.ReduceOp.MIN)
        # ---- timing (NCCL in place on ny: its values grow, which does not change what it moves)
        ms_b2r = _max_over_ranks(dist, _timed(ours, iters, 3, comm.barrier), dev)
        ms_nccl = _max_over_ranks(dist, _timed(theirs, iters, 3, comm.barrier), dev)
        factor = 2.0 * (world - 1) / world if coll == "allreduce" else 1.0  # nccl-tests bus-bandwidth convention
        bus_b2r = nbytes / (ms_b2r * 1e-3) * 1e-9 * factor
        bus_nccl = nbytes / (ms_nccl * 1e-3) * 1e-9 * factor
        return iters, {"coll": coll, "dtype": str(dtype).replace("torch.", ""), "op": "sum", "bytes": nbytes, "iters": iters,
                       "algo": comm.select_algo(nbytes, dtype, "sum", allreduce=(coll == "allreduce")),
                       "b2r_us": ms_b2r * 1e3, "nccl_us": ms_nccl * 1e3, "b2r_busbw_gbs": bus_b2r, "nccl_busbw_gbs": bus_nccl,
                       "ratio": bus_b2r / bus_nccl if bus_nccl > 0 else None,
                       "b2r_check": "PASSED" if flags[0].item() else "FAILED", "nccl_check": "PASSED" if flags[1].item() else "FAILED"}

    for coll, dtype, nbytes in cases:
        iters, row = measure(coll, dtype, nbytes)
        launches += iters
        rows.append(row)
    # One mid-size row after the ones above (a failure here leaves them standing): reduce-to-root of 4 MiB, where AUTO takes the
    # posted reduce from 4 GPUs up. Both arms work out of L2 at this size (send + recv = 8 MiB), and each call is long enough
    # (~20 us) for the host side of either library to stay ahead of the GPU.
    mid_error = None
    try:
        iters, row = measure("reduce", torch.float32, min(4 << 20, big))
        row["l2_resident"] = True
        launches += iters
        rows.append(row)
    except Exception as e:  # noqa: BLE001
        mid_error = f"{type(e).__name__}: {e}"
    return {"rows": rows, "timed_launches_per_gpu": launches, "mid_size_row_error": mid_error,
            "buffers": "b2r: symmetric-heap tensors (zero copy); NCCL: ordinary torch tensors, in place",
            "busbw": "allreduce: algbw x 2(P-1)/P; reduce: algbw (nccl-tests convention)",
            "timing": "CUDA events after a device-side barrier, 3 warm-up calls, max over ranks; messages >= 64 MiB per rank (> L2 with send+recv), except the row marked l2_resident"}


def config4_section(comm, dist, dev, rank: int, world: int, total_elems: int = 1 << 27, steps: int = 20) -> dict:
    """1 GiB of fp64 sharded over the GPUs -> one scalar at rank 0."""
    import cuda_mpi_reductions_b200 as b2r
    from .timing import rotating_copies

    n = total_elems // world
    nbytes = n * 8
    copies = rotating_copies(nbytes)
    bigbuf = torch.empty(copies * n, dtype=torch.float64, device=dev)
    inputs = [bigbuf[c * n:(c + 1) * n] for c in range(copies)]
    for c, t in enumerate(inputs):
        b2r.fill_device(n, torch.float64, seed=5000 + 100 * rank + c, mode="sdk", out=t)
    out = torch.zeros((), dtype=torch.float64, device=dev)
    cfg = b2r.launch_config(pdl=b2r.PDL_INPUT_READY)
    start, _ = comm.prepare_reduce_scalar_sharded(inputs, "sum", root=0, out=out, cfg=cfg)
    tout = torch.zeros((), dtype=torch.float64, device=dev)
    counter = [0]

    def ours():
        start(counter[0])
        counter[0] += 1

    def theirs():
        torch.sum(inputs[counter[0] % copies], dim=0, out=tout)
        dist.reduce(tout, dst=0, op=dist.ReduceOp.SUM)
        counter[0] += 1

    # both arms on input 0, compared with each other and with a CPU oracle of rank 0's shard combined by NCCL
    counter[0] = 0
    ours()
    comm.check()
    counter[0] = 0
    theirs()
    torch.cuda.synchronize()
    local = b2r.reduce_oracle(inputs[0].cpu(), "sum").to(dev)
    dist.all_reduce(local, op=dist.ReduceOp.SUM)
    ok = True
    if rank == 0:
        w = local.item()
        ok = abs(out.item() - w) <= 1e-12 * abs(w) * max(1, world) and abs(tout.item() - w) <= 1e-9 * abs(w)
    flag = torch.tensor([int(ok)], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    ms_b2r = _max_over_ranks(dist, _timed(ours, steps, 3, comm.barrier), dev)
    ms_lib = _max_over_ranks(dist, _timed(theirs, steps, 3, comm.barrier), dev)
    total_bytes = total_elems * 8
    return {"what": "BASELINE config 4: 1 GiB fp64 sharded over the GPUs, SUM to one scalar at rank 0", "total_bytes": total_bytes,
            "elements_per_gpu": n, "steps": steps, "cache_control": f"{copies} rotating shards of {nbytes >> 20} MiB per GPU",
            "b2r_ms": ms_b2r, "b2r_gbs": total_bytes / (ms_b2r * 1e-3) * 1e-9, "library_ms": ms_lib,
            "library_gbs": total_bytes / (ms_lib * 1e-3) * 1e-9, "ratio": ms_lib / ms_b2r,
            "library": "torch.sum per GPU + torch.distributed.reduce (ncclReduce) of the scalars",
            "self_check": "PASSED" if flag.item() else "FAILED", "timed_launches_per_gpu": steps}

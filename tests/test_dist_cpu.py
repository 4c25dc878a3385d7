"""CPU-only, real processes: the Python collective API over torch.distributed with the gloo backend (world_size 2..4).
gloo only carries the shared-memory segment's name; the reductions are done by the library's host algorithms — the same
slice / tree / chain logic and combination order as the device kernels — and are checked against gloo's own all_reduce and
the element-wise oracle."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import cuda_mpi_reductions_b200 as b2r
    from cuda_mpi_reductions_b200.parallel import HostCommunicator

    comm = HostCommunicator.from_torch_distributed(window_bytes=1 << 20)
    fails = []
    try:
        for dtype in (torch.int32, torch.float32, torch.float64, torch.bfloat16):
            for count in (1, 1000, 300001):  # 300001 fp64 = 2.4 MB > the 1 MiB window: exercises chunking
                x = b2r.fill_host(count, dtype, seed=10 * rank + count, mode="signed")
                gathered = [torch.empty_like(x) for _ in range(world)]
                dist.all_gather(gathered, x)
                for op in ("sum", "min", "max"):
                    if dtype == torch.bfloat16:
                        acc = gathered[0].float()
                        for g in gathered[1:]:
                            acc = acc + g.float() if op == "sum" else (torch.minimum(acc, g.float()) if op == "min" else torch.maximum(acc, g.float()))
                        want = acc.bfloat16()
                    else:
                        want = gathered[0].clone()
                        for g in gathered[1:]:
                            b2r.combine_oracle(want, g, op)
                    for algo in ("oneshot", "twoshot", "tree"):
                        got = comm.allreduce(x, op, algo=algo)
                        exact = algo != "tree" or op != "sum" or dtype == torch.int32
                        ok = torch.equal(got, want) if exact else torch.allclose(got.double(), want.double(), rtol=2e-2, atol=0.1)
                        if not ok:
                            fails.append(("allreduce", str(dtype), count, op, algo))
                    for algo in ("twoshot", "ring"):
                        root = count % world
                        red = comm.reduce(x, op, root=root, algo=algo)
                        if rank == root:
                            exact = algo != "ring" or op != "sum" or dtype == torch.int32
                            ok = torch.equal(red, want) if exact else torch.allclose(red.double(), want.double(), rtol=2e-2, atol=0.1)
                            if not ok:
                                fails.append(("reduce", str(dtype), count, op, algo))
                        else:
                            assert red is None
            if dtype != torch.bfloat16:  # agree with gloo's own all_reduce
                y = x.clone()
                dist.all_reduce(y, op=dist.ReduceOp.MAX)
                if not torch.equal(comm.allreduce(x, "max"), y):
                    fails.append(("vs-gloo", str(dtype)))
        comm.barrier()
    finally:
        comm.close()
        dist.destroy_process_group()
    q.put((rank, fails))


@pytest.mark.parametrize("world", [2, 3, 4])
def test_host_communicator_over_gloo(world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert sorted(r for r, _ in results) == list(range(world))
    for rank, fails in results:
        assert not fails, f"rank {rank}: {fails[:5]}"

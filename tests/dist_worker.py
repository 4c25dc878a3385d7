"""Worker for the multi-process tests: launched by torchrun (one process per GPU) or by test_dist_cpu.py (gloo, CPU).
GPU mode checks the real NVLink path: Communicator.from_torch_distributed() -> VMM export/import over the Unix-socket
bootstrap -> fused collectives, against torch.distributed (NCCL) and the CPU oracle. Exit code 0 = all checks passed."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_mpi_reductions_b200 as b2r  # noqa: E402
from cuda_mpi_reductions_b200.parallel import Communicator  # noqa: E402


# absolute slack per rank for reordered fp sums (tree / NVLS): the tree keeps bf16 partials between rounds, values are in [-1, 1]
ATOL = {torch.int32: 0, torch.float32: 1e-5, torch.float64: 1e-13, torch.bfloat16: 2.0**-7}


def check_posted(comm, rank, world, fails):
    # posted reduce on real NVLink: a back-to-back series with rotating roots, changing sizes and no host synchronisation (slot
    # parity, acknowledgements of the root two calls back), interleaved with barrier-framed algorithms that use the same work
    # region; with a switch the staged copies are read by multimem (fp32 min travels as integer keys: exact)
    for dtype, op in ((torch.int32, "sum"), (torch.float32, "min"), (torch.float64, "max"), (torch.bfloat16, "max")):
        sizes = (7, 70_001, (1 << 20) + 5, (3 << 20) + 1)
        xs, wants = {}, {}
        for n in sizes:
            xs[n] = b2r.fill_device(n, dtype, seed=600 + rank + n % 97, mode="signed")
            gathered = [torch.empty_like(xs[n]) for _ in range(world)]
            dist.all_gather(gathered, xs[n])
            w = gathered[0].cpu().clone()
            for g in gathered[1:]:
                if dtype == torch.bfloat16:
                    w = torch.maximum(w.float(), g.cpu().float()).bfloat16()
                else:
                    b2r.combine_oracle(w, g.cpu(), op)
            wants[n] = w
        pending = []
        for i in range(60):
            n = sizes[(i * 3) % len(sizes)]
            root = (i * 5 + i // 7) % world
            algo = "posted" if i % 10 != 9 else ("push", "tree", "twoshot")[(i // 10) % 3]
            pending.append((n, root, comm.reduce(xs[n], op, root=root, algo=algo), algo))
        comm.check()
        for n, root, red, algo in pending:
            if rank == root and not torch.equal(red.cpu(), wants[n]):
                fails[0] += 1
                print(f"[rank {rank}] posted series MISMATCH dtype={dtype} op={op} count={n} algo={algo}", flush=True)


def finish(comm, dev, rank, fails):
    comm.barrier()
    comm.check()
    t = torch.tensor([fails], device=dev)
    dist.all_reduce(t)
    comm.close()
    dist.destroy_process_group()
    if rank == 0:
        print("DIST_WORKER " + ("PASSED" if t.item() == 0 else f"FAILED ({t.item()} mismatches)"), flush=True)
    sys.exit(0 if t.item() == 0 else 1)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", os.environ["RANK"]))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    comm = Communicator.from_torch_distributed(heap_bytes=256 << 20, timeout_ms=20000)
    assert comm.rank == rank and comm.size == world
    if rank == 0:
        print(f"world={world} multicast={comm.has_multicast}", flush=True)
    algos = ["oneshot", "twoshot", "tree", "push"] + (["nvls"] if comm.has_multicast else [])
    fails = 0
    if os.environ.get("B2R_DIST_ONLY") == "posted":  # the short form used while tuning: only the posted series
        box = [0]
        check_posted(comm, rank, world, box)
        finish(comm, dev, rank, box[0])
    for dtype in (torch.int32, torch.float32, torch.float64, torch.bfloat16):
        for count in (1, 1000, 65537, (1 << 22) + 3):
            x = b2r.fill_device(count, dtype, seed=100 + rank, mode="signed")
            gathered = [torch.empty_like(x) for _ in range(world)]
            dist.all_gather(gathered, x)
            hs = [g.cpu() for g in gathered]
            for op in ("sum", "min", "max"):
                for algo in algos:
                    if algo == "nvls" and dtype in (torch.float32, torch.float64) and op != "sum":
                        continue
                    if dtype == torch.bfloat16:
                        acc = hs[0].float()
                        for h in hs[1:]:
                            acc = acc + h.float() if op == "sum" else (torch.minimum(acc, h.float()) if op == "min" else torch.maximum(acc, h.float()))
                        want = acc.bfloat16()
                    else:
                        want = hs[0].clone()
                        for h in hs[1:]:
                            b2r.combine_oracle(want, h, op)
                    out = comm.allreduce(x, op, algo=algo)
                    comm.check()
                    exact = algo in ("oneshot", "twoshot", "push") or op != "sum" or dtype == torch.int32
                    got = out.cpu()
                    ok = torch.equal(got, want) if exact else torch.allclose(got.double(), want.double(), rtol=2e-2 if dtype == torch.bfloat16 else 1e-5, atol=ATOL[dtype] * world)
                    if not ok:
                        fails += 1
                        print(f"[rank {rank}] allreduce MISMATCH dtype={dtype} count={count} op={op} algo={algo}", flush=True)
                    root = (count + rank * 0) % world
                    for ralgo in ([algo, "ring", "posted"] if algo == "twoshot" else [algo]):
                        red = comm.reduce(x, op, root=root, algo=ralgo)
                        comm.check()
                        if rank == root:
                            rexact = ralgo in ("oneshot", "twoshot", "push") or op != "sum" or dtype == torch.int32
                            got = red.cpu()
                            ok = torch.equal(got, want) if rexact else torch.allclose(got.double(), want.double(), rtol=2e-2 if dtype == torch.bfloat16 else 1e-5, atol=ATOL[dtype] * world)
                            if not ok:
                                fails += 1
                                print(f"[rank {rank}] reduce MISMATCH dtype={dtype} count={count} op={op} algo={ralgo} root={root}", flush=True)
            # scalar sharded
            for op in ("sum", "max"):
                whole = torch.cat(hs)
                want = b2r.reduce_oracle(whole, op)
                for allflag in (False, True):
                    o = comm.reduce_scalar_sharded(x, op, root=world - 1, all=allflag)
                    comm.check()
                    if o is not None:
                        g = o.cpu()
                        tol = {torch.int32: 0, torch.float32: 1e-5, torch.float64: 1e-12, torch.bfloat16: 2.0**-7}[dtype]
                        if abs(g.double().item() - want.double().item()) > tol * max(1.0, abs(want.double().item())):
                            fails += 1
                            print(f"[rank {rank}] scalar MISMATCH dtype={dtype} count={count} op={op} got={g.item()} want={want.item()}", flush=True)
    # end-to-end flavour: shards in pinned host memory, native chunked pipeline + cross-GPU combine, every root, both flavours
    for dtype in (torch.int32, torch.float64, torch.bfloat16):
        for count in (1, 3_000_001, (5 << 20) + 7):
            h = b2r.fill_host(count, dtype, seed=300 + rank, mode="sdk", pin_memory=True)
            d = h.cuda()
            gathered = [torch.empty_like(d) for _ in range(world)]
            dist.all_gather(gathered, d)
            whole = torch.cat([g.cpu() for g in gathered])
            for op in ("sum", "min"):
                want = b2r.reduce_oracle(whole, op)
                for root, allflag in ((0, False), (world - 1, False), (0, True)):
                    o = comm.reduce_scalar_sharded_from_host(h, op, root=root, all=allflag)
                    comm.check()
                    if (o is None) != (not allflag and rank != root):
                        fails += 1
                        print(f"[rank {rank}] from_host returned a result on the wrong ranks", flush=True)
                    if o is not None:
                        tol = {torch.int32: 0, torch.float64: 1e-12, torch.bfloat16: 2.0**-7}[dtype]
                        if abs(o.double().item() - want.double().item()) > tol * max(1.0, abs(want.double().item())):
                            fails += 1
                            print(f"[rank {rank}] from_host MISMATCH dtype={dtype} count={count} op={op} got={o.item()} want={want.item()}", flush=True)
    # the call number of the scalar-sharded reduce lives on the device: a captured graph of calls can be replayed
    xs = [b2r.fill_device(1 << 20, torch.int32, seed=500 + 10 * rank + k, mode="sdk") for k in range(4)]
    wants = []
    for x in xs:
        t = x.sum(dtype=torch.int64)
        dist.all_reduce(t)
        wants.append(int(t.item()))
    outs = [torch.zeros((), dtype=torch.int32, device=dev) for _ in xs]
    side = torch.cuda.Stream(dev)
    with torch.cuda.stream(side):
        cfg = b2r.launch_config(pdl=b2r.PDL_INPUT_READY)
        for x, o in zip(xs, outs):
            comm.reduce_scalar_sharded(x, "sum", root=0, out=o, cfg=cfg)  # warm-up outside the capture
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for x, o in zip(xs, outs):
                comm.reduce_scalar_sharded(x, "sum", root=0, out=o, cfg=cfg)
        for rep in range(6):  # 24 calls: beyond the 16 mailbox slots, so slot reuse and the acknowledgements are exercised
            for o in outs:
                o.zero_()
            g.replay()
            torch.cuda.synchronize()
            if rank == 0 and [o.item() for o in outs] != wants:
                fails += 1
                print(f"[rank 0] graph replay {rep} MISMATCH {[o.item() for o in outs]} vs {wants}", flush=True)
    comm.check()
    # the experimental-no-more paths on real NVLink: LL (small), ring allreduce, pipelined ring
    for dtype in (torch.int32, torch.float32):
        for count in (5, 1000, comm.ll_max_bytes // 4):
            x = b2r.fill_device(count, dtype, seed=400 + rank, mode="signed")
            gathered = [torch.empty_like(x) for _ in range(world)]
            dist.all_gather(gathered, x)
            want = gathered[0].cpu().clone()
            for g in gathered[1:]:
                b2r.combine_oracle(want, g.cpu(), "sum")
            for i in range(6):
                out = comm.allreduce(x, "sum", algo="ll")
                red = comm.reduce(x, "sum", root=i % world, algo="ll")
                comm.check()
                if not torch.equal(out.cpu(), want) or (rank == i % world and not torch.equal(red.cpu(), want)):
                    fails += 1
                    print(f"[rank {rank}] LL MISMATCH dtype={dtype} count={count}", flush=True)
    fails_box = [0]
    check_posted(comm, rank, world, fails_box)
    fails += fails_box[0]
    # symmetric zero-copy + in place + stress of small back-to-back collectives
    s = comm.symm_empty(1 << 20, torch.float32)
    s.fill_(float(rank + 1))
    for i in range(200):
        comm.allreduce(s, "max", out=s, algo=algos[i % len(algos)] if algos[i % len(algos)] != "nvls" else "twoshot")
    comm.check()
    if not bool((s == world).all()):
        fails += 1
        print(f"[rank {rank}] in-place symmetric stress MISMATCH", flush=True)
    finish(comm, dev, rank, fails)


if __name__ == "__main__":
    main()

"""GPU, >= 2 devices: the real NVLink paths. (1) one process driving all GPUs (b2r_comm_init_all), (2) one process per
GPU under torchrun (b2r_comm_init_rank: VMM handles exchanged as file descriptors). Skipped on a one-GPU box, where the
same kernels are covered by the loopback tests."""
import os
import subprocess
import sys

import pytest
import torch

import cuda_mpi_reductions_b200 as b2r
from cuda_mpi_reductions_b200.parallel import Communicator, host_reduce_threads

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_single_process_all_devices():
    n = torch.cuda.device_count()
    comms = Communicator.init_all(list(range(n)), heap_bytes=128 << 20, timeout_ms=20000)
    try:
        algos = ["oneshot", "twoshot", "tree", "push"] + (["nvls"] if comms[0].has_multicast else [])
        for dtype in (torch.int32, torch.float32, torch.bfloat16):
            for count in (1, 4099, (1 << 21) + 1):
                xs_h = [b2r.fill_host(count, dtype, seed=7 * r + count, mode="signed") for r in range(n)]
                xs = [x.to(f"cuda:{r}") for r, x in enumerate(xs_h)]
                outs = [torch.empty_like(x) for x in xs]
                for d in range(n):
                    torch.cuda.synchronize(d)
                for algo in algos:
                    Communicator.group_allreduce(comms, xs, "sum", algo=algo, outs=outs)
                    Communicator.group_check(comms)
                    want = host_reduce_threads(xs_h, "sum", -1, "tree" if algo == "tree" else "oneshot")[0]
                    for r in range(n):
                        got = outs[r].cpu()
                        if algo == "nvls" and dtype != torch.int32:
                            assert torch.allclose(got.double(), want.double(), rtol=2e-2 if dtype == torch.bfloat16 else 1e-5, atol=1e-5)
                        else:
                            assert torch.equal(got, want), (algo, dtype, count, r)
                res = Communicator.group_reduce_scalar(comms, xs, "sum", root=n - 1, all=True)
                Communicator.group_check(comms)
                want = b2r.reduce_oracle(torch.cat(xs_h), "sum")
                for r in range(n):
                    assert abs(res[r].cpu().double().item() - want.double().item()) <= 1e-2 * max(1.0, abs(want.double().item()))
    finally:
        for c in comms:
            c.close()


def test_one_process_per_gpu_torchrun():
    n = min(torch.cuda.device_count(), 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(REPO, "tests", "dist_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    os.makedirs(os.path.join(REPO, "gpurun_out"), exist_ok=True)
    with open(os.path.join(REPO, "gpurun_out", f"dist_worker_{n}.log"), "w") as f:
        f.write(res.stdout + "\n==== stderr ====\n" + res.stderr)
    lines = [l for l in (res.stdout + res.stderr).splitlines() if "MISMATCH" in l or "Error" in l or "error" in l]
    tail = "\n".join(lines[:40]) + "\n" + res.stdout[-1500:]
    assert res.returncode == 0 and "DIST_WORKER PASSED" in res.stdout, tail
